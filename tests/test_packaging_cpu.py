"""Packaging tail: the .deb writer and the installer script, on temporary directories."""
import io
import os
import subprocess
import sys
import tarfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "packaging"))


def _ar_members(blob: bytes):
    assert blob[:8] == b"!<arch>\n"
    pos, out = 8, {}
    while pos < len(blob):
        hdr = blob[pos:pos + 60]
        name = hdr[:16].decode().strip()
        size = int(hdr[48:58].decode().strip())
        assert hdr[58:60] == b"`\n"
        out[name] = blob[pos + 60:pos + 60 + size]
        pos += 60 + size + (size % 2)
    return out


def test_deb_has_debian_binary_control_and_payload(tmp_path):
    import build_deb

    deb = build_deb.build(tmp_path)
    assert deb.name.startswith("uda-b200_1.0.") and deb.name.endswith("_amd64.deb")
    m = _ar_members(deb.read_bytes())
    assert list(m) == ["debian-binary", "control.tar.gz", "data.tar.gz"] and m["debian-binary"] == b"2.0\n"
    with tarfile.open(fileobj=io.BytesIO(m["control.tar.gz"]), mode="r:gz") as tf:
        control = tf.extractfile("./control").read().decode()
        assert "./postinst" in tf.getnames()
    assert "Package: uda-b200" in control and "Architecture: amd64" in control and "Version: 1.0." in control
    with tarfile.open(fileobj=io.BytesIO(m["data.tar.gz"]), mode="r:gz") as tf:
        names = set(tf.getnames())
    assert "./usr/lib64/uda/uda_b200/models/terasort.py" in names
    assert "./usr/lib64/uda/java/shared/com/mellanox/hadoop/mapred/UdaBridge.java" in names
    assert "./usr/lib64/uda/install-uda.sh" in names and "./usr/share/doc/uda-b200/README.md" in names
    if (ROOT / "uda_b200" / "lib" / "libuda.so").exists():
        assert "./usr/lib64/uda/libuda.so" in names
    if subprocess.run(["which", "dpkg-deb"], capture_output=True).returncode == 0:   # the real tool accepts it
        p = subprocess.run(["dpkg-deb", "--info", str(deb)], capture_output=True, text=True)
        assert p.returncode == 0 and "uda-b200" in p.stdout, p.stderr


def test_installer_writes_env_and_hooks_hadoop_once(tmp_path):
    uda_dir, conf = tmp_path / "opt" / "uda", tmp_path / "hadoop" / "conf"
    conf.mkdir(parents=True)
    henv = conf / "hadoop-env.sh"
    henv.write_text("# hadoop-env\nexport JAVA_HOME=/usr\n")
    cmd = ["bash", str(ROOT / "packaging" / "install-uda.sh"), f"--uda-dir={uda_dir}", f"--prefix={tmp_path / 'root'}",
           f"--uda-hadoop-conf-dir={conf}"]
    for _ in range(2):      # the second run must not add a second source line
        p = subprocess.run(cmd, capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
    env = (uda_dir / "uda-env.sh").read_text()
    assert f"HADOOP_CLASSPATH=$HADOOP_CLASSPATH:{uda_dir}/uda.jar" in env and f"UDA_KERNELS_LIB={uda_dir}/libuda_kernels.so" in env
    assert henv.read_text().count("uda-env.sh ]; then") == 1
    assert (uda_dir / "uda_b200" / "config.py").exists() and (uda_dir / "java" / "README.md").exists()
    # the hook works: sourcing hadoop-env.sh puts the jar on the class path
    out = subprocess.run(["bash", "-c", f"HADOOP_CLASSPATH=x; . {henv}; echo $HADOOP_CLASSPATH"], capture_output=True, text=True).stdout
    assert out.strip() == f"x:{uda_dir}/uda.jar"
    p = subprocess.run(["bash", str(ROOT / "packaging" / "install-uda.sh")], capture_output=True, text=True)
    assert p.returncode != 0 and "Missing param" in p.stderr


def test_scheduler_case_wrapper_maps_cases_to_the_regression_driver():
    """cases/: the wrapper turns a case's arguments into the regression driver's command line (reference: cases/uda_wrapper.py)."""
    import importlib.util
    import xml.etree.ElementTree as ET

    spec = importlib.util.spec_from_file_location("uda_b200_wrapper", ROOT / "cases" / "uda_b200_wrapper.py")
    w = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(w)
    cmd = w.get_command("cpu", "/tmp/x")
    assert cmd[1].endswith("scripts/regression/auto_tester.py") and "--matrix" in cmd and cmd[cmd.index("--matrix") + 1].endswith("matrix_cpu.csv")
    assert "--matrix" not in w.get_command("all", "/tmp/x")
    tree = ET.parse(ROOT / "cases" / "uda_b200.cases")
    assert tree.getroot().tag == "CASEDEF" and len(tree.getroot().findall("case")) == 2
    assert tree.getroot().find("global/Case/wrapper").text == "./uda_b200_wrapper.py"
