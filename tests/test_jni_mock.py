"""Runs the mock-JVM test program against libuda.so (see tests/native/test_jni_mock.cc)."""
import subprocess


def test_jni_bridge_against_mock_jvm(built, tmp_path):
    exe, lib = built["test_jni_mock"], built["runtime"]
    p = subprocess.run([str(exe), str(lib), str(tmp_path)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "all checks passed" in p.stdout


def test_jni_bridge_with_the_device_merge(built, tmp_path):
    """The same flow with the JNI consumer on merge approach 3, against the host-only mock of the kernel library: the
    merged stream reaches dataFromUda through per-call direct buffers cut out of the pinned delivery windows."""
    import os
    from pathlib import Path
    exe, lib = built["test_jni_mock"], built["runtime"]
    mock = Path(exe).parent / "libuda_kernels_mock.so"
    if not mock.exists():
        mock = Path(__file__).resolve().parent.parent / "build" / "bin" / "libuda_kernels_mock.so"
    p = subprocess.run([str(exe), str(lib), str(tmp_path), "3"], capture_output=True, text=True, timeout=120,
                       env={**os.environ, "UDA_KERNELS_LIB": str(mock)})
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "all checks passed" in p.stdout
