"""Runs the mock-JVM test program against libuda.so (see tests/native/test_jni_mock.cc)."""
import subprocess


def test_jni_bridge_against_mock_jvm(built, tmp_path):
    exe, lib = built["test_jni_mock"], built["runtime"]
    p = subprocess.run([str(exe), str(lib), str(tmp_path)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "all checks passed" in p.stdout
