import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box via gpurun)")


def _has_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # UDA_B200_RUN_GPU_MARKED=1: a test run that replaces the kernel library by the host-only mock
    # (tests/native/mock/mock_kernels.cc via UDA_KERNELS_LIB) may run GPU-marked tests of the native runtime without a GPU;
    # tests/test_device_merge_mock_cpu.py does that for tests/test_device_merge_gpu.py in a subprocess
    if _has_gpu() or os.environ.get("UDA_B200_RUN_GPU_MARKED") == "1":
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """Make sure the native libraries exist (cheap when up to date)."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import build as _b

    if os.environ.get("UDA_B200_SKIP_BUILD"):
        # sanitizer runs preload a runtime into this process; the compilers must not inherit it, and the libraries
        # under test were built beforehand
        root = _b.ROOT
        return {"kernels": _b.LIB_DIR / "libuda_kernels.so", "runtime": _b.LIB_DIR / "libuda.so",
                "test_jni_mock": root / "build" / "bin" / "test_jni_mock", "reader_bench": root / "build" / "bin" / "reader_bench"}
    return _b.build()
