"""uda_shuffle_push_static: the push entry point for callers that may not allocate device memory.

Some users of this library cannot hand it scratch space: a shuffle provider loaded as a JNI plugin
into a container whose GPU memory another runtime has already claimed, or a long-running service
that fixes all of its buffers at start-up and forbids cudaMalloc afterwards (it synchronises the
device and can fail once memory is fragmented). For them the library carries its own small scratch
array. This test drives that entry point through the plain C ABI (ctypes on libuda_kernels.so, no
Python wrapper in between) and checks, on two calls in a row, that it moves exactly the same bytes
and publishes exactly the same landed-byte counters as uda_shuffle_push does with caller scratch.
The second call uses a different, smaller descriptor table, so it also shows that the table left in
the shared array by the first call does no harm.

The file name sorts last on purpose: a run that stops at the first failure has then already run
every other test.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

PUSH_DESC = np.dtype([("src", "<u8"), ("dst", "<u8"), ("bytes", "<i8"), ("flag", "<u8")])


def _load():
    from uda_b200._lib import lib_path

    lib = C.CDLL(str(lib_path("libuda_kernels.so")))
    vp, ll, i = C.c_void_p, C.c_longlong, C.c_int
    lib.uda_shuffle_push_static.restype = i
    lib.uda_shuffle_push_static.argtypes = [vp, i, i, vp]
    lib.uda_shuffle_push.restype = i
    lib.uda_shuffle_push.argtypes = [vp, i, vp, i, vp]
    lib.uda_push_scratch_bytes.restype = ll
    lib.uda_push_scratch_bytes.argtypes = [i]
    lib.uda_shuffle_push_static_max_segments.restype = i
    lib.uda_error_string.restype = C.c_char_p
    lib.uda_error_string.argtypes = [i]
    return lib


def _case(dev, sizes, seed):
    """Source bytes, two zeroed destinations (+flags) and the descriptor tables that target each."""
    rng = np.random.RandomState(seed)
    total = sum(sizes) + 32 * len(sizes) + 64
    src = torch.from_numpy(rng.randint(0, 256, size=total, dtype=np.uint8)).to(dev)
    outs = []
    for _ in range(2):
        dst = torch.zeros(total, dtype=torch.uint8, device=dev)
        flags = torch.zeros(len(sizes), dtype=torch.int64, device=dev)
        d = np.zeros(len(sizes), dtype=PUSH_DESC)
        so = do = 0
        for j, s in enumerate(sizes):
            so = (so + 15) // 16 * 16 + (j * 5) % 16
            do = (do + 15) // 16 * 16 + so % 16
            d[j] = (src.data_ptr() + so, dst.data_ptr() + do, s, flags.data_ptr() + 8 * j)
            so += s
            do += s
        table = torch.from_numpy(d.view(np.uint8).reshape(-1).copy()).to(dev)
        outs.append((dst, flags, table))
    return src, outs


def test_push_static_matches_push_twice(built):
    lib = _load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.uda_shuffle_push_static_max_segments() == 16384
    for call, sizes in enumerate([[1_000_003, 0, 17, 65536, 65537, 4096, 333_333, 1] * 6, [70_001, 5, 131072]]):
        src, ((dst_a, flags_a, tab_a), (dst_b, flags_b, tab_b)) = _case(dev, sizes, seed=call)
        scratch = torch.empty(int(lib.uda_push_scratch_bytes(len(sizes))), dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        rc = lib.uda_shuffle_push(tab_a.data_ptr(), len(sizes), scratch.data_ptr(), 0, stream)
        assert rc == 0, lib.uda_error_string(rc)
        rc = lib.uda_shuffle_push_static(tab_b.data_ptr(), len(sizes), 0, stream)
        assert rc == 0, lib.uda_error_string(rc)
        torch.cuda.synchronize()
        # byte-exact: the same tolerance (none) every other copy test in this suite uses
        assert torch.equal(dst_a, dst_b), f"call {call}: pushed bytes differ"
        assert torch.equal(flags_a, flags_b), f"call {call}: landed counters differ"
        assert flags_b.cpu().numpy().tolist() == sizes
        assert int(dst_b.to(torch.int64).sum().item()) > 0
        del src


def test_push_static_rejects_oversize(built):
    lib = _load()
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    dummy = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = lib.uda_shuffle_push_static(dummy.data_ptr(), 16384 + 1, 0, stream)
    assert rc == -2 and b"maximum" in lib.uda_error_string(rc)
    assert lib.uda_shuffle_push_static(dummy.data_ptr(), -1, 0, stream) == -1
    assert lib.uda_shuffle_push_static(dummy.data_ptr(), 0, 0, stream) == 0
