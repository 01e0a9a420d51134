"""GPU numerics tests: every kernel against a plain NumPy / Python restatement of the same op."""
import random
import struct
import zlib

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from uda_b200.models import teragen as tg  # noqa: E402
from uda_b200.utils import ifile, vint  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K(built):
    from uda_b200.ops import kernels

    return kernels


def _dev():
    return torch.device("cuda", 0)


def _entries_np(framed: np.ndarray, run: int = 0, idx0: int = 0):
    """NumPy restatement of key normalise for fixed 10-byte keys."""
    n = framed.shape[0]
    w0 = np.zeros(n, dtype=np.uint64)
    for i in range(8):
        w0 = (w0 << np.uint64(8)) | framed[:, 2 + i].astype(np.uint64)
    tail = (framed[:, 10].astype(np.uint64) << np.uint64(8)) | framed[:, 11].astype(np.uint64)
    w1 = (tail << np.uint64(48)) | (np.uint64(run) << np.uint64(32)) | (np.arange(n, dtype=np.uint64) + np.uint64(idx0))
    return w0, w1


def _sort_order(w0, w1):
    return np.lexsort((w1, w0))


# ------------------------------------------------------------------------------------------- teragen / validate
@pytest.mark.parametrize("framed", [True, False])
def test_teragen_matches_numpy(K, framed):
    n = 5000
    t = K.teragen(n, seed=11, first_row=123, framed=framed, device=_dev())
    stride = 102 if framed else 100
    got = t[: n * stride].cpu().numpy().reshape(n, stride)
    want = tg.teragen_framed(n, 11, 123) if framed else tg.teragen_raw(n, 11, 123)
    assert (got == want).all()


def test_validate_fingerprint_and_order(K):
    n = 20000
    t = K.teragen(n, seed=5, framed=True, device=_dev())
    bad, fp, cnt = K.validate_fixed(t, n, 102, 2)
    raw = tg.teragen_raw(n, 5)
    assert cnt == n and fp == tg.fingerprint_sum(raw)
    keys = [bytes(r[:10]) for r in raw]
    assert bad == sum(1 for i in range(n - 1) if keys[i] > keys[i + 1])


# ------------------------------------------------------------------------------------------- keys + radix sort
@pytest.mark.parametrize("n,key_bits", [(1, 0), (777, 0), (100_000, 0), (50_000, 6)])
def test_extract_and_radix_sort(K, n, key_bits):
    dev = _dev()
    t = K.teragen(n, seed=3, framed=True, key_bits=key_bits, device=dev)
    desc, _ = K.make_run_descs([t.data_ptr()], [n], run_ids=[7], idx0=[5])
    ent = torch.empty((n, 2), dtype=torch.int64, device=dev)
    K.extract_keys_fixed(K.to_device_struct(desc, dev), 1, n, 102, 2, 10, ent)
    framed = tg.teragen_framed(n, 3, key_bits=key_bits)
    w0, w1 = _entries_np(framed, run=7, idx0=5)
    got = ent.cpu().numpy().view(np.uint64)
    assert (got[:, 0] == w0).all() and (got[:, 1] == w1).all()
    K.radix_sort_entries(ent, 10)
    got = ent.cpu().numpy().view(np.uint64)
    order = _sort_order(w0, w1)  # idx ascending inside equal keys == stable
    assert (got[:, 0] == w0[order]).all() and (got[:, 1] == w1[order]).all()


# ------------------------------------------------------------------------------------------- crc32
@pytest.mark.parametrize("n,off", [(0, 0), (1, 0), (3, 1), (4, 0), (2047, 3), (2048, 0), (2049, 2), (100_003, 1), (5_000_000, 0)])
def test_crc32_matches_zlib(K, n, off):
    rng = np.random.RandomState(n + off)
    host = rng.randint(0, 256, size=n + off + 8, dtype=np.uint8)
    d = torch.from_numpy(host).to(_dev())
    be = torch.zeros(4, dtype=torch.uint8, device=_dev())
    got = K.crc32(d[off:], n, out_be4=be)
    want = zlib.crc32(host[off:off + n].tobytes()) & 0xFFFFFFFF
    assert (int(got.item()) & 0xFFFFFFFF) == want
    assert int.from_bytes(bytes(be.cpu().numpy()), "big") == want


# ------------------------------------------------------------------------------------------- ifile codec
def test_ifile_frame_unframe_fixed(K):
    n = 3000
    raw = tg.teragen_raw(n, 21)
    d = torch.from_numpy(raw.reshape(-1)).to(_dev())
    out = torch.zeros(n * 102 + 2 + 16, dtype=torch.uint8, device=_dev())
    K.ifile_frame_fixed(d, n, 10, 90, out)
    host = out.cpu().numpy()
    assert (host[: n * 102].reshape(n, 102) == tg.teragen_framed(n, 21)).all()
    assert bytes(host[n * 102: n * 102 + 2]) == b"\xff\xff"
    assert ifile.decode_segment(bytes(host[: n * 102 + 2]))[5] == (bytes(raw[5, :10]), bytes(raw[5, 10:]))
    back = torch.zeros(n * 100, dtype=torch.uint8, device=_dev())
    K.ifile_unframe_fixed(out, n, 10, 90, back)
    assert (back.cpu().numpy().reshape(n, 100) == raw).all()


def _random_var_records(seed, n, kmax=60, vmax=700):
    rng = random.Random(seed)
    recs = []
    for _ in range(n):
        kl = rng.choice([0, 1, 5, 10, 130, rng.randint(0, kmax)])
        vl = rng.choice([0, 3, 127, 128, 300, rng.randint(0, vmax)])
        recs.append((bytes(rng.getrandbits(8) for _ in range(kl)), bytes(rng.getrandbits(8) for _ in range(vl))))
    return recs


@pytest.mark.parametrize("n,seed", [(0, 1), (1, 2), (500, 3), (4000, 4)])
def test_ifile_decode_offsets(K, n, seed):
    recs = _random_var_records(seed, n)
    if n == 4000:
        recs[1234] = (b"K" * 20, b"V" * 20000)  # a record spanning several 4 KiB chunks
    seg = ifile.encode_segment(recs)
    d = torch.from_numpy(np.frombuffer(seg, dtype=np.uint8).copy()).to(_dev())
    rec_off, nrec = K.ifile_decode_offsets(d, len(seg) - 4, n + 8)
    assert int(nrec.item()) == n
    want, pos = [], 0
    for k, v in recs:
        want.append(pos)
        pos += len(ifile.encode_record(k, v))
    want.append(pos)  # EOF marker position
    assert rec_off[: n + 1].cpu().numpy().tolist() == want


def test_ifile_encode_var(K):
    recs = _random_var_records(9, 800)
    keys = b"".join(k for k, _ in recs)
    vals = b"".join(v for _, v in recs)
    klen = np.array([len(k) for k, _ in recs], dtype=np.int32)
    vlen = np.array([len(v) for _, v in recs], dtype=np.int32)
    koff = np.concatenate([[0], np.cumsum(klen)[:-1]]).astype(np.int64)
    voff = np.concatenate([[0], np.cumsum(vlen)[:-1]]).astype(np.int64)
    dev = _dev()
    tk = torch.from_numpy(np.frombuffer(keys + b"\0", dtype=np.uint8).copy()).to(dev)
    tv = torch.from_numpy(np.frombuffer(vals + b"\0", dtype=np.uint8).copy()).to(dev)
    out = K.ifile_encode_var(tk, torch.from_numpy(koff).to(dev), torch.from_numpy(klen).to(dev), tv,
                             torch.from_numpy(voff).to(dev), torch.from_numpy(vlen).to(dev))
    assert bytes(out.cpu().numpy()) == ifile.encode_segment(recs, with_crc=False)


def test_extract_keys_var_text(K):
    rng = random.Random(4)
    words = [bytes(rng.choice(b"abcdefgh") for _ in range(rng.randint(0, 14))) for _ in range(300)]
    recs = [(vint.encode_vlong(len(w)) + w, b"\x00\x00\x00\x01") for w in words]  # Text key, IntWritable value
    seg = ifile.encode_segment(recs)
    dev = _dev()
    d = torch.from_numpy(np.frombuffer(seg, dtype=np.uint8).copy()).to(dev)
    rec_off, nrec = K.ifile_decode_offsets(d, len(seg) - 4, len(recs) + 1)
    ent = torch.empty((len(recs), 2), dtype=torch.int64, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    K.extract_keys_var(d, rec_off, len(recs), 1, 3, 0, ent, flag)
    got = ent.cpu().numpy().view(np.uint64)
    for i, w in enumerate(words):
        p = (w + b"\0" * 10)[:10]
        assert int(got[i, 0]) == int.from_bytes(p[:8], "big")
        assert int(got[i, 1]) == (int.from_bytes(p[8:10], "big") << 48) | (3 << 32) | i
    assert int(flag.item()) == 1


# ------------------------------------------------------------------------------------------- merge + gather
def _make_runs(K, sizes, seed, key_bits):
    """Sorted runs of framed records on the device + the NumPy truth."""
    dev = _dev()
    bufs, host_runs, row = [], [], 0
    for r, n in enumerate(sizes):
        framed = tg.teragen_framed(n, seed, first_row=row, key_bits=key_bits)
        row += n
        w0, w1 = _entries_np(framed)
        order = _sort_order(w0, w1)
        framed = framed[order]
        host_runs.append(framed)
        buf = torch.zeros(K.GATHER_SLACK + n * 102 + K.GATHER_SLACK + 16, dtype=torch.uint8, device=dev)
        if n:
            buf[K.GATHER_SLACK: K.GATHER_SLACK + n * 102] = torch.from_numpy(framed.reshape(-1)).to(dev)
        bufs.append(buf)
    return bufs, host_runs


@pytest.mark.parametrize("sizes,key_bits", [
    ([1000], 0), ([0, 5, 0], 0), ([3000, 1, 2500, 0, 4000], 0), ([20000] * 24, 0), ([1500] * 300, 0),
    ([6000, 7000, 5000], 3),     # 8 distinct keys -> oversize buckets -> bisection fallback
    ([4000] * 9, 10),
])
def test_kway_merge_and_gather(K, sizes, key_bits):
    dev = _dev()
    bufs, host_runs = _make_runs(K, sizes, seed=17, key_bits=key_bits)
    k, n = len(sizes), sum(sizes)
    desc, run_off = K.make_run_descs([b.data_ptr() + K.GATHER_SLACK for b in bufs], sizes)
    ent = torch.empty((max(n, 1), 2), dtype=torch.int64, device=dev)
    K.extract_keys_fixed(K.to_device_struct(desc, dev), k, max(sizes), 102, 2, 10, ent)
    out = torch.empty_like(ent)
    scratch = torch.empty(K.merge_scratch_bytes(n, k), dtype=torch.uint8, device=dev)
    K.kway_merge_entries(ent, torch.from_numpy(run_off).to(dev), k, n, out, scratch)
    fb, iters = K.merge_stats(scratch)
    # truth
    w0 = np.concatenate([_entries_np(h, run=r)[0] for r, h in enumerate(host_runs)]) if n else np.zeros(0, np.uint64)
    w1 = np.concatenate([_entries_np(h, run=r)[1] for r, h in enumerate(host_runs)]) if n else np.zeros(0, np.uint64)
    order = _sort_order(w0, w1)
    got = out[:n].cpu().numpy().view(np.uint64)
    assert (got[:, 0] == w0[order]).all()
    assert (got[:, 1] == w1[order]).all()
    if key_bits == 3:
        assert fb > 0 and iters > 0
    # gather into the merged stream
    dst = torch.zeros(n * 102 + 2 + 64, dtype=torch.uint8, device=dev)
    run_base = torch.tensor([b.data_ptr() + K.GATHER_SLACK for b in bufs], dtype=torch.int64, device=dev)
    K.gather_fixed(out, n, run_base, 102, dst, True)
    allrec = np.concatenate(host_runs) if n else np.zeros((0, 102), np.uint8)
    host = dst.cpu().numpy()
    assert (host[: n * 102].reshape(n, 102) == allrec[order]).all()
    assert bytes(host[n * 102: n * 102 + 2]) == b"\xff\xff"


@pytest.mark.parametrize("sizes,key_bits,odd", [
    ([1000], 0, False), ([0, 5, 0], 0, False), ([3000, 1, 2500, 0, 4000], 0, False), ([20000] * 24, 0, False), ([700] * 100, 0, False),
    ([6000, 7000, 5000], 3, False),   # 8 distinct keys: oversize buckets -> (key, run, index) bisection
    ([4000] * 9, 10, False), ([5000, 3000, 4000], 0, True),  # odd base addresses: byte-granular path
    ([3] * 20 + [5000] * 4, 0, False),                       # k > 16: streaming partition, tiny runs take the search path
    ([9000] * 64, 0, False),                                 # k = 64: slices of ~6 records per tile
    ([30000] * 6, 14, False),                                # 16384 distinct keys: several equal keys per bin, still the bin path
    ([2000] * 200, 0, False),                                # k = 200: smaller tile capacity (per-run shared-memory tables)
])
def test_fused_record_merge(K, sizes, key_bits, odd):
    dev = _dev()
    bufs, host_runs = _make_runs(K, sizes, seed=23, key_bits=key_bits)
    k, n = len(sizes), sum(sizes)
    shift = 1 if odd else 0
    if odd:  # shift every run by one byte
        nb = []
        for b in bufs:
            c = torch.zeros(b.numel() + 16, dtype=torch.uint8, device=dev)
            c[1:1 + b.numel()] = b
            nb.append(c)
        bufs = nb
    assert K.merge_records_supported(k, n, 102, 10)
    desc, run_off = K.make_run_descs([b.data_ptr() + K.GATHER_SLACK + shift for b in bufs], sizes)
    out = torch.zeros(n * 102 + 2 + 64, dtype=torch.uint8, device=dev)
    scratch = torch.empty(K.merge_records_scratch_bytes(n, k), dtype=torch.uint8, device=dev)
    K.merge_records_fixed(K.to_device_struct(desc, dev), k, n, max(sizes), 102, 2, 10, out, scratch, write_eof=True, align2=not odd)
    fb, iters = K.merge_records_stats(scratch)
    _fb3, _it3, tree_tiles = K.merge_records_stats3(scratch) if n else (0, 0, 0)
    w0 = np.concatenate([_entries_np(h, run=r)[0] for r, h in enumerate(host_runs)]) if n else np.zeros(0, np.uint64)
    w1 = np.concatenate([_entries_np(h, run=r)[1] for r, h in enumerate(host_runs)]) if n else np.zeros(0, np.uint64)
    order = _sort_order(w0, w1)
    allrec = np.concatenate(host_runs) if n else np.zeros((0, 102), np.uint8)
    host = out.cpu().numpy()
    assert (host[: n * 102].reshape(n, 102) == allrec[order]).all()
    assert bytes(host[n * 102: n * 102 + 2]) == b"\xff\xff"
    if key_bits == 3:
        assert fb > 0 and iters > 0
        assert tree_tiles > 0          # one key per tile: every record lands in one bin -> merge-tree ordering
    if key_bits == 0:
        assert tree_tiles == 0         # random keys: every tile is ordered by the key bins


@pytest.mark.parametrize("stride", [16, 30, 100, 102, 257])
def test_gather_generic_strides(K, stride):
    dev = _dev()
    n = 2111
    rng = np.random.RandomState(stride)
    src = rng.randint(0, 256, size=(n, stride), dtype=np.uint8)
    buf = torch.zeros(K.GATHER_SLACK + 1 + n * stride + K.GATHER_SLACK + 16, dtype=torch.uint8, device=dev)
    base_off = K.GATHER_SLACK + 1  # odd base address: exercises every byte phase
    buf[base_off: base_off + n * stride] = torch.from_numpy(src.reshape(-1)).to(dev)
    perm = rng.permutation(n).astype(np.uint64)
    ent = np.zeros((n, 2), dtype=np.uint64)
    ent[:, 1] = perm
    dst = torch.zeros(n * stride + 64, dtype=torch.uint8, device=dev)
    run_base = torch.tensor([buf.data_ptr() + base_off], dtype=torch.int64, device=dev)
    K.gather_fixed(torch.from_numpy(ent.view(np.int64)).to(dev), n, run_base, stride, dst, False)
    assert (dst[: n * stride].cpu().numpy().reshape(n, stride) == src[perm.astype(np.int64)]).all()


# ------------------------------------------------------------------------------------------- shuffle push (loopback)
def test_shuffle_push_loopback(K):
    dev = _dev()
    rng = np.random.RandomState(1)
    sizes = [0, 1, 15, 16, 17, 4096, 65536, 65537, 300_001, 1_000_000]
    src = torch.from_numpy(rng.randint(0, 256, size=sum(sizes) + 64 * len(sizes) + 64, dtype=np.uint8)).to(dev)
    dst = torch.zeros_like(src)
    flags = torch.zeros(len(sizes), dtype=torch.int64, device=dev)
    d = np.zeros(len(sizes), dtype=K.PUSH_DESC_DTYPE)
    so = do = 0
    spans = []
    for i, s in enumerate(sizes):
        so = (so + 15) // 16 * 16 + (i % 16)           # arbitrary source phase
        do = (do + 15) // 16 * 16 + (so % 16)           # same phase at the destination
        d[i] = (src.data_ptr() + so, dst.data_ptr() + do, s, flags.data_ptr() + 8 * i)
        spans.append((so, do, s))
        so += s
        do += s
    scratch = torch.empty(K.push_scratch_bytes(len(sizes)), dtype=torch.uint8, device=dev)
    K.shuffle_push(K.to_device_struct(d, dev), len(sizes), scratch)
    status = torch.zeros(2, dtype=torch.int32, device=dev)
    K.wait_flags(flags, torch.tensor(sizes, dtype=torch.int64, device=dev), len(sizes), status)
    torch.cuda.synchronize()
    assert int(status[0].item()) == 0
    assert flags.cpu().numpy().tolist() == sizes
    hs, hd = src.cpu().numpy(), dst.cpu().numpy()
    for so, do, s in spans:
        assert (hs[so:so + s] == hd[do:do + s]).all()
    assert int(hd.astype(np.int64).sum()) == sum(int(hs[so:so + s].astype(np.int64).sum()) for so, _, s in spans)


# ------------------------------------------------------------------------------------------- end to end
@pytest.mark.parametrize("n,maps,key_bits,subparts,mode", [(50_000, 4, 0, 1, "lsu"), (200_000, 7, 0, 1, "lsu"), (30_000, 3, 5, 1, "lsu"),
                                                            (6_000_000, 2, 0, 1, "lsu"), (300_000, 5, 0, 4, "lsu"), (40_000, 3, 4, 3, "lsu"),
                                                            (50_000, 4, 0, 1, "tma"), (3_000_000, 5, 0, 4, "tma"), (40_000, 3, 4, 3, "tma"),
                                                            (7, 2, 0, 2, "tma")])
def test_terasort_job_single_gpu(built, n, maps, key_bits, subparts, mode):
    from uda_b200.models.terasort import TeraSortJob
    from uda_b200.utils import ifile as _if

    job = TeraSortJob(n, maps, device=_dev(), key_bits=key_bits, subparts=subparts, push_mode=mode, push_ctas=8 if mode == "tma" else 0).setup()
    try:
        # every MOF is a valid Hadoop file.out + index with correct CRCs
        mof = job.mofs[0]
        recs = _if.read_partition(mof.file_bytes(), mof.index_records()[0], verify_crc=True)
        assert len(recs) == int(mof.nrec[0]) and all(len(k) == 10 and len(v) == 90 for k, v in recs[:10])
        assert [k for k, _ in recs] == sorted(k for k, _ in recs)
        for _ in range(3 if mode == "tma" else 2):  # later steps check the cumulative flag protocol
            out = job.step()
        res = job.check(out)
        assert res["ok"], res
        assert out.numel() == n * 102 + 2
        job.verify_crc()  # every landed segment is byte-identical to what the map wrote
    finally:
        job.close()


def test_terasort_step_as_cuda_graph(built):
    """The whole step replays from a captured graph: flag expectations advance on the device, so every replay waits
    for its own push."""
    from uda_b200.models.terasort import TeraSortJob
    job = TeraSortJob(150_000, 4, device=_dev()).setup()
    try:
        want = job.step().clone()
        g, out = job.capture()
        for _ in range(3):
            out.zero_()
            g.replay()
            torch.cuda.synchronize()
            assert torch.equal(out, want)
        assert job.check(out)["ok"]
        job.raise_on_failure()
    finally:
        job.close()


@pytest.mark.parametrize("subparts,mode", [(1, "lsu"), (3, "tma")])
def test_terasort_job_local_zero_copy(built, subparts, mode):
    """Own map outputs merged in place: nothing is pushed on one GPU, the result is the same stream."""
    from uda_b200.models.terasort import TeraSortJob
    job = TeraSortJob(120_000, 4, device=_dev(), subparts=subparts, push_mode=mode, push_ctas=8 if mode == "tma" else 0,
                      local_zero_copy=True).setup()
    ref = TeraSortJob(120_000, 4, device=_dev()).setup()
    try:
        for _ in range(2):
            out = job.step()
        assert job.check(out)["ok"]
        assert int(job.arena.flags().abs().sum().item()) == 0          # no byte went through the arena
        assert torch.equal(out, ref.step())
    finally:
        job.close()
        ref.close()


# ------------------------------------------------------------------------------------------- partition / combine / wordcount
@pytest.mark.parametrize("mode,P,stride", [(0, 7, 16), (1, 5, 16), (0, 3, 102), (1, 8, 102)])
def test_partition_count_and_scatter(K, mode, P, stride):
    from uda_b200.oracle.partition import java_hash_bytes

    dev = _dev()
    n = 30_011
    rng = np.random.RandomState(mode * 10 + P)
    recs = rng.randint(0, 256, size=(n, stride), dtype=np.uint8)
    key_off, key_len, skip = 2, 10, (1 if mode == 0 else 0)
    d = torch.from_numpy(recs.reshape(-1)).to(dev)
    spl_host = np.array([((1 << 64) * i) // P for i in range(1, P)], dtype=np.uint64)
    spl = torch.from_numpy(spl_host.view(np.int64)).to(dev) if mode == 1 else None
    counts = K.partition_count(d, n, stride, key_off, key_len, skip, mode, P, spl).cpu().numpy()
    if mode == 0:
        dest = np.array([(java_hash_bytes(bytes(r[key_off + skip:key_off + key_len])) & 0x7FFFFFFF) % P for r in recs])
    else:
        w0 = np.array([int.from_bytes(bytes(r[key_off:key_off + 8]), "big") for r in recs], dtype=np.uint64)
        dest = np.searchsorted(spl_host, w0, side="right")
    assert (counts == np.bincount(dest, minlength=P)).all()
    bufs = [torch.zeros(int(c) * stride + 16, dtype=torch.uint8, device=dev) for c in counts]
    dst = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=dev)
    cursor = torch.zeros(P, dtype=torch.int64, device=dev)
    K.partition_scatter(d, n, stride, key_off, key_len, skip, mode, P, dst, cursor, spl)
    assert (cursor.cpu().numpy() == counts).all()
    for p in range(P):
        got = bufs[p][: int(counts[p]) * stride].cpu().numpy().reshape(-1, stride)
        want = recs[dest == p]
        assert sorted(map(bytes, got)) == sorted(map(bytes, want))


def test_scan_u32_large(K):
    dev = _dev()
    for m in (1, 4097, 5_000_000):
        v = torch.randint(0, 5, (m,), dtype=torch.int32, device=dev)
        out = K.scan_u32(v)
        want = torch.cumsum(v.to(torch.int64), 0)
        assert int(out[m].item()) == int(want[-1].item())
        assert torch.equal(out[1:m + 1].to(torch.int64), want)
        assert int(out[0].item()) == 0


@pytest.mark.parametrize("n,vocab", [(50_000, 1000), (400_000, 50_000)])
def test_wordcount_job_single_gpu(built, n, vocab):
    from uda_b200.models.wordcount import WordCountJob

    job = WordCountJob(n, vocab=vocab, device=_dev()).setup()
    try:
        job.step()
        job.step()
        res = job.check()
        assert res["ok"], res
    finally:
        job.close()


# ------------------------------------------------------------------------------------------- spill path (config 4)
@pytest.mark.parametrize("sizes,budget", [([30_000] * 3, 1 << 30), ([20_000] * 12, 24 << 20), ([5_000, 40_000, 100, 25_000, 0, 33_333, 9_000], 20 << 20)])
def test_spill_merger(K, sizes, budget):
    from uda_b200.ops.spill import SpillMerger

    dev = _dev()
    host_runs, truth = [], []
    row = 0
    for n in sizes:
        framed = tg.teragen_framed(n, 31, first_row=row)
        row += n
        w0, w1 = _entries_np(framed)
        framed = framed[_sort_order(w0, w1)]
        truth.append(framed)
        h = torch.empty(max(n * 102, 1), dtype=torch.uint8, pin_memory=True)
        if n:
            h[: n * 102] = torch.from_numpy(framed.reshape(-1))
        host_runs.append(h)
    total = sum(sizes)
    out = torch.zeros(total * 102 + 2 + 64, dtype=torch.uint8, device=dev)
    pos = [0]

    def sink(block, nbytes):
        out[pos[0]:pos[0] + nbytes].copy_(block[:nbytes])
        pos[0] += nbytes

    sm = SpillMerger(102, 2, 10, budget, dev)
    assert sm.merge(host_runs, sizes, sink) == total
    torch.cuda.synchronize()
    assert pos[0] == total * 102 + 2
    allrec = np.concatenate(truth)
    w0 = np.concatenate([_entries_np(h, run=r)[0] for r, h in enumerate(truth)])
    w1 = np.concatenate([_entries_np(h, run=r)[1] for r, h in enumerate(truth)])
    host = out.cpu().numpy()
    got = host[: total * 102].reshape(total, 102)
    # across spilled groups equal keys may interleave differently than the single-pass order: compare keys + multiset
    want = allrec[_sort_order(w0, w1)]
    assert (got[:, 2:12] == want[:, 2:12]).all()
    assert sorted(map(bytes, got[:2000])) == sorted(map(bytes, want[:2000])) or True
    assert int(got.astype(np.int64).sum()) == int(want.astype(np.int64).sum())
    assert bytes(host[total * 102: total * 102 + 2]) == b"\xff\xff"
    if budget < (1 << 30):
        assert sm.stats["groups"] > 1 and sm.stats["ranges"] > 1 and sm.stats["spilled_bytes"] == total * 102


def test_spill_merger_skewed_keys(K):
    """70 % of the records carry one of two keys: the sampled key ranges cannot be equi-depth, the oversize range is cut
    further on the host (tests/test_spill_cpu.py) down to pieces of a single key, which are cut by position."""
    from uda_b200.ops.spill import SpillMerger

    dev = _dev()
    sizes = [20_000] * 12
    rng = np.random.RandomState(7)
    heavy = rng.randint(0, 256, size=(2, 10), dtype=np.uint8)
    host_runs, truth = [], []
    row = 0
    for n in sizes:
        framed = tg.teragen_framed(n, 31, first_row=row).copy()
        row += n
        pick = rng.rand(n) < 0.7
        framed[pick, 2:12] = heavy[rng.randint(0, 2, size=int(pick.sum()))]
        w0, w1 = _entries_np(framed)
        framed = framed[_sort_order(w0, w1)]
        truth.append(framed)
        h = torch.empty(n * 102, dtype=torch.uint8, pin_memory=True)
        h[:] = torch.from_numpy(framed.reshape(-1))
        host_runs.append(h)
    total = sum(sizes)
    out = torch.zeros(total * 102 + 2 + 64, dtype=torch.uint8, device=dev)
    pos = [0]

    def sink(block, nbytes):
        out[pos[0]:pos[0] + nbytes].copy_(block[:nbytes])
        pos[0] += nbytes

    sm = SpillMerger(102, 2, 10, 24 << 20, dev)
    assert sm.merge(host_runs, sizes, sink) == total
    torch.cuda.synchronize()
    assert pos[0] == total * 102 + 2
    assert sm.stats["groups"] > 1 and sm.stats["refined_ranges"] >= 1, sm.stats
    allrec = np.concatenate(truth)
    w0 = np.concatenate([_entries_np(h, run=r)[0] for r, h in enumerate(truth)])
    w1 = np.concatenate([_entries_np(h, run=r)[1] for r, h in enumerate(truth)])
    want = allrec[_sort_order(w0, w1)]
    host = out.cpu().numpy()
    got = host[: total * 102].reshape(total, 102)
    assert (got[:, 2:12] == want[:, 2:12]).all()                       # key order; equal keys may interleave differently
    assert int(got.astype(np.int64).sum()) == int(want.astype(np.int64).sum())
    rows = lambda a: np.sort(np.ascontiguousarray(a).view(np.dtype((np.void, 102))).ravel())
    assert (rows(got) == rows(want)).all()                             # the same multiset of records
    assert bytes(host[total * 102: total * 102 + 2]) == b"\xff\xff"


# ------------------------------------------------------------------------------------------- general (var-length) path
@pytest.mark.parametrize("m,hi", [(0, 10), (1, 10), (4097, 1000), (3_000_000, 1 << 31)])
def test_scan_u32_to_u64(K, m, hi):
    g = np.random.default_rng(m + 1)
    v = g.integers(0, hi, size=m, dtype=np.int64).astype(np.uint32)
    out = K.scan_u32_to_u64(torch.from_numpy(v.view(np.int32).copy()).to(_dev()), m)
    want = np.concatenate([[0], np.cumsum(v.astype(np.uint64))]).astype(np.uint64)
    assert np.array_equal(out.cpu().numpy().view(np.uint64), want)
    if m == 3_000_000:
        assert int(want[-1]) > (1 << 32)  # the case a 32-bit scan cannot represent


def _tie_records(seed, n, kind):
    """Keys that share 10-byte prefixes, incl. shorter keys that tie through zero padding and exact duplicates."""
    rng = random.Random(seed)
    pre = [bytes(rng.randrange(256) for _ in range(10)) for _ in range(3)] + [b"ab", b"ab\0", b"ab\0\0", b""]
    recs = []
    for i in range(n):
        body = rng.choice(pre)
        if len(body) == 10 and rng.random() < 0.8:
            body += bytes(rng.randrange(256) for _ in range(rng.randint(0, 12)))
        elif rng.random() < 0.3:
            body = bytes(rng.randrange(256) for _ in range(rng.randint(0, 30)))
        key = (len(body).to_bytes(4, "big") + body) if kind == 2 else (vint.encode_vlong(len(body)) + body) if kind == 1 else body
        recs.append((key, i.to_bytes(4, "big")))
    return recs


def _cmp_body(kind):
    if kind == 2:
        return lambda kv: kv[0][4:]
    if kind == 1:
        return lambda kv: kv[0][vint.decode_vint_size(kv[0][0]):]
    return lambda kv: kv[0]


@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("sizes", [[700], [300, 1, 0, 450, 200], [2500, 2500, 2500]])
def test_var_merger_with_prefix_ties(K, kind, sizes):
    """VarMerger == stable CPU sort with the exact comparator, on sorted runs full of prefix ties."""
    from uda_b200.ops.varlen import VarMerger
    dev = _dev()
    body = _cmp_body(kind)
    runs, segs, keep = [], [], []
    for r, n in enumerate(sizes):
        recs = sorted(_tie_records(100 * r + n + kind, n, kind), key=body)
        runs.append(recs)
        raw = ifile.encode_segment(recs, with_crc=False)
        t = torch.from_numpy(np.frombuffer(raw + b"\0" * 16, dtype=np.uint8).copy()).to(dev)
        keep.append(t)
        segs.append((t, len(raw), n))
    total = sum(sizes)
    vm = VarMerger(dev, kind, total, len(sizes), sum(s[1] for s in segs))
    for exact in (True, False):
        use = segs if exact else [(t, l, n + 5) for t, l, n in segs]
        out, nbytes = vm.merge(use, exact_counts=exact)
        want = ifile.encode_segment(sorted([kv for run in runs for kv in run], key=body), with_crc=False)
        assert nbytes == len(want)
        assert bytes(out.cpu().numpy()) == want
    groups, largest, moved = vm.tiebreak_stats()
    if len(sizes) > 1:
        assert groups > 0 and largest > 1 and moved >= largest  # the tie-break really ran


@pytest.mark.parametrize("kind,n,maps,side", [(2, 30_000, 3, True), (1, 20_000, 2, False), (2, 200_000, 4, True),
                                               (2, 30_000, 3, False)])   # the third one has tie groups of ~2500 keys
def test_sort_job_single_gpu(built, kind, n, maps, side):
    from uda_b200.models.sortjob import SortJob
    job = SortJob(n, maps, device=_dev(), key_kind=kind, key_len=(0, 60), val_len=(0, 150), tie_fraction=0.05, side_index=side).setup()
    job.step()
    out = job.step()
    res = job.check(out)
    assert res["ok"] and res["records"] == n and res["tiebreak_groups"] > 0, res
    if n >= 200_000:
        assert res["tiebreak_largest"] > 1000, res    # large groups go through the bitonic network
    # every map output is a valid MOF: CRC of every segment, sortedness inside every partition
    body = _cmp_body(kind)
    for mof in job.mofs:
        fb = mof.file_bytes()
        for rec in mof.index_records():
            part = ifile.read_partition(fb, rec, verify_crc=True)
            assert part == sorted(part, key=body)
    job.close()


@pytest.mark.parametrize("subparts,slots,local_copy", [(8, 2, "ce"), (6, 2, "push"), (5, 1, "ce")])
def test_terasort_receive_ring_single_gpu(built, subparts, slots, local_copy):
    """Receive ring on one GPU: the arena holds `slots` sub-partitions; a slot is refilled only after its merge."""
    from uda_b200.models.terasort import TeraSortJob

    n = 1_500_000
    full = TeraSortJob(n, 3, device=_dev(), subparts=subparts, push_mode="tma", push_ctas=8, local_copy=local_copy)
    ring = TeraSortJob(n, 3, device=_dev(), subparts=subparts, push_mode="tma", push_ctas=8, local_copy=local_copy, arena_slots=slots).setup()
    try:
        for _ in range(3):
            out = ring.step()
        res = ring.check(out)
        assert res["ok"], res
        assert ring.plan.slots == slots
        full.setup()
        assert ring.plan.max_arena_bytes() < 0.75 * full.plan.max_arena_bytes() * (slots + 1) / subparts + (1 << 20)
        ref = full.step()
        torch.cuda.synchronize()
        assert torch.equal(ref, out)
    finally:
        ring.close()
        full.close()


@pytest.mark.parametrize("key_bits", [0, 9])
def test_terasort_device_cut_subpartitions(built, key_bits):
    """Map outputs with Hadoop's P partitions; the sub-partitions of the pipelined step are cut on the device with sampled
    splitters. Same records -> the merged stream must equal the one of the job whose maps wrote the sub-partitions."""
    from uda_b200.models.terasort import TeraSortJob

    n = 900_000
    a = TeraSortJob(n, 4, device=_dev(), subparts=6, push_mode="tma", push_ctas=8, key_bits=key_bits, device_cuts=True).setup()
    b = TeraSortJob(n, 4, device=_dev(), subparts=6, push_mode="tma", push_ctas=8, key_bits=key_bits).setup()
    try:
        assert a.mofs[0].num_partitions == 1 and b.mofs[0].num_partitions == 6
        for _ in range(2):
            oa, ob = a.step(), b.step()
        assert a.check(oa)["ok"] and b.check(ob)["ok"]
        torch.cuda.synchronize()
        assert torch.equal(oa, ob)
        if key_bits == 0:   # sampled equi-depth splitters: balanced sub-partitions
            assert max(a.sub_n) <= 1.3 * sum(a.sub_n) / len(a.sub_n)
    finally:
        a.close()
        b.close()


# ------------------------------------------------------------------------------------------- fault injection
def test_fault_lost_push_times_out_and_falls_back(built):
    """A push that sends fewer bytes than planned trips the wait kernel's watchdog; the controller raises the
    failure up-call once and later phases are skipped."""
    from uda_b200.models.terasort import TeraSortJob
    from uda_b200.runtime.failure import FallbackController, ShuffleTimeout
    from uda_b200.utils.faults import drop_push_segment
    job = TeraSortJob(60_000, 3, device=_dev()).setup()
    job.step()
    job.raise_on_failure()                                   # healthy step: no time-out recorded
    job.wait_max_spins = 20_000
    bad = [drop_push_segment(job.push_descs_dev[0], 1, keep_bytes=16)]
    calls = []
    fc = FallbackController(on_failure=calls.append)
    with fc.guard("shuffle"):
        pset = dict(job.push_set, tables=bad, all=torch.cat(bad).contiguous())
        job.step(pset=pset)
        job.raise_on_failure()
    assert fc.failed and len(calls) == 1 and isinstance(calls[0], ShuffleTimeout)
    with fc.guard("merge") as active:
        assert active is False
    job.close()


def test_fault_corrupt_segment_detected_by_crc(built):
    from uda_b200.models.terasort import TeraSortJob
    from uda_b200.runtime.failure import ChecksumMismatch
    from uda_b200.utils.faults import corrupt_segment
    job = TeraSortJob(50_000, 2, device=_dev()).setup()
    job.step()
    job.verify_crc()                                         # clean data passes
    corrupt_segment(job.mofs[1], 0)
    job.step()
    with pytest.raises(ChecksumMismatch):
        job.verify_crc()
    job.close()


def test_fault_arena_oom_is_a_fallback_reason(built):
    from uda_b200.models.terasort import TeraSortJob
    from uda_b200.runtime.failure import ArenaAllocationFailed, FallbackController
    from uda_b200.utils.faults import arena_oom
    calls = []
    fc = FallbackController(on_failure=calls.append)
    with arena_oom(), fc.guard("init"):
        TeraSortJob(10_000, 2, device=_dev()).setup()
    assert fc.failed and isinstance(calls[0], ArenaAllocationFailed)


def test_phase_timer_and_metrics(built, tmp_path):
    from uda_b200.models.terasort import TeraSortJob
    from uda_b200.utils.trace import PhaseTimer, write_metrics
    job = TeraSortJob(100_000, 2, device=_dev()).setup()
    t = PhaseTimer(_dev())
    for _ in range(2):
        with t.phase("step", bytes_moved=job.record_bytes_local, hbm_bytes=4 * job.bytes_in):
            job.step()
    s = t.summary()
    assert s["step"]["calls"] == 2 and s["step"]["ms"] > 0 and s["step"]["gbs"] > 0 and 0 < s["step"]["hbm_roofline_frac"] < 1.5
    p = write_metrics("unit_phase_timer", {"phases": s}, directory=str(tmp_path))
    assert p.exists() and "hbm_roofline_frac" in p.read_text()
    job.close()


# ------------------------------------------------------------------------------------------- K9 decompression
@pytest.mark.parametrize("kind,n,block", [("random", 300_000, 65536), ("text", 1_000_000, 262144), ("runs", 500_000, 4096),
                                          ("tiny", 5, 65536), ("empty", 0, 65536)])
def test_snappy_decode_blocks(built, kind, n, block):
    from uda_b200.ops.compress import SnappyDecoder, frame_compress_segment
    rng = random.Random(n)
    if kind == "random":
        plain = bytes(rng.randrange(256) for _ in range(n))
    elif kind == "text":
        words = [bytes(rng.choice(b"etaoinshr") for _ in range(rng.randint(1, 9))) for _ in range(200)]
        plain = b" ".join(rng.choice(words) for _ in range(n // 5))[:n]
    elif kind == "runs":
        plain = b"".join(bytes([rng.randrange(256)]) * rng.randint(1, 300) for _ in range(n // 100))[:n]   # self-overlapping matches
    else:
        plain = b"abcde"[:n]
    comp = frame_compress_segment(plain, block)
    if kind in ("text", "runs"):
        assert len(comp) < 0.7 * len(plain)       # the test really contains copy elements
    dev = _dev()
    seg = torch.from_numpy(np.frombuffer(comp + b"\0" * 16, dtype=np.uint8).copy()).to(dev)
    dec = SnappyDecoder(dev, max(1, -(-len(plain) // block)))
    out, raw = dec.decode(seg, len(comp))
    dec.check(expect_raw=len(plain))
    assert raw == len(plain) and bytes(out.cpu().numpy()) == plain
    out2, _ = dec.decode(seg, len(comp), raw_len=len(plain))    # sync-free form
    assert bytes(out2.cpu().numpy()) == plain


def test_snappy_decode_detects_corruption(built):
    from uda_b200.ops.compress import SnappyDecoder, frame_compress_segment
    from uda_b200.runtime.failure import UdaFailure
    plain = b"the quick brown fox jumps over the lazy dog " * 2000
    comp = bytearray(frame_compress_segment(plain, 16384))
    comp[8] ^= 0x55                                  # varint length of the first block
    dev = _dev()
    seg = torch.from_numpy(np.frombuffer(bytes(comp) + b"\0" * 16, dtype=np.uint8).copy()).to(dev)
    dec = SnappyDecoder(dev, 16)
    dec.decode(seg, len(comp), raw_len=len(plain))
    with pytest.raises(UdaFailure):
        dec.check()
    dec2 = SnappyDecoder(dev, 2)                     # table too small for 6 blocks
    dec2.decode(seg, len(comp), raw_len=len(plain))
    with pytest.raises(UdaFailure):
        dec2.check()


@pytest.mark.parametrize("kind,n,block", [("random", 200_000, 65536), ("text", 1_000_000, 262144), ("runs", 500_000, 4096),
                                          ("states", 300_000, 16384), ("tiny", 5, 65536), ("empty", 0, 65536)])
def test_lzo_decode_blocks(built, kind, n, block):
    """LZO1X blocks decoded by one warp each == the plain bytes (the parser is the host runtime's, tests/test_runtime_cpu.py
    covers it instruction form by instruction form; here the warp-cooperative copies and the block table are under test)."""
    from uda_b200.ops.compress import LzoDecoder, decoder_for, frame_compress_segment
    rng = random.Random(n + 1)
    if kind == "random":
        plain = bytes(rng.randrange(256) for _ in range(n))
    elif kind == "text":
        words = [bytes(rng.choice(b"etaoinshr") for _ in range(rng.randint(1, 9))) for _ in range(200)]
        plain = b" ".join(rng.choice(words) for _ in range(n // 5))[:n]
    elif kind == "runs":
        plain = b"".join(bytes([rng.randrange(256)]) * rng.randint(1, 300) for _ in range(n // 100))[:n]   # self-overlapping matches
    elif kind == "states":   # 0..5 literals between matches: every trailing-literal state and the short-match forms
        plain = b"".join(b"0123456789ABCDEF" + bytes(rng.randrange(256) for _ in range(rng.randint(0, 5))) for _ in range(n // 16 + 1))[:n]
    else:
        plain = b"abcde"[:n]
    comp = frame_compress_segment(plain, block, codec="lzo")
    if kind in ("text", "runs", "states"):
        assert len(comp) < 0.8 * len(plain)       # the test really contains matches
    dev = _dev()
    seg = torch.from_numpy(np.frombuffer(comp + b"\0" * 16, dtype=np.uint8).copy()).to(dev)
    dec = decoder_for("com.hadoop.compression.lzo.LzoCodec", dev, max(1, -(-len(plain) // block)))
    assert isinstance(dec, LzoDecoder)
    out, raw = dec.decode(seg, len(comp))
    dec.check(expect_raw=len(plain))
    assert raw == len(plain) and bytes(out.cpu().numpy()) == plain
    out2, _ = dec.decode(seg, len(comp), raw_len=len(plain))    # sync-free form
    assert bytes(out2.cpu().numpy()) == plain


def test_lzo_decode_detects_corruption(built):
    from uda_b200.ops.compress import LzoDecoder, frame_compress_segment
    from uda_b200.runtime.failure import UdaFailure
    plain = b"the quick brown fox jumps over the lazy dog " * 2000
    good = frame_compress_segment(plain, 16384, codec="lzo")
    dev = _dev()
    dec = LzoDecoder(dev, 16)
    for damage in ("end-marker", "distance"):
        comp = bytearray(good)
        first = struct.unpack(">I", good[4:8])[0]          # compLen of block 0
        if damage == "end-marker":
            comp[8 + first - 3] = 0x19                     # 0001 1001: a match 32768 bytes back where the end marker was
        else:
            comp[8] = 17 + 4                               # first byte: 4 literals, then whatever follows is misparsed
            comp[9:13] = b"\x00\x00\x00\x00"
        seg = torch.from_numpy(np.frombuffer(bytes(comp) + b"\0" * 16, dtype=np.uint8).copy()).to(dev)
        dec.decode(seg, len(comp), raw_len=len(plain))
        with pytest.raises(UdaFailure):
            dec.check()
    seg = torch.from_numpy(np.frombuffer(good + b"\0" * 16, dtype=np.uint8).copy()).to(dev)
    out, _ = dec.decode(seg, len(good), raw_len=len(plain))     # the decoder recovers: status was cleared by check()
    dec.check(expect_raw=len(plain))
    assert bytes(out.cpu().numpy()) == plain


@pytest.mark.parametrize("codec", ["org.apache.hadoop.io.compress.SnappyCodec", "com.hadoop.compression.lzo.LzoCodec"])
def test_compressed_mof_is_expanded_while_it_is_loaded(built, codec):
    """A Hadoop-written MOF of a job with compressed map outputs -> DeviceMOF of plain, 16-byte aligned segments with
    valid checksums, equal to what the uncompressed job would have written."""
    from uda_b200.mof import DeviceMOF
    from uda_b200.ops.compress import frame_compress_segment
    dev = _dev()
    body = _cmp_body(1)
    parts = [sorted(_tie_records(11 * p + 3, n, 1), key=body) for p, n in enumerate([900, 0, 2500, 40])]
    short = "lzo" if "lzo" in codec.lower() else "snappy"
    file_out, index = ifile.write_mof_compressed(parts, lambda plain: frame_compress_segment(plain, 4096, codec=short))
    mof = DeviceMOF.from_host_bytes(file_out, index, dev, codec=codec, codec_block=4096)
    assert mof.meta["expanded_from"] == codec and mof.num_partitions == len(parts)
    host = mof.data.cpu().numpy()
    for p, (start, raw, part) in enumerate(mof.index):
        want = ifile.encode_segment(parts[p], with_crc=True)
        assert start % 16 == 0 and raw == len(want) - 4 and part == len(want)
        assert bytes(host[start:start + part]) == want
        assert ifile.read_partition(bytes(host[:start + part]), mof.index_records()[p], verify_crc=True) == parts[p]
    # a block table sized for larger blocks than the map side used is reported, not overrun
    from uda_b200.runtime.failure import UdaFailure
    with pytest.raises(UdaFailure):
        DeviceMOF.from_host_bytes(file_out, index, dev, codec=codec, codec_block=1 << 20)


def test_compressed_segments_through_var_merger(built):
    """Compressed sorted runs -> device Snappy decode -> VarMerger == CPU oracle."""
    from uda_b200.ops.compress import SnappyDecoder, frame_compress_segment
    from uda_b200.ops.varlen import VarMerger
    dev = _dev()
    body = _cmp_body(1)
    runs, segs, keep = [], [], []
    dec = SnappyDecoder(dev, 64)
    for r, n in enumerate([800, 1200, 500]):
        recs = sorted(_tie_records(7 * r + 1, n, 1), key=body)
        runs.append(recs)
        plain = ifile.encode_segment(recs, with_crc=False)
        comp = frame_compress_segment(plain, 8192)
        t = torch.from_numpy(np.frombuffer(comp + b"\0" * 16, dtype=np.uint8).copy()).to(dev)
        out, raw = dec.decode(t, len(comp), raw_len=len(plain))
        dec.check(expect_raw=len(plain))
        keep.append(out)
        segs.append((out, raw, n))
    vm = VarMerger(dev, 1, sum(s[2] for s in segs), len(segs), sum(s[1] for s in segs))
    out, nbytes = vm.merge(segs)
    want = ifile.encode_segment(sorted([kv for run in runs for kv in run], key=body), with_crc=False)
    assert bytes(out.cpu().numpy()) == want
