"""The NCCL + torch.sort comparison arm (baseline/nccl_sort.py): its key packing and two-pass ordering, on CPU tensors."""
import numpy as np
import torch

from baseline.nccl_sort import pack_keys


def test_pack_keys_orders_like_the_key_bytes():
    g = torch.Generator().manual_seed(3)
    recs = torch.randint(0, 256, (5000, 102), dtype=torch.uint8, generator=g)
    recs[:200, 2:10] = recs[0, 2:10]                        # equal 8-byte prefixes: the tail decides
    recs[200:260, 2:12] = recs[200, 2:12]                   # fully equal keys: stable order decides
    recs[300, 2:12] = 0
    recs[301, 2:12] = 255
    w0, tail = pack_keys(recs)
    keys = recs[:, 2:12].numpy()
    # straightforward restatement
    want_w0 = np.zeros(len(keys), dtype=np.uint64)
    for b in range(8):
        want_w0 = (want_w0 << np.uint64(8)) | keys[:, b].astype(np.uint64)
    assert ((w0.numpy().view(np.uint64) ^ np.uint64(1 << 63)) == want_w0).all()
    assert (tail.numpy() == ((keys[:, 8].astype(np.int32) << 8) | keys[:, 9])).all()
    # the arm's two stable passes give the order of the raw key bytes, ties in input order
    i1 = torch.sort(tail, stable=True).indices
    perm = i1[torch.sort(w0[i1], stable=True).indices].numpy()
    want = sorted(range(len(keys)), key=lambda i: (keys[i].tobytes(), i))
    assert perm.tolist() == want


def _bench(impl, *extra):
    import json, subprocess, sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    p = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", impl, "--steps", "2", "--warmup", "1", "--records-per-gpu", "400000",
                        "--maps-per-gpu", "4", *extra], capture_output=True, text=True, timeout=600, cwd=str(root))
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, p.stdout                      # exactly one JSON line on stdout
    return json.loads(lines[0])


def test_bench_cpu_arm_prints_the_contract_line(built):
    d = _bench("cpu")
    assert d["impl"] == "cpu" and d["unit"] == "GB/s" and d["value"] > 0 and d["steps"] == 2 and d["validated"] is True
    assert d["gpu_launches"] == 0 and d["e2e"]["value"] == d["value"] and d["config"]["maps_per_gpu"] == 4


def test_bench_reference_arm_prints_the_contract_line(built):
    d = _bench("reference")
    if "unavailable" in d:                                # no reference tree on this machine: the arm says so and exits 0
        assert d["impl"] == "reference" and isinstance(d["unavailable"], str)
        return
    assert d["impl"] == "reference" and d["value"] > 0 and d["steps"] == 2 and d["validated"] is True and d["gpu_launches"] == 0
    assert d["config"]["runs_per_reducer"] == 4 and "unmodified" in d["config"]["reference_code"]


def test_reference_arm_sources_are_byte_identical_to_the_reference():
    """The reference arm compiles the reference's own files: every copied file still hashes to its manifest entry, and to the
    file in the reference tree when that tree is mounted."""
    import hashlib
    import os
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    man = root / "baseline" / "_ref" / "MANIFEST.txt"
    if not man.exists():
        import pytest
        pytest.skip("reference arm not built on this machine")
    ref = Path(os.environ.get("UDA_REFERENCE_DIR", "/root/reference"))
    entries = [l.split(None, 1) for l in man.read_text().splitlines() if l.strip()]
    assert len(entries) >= 12
    checked_against_tree = 0
    for digest, rel in entries:
        rel = rel.strip()
        copy = root / "baseline" / "_ref" / rel
        assert hashlib.sha256(copy.read_bytes()).hexdigest() == digest, rel
        orig = ref / rel
        if orig.exists():
            assert hashlib.sha256(orig.read_bytes()).hexdigest() == digest, rel
            checked_against_tree += 1
    assert checked_against_tree == len(entries) or not ref.exists()
    assert any(rel.strip().endswith("Merger/StreamRW.cc") for _, rel in entries) and any("MergeQueue.h" in rel for _, rel in entries)
