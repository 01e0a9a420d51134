"""CPU restatements used as oracles and data sources: teragen, TeraValidate checks, the WordCount record generator and its
Java-compatible hash partitioner, measured-peaks lookup, metrics writer, max-over-ranks reduction (gloo, 2 ranks)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

from uda_b200.models import teragen as tg
from uda_b200.oracle.partition import hash_partition, java_hash_bytes
from uda_b200.oracle.validate import record_checksum, teravalidate

ROOT = Path(__file__).resolve().parent.parent


def test_teragen_is_a_pure_function_of_seed_and_row():
    a = tg.teragen_raw(1000, 31, first_row=0)
    b = tg.teragen_raw(400, 31, first_row=600)
    assert a.shape == (1000, 100) and a.dtype == np.uint8
    assert (a[600:] == b).all()                                   # any slice of the table can be generated on its own
    assert not (tg.teragen_raw(1000, 32)[:, :10] == a[:, :10]).all()
    rows = np.zeros(1000, dtype=np.uint64)
    for j in range(10, 18):
        rows = (rows << np.uint64(8)) | a[:, j].astype(np.uint64)
    assert (rows == np.arange(1000, dtype=np.uint64)).all()       # bytes 10..17 of the value carry the row number
    assert len({bytes(k) for k in a[:, :10]}) == 1000
    f = tg.teragen_framed(1000, 31)
    assert f.shape == (1000, 102) and (f[:, 0] == 10).all() and (f[:, 1] == 90).all() and (f[:, 2:] == a).all()


def test_teragen_key_bits_limits_the_distinct_keys():
    a = tg.teragen_raw(5000, 7, key_bits=3)
    keys = {bytes(k) for k in a[:, :10]}
    assert len(keys) <= 8 and all(k[1:] == b"\0" * 9 for k in keys)


def test_fingerprint_is_order_independent_and_content_sensitive():
    a = tg.teragen_raw(2000, 5)
    perm = np.random.default_rng(0).permutation(2000)
    assert tg.fingerprint_sum(a) == tg.fingerprint_sum(a[perm])
    b = a.copy()
    b[17, 99] ^= 1
    assert tg.fingerprint_sum(a) != tg.fingerprint_sum(b)
    assert tg.fingerprint_sum(a[:1999]) != tg.fingerprint_sum(a)


def test_teravalidate_accepts_sorted_partitions_and_names_the_fault():
    a = tg.teragen_raw(3000, 11)
    recs = sorted((bytes(r[:10]), bytes(r[10:])) for r in a)
    parts = [recs[:1000], recs[1000:1800], recs[1800:]]
    teravalidate(parts, record_checksum(recs))
    swapped = [parts[1], parts[0], parts[2]]
    with pytest.raises(AssertionError, match="misordered"):
        teravalidate(swapped, record_checksum(recs))
    with pytest.raises(AssertionError, match="multiset"):
        teravalidate([recs[:1000], recs[1000:1800], recs[1800:-1]], record_checksum(recs))
    dup = [recs[:1000], recs[1000:1800], recs[1800:-1] + [recs[-2]]]   # same count and bytes, one record replaced
    with pytest.raises(AssertionError, match="multiset"):
        teravalidate(dup, record_checksum(recs))


def test_wordcount_records_and_hash_partitioner_match_the_oracle():
    from uda_b200.models import wordcount as wc
    ids = wc.zipf_word_ids(20000, 5000, 1.1, 3, torch.device("cpu"))
    assert ids.min() >= 0 and ids.max() < 5000
    counts = torch.bincount(ids, minlength=5000)
    assert counts[0] > counts[10] > counts[1000]                   # Zipf: frequency falls with the rank
    assert (wc.zipf_word_ids(20000, 5000, 1.1, 3, torch.device("cpu")) == ids).all()
    recs = wc.encode_records(ids[:50], torch.arange(50) * 1000 + 1).numpy()
    assert recs.shape == (50, 16)
    for i in range(50):
        r = recs[i]
        assert r[0] == 10 and r[1] == 4 and r[2] == 9              # VInt(10) VInt(4), Text length 9
        assert bytes(r[3:12]) == b"%09d" % int(ids[i])
        assert int.from_bytes(bytes(r[12:16]), "big") == i * 1000 + 1
    # the vectorised Java hash of the 9 digits == WritableComparator.hashBytes of the same bytes
    h = wc.java_hash_of_word_ids(ids[:200].numpy())
    for i in range(200):
        word = b"%09d" % int(ids[i])
        assert int(h[i]) == java_hash_bytes(word) & 0xFFFFFFFF
        assert (int(h[i]) & 0x7FFFFFFF) % 8 == hash_partition(word, 8)


def test_measured_peaks_fallback_and_file(tmp_path, monkeypatch):
    from uda_b200.utils import trace
    monkeypatch.setattr(trace, "_ROOT", tmp_path)
    assert trace.measured_peaks() == {"hbm_gbs": 6541.5, "nvlink_gbs": 770.0}
    (tmp_path / "MEASURED_PEAKS.json").write_text(json.dumps({"hbm_copy_gbs": 6400.0, "nvlink_p2p_gbs": 760.0}))
    assert trace.measured_peaks() == {"hbm_gbs": 6400.0, "nvlink_gbs": 760.0}
    (tmp_path / "MEASURED_PEAKS.json").write_text("not json")
    assert trace.measured_peaks()["hbm_gbs"] == 6541.5
    p = trace.write_metrics("unit", {"a": 1}, directory=str(tmp_path / "out"))
    doc = json.loads(p.read_text())
    assert doc["a"] == 1 and "written_at" in doc


def test_reduce_max_over_ranks_two_processes(tmp_path):
    script = tmp_path / "r.py"
    script.write_text(f"""
import sys, json
sys.path.insert(0, {str(ROOT)!r})
import torch, torch.distributed as dist
from uda_b200.utils.trace import reduce_max_over_ranks
dist.init_process_group("gloo")
r = dist.get_rank()
s = reduce_max_over_ranks({{"push": {{"ms": 1.0 + r, "calls": 1}}, "merge": {{"ms": 5.0 - r, "calls": 2}}}}, torch.device("cpu"))
if r == 0:
    print("RES " + json.dumps(s))
dist.destroy_process_group()
""")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", str(script)], capture_output=True, text=True, timeout=300,
                       env={**os.environ, "MASTER_ADDR": "127.0.0.1"})
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    line = next(l for l in p.stdout.splitlines() if l.startswith("RES "))
    s = json.loads(line[4:])
    assert s["push"]["ms_max_over_ranks"] == 2.0 and s["merge"]["ms_max_over_ranks"] == 5.0 and s["push"]["ms"] == 1.0
