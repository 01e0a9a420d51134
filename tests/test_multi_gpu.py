"""Multi-GPU tests (need >= 2 visible GPUs; run with `gpurun --gpus N`). One process per GPU via torchrun."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(n, script, *args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + n), os.path.join(ROOT, "tests", "scripts", script), *map(str, args)]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


@pytest.mark.parametrize("world,subparts,mode", [(2, 1, "lsu"), (2, 3, "lsu"), (4, 1, "lsu"), (8, 1, "lsu"), (8, 4, "lsu"),
                                                 (2, 1, "tma"), (2, 4, "tma"), (8, 4, "tma")])
def test_terasort_multi_rank(built, world, subparts, mode):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    p = _torchrun(world, "mp_terasort.py", 300_000 if mode == "lsu" else 2_000_000, 3, 3, subparts, mode)
    text = p.stdout + "\n" + p.stderr
    assert p.returncode == 0, text[-6000:]
    assert text.count("MPRESULT") == world and text.count('"ok": true') == world, text[-6000:]


@pytest.mark.parametrize("world,subparts,slots,local_copy", [(2, 8, 2, "ce"), (2, 8, 2, "push"), (2, 12, 3, "ce"), (8, 12, 3, "ce")])
def test_terasort_receive_ring(built, world, subparts, slots, local_copy):
    """Bounded receive memory: the arena holds `slots` of the `subparts` sub-partitions (a quarter of the partition),
    senders wait for credits before they refill a slot. Several steps, so slots and credits are reused across steps; the
    output must be the full sorted partition on every rank."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    p = _torchrun(world, "mp_terasort.py", 2_000_000, 3, 4, subparts, "tma", 0, slots, local_copy)
    text = p.stdout + "\n" + p.stderr
    assert p.returncode == 0, text[-6000:]
    assert text.count("MPRESULT") == world and text.count('"ok": true') == world, text[-6000:]
    assert text.count(f'"slots": {slots}') == world


def test_terasort_local_zero_copy_two_ranks(built):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    p = _torchrun(2, "mp_terasort.py", 500_000, 3, 3, 4, "tma", 1)
    text = p.stdout + "\n" + p.stderr
    assert p.returncode == 0, text[-6000:]
    assert text.count("MPRESULT") == 2 and text.count('"ok": true') == 2, text[-6000:]


@pytest.mark.parametrize("world", [2, 8])
def test_wordcount_multi_rank(built, world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    p = _torchrun(world, "mp_wordcount.py", 300_000, 30_000)
    text = p.stdout + "\n" + p.stderr
    assert p.returncode == 0, text[-6000:]
    assert text.count("MPRESULT") == world and text.count('"ok": true') == world, text[-6000:]


@pytest.mark.parametrize("world", [2, 8])
def test_side_table_broadcast(built, world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    p = _torchrun(world, "mp_bcast.py")
    text = p.stdout + "\n" + p.stderr
    assert p.returncode == 0, text[-6000:]
    assert text.count('"ok": true') == world, text[-6000:]


@pytest.mark.parametrize("world,kind,side", [(2, 2, 1), (2, 1, 0), (8, 2, 1)])
def test_sort_job_multi_rank(built, world, kind, side):
    """Variable-length records, BytesWritable (2) / Text (1) keys, 5 % of the keys tie on the sort prefix."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    p = _torchrun(world, "mp_sort.py", 40_000, kind, side)
    text = p.stdout + "\n" + p.stderr
    assert p.returncode == 0, text[-6000:]
    assert text.count("MPRESULT") == world and text.count('"ok": true') == world, text[-6000:]
