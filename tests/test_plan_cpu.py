"""CPU tests of the host-side shuffle plan (layout determinism, descriptor coverage, peer rotation)."""
import numpy as np

from uda_b200.parallel.shuffle import ShufflePlan


def _index(world, maps, seed=0, stride=102):
    rng = np.random.RandomState(seed)
    idx = np.zeros((world, maps, world, 3), dtype=np.int64)
    for s in range(world):
        for m in range(maps):
            cur = 0
            for p in range(world):
                n = int(rng.randint(0, 50))
                body = n * stride
                idx[s, m, p] = (cur, body + 2, body + 6)
                cur = (cur + body + 6 + 15) // 16 * 16
    return idx


def test_layout_non_overlapping_and_phase_preserving():
    plan = ShufflePlan(_index(4, 3), record_stride=102)
    assert plan.nruns == 12
    for r in range(4):
        lay = plan.layouts[r]
        order = np.argsort(lay.run_offsets)
        ends = lay.run_offsets[order] + lay.run_bytes[order]
        assert (lay.run_offsets[order][1:] >= ends[:-1]).all()
        assert lay.total_bytes >= ends.max() + 32
        for s in range(4):
            for m in range(3):
                rid = plan.run_id(s, m)
                assert lay.run_offsets[rid] % 16 == plan.index[s, m, r, 0] % 16
                assert lay.run_nrec[rid] == (plan.index[s, m, r, 1] - 2) // 102


def test_push_descs_cover_everything_and_rotate_peers():
    world, maps = 4, 2
    plan = ShufflePlan(_index(world, maps, seed=3), record_stride=102)
    peer_data = [1_000_000 * (p + 1) for p in range(world)]
    peer_flag = [50_000_000 + 10_000 * p for p in range(world)]
    for rank in range(world):
        d = plan.push_descs(rank, [10_000 * (m + 1) for m in range(maps)], peer_data, peer_flag)
        assert len(d) == world * maps
        # first block of descriptors targets rank+1, the last block targets the rank itself
        first_dst = int(d[0]["dst"]) // 1_000_000 - 1
        last_dst = int(d[-1]["dst"]) // 1_000_000 - 1
        assert first_dst == (rank + 1) % world and last_dst == rank
        assert int(d["bytes"].sum()) == int(plan.index[rank, :, :, 2].sum())
        flags = {int(f) for f in d["flag"]}
        assert len(flags) == world * maps
    rd, off, base, nbytes = plan.run_descs(1, 7_000_000)
    assert off[-1] == rd["nrec"].sum() and (base == 7_000_000 + plan.layouts[1].run_offsets).all()
    assert (nbytes == plan.layouts[1].run_bytes).all()
