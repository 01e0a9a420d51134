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


def _var_index(world, maps, seed=0):
    """Index tables with arbitrary (non-stride) segment sizes, 16-byte aligned starts."""
    rng = np.random.RandomState(seed)
    idx = np.zeros((world, maps, world, 3), dtype=np.int64)
    for s in range(world):
        for m in range(maps):
            cur = 0
            for p in range(world):
                body = int(rng.randint(0, 5000))
                idx[s, m, p] = (cur, body + 2, body + 6)
                cur = (cur + body + 6 + 15) // 16 * 16
    return idx


def test_variable_length_plan_and_sub_partitions():
    """stride 0 = variable-length job: no record counts, byte-exact coverage; sub-partitions share one arena and
    every (sub, run) owns its own landed-byte counter slot."""
    world, maps, C = 3, 2, 4
    rng = np.random.RandomState(5)
    idx = np.zeros((world, maps, world * C, 3), dtype=np.int64)
    for s in range(world):
        for m in range(maps):
            cur = 0
            for lp in range(world * C):
                body = int(rng.randint(0, 3000))
                idx[s, m, lp] = (cur, body + 2, body + 6)
                cur = (cur + body + 6 + 15) // 16 * 16
    plan = ShufflePlan(idx, record_stride=0, subparts=C)
    assert plan.P == world and plan.nruns == world * maps and len(plan.layouts) == world * C
    for r in range(world):
        spans = []
        for j in range(C):
            lay = plan.layout(r, j)
            assert (lay.run_nrec == -1).all()
            for rid in range(plan.nruns):
                spans.append((int(lay.run_offsets[rid]), int(lay.run_offsets[rid] + lay.run_bytes[rid])))
        spans.sort()
        assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))          # nothing overlaps inside the arena
        assert spans[-1][1] + 32 <= plan.arena_bytes[r]
    flags = set()
    total = 0
    for j in range(C):
        d = plan.push_descs(0, [1 << 20, 2 << 20], [10 << 20, 20 << 20, 30 << 20], [1 << 30, 2 << 30, 3 << 30], j)
        total += int(d["bytes"].sum())
        flags |= {int(f) for f in d["flag"]}
        # the descriptors addressed to the rank itself come last: that is what local_zero_copy relies on
        assert all(int(x) // (10 << 20) == 1 for x in d["dst"][-maps:])
    assert total == int(idx[0, :, :, 2].sum()) and len(flags) == C * world * maps


def test_plan_is_identical_on_every_rank():
    """Every rank derives layouts and peer offsets from the same all-gathered tables: no addresses are exchanged."""
    idx = _var_index(4, 3, seed=9)
    a, b = ShufflePlan(idx.copy(), 0), ShufflePlan(idx.copy(), 0)
    assert a.arena_bytes == b.arena_bytes
    for la, lb in zip(a.layouts, b.layouts):
        assert (la.run_offsets == lb.run_offsets).all() and (la.run_bytes == lb.run_bytes).all()
    # what rank 2 pushes to rank 1 lands exactly where rank 1 expects run (2, m)
    d = a.push_descs(2, [0, 0, 0], [0, 1 << 32, 0, 0], [0, 0, 0, 0])
    to1 = [x for x in d if int(x["dst"]) >> 32 == 1]
    lay1 = b.layout(1)
    assert sorted(int(x["dst"]) - (1 << 32) for x in to1) == sorted(int(lay1.run_offsets[a.run_id(2, m)]) for m in range(3))


def _sub_index(world, maps, subparts, seed=0, stride=102):
    rng = np.random.RandomState(seed)
    idx = np.zeros((world, maps, world * subparts, 3), dtype=np.int64)
    for s in range(world):
        for m in range(maps):
            cur = 0
            for p in range(world * subparts):
                body = int(rng.randint(0, 80)) * stride
                idx[s, m, p] = (cur, body + 2, body + 6)
                cur = (cur + body + 6 + 15) // 16 * 16
    return idx


def test_receive_ring_layout_reuses_slots_and_bounds_the_arena():
    """`slots < subparts`: sub-partition j lives in slot j % slots; sub-partitions that share a slot overlap in memory (they
    never coexist), sub-partitions in different slots never do, and the arena shrinks to about slots / subparts."""
    world, maps, C, S = 3, 2, 8, 2
    idx = _sub_index(world, maps, C, seed=5)
    full = ShufflePlan(idx, record_stride=102, subparts=C)
    ring = ShufflePlan(idx, record_stride=102, subparts=C, slots=S)
    assert full.slots == C and ring.slots == S
    assert ring.max_arena_bytes() < 0.5 * full.max_arena_bytes()
    for r in range(world):
        spans = []
        for j in range(C):
            lay = ring.layout(r, j)
            lo = int(lay.run_offsets.min())
            hi = int((lay.run_offsets + lay.run_bytes).max())
            spans.append((lo, hi))
            assert hi + 32 <= ring.arena_bytes[r]
            # the same run sizes and record counts as in the full plan, phase preserved
            assert (lay.run_bytes == full.layout(r, j).run_bytes).all() and (lay.run_nrec == full.layout(r, j).run_nrec).all()
            for s in range(world):
                for m in range(maps):
                    assert lay.run_offsets[ring.run_id(s, m)] % 16 == idx[s, m, r * C + j, 0] % 16
        for a in range(C):
            for b in range(a + 1, C):
                disjoint = spans[a][1] <= spans[b][0] or spans[b][1] <= spans[a][0]
                if a % S != b % S:
                    assert disjoint, (a, b, spans[a], spans[b])
    # slots >= subparts or 0 means "whole partition"
    assert ShufflePlan(idx, record_stride=102, subparts=C, slots=C + 3).slots == C
    assert ShufflePlan(idx, record_stride=102, subparts=C, slots=0).max_arena_bytes() == full.max_arena_bytes()


def test_receive_ring_descriptors_point_into_the_slots():
    world, maps, C, S = 2, 3, 6, 2
    idx = _sub_index(world, maps, C, seed=9)
    ring = ShufflePlan(idx, record_stride=102, subparts=C, slots=S)
    base = [1 << 40, 2 << 40]
    flags = [3 << 40, 4 << 40]
    for rank in range(world):
        for j in range(C):
            d = ring.push_descs(rank, [100_000 * (m + 1) for m in range(maps)], base, flags, j)
            for i in range(len(d)):
                p = (rank + 1 + i // maps) % world
                off = int(d[i]["dst"]) - base[p]
                assert 0 <= off and off + int(d[i]["bytes"]) <= ring.arena_bytes[p]
                # landed-byte counters are per (sub-partition, run): never recycled with the slots
                assert int(d[i]["flag"]) == flags[p] + 8 * (j * ring.nruns + ring.run_id(rank, i % maps))
