"""Host-side range refinement of the spill merger (skewed keys): pure numpy, no GPU."""
import random

import numpy as np
import pytest

from uda_b200.ops.spill import split_range

S, KOFF, KLEN = 24, 2, 10


def _group(rng, n, heavy, frac):
    """n sorted records of stride S; `frac` of them carry one of the `heavy` keys."""
    keys = []
    for _ in range(n):
        if heavy and rng.random() < frac:
            keys.append(rng.choice(heavy))
        else:
            keys.append(bytes(rng.randrange(256) for _ in range(KLEN)))
    keys.sort()
    buf = np.zeros(n * S, dtype=np.uint8)
    for i, k in enumerate(keys):
        buf[i * S + KOFF:i * S + KOFF + KLEN] = np.frombuffer(k, dtype=np.uint8)
        buf[i * S + KOFF + KLEN:i * S + S] = rng.randrange(256)     # payload
    return buf


def _keys(buf, a, b):
    return [buf[i * S + KOFF:i * S + KOFF + KLEN].tobytes() for i in range(a, b)]


@pytest.mark.parametrize("sizes,heavy_n,frac,cap", [
    ([400, 300, 500], 0, 0.0, 100),          # no skew, small capacity: pure median halving
    ([400, 300, 500], 2, 0.8, 100),          # two heavy keys hold 80 %: one-key parts are cut by position
    ([1000], 1, 1.0, 64),                    # a single key everywhere, one group
    ([50, 0, 70, 3], 1, 0.5, 1000),          # fits: returned unchanged
    ([200, 200], 3, 0.95, 1),                # capacity of one record
    ([0, 0], 0, 0.0, 10),                    # nothing
])
def test_split_range_partitions_in_key_order(sizes, heavy_n, frac, cap):
    rng = random.Random(len(sizes) * 1000 + cap)
    heavy = [bytes(rng.randrange(256) for _ in range(KLEN)) for _ in range(heavy_n)]
    if heavy_n >= 2:
        heavy[1] = heavy[0][:-1] + bytes([(heavy[0][-1] + 1) % 256])     # neighbours in key space
    groups = [_group(rng, n, heavy, frac) for n in sizes]
    tasks = split_range(groups, [(g, 0, n) for g, n in enumerate(sizes)], S, KOFF, KLEN, cap)
    total = sum(sizes)
    if total == 0:
        assert tasks == []
        return
    if total <= cap:
        assert len(tasks) == 1
    # every record is covered exactly once, pieces of one group are consecutive and in order
    nxt = [0] * len(sizes)
    prev_max = None
    for t in tasks:
        assert t and sum(b - a for _, a, b in t) <= cap
        assert len({g for g, _, _ in t}) == len(t)                       # a group appears once per task
        ks = []
        for g, a, b in t:
            assert a == nxt[g] and b > a
            nxt[g] = b
            ks += _keys(groups[g], a, b)
        if prev_max is not None:
            assert min(ks) >= prev_max                                    # key order across tasks
        prev_max = max(ks)
    assert nxt == list(sizes)
    # merging every task and concatenating is the globally sorted key sequence
    merged = []
    for t in tasks:
        merged += sorted(k for g, a, b in t for k in _keys(groups[g], a, b))
    assert merged == sorted(k for g, n in enumerate(sizes) for k in _keys(groups[g], 0, n))


def test_split_range_sub_slices():
    """The range to split is a window of every group, not the whole group."""
    rng = random.Random(5)
    groups = [_group(rng, 300, [], 0.0) for _ in range(3)]
    window = [(0, 50, 250), (1, 0, 120), (2, 299, 300)]
    # a window must be a key range: cut all groups at two keys
    lo, hi = sorted([_keys(groups[0], 60, 61)[0], _keys(groups[0], 200, 201)[0]])
    window = []
    for g in range(3):
        ks = _keys(groups[g], 0, 300)
        a = next((i for i, k in enumerate(ks) if k >= lo), 300)
        b = next((i for i, k in enumerate(ks) if k >= hi), 300)
        window.append((g, a, b))
    tasks = split_range(groups, window, S, KOFF, KLEN, 37)
    got = [k for t in tasks for k in sorted(k for g, a, b in t for k in _keys(groups[g], a, b))]
    assert got == sorted(k for g, a, b in window for k in _keys(groups[g], a, b))
    assert all(sum(b - a for _, a, b in t) <= 37 for t in tasks)


def test_split_range_property():
    """Random group counts, sizes, duplicate densities and capacities: every record exactly once, every part within the
    capacity, parts in key order."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(seed=st.integers(0, 10**6), ngroups=st.integers(1, 6), cap=st.integers(1, 400), alphabet=st.integers(1, 40))
    def run(seed, ngroups, cap, alphabet):
        rng = random.Random(seed)
        sizes = [rng.randint(0, 300) for _ in range(ngroups)]
        pool = [bytes(rng.randrange(256) for _ in range(KLEN)) for _ in range(alphabet)]     # few distinct keys: heavy duplicates
        groups = []
        for n in sizes:
            keys = sorted(rng.choice(pool) for _ in range(n))
            buf = np.zeros(n * S, dtype=np.uint8)
            for i, k in enumerate(keys):
                buf[i * S + KOFF:i * S + KOFF + KLEN] = np.frombuffer(k, dtype=np.uint8)
            groups.append(buf)
        tasks = split_range(groups, [(g, 0, n) for g, n in enumerate(sizes)], S, KOFF, KLEN, cap)
        nxt = [0] * ngroups
        prev_max = None
        for t in tasks:
            assert t and sum(b - a for _, a, b in t) <= cap
            ks = []
            for g, a, b in t:
                assert a == nxt[g] and b > a
                nxt[g] = b
                ks += _keys(groups[g], a, b)
            assert prev_max is None or min(ks) >= prev_max
            prev_max = max(ks)
        assert nxt == sizes

    run()
