"""The dataflow schedule of a sharded scan (csrc/capi_dataflow.cu.inc), host side only: every (stage, view) task is
placed exactly once, a task sits `lag` rounds after every task it reads from (ACMMP.cpp:608-635: own view + sources of
the previous stage), no task overwrites a record generation that a task of a later round still reads, and the round
count beats the staged schedule's stage barrier (49 views on 8 GPUs: 37 rounds instead of 6 x 7)."""
import numpy as np
import pytest

from acmmp_b200 import host


def ring(n, k=10):
    return [[(v + d) % n for d in list(range(-k // 2, 0)) + list(range(1, k // 2 + 1))] for v in range(n)]


def nearest(n, k=10):
    return [sorted(range(n), key=lambda u: (abs(u - v), u))[1:k + 1] for v in range(n)]


def random_pairs(n, k, seed):
    rng = np.random.default_rng(seed)
    return [list(rng.choice([u for u in range(n) if u != v], size=min(k, n - 1), replace=False)) for v in range(n)]


def check_plan(pairs, levels, world, photometric_only=False):
    n = len(pairs)
    per_level = 1 if photometric_only else 3
    rounds, ranks, n_rounds, lag = host.plan_dataflow(pairs, levels, world, photometric_only)
    stages = levels * per_level
    assert len(rounds) == stages * n and rounds.min() == 0 and rounds.max() == n_rounds - 1
    assert 1 <= lag <= 3 and (world > 1 or lag == 1)
    slots = set(zip(rounds.tolist(), ranks.tolist()))
    assert len(slots) == len(rounds) and ranks.min() >= 0 and ranks.max() < world      # one task per (round, rank)
    users = [[u for u in range(n) if v in pairs[u]] for v in range(n)]
    for g in range(1, stages):
        kind = 0 if photometric_only else g % 3
        prev_kind = 0 if photometric_only else (g - 1) % 3
        for v in range(n):
            r = rounds[g * n + v]
            reads = [v] + (list(pairs[v]) if kind != 0 else [])
            for u in reads:                                                              # true dependencies
                assert rounds[(g - 1) * n + u] + lag <= r, (g, v, u)
            if prev_kind != 0:                                                           # readers of the generation (g, v) overwrites
                for u in users[v]:
                    assert rounds[(g - 1) * n + u] <= r, (g, v, u)
    # work, not task count, is what is even across ranks: within one costliest task of the mean
    cost = np.array([(4.0 ** (g // per_level)) * (2.0 if (not photometric_only and g % 3 == 0) else 1.0) for g in range(stages) for _ in range(n)])
    load = np.array([cost[ranks == r].sum() for r in range(world)])
    assert load.max() - load.min() <= cost.max() + 1e-6, load
    return n_rounds


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_plan_is_valid_and_dense_for_the_headline_shape(world):
    for pairs in (ring(49), nearest(49)):
        n_rounds = check_plan(pairs, 2, world)
        assert n_rounds == -(-6 * 49 // world)                     # no round lost to a stage barrier
        assert n_rounds <= 6 * -(-49 // world)


def test_plan_other_shapes():
    assert check_plan(nearest(300, 20), 2, 8) == 225               # configs[4]
    assert check_plan(nearest(11, 10), 3, 8) <= 16                 # configs[3]: every view pairs with every other: 2 rounds per stage,
    assert check_plan(nearest(11, 10), 3, 1) == 99                 #             geom-1 -> next level's photometric stage overlap
    assert check_plan(ring(49), 1, 8, photometric_only=True) == 7  # configs[1]: independent tasks
    for seed in range(4):                                          # asymmetric pair graphs
        check_plan(random_pairs(23, 6, seed), 2, 4)
        check_plan(random_pairs(9, 3, seed), 3, 8)
    assert check_plan([[1], [0]], 1, 2) == 3                       # 2 views on 2 ranks: the three stages stay a chain


def test_plan_rejects_bad_arguments():
    from acmmp_b200 import lib as L
    import ctypes as C
    lib = L.load()
    off = (C.c_int32 * 3)(0, 1, 2)
    ids = (C.c_int32 * 2)(1, 7)
    assert lib.acmmp_plan_dataflow(2, off, ids, 1, 0, 2, None, None, None, None, None) != 0
    ids = (C.c_int32 * 2)(1, 0)
    assert lib.acmmp_plan_dataflow(2, off, ids, 0, 0, 2, None, None, None, None, None) != 0
    assert lib.acmmp_plan_dataflow(2, off, ids, 1, 0, 0, None, None, None, None, None) != 0
