"""CPU tests of the CI-test pooling behind obn_ci_tests: the planner and the
kernel's marginalisation passes, both fetched through the C ABI and replayed in numpy against
tables counted directly by the checker.  The device path is covered by tests/test_gpu_kernels.py."""
import itertools

import numpy as np
import pytest

import oracle
from helpers import ranked
from openbnsl_b200 import capi, networks as nw


def frontier(rng, n_vars, n_pairs, depth, pool_size):
    """A PC-like batch: for each pair, every `depth`-subset of a random neighbourhood (pc.cpp:27-47)."""
    xs, ys, tuples = [], [], []
    for _ in range(n_pairs):
        x, y = (int(v) for v in rng.choice(n_vars, size=2, replace=False))
        others = [v for v in range(n_vars) if v not in (x, y)]
        neigh = sorted(int(v) for v in rng.choice(others, size=min(pool_size, len(others)), replace=False))
        for z in itertools.combinations(neigh, depth):
            xs.append(x); ys.append(y); tuples.append(sorted(list(z) + [x, y]))
    return xs, ys, tuples


def kernel_order(pool, x, y):
    return [v for v in pool if v not in (x, y)] + [min(x, y), max(x, y)]


def dense_in_order(ods, cards, order):
    """Table over `order` (first variable most significant), from the checker's ascending-id table."""
    asc = sorted(order)
    t = np.asarray(oracle.count_table(ods, asc)).reshape([cards[v] for v in asc])
    return np.transpose(t, [asc.index(v) for v in order]).reshape(-1)


@pytest.mark.parametrize("shuffled", [False, True])
@pytest.mark.parametrize("netf,depth,pool_size,max_cells", [(nw.alarm, 1, 5, 4096), (nw.alarm, 2, 6, 4096),
                                                           (nw.alarm, 3, 5, 1024), (nw.munin1_like, 1, 4, 2048),
                                                           (nw.asia, 2, 6, 64)])
def test_plan_and_marginalisation_reproduce_direct_tables(netf, depth, pool_size, max_cells, shuffled):
    net = netf()
    codes, cards = ranked(net.sample(1500, 3))
    ods = oracle.Dataset(codes, cards)
    rng = np.random.default_rng(depth * 100 + pool_size)
    xs, ys, tuples = frontier(rng, net.n, 12, depth, pool_size)
    if shuffled:        # the tests of a pair no longer arrive back to back: the pair table, not the fast path, finds the pool
        order = rng.permutation(len(tuples))
        xs, ys, tuples = [xs[i] for i in order], [ys[i] for i in order], [tuples[i] for i in order]
    pools, pool_of, keep = capi.ci_pool_plan(cards, tuples, xs, ys, max_cells)
    assert len(pools) <= len(tuples)
    shared = 0
    members = {}
    for t, q in enumerate(pool_of):
        members.setdefault(int(q), []).append(t)
    assert sorted(members) == list(range(len(pools)))                  # every pool has a member, ids are dense
    for q, pool in enumerate(pools):
        x, y = min(xs[members[q][0]], ys[members[q][0]]), max(xs[members[q][0]], ys[members[q][0]])
        assert pool == sorted(set(pool)) and x in pool and y in pool
        order = kernel_order(pool, x, y)
        zs = order[:-2]
        cells = int(np.prod([cards[v] for v in pool]))
        if len(members[q]) > 1:
            assert cells <= max_cells
            shared += len(members[q])
        joint = dense_in_order(ods, cards, order)
        for t in members[q]:
            assert (min(xs[t], ys[t]), max(xs[t], ys[t])) == (x, y)       # one pair per pool
            kept = [zs[i] for i in range(len(zs)) if (int(keep[t]) >> i) & 1]
            assert sorted(kept + [x, y]) == tuples[t]                   # the mask names exactly the test's Z
            got = capi.ci_pool_marginalise(joint, [cards[v] for v in zs], int(keep[t]))
            want = dense_in_order(ods, cards, kept + [x, y])
            assert got.shape == want.shape and (got == want).all(), (q, t)
    assert shared >= len(tuples) // 2 or max_cells <= 64                 # the frontier shape does pool


def test_plan_edge_cases():
    cards = [2, 3, 1, 4, 2, 3]
    # empty batch
    pools, pool_of, keep = capi.ci_pool_plan(cards, [], [], [])
    assert pools == [] and len(pool_of) == 0
    # x > y is normalised, tests of different pairs never share, order of first appearance is kept
    tuples = [[0, 1, 3], [0, 1, 4], [1, 3, 4], [0, 1, 5], [0, 1]]
    xs, ys = [1, 0, 4, 0, 1], [0, 1, 3, 1, 0]
    pools, pool_of, keep = capi.ci_pool_plan(cards, tuples, xs, ys, 4096)
    assert pools == [[0, 1, 3, 4, 5], [1, 3, 4]]
    assert pool_of.tolist() == [0, 0, 1, 0, 0]
    assert keep.tolist() == [0b001, 0b010, 0b1, 0b100, 0]
    # a cap below the union splits the pool; a test above the cap stays alone
    pools, pool_of, keep = capi.ci_pool_plan(cards, tuples, xs, ys, 2 * 3 * 4 * 2)
    assert pools == [[0, 1, 3, 4], [1, 3, 4], [0, 1, 5]] and pool_of.tolist() == [0, 0, 1, 2, 2]
    pools, pool_of, keep = capi.ci_pool_plan(cards, tuples, xs, ys, 7)
    assert len(pools) == 5 and pool_of.tolist() == [0, 1, 2, 3, 4]       # tables above the cap stay alone, never open a pool
    # at most max_z conditioning variables per pool
    pools, pool_of, keep = capi.ci_pool_plan(cards, tuples, xs, ys, 4096, max_z=2)
    assert pools == [[0, 1, 3, 4], [1, 3, 4], [0, 1, 5]] and pool_of.tolist() == [0, 0, 1, 2, 2]
    # at most 32 members per pool (the kernel keeps their statistics in shared memory)
    many = [[0, 1, 3]] * 70
    pools, pool_of, keep = capi.ci_pool_plan(cards, many, [0] * 70, [1] * 70, 4096)
    assert len(pools) == 3 and np.bincount(pool_of).tolist() == [32, 32, 6]
    # a repeated id is passed through untouched
    pools, pool_of, keep = capi.ci_pool_plan(cards, [[0, 1, 3, 3], [0, 1, 4]], [0, 0], [1, 1], 4096)
    assert pools == [[0, 1, 3, 3], [0, 1, 4]] and keep.tolist() == [0xffffffff, 1]
    # the batch checks of obn_ci_tests, same classes and messages
    with pytest.raises(ValueError, match="x and y must differ"):
        capi.ci_pool_plan(cards, [[0, 1]], [0], [0])
    with pytest.raises(ValueError, match="x or y not contained"):
        capi.ci_pool_plan(cards, [[0, 1]], [0], [2])
    with pytest.raises(ValueError, match="must be sorted"):
        capi.ci_pool_plan(cards, [[1, 0]], [0], [1])
    with pytest.raises(IndexError):
        capi.ci_pool_plan(cards, [[0, 9]], [0], [9])


def test_marginalise_one_state_digits_and_identity():
    rng = np.random.default_rng(0)
    cards = [3, 1, 2, 1, 4]                       # conditioning digits (kernel order), then x, y
    rx, ry = 2, 3
    joint = rng.integers(0, 50, size=int(np.prod(cards)) * rx * ry).astype(np.int32)
    full = joint.reshape(cards + [rx, ry])
    for keep in range(32):
        got = capi.ci_pool_marginalise(joint, cards, keep)
        axes = tuple(i for i in range(5) if not (keep >> i) & 1)
        want = full.sum(axis=axes).reshape(-1)
        assert (got == want).all(), keep
    assert (capi.ci_pool_marginalise(joint, cards, 0xffffffff) == joint).all()
    # buffer rule of the kernel: the first pass output is <= cells/2, the second <= cells/4 (one-state digits skipped)
    sizes, cur = [], joint.size
    for i, c in enumerate(cards):
        if c > 1:
            cur //= c
            sizes.append(cur)
    assert sizes[0] <= joint.size // 2 and sizes[1] <= joint.size // 4
