"""CPU tests of the table sharing behind obn_ci_tests / obn_pc_round (obn_ci_groups.inl, DESIGN.md
3.5): the host planner through the C ABI's diagnostics entry point (needs no device), and the
arithmetic it relies on -- one table in reference key order serves every (x, y | rest) question over
its variable set -- restated in numpy against the oracle."""
import itertools

import numpy as np
import pytest

import oracle
from helpers import ranked
from openbnsl_b200 import capi
from openbnsl_b200 import networks as nw


def star_frontier(rng, n_vars, n_centres, depth, degree):
    """The shape of a PC depth: for a centre x with neighbours N, every (x, y | Z), y in N, Z a
    depth-subset of N - {y} (pc.cpp:27-47)."""
    xs, ys, tuples = [], [], []
    for _ in range(n_centres):
        x = int(rng.integers(n_vars))
        neigh = sorted(int(v) for v in rng.choice([v for v in range(n_vars) if v != x], size=degree, replace=False))
        for y in neigh:
            for z in itertools.combinations([v for v in neigh if v != y], depth):
                xs.append(x); ys.append(y); tuples.append(sorted(list(z) + [x, y]))
    return xs, ys, tuples


@pytest.mark.parametrize("depth,degree", [(0, 4), (1, 4), (2, 5), (3, 5)])
@pytest.mark.parametrize("shuffled", [False, True])
def test_plan_groups_by_variable_set(depth, degree, shuffled):
    net = nw.alarm()
    cards = [int(c) for c in net.cards]
    rng = np.random.default_rng(10 * depth + degree)
    xs, ys, tuples = star_frontier(rng, net.n, 6, depth, degree)
    if shuffled:
        perm = rng.permutation(len(tuples))
        xs, ys, tuples = [xs[i] for i in perm], [ys[i] for i in perm], [tuples[i] for i in perm]
    groups, group_of, digits = capi.ci_group_plan(cards, tuples, xs, ys)
    assert len(group_of) == len(tuples)
    seen = {}
    for t, q in enumerate(group_of):
        assert groups[q] == tuples[t]                                   # a test's group IS its variable set
        px, py = digits[t]
        assert groups[q][px] == min(xs[t], ys[t]) and groups[q][py] == max(xs[t], ys[t]) and px < py   # citest.h:48
        seen.setdefault(q, []).append(t)
    assert sorted(seen) == list(range(len(groups)))                     # dense ids, every group has a member
    firsts = [min(m) for q, m in sorted(seen.items())]
    assert firsts == sorted(firsts)                                     # groups in order of their first test
    # distinct sets <-> groups (no group is full here), and the frontier shape shares depth+1 ways
    assert len(groups) == len({tuple(t) for t in tuples})
    if depth:
        assert len(tuples) >= (depth + 1) * len(groups) * 0.99


def test_plan_limits_and_errors():
    cards = [2, 3, 4, 2, 3, 2]
    groups, group_of, digits = capi.ci_group_plan(cards, [], [], [])
    assert groups == [] and len(group_of) == 0
    # at most 32 members per group (the kernel keeps their statistics in shared memory)
    many = [[0, 1, 2]] * 70
    groups, group_of, _ = capi.ci_group_plan(cards, many, [0] * 70, [1] * 70)
    assert len(groups) == 3 and np.bincount(group_of).tolist() == [32, 32, 6]
    # a repeated id is a legal table (the reference counts it): x and y are the FIRST occurrences (citest.h:63-64)
    groups, group_of, digits = capi.ci_group_plan(cards, [[2, 2, 3], [2, 3, 3]], [2, 3], [3, 2])
    assert groups == [[2, 2, 3], [2, 3, 3]] and digits == [(0, 2), (0, 1)]
    with pytest.raises(ValueError, match="x and y must differ"):
        capi.ci_group_plan(cards, [[0, 1]], [0], [0])
    with pytest.raises(ValueError, match="not contained"):
        capi.ci_group_plan(cards, [[0, 1]], [0], [2])
    with pytest.raises(ValueError, match="sorted"):
        capi.ci_group_plan(cards, [[1, 0]], [0], [1])
    with pytest.raises(IndexError):
        capi.ci_group_plan(cards, [[0, 9]], [0], [9])


def strided_citest(table, cards, px, py, kind):
    """What the kernel's epilogue does (obn_kernels.cuh: CiGeom): walk the slices of (x at digit px,
    y at digit py | every other digit) directly in the dense reference-order table."""
    t = np.asarray(table).reshape(cards)
    t = np.moveaxis(t, (px, py), (-2, -1)).reshape(-1, cards[px], cards[py])
    stat, dof = 0.0, 0
    for s in t:
        n = s.sum()
        if n == 0:
            continue
        rs, cs = s.sum(1), s.sum(0)
        r_obs, c_obs = np.nonzero(rs)[0], np.nonzero(cs)[0]
        d = (len(r_obs) - 1) * (len(c_obs) - 1)
        dof += d
        for i in r_obs:
            for j in c_obs:
                obs, ex = float(s[i, j]), rs[i] * cs[j] / n
                if d == 1:
                    diff = ex - obs
                    obs += min(0.5, abs(diff)) * np.sign(diff)
                if kind == 0:
                    stat += (obs - ex) ** 2 / ex
                elif obs > 0:
                    stat += 2.0 * obs * np.log(obs / ex)
    return stat, dof


@pytest.mark.parametrize("kind", [0, 1])
def test_one_table_serves_every_role_assignment(kind):
    net = nw.alarm()
    codes, cards = ranked(net.sample(4000, 3))
    ods = oracle.Dataset(codes, cards)
    rng = np.random.default_rng(5)
    for _ in range(6):
        k = int(rng.integers(2, 6))
        vs = sorted(int(v) for v in rng.choice(net.n, size=k, replace=False))
        table = oracle.count_table(ods, vs)
        cs = [cards[v] for v in vs]
        for px, py in itertools.combinations(range(k), 2):
            z = [v for i, v in enumerate(vs) if i not in (px, py)]
            want_stat, want_dof, _ = oracle.citest(ods, vs[px], vs[py], z, kind)
            got_stat, got_dof = strided_citest(table, cs, px, py, kind)
            assert got_dof == want_dof
            assert got_stat == pytest.approx(want_stat, rel=1e-9, abs=1e-9)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_plan_deals_every_variable_set_to_exactly_one_rank(world):
    """A sharded PC round (obn_pc_round, shard != 0): every rank plans the same batch and keeps the variable sets
    whose hash falls to it.  The ranks must partition the tests, no set may be split between ranks (the table
    sharing survives the split), and a rank's groups must be the unsharded groups of its own tests."""
    net = nw.hepar2_like()
    cards = [int(c) for c in net.cards]
    rng = np.random.default_rng(world)
    xs, ys, tuples = star_frontier(rng, net.n, 12, 2, 6)
    n = len(tuples)
    g_all, of_all, dig_all = capi.ci_group_plan(cards, tuples, xs, ys)
    owner = np.full(n, -1)
    sizes = []
    for rank in range(world):
        groups, group_of, digits = capi.ci_group_plan(cards, tuples, xs, ys, rank, world)
        mine = np.flatnonzero(group_of != 0xffffffff)
        assert (owner[mine] == -1).all()                       # nobody else took these tests
        owner[mine] = rank
        sizes.append(len(mine))
        assert sorted(set(group_of[mine].tolist())) == list(range(len(groups)))
        for t in mine.tolist():
            assert groups[group_of[t]] == list(tuples[t])      # its group is its variable set
            assert digits[t] == dig_all[t]                     # and the roles are those of the unsharded plan
        # the same sets are grouped the same way as without sharding: one group per distinct set here (< 32 tests each)
        assert len(groups) == len({tuple(tuples[t]) for t in mine.tolist()})
    assert (owner >= 0).all()                                  # every test has an owner
    by_set = {}
    for t in range(n):
        by_set.setdefault(tuple(tuples[t]), set()).add(int(owner[t]))
    assert all(len(r) == 1 for r in by_set.values())           # a variable set lives on one rank
    assert min(sizes) > 0.4 * n / world                        # and the deal is not lopsided
    with pytest.raises(ValueError):
        capi.ci_group_plan(cards, tuples, xs, ys, world, world)
