"""GPU parity at BASELINE.json's FULL sizes (configs 3, 4, 5): 1M rows, 24 / 70 /
186 variables, cardinalities up to 21.  The reference cannot run these configs
end to end, so parity is checked
  * bit-exactly against numpy's bincount of the mixed-radix key (an independent
    count oracle that is fast at 1M rows), over every table-size class up to
    2^24 cells,
  * against the C oracle on a sample of tuples (statistics 1e-9 relative, dof
    exact, exact-order BDeu bit-exact),
  * through size-independent properties: counts sum to N, marginalising a table
    gives the table of the sub-tuple, a CI test is symmetric in (x, y), and the
    per-subset scores agree with the per-child scores."""
import numpy as np
import pytest

import oracle
from openbnsl_b200 import networks as nw

pytestmark = pytest.mark.gpu
N_ROWS = 1_000_000


def _ranked(codes):
    cards = []
    for v in range(codes.shape[1]):
        u = np.unique(codes[:, v])
        if len(u) != int(u.max()) + 1:
            lut = np.zeros(256, dtype=np.uint8)
            lut[u] = np.arange(len(u), dtype=np.uint8)
            codes[:, v] = lut[codes[:, v]]
        cards.append(len(u))
    return codes, cards


def _np_counts(codes, cards, t):
    key = np.zeros(codes.shape[0], dtype=np.int64)
    cells = 1
    for v in t:                                   # first variable most significant (contingency_table.h:76-85)
        key = key * cards[v] + codes[:, v]
        cells *= cards[v]
    return np.bincount(key, minlength=cells)


class Big:
    def __init__(self, gpu_ctx, net, seed):
        codes, cards = _ranked(net.sample(N_ROWS, seed))
        self.codes, self.cards, self.n = codes, cards, net.n
        self.ods = oracle.Dataset(codes, cards)
        self.ds = gpu_ctx.upload(codes, cards)


@pytest.fixture(scope="module")
def hepar(gpu_ctx):
    return Big(gpu_ctx, nw.hepar2_like(), 4)


@pytest.fixture(scope="module")
def munin(gpu_ctx):
    return Big(gpu_ctx, nw.munin1_like(), 5)


@pytest.fixture(scope="module")
def bin24(gpu_ctx):
    return Big(gpu_ctx, nw.random_binary(24), 3)


def _check_counts(big, tuples):
    got = big.ds.count_tables(tuples)
    for t, g in zip(tuples, got):
        assert int(g.astype(np.int64).sum()) == N_ROWS, t
        assert np.array_equal(g.astype(np.int64), _np_counts(big.codes, big.cards, t)), t
    return got


def _check_ci(big, xs, ys, zs, kind):
    stat, dof, p = big.ds.ci_tests(xs, ys, zs, kind)
    s2, d2, p2 = big.ds.ci_tests(ys, xs, zs, kind)               # symmetric in (x, y): citest.h:48
    assert np.array_equal(dof, d2) and np.array_equal(stat, s2) and np.array_equal(p, p2)
    for i in range(len(xs)):
        s, d, pv = oracle.citest(big.ods, xs[i], ys[i], zs[i], kind)
        assert int(dof[i]) == d, (xs[i], ys[i], zs[i])
        assert stat[i] == pytest.approx(s, rel=1e-9, abs=1e-9)
        assert p[i] == pytest.approx(pv, rel=1e-9, abs=1e-12)


def test_c4_hepar2_like_counts_and_marginals(hepar):
    rng = np.random.default_rng(0)
    tuples = [sorted(rng.choice(hepar.n, size=k, replace=False).tolist()) for k in (1, 2, 3, 4, 5, 6) for _ in range(4)]
    got = _check_counts(hepar, tuples)
    for t, g in zip(tuples, got):                                # drop the last variable = sum over its digit
        if len(t) < 2:
            continue
        sub = hepar.ds.count_tables([t[:-1]])[0]
        assert np.array_equal(g.reshape(-1, hepar.cards[t[-1]]).sum(axis=1), sub), t


@pytest.mark.parametrize("kind", [0, 1])
def test_c4_hepar2_like_ci(hepar, kind):
    rng = np.random.default_rng(1 + kind)
    xs, ys, zs = [], [], []
    for nz in (0, 1, 2, 3, 4):
        for _ in range(4):
            v = rng.choice(hepar.n, size=nz + 2, replace=False).tolist()
            xs.append(v[0]); ys.append(v[1]); zs.append(v[2:])
    _check_ci(hepar, xs, ys, zs, kind)


def test_c5_munin1_like_high_cardinality(munin):
    order = np.argsort(munin.cards)[::-1]
    top = [int(v) for v in order[:4]]                            # the widest variables (cards up to 21)
    assert munin.cards[top[0]] >= 15
    rng = np.random.default_rng(2)
    tuples = [sorted(top[:2]), sorted(top[:3]), sorted(top[:4])]     # 441 .. ~190k cells: smem and global tiers
    tuples += [sorted(rng.choice(munin.n, size=k, replace=False).tolist()) for k in (2, 3, 4, 5) for _ in range(3)]
    _check_counts(munin, tuples)
    xs = [top[0], top[0], top[1], int(order[10])]
    ys = [top[1], top[2], top[2], int(order[40])]
    zs = [[], [top[1]], [top[0], top[3]], [int(order[20]), int(order[30]), int(order[50])]]
    _check_ci(munin, xs, ys, zs, 0)
    _check_ci(munin, xs, ys, zs, 1)


def test_c3_binary24_tables_up_to_2_pow_24_cells(bin24):
    rng = np.random.default_rng(3)
    tuples = [sorted(rng.choice(24, size=k, replace=False).tolist()) for k in (0, 1, 6, 8, 12, 14, 16, 20)]
    tuples.append(list(range(24)))                               # the lattice's top: 2^24 cells
    _check_counts(bin24, tuples)


def test_c3_binary24_scores(bin24):
    rng = np.random.default_rng(4)
    tuples = [sorted(rng.choice(24, size=k, replace=False).tolist()) for k in (1, 2, 4, 6, 10) for _ in range(2)]
    for exact in (False, True):
        sub, offs = bin24.ds.bdeu_subset_scores(tuples, 1.0, exact_order=exact)
        child = [t[len(t) // 2] for t in tuples]
        one = bin24.ds.bdeu_scores(tuples, child, 1.0, exact_order=exact)
        for i, t in enumerate(tuples):
            j = len(t) // 2
            assert sub[int(offs[i]) + j] == one[i]               # same table, same epilogue
            want = oracle.bdeu_local_score(bin24.ods, child[i], [v for v in t if v != child[i]], 1.0)
            if exact:
                assert one[i] == want, t
            else:
                assert one[i] == pytest.approx(want, rel=1e-9)


def _seam_cases():
    import json, os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pc_fullsize_seam.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("case", _seam_cases(), ids=lambda c: c["config"])
def test_full_size_pc_matches_reference_seam(case):
    """BASELINE configs 4 and 5 at FULL size (1M rows): the PDAG learned on the GPU must equal, arc for arc, the one the
    same host driver learns when every CI decision is taken by the UNMODIFIED reference functions
    (tests/golden/make_fullsize_seam.py: ContingencyTable<true> + citest<true> under OpenMP, 5 and 10 minutes of CPU).
    Covers what the sampled checks above cannot: every decision of the depth >= 2 frontiers, the cardinality-21 tables
    and the global tier inside a real run, with early-exit rounds, whole-depth rounds and the overlapped driver."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    import openbnsl_b200
    L = openbnsl_b200.openbnsllib
    c = bench.CONFIGS[case["config"]]
    net, codes, cards = bench.make_data(case["config"])
    assert cards == case["cards"]
    assert int(np.frombuffer(codes.tobytes(), dtype=np.uint8).astype(np.uint64).sum()) == case["codes_sha"]
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    T = L.citest.ChiSquare(c["level"]) if c["ci_test"] == "chi2" else L.citest.GSquare(c["level"])
    for kw in ({}, {"target_batch": 100000000, "overlap": False}):
        g = L.structure_learning.pc(df, T, c["max_cond_vars"], True, **kw)
        arcs = [[p, v] for p in range(net.n) for v in range(net.n) if g.has_edge(p, v)]
        assert arcs == case["arcs"], kw
