"""GPU parity of the batched kernels against the oracle, through the C ABI."""
import itertools

import numpy as np
import pytest

import oracle
from openbnsl_b200 import networks as nw

pytestmark = pytest.mark.gpu


def _mk(net, n_rows, seed):
    codes = net.sample(n_rows, seed)
    return codes, oracle.Dataset(codes, net.cards)


@pytest.fixture(scope="module")
def asia(gpu_ctx):
    net = nw.asia()
    codes, ods = _mk(net, 10000, 0)
    return net, ods, gpu_ctx.upload(codes, net.cards)


@pytest.fixture(scope="module")
def alarm(gpu_ctx):
    net = nw.alarm()
    codes, ods = _mk(net, 20011, 1)      # not a multiple of 16: exercises the tail
    return net, ods, gpu_ctx.upload(codes, ods.cards)


def test_counts_all_subsets_asia(asia):
    net, ods, ds = asia
    tuples = [list(c) for k in range(0, 9) for c in itertools.combinations(range(8), k)]
    got = ds.count_tables(tuples)
    for t, g in zip(tuples, got):
        want = oracle.count_table(ods, t)
        assert g.sum() == ods.n_rows
        assert np.array_equal(g.astype(np.int64), want), t


def test_counts_tiers_alarm(alarm):
    net, ods, ds = alarm
    rng = np.random.default_rng(0)
    tuples = []
    for k in (1, 2, 3, 4, 5, 6, 7, 8, 10, 12):
        for _ in range(6):
            tuples.append(sorted(rng.choice(37, size=k, replace=False).tolist()))
    got = ds.count_tables(tuples)
    sizes = set()
    for t, g in zip(tuples, got):
        want = oracle.count_table(ods, t)
        sizes.add(len(want))
        assert np.array_equal(g.astype(np.int64), want), t
    assert max(sizes) > 70000 and min(sizes) <= 4      # private, shared and global tiers all hit


@pytest.mark.parametrize("kind", [0, 1])
def test_ci_random_alarm(alarm, kind):
    net, ods, ds = alarm
    rng = np.random.default_rng(10 + kind)
    xs, ys, zs = [], [], []
    for nz in (0, 1, 2, 3, 4, 5, 7):
        for _ in range(12):
            v = rng.choice(37, size=nz + 2, replace=False).tolist()
            xs.append(v[0]); ys.append(v[1]); zs.append(v[2:])
    stat, dof, p = ds.ci_tests(xs, ys, zs, kind)
    for i in range(len(xs)):
        s, d, pv = oracle.citest(ods, xs[i], ys[i], zs[i], kind)
        assert int(dof[i]) == d, (xs[i], ys[i], zs[i])
        assert stat[i] == pytest.approx(s, rel=1e-9, abs=1e-9)
        assert p[i] == pytest.approx(pv, rel=1e-9, abs=1e-12)
        for level in (0.01, 0.05):
            if abs(pv - level) > 1e-9:
                assert (p[i] >= level) == (pv >= level)


def test_exact_order_table_cache_evicts_instead_of_failing(gpu_ctx, alarm):
    """The lgamma tables of the exact-order epilogue are cached on the ctx up to a cap.  A ctx that has served many
    ess values must not fail (or silently change precision) because of what EARLIER calls left in the cache: it evicts
    and rebuilds.  Only a single batch that needs more than the cap by itself is refused (exact_order=1) or answered
    in fast mode and counted in the profile (exact_order=2)."""
    net, ods, ds = alarm                         # 20011 rows: one table = 160 KB
    tuples, child = [[0, 1, 2], [3, 4], [5, 9, 22]], [1, 3, 22]
    gpu_ctx.set_option("lut_max_mb", 2)          # room for thirteen tables; one call below needs up to nine
    try:
        gpu_ctx.profile_enable(True)
        gpu_ctx.profile_read(True)
        for i in range(12):                      # every call brings new tables; the cache overflows several times
            ess = 0.5 + 0.37 * i
            got = ds.bdeu_scores(tuples, child, ess, exact_order=1)
            for t, c, g in zip(tuples, child, got):
                assert g == oracle.bdeu_local_score(ods, c, [v for v in t if v != c], ess), (i, t)
        assert gpu_ctx.profile_read(True).exact_fallbacks == 0
        # one batch that cannot fit by itself
        rng = np.random.default_rng(2)
        many = [sorted(rng.choice(37, size=k, replace=False).tolist()) for k in (1, 2, 3, 4, 5, 6) for _ in range(4)]
        ch = [t[0] for t in many]
        with pytest.raises(RuntimeError, match="lgamma tables"):
            ds.bdeu_scores(many, ch, 1.0, exact_order=1)
        got = ds.bdeu_scores(many, ch, 1.0, exact_order=2)
        assert gpu_ctx.profile_read(True).exact_fallbacks == 1
        for t, c, g in zip(many, ch, got):
            assert g == pytest.approx(oracle.bdeu_local_score(ods, c, [v for v in t if v != c], 1.0), rel=1e-9)
    finally:
        gpu_ctx.set_option("lut_max_mb", 2048)
        gpu_ctx.profile_enable(False)


@pytest.mark.parametrize("exact", [False, True])
def test_bdeu_random(asia, alarm, exact):
    for net, ods, ds in (asia, alarm):
        rng = np.random.default_rng(5)
        tuples, child = [], []
        for npar in (0, 1, 2, 3, 4, 5):
            for _ in range(8):
                v = rng.choice(ods.n_vars, size=npar + 1, replace=False).tolist()
                child.append(v[0]); tuples.append(sorted(v))
        got = ds.bdeu_scores(tuples, child, 1.0, exact_order=exact)
        sub, offs = ds.bdeu_subset_scores(tuples, 1.0, exact_order=exact)
        for i, (t, c) in enumerate(zip(tuples, child)):
            want = oracle.bdeu_local_score(ods, c, [v for v in t if v != c], 1.0)
            if exact:
                assert got[i] == want, (t, c)
            else:
                assert got[i] == pytest.approx(want, rel=1e-9)
            for j, cc in enumerate(t):
                w2 = oracle.bdeu_local_score(ods, cc, [v for v in t if v != cc], 1.0)
                g2 = sub[int(offs[i]) + j]
                if exact:
                    assert g2 == w2
                else:
                    assert g2 == pytest.approx(w2, rel=1e-9)


def test_row_splits_small_batch(gpu_ctx):
    """A batch smaller than the SM count is split over rows and merged through global scratch."""
    net = nw.alarm()
    codes, ods = _mk(net, 300007, 3)
    ds = gpu_ctx.upload(codes, ods.cards)
    tuples = [[0, 5], [1, 2, 3], [4, 9, 20, 30], [2, 7, 11, 13, 17, 19, 23]]
    for rep in range(2):       # twice: scratch must come back zeroed
        got = ds.count_tables(tuples)
        for t, g in zip(tuples, got):
            assert np.array_equal(g.astype(np.int64), oracle.count_table(ods, t)), (rep, t)
    stat, dof, p = ds.ci_tests([0, 1], [5, 3], [[], [2]], 1)
    for i, (x, y, z) in enumerate([(0, 5, []), (1, 3, [2])]):
        s, d, pv = oracle.citest(ods, x, y, z, 1)
        assert int(dof[i]) == d and stat[i] == pytest.approx(s, rel=1e-9, abs=1e-9)


def test_errors(asia):
    net, ods, ds = asia
    with pytest.raises(ValueError, match="sorted"):
        ds.count_tables([[2, 1]])
    with pytest.raises(ValueError, match="differ"):
        ds.ci_tests([1], [1], [[2]], 0)
    with pytest.raises(IndexError):
        ds.count_tables([[0, 99]])
    with pytest.raises(ValueError, match="child_var"):
        ds.bdeu_scores([[0, 1]], [5])
