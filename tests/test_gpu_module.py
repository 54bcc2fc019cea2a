"""GPU parity of the drop-in surface (pybind11 module -> C++ drivers -> C ABI ->
sm_100a kernels) against the golden fixtures of the unmodified reference and
against the oracle at the benchmark's full size.  Mirrors the reference's own
tests: pytest/base/test_contingency_table.py, pytest/score/test_local_score.py,
pytest/citest/test_citest.py, pytest/structure_learning/test_{exhaustive_search,pc}.py."""
import numpy as np
import pytest

import oracle
import openbnsl_b200
from helpers import GOLDEN_SETS, close, load_golden, pdag_to_adj, ranked
from openbnsl_b200 import networks as nw
from openbnsl_b200 import structural_distance as sd

pytestmark = pytest.mark.gpu
L = openbnsl_b200.openbnsllib


@pytest.fixture(scope="module", params=GOLDEN_SETS)
def gold(request):
    g, net, codes, cards = load_golden(request.param)
    import types
    frame = nw.to_reference_frame(codes, net.names, cards)      # str cells, like the reference's input
    df = L.base.DataframeWrapper(frame)
    assert list(df.num_of_values) == cards
    return g, net, df


def test_golden_tables(gold):
    g, net, df = gold
    for t in g["tables"]:
        ct = L.base.ContingencyTable(t["var_ids"], df)
        assert {str(k): v for k, v in ct.counts.items()} == t["counts"]
        assert sum(ct.counts.values()) == df.num_of_datapoints
        assert list(ct.var_ids) == t["var_ids"]
        assert list(ct.cardinalities) == [g["cards"][v] for v in t["var_ids"]]


def test_golden_scores_bit_exact(gold):
    g, net, df = gold
    for s in g["scores"]:
        ct = L.base.ContingencyTable(sorted(s["parents"] + [s["child"]]), df)
        got = L.score.calculate_local_score(s["child"], s["parents"], ct, L.score.BDeu(s["ess"]))
        assert got == s["score"], s          # exact-order mode: bit-identical to the reference


def test_golden_ci(gold):
    g, net, df = gold
    for c in g["ci"]:
        ct = L.base.ContingencyTable(sorted(c["z"] + [c["x"], c["y"]]), df)
        T = L.citest.ChiSquare(0.05) if c["kind"] == 0 else L.citest.GSquare(0.05)
        stat, dof, p = L.citest.citest_statistic(c["x"], c["y"], c["z"], ct, T)
        assert dof == c["dof"]
        if c["stat"] is not None:
            assert close(stat, c["stat"], 1e-9, 1e-9), c
        assert close(p, c["p"], 1e-9, 1e-12), c
        if abs(c["p"] - 0.05) > 1e-9:
            assert L.citest.citest(c["x"], c["y"], c["z"], ct, T) == c["independent_at_0.05"]
        # argument order must not matter (citest.h:48)
        assert L.citest.citest_statistic(c["y"], c["x"], c["z"], ct, T)[1] == c["dof"]


def test_golden_drivers(gold):
    g, net, df = gold
    for e in g.get("exhaustive_search", []):
        got = L.structure_learning.exhaustive_search(df, L.score.BDeu(e["ess"]), e["max_parents"])
        assert pdag_to_adj(got, net.n).tolist() == e["adj"], e["max_parents"]      # identical DAG
    for e in g.get("pc", []):
        T = L.citest.ChiSquare(e["level"]) if e["kind"] == 0 else L.citest.GSquare(e["level"])
        got = L.structure_learning.pc(df, T, e["max_cond_vars"], e["stable"])
        assert pdag_to_adj(got, net.n).tolist() == e["adj"], (e["kind"], e["stable"])


def test_errors_on_gpu(gold):
    g, net, df = gold
    ct = L.base.ContingencyTable([0, 1], df)
    with pytest.raises(ValueError, match="differ"):
        L.citest.citest(1, 1, [], ct, L.citest.GSquare())
    with pytest.raises(ValueError, match="not contained"):
        L.citest.citest(0, 2, [], ct, L.citest.GSquare())
    with pytest.raises(ValueError, match="child_var"):
        L.score.calculate_local_score(2, [0], ct, L.score.BDeu())
    with pytest.raises(ValueError, match="Unsupported score type"):
        L.score.calculate_local_score(0, [1], ct, L.score.K2())


# ---------------------------------------------------------------------------
# full-size configurations (BASELINE.json configs[0] and configs[1])
# ---------------------------------------------------------------------------
def test_config0_asia_es_full_lattice_fast_and_exact():
    net = nw.asia()
    codes, cards = ranked(net.sample(10000, 0))
    ods = oracle.Dataset(codes, cards)
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    want, gs, ch, ls = oracle.exhaustive_search(ods, 1.0, 7, want_tables=True)
    for exact in (1, 0):
        got = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), 7, exact_order=exact)
        st = L.structure_learning.last_es_stats()
        assert st["tables"] == 256 and st["scores"] == 1024
        if exact:
            assert (pdag_to_adj(got, 8) == want).all()
            assert st["best_score"] == gs[255]                     # bit-identical network score
        else:
            assert close(st["best_score"], gs[255], 1e-9)
    # the learned structure is the true Asia CPDAG skeleton
    und = {(min(a, b), max(a, b)) for a, b in net.edges()}
    got_und = {(min(a, b), max(a, b)) for a in range(8) for b in range(8) if want[a, b]}
    assert got_und == und - {(net.names.index("asia"), net.names.index("tub"))} or got_und == und


def load_big_pc():
    import json, os
    from helpers import GOLDEN_DIR
    with open(os.path.join(GOLDEN_DIR, "pc_big.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("case", load_big_pc(), ids=lambda c: c["name"])
@pytest.mark.parametrize("target_batch", [0, 100000000])
def test_pc_big_networks_match_reference(case, target_batch):
    """Down-scaled stand-ins of BASELINE configs 4 and 5 (the benchmark's networks at 20 000 rows, where the
    serial reference finishes): the STOCK reference pc() made the fixture (tests/golden/make_golden.py --big-pc);
    depth >= 2 frontiers, cardinalities up to 21 and every table tier run inside a real PC call here and the
    PDAG must be identical arc for arc -- with the default rounds (early exit) and with whole-depth rounds
    (no early exit, maximal table sharing)."""
    net = getattr(nw, case["network"])()
    codes, cards = ranked(net.sample(case["n_rows"], case["seed"]))
    assert cards == case["cards"]
    assert int(np.frombuffer(codes.tobytes(), dtype=np.uint8).astype(np.uint64).sum()) == case["codes_sha"]
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    T = L.citest.ChiSquare(case["level"]) if case["kind"] == 0 else L.citest.GSquare(case["level"])
    for overlap in (False, True):
        g = L.structure_learning.pc(df, T, case["max_cond_vars"], case["stable"], target_batch=target_batch, overlap=overlap)
        arcs = [[p, v] for p in range(net.n) for v in range(net.n) if g.has_edge(p, v)]
        assert arcs == case["arcs"]


def test_config1_alarm_pc_full_size():
    net = nw.alarm()
    codes, cards = ranked(net.sample(100_000, 0))
    ods = oracle.Dataset(codes, cards)
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    for stable in (True, False):
        got = L.structure_learning.pc(df, L.citest.GSquare(0.05), 4, stable, record_tests=True)
        want = oracle.pc(ods, 1, 0.05, 4, stable)
        assert (pdag_to_adj(got, net.n) == want).all(), stable
    # size-independent properties on the recorded frontier of the last run
    var_ids, offsets, x, y = L.structure_learning.last_pc_tests()
    assert len(x) == L.structure_learning.last_pc_stats()["total_tests"] > 1000
    from openbnsl_b200 import capi
    ctx = capi.Context(0)
    ds = ctx.upload(codes, cards)
    pick = np.random.default_rng(0).choice(len(x), size=200, replace=False)
    tuples = [[int(v) for v in var_ids[int(offsets[t]):int(offsets[t + 1])]] for t in pick]
    for t, c in zip(tuples, ds.count_tables(tuples)):
        assert int(c.sum()) == 100_000
        marg = oracle.count_table(ods, t[:1])                     # marginalising must give the 1-var table
        assert np.array_equal(c.reshape(len(marg), -1).sum(axis=1), marg)
    stat, dof, p = ds.ci_tests(x[pick], y[pick], [[v for v in t if v not in (int(a), int(b))]
                                                  for t, a, b in zip(tuples, x[pick], y[pick])], 1)
    for i, t in enumerate(tuples):
        a, b = int(x[pick][i]), int(y[pick][i])
        s, d, pv = oracle.citest(ods, a, b, [v for v in t if v not in (a, b)], 1)
        assert d == int(dof[i]) and close(stat[i], s, 1e-9, 1e-9) and close(p[i], pv, 1e-9, 1e-12)
    ctx.close()


@pytest.mark.parametrize("max_parents,exact", [(11, 2), (3, 2), (11, 0)])
def test_exhaustive_search_device_dp_equals_host_dp(max_parents, exact):
    """The HBM-resident DP (obn_es_*) and the host DP over GPU scores pick the same DAG and score."""
    net = nw.random_binary(12, seed=7)
    codes, cards = ranked(net.sample(20000, 2))
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    a = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, exact_order=exact)
    sa = L.structure_learning.last_es_stats()
    b = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, exact_order=exact, host_dp=True)
    sb = L.structure_learning.last_es_stats()
    assert np.array_equal(pdag_to_adj(a, net.n), pdag_to_adj(b, net.n))
    assert sa["best_score"] == sb["best_score"]
    assert sa["tables"] == sb["tables"] and sa["scores"] == sb["scores"]
    if exact:                                   # bit-exact scores => the oracle's DAG, not just its CPDAG
        ods = oracle.Dataset(codes, cards)
        assert np.array_equal(pdag_to_adj(a, net.n), oracle.exhaustive_search(ods, 1.0, max_parents))


@pytest.mark.parametrize("which,max_parents", [("mixed", 9), ("mixed", 3), ("binary", 11)])
def test_exhaustive_search_lattice_mode(which, max_parents):
    """Fast mode: every local score is F(pa + child) - F(pa) with F from the marginalisation lattice
    over the joint table.  The optimum must equal the exact-order path's within 1e-9 relative, the
    DAG it returns must really have that score (recomputed with the oracle's BDeu), and -- the
    reference's own criterion, pytest/structure_learning/test_exhaustive_search.py:26-30 over
    helpers/structural_distance.py:95-96 -- its CPDAG must equal the exact-order DAG's CPDAG."""
    if which == "mixed":
        net = nw.random_network(10, 14, 3, [2, 3, 4, 2, 3, 2, 4, 2, 3, 2], seed=11)
    else:
        net = nw.random_binary(12, seed=9)
    codes, cards = ranked(net.sample(6000, 3))
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    g_exact = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, exact_order=1, lattice=False)
    want = L.structure_learning.last_es_stats()
    assert not want["lattice"]
    g = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, exact_order=0, lattice=True)
    got = L.structure_learning.last_es_stats()
    assert got["lattice"]
    assert got["tables"] == want["tables"] and got["scores"] == want["scores"]
    assert got["best_score"] == pytest.approx(want["best_score"], rel=1e-9)
    ods = oracle.Dataset(codes, cards)
    adj = pdag_to_adj(g, net.n)
    total = 0.0
    for v in range(net.n):
        pa = [int(p) for p in np.nonzero(adj[:, v])[0]]
        assert len(pa) <= max_parents
        total += oracle.bdeu_local_score(ods, v, pa, 1.0)
    assert total == pytest.approx(got["best_score"], rel=1e-9)
    assert np.array_equal(sd.pdag2cpdag(adj), sd.pdag2cpdag(pdag_to_adj(g_exact, net.n)))
    assert np.array_equal(pdag_to_adj(g_exact, net.n), oracle.exhaustive_search(ods, 1.0, max_parents))


@pytest.mark.parametrize("cards,max_l,no_lut", [
    ([2] * 13, 20, 0),                      # one block: h = 0, 5 mid + 8 bottom variables
    ([2] * 13, 10, 0),                      # 8 blocks of 2 mid + 8 bottom variables
    ([3, 2, 4, 2, 3] + [2] * 8, 11, 0),     # mixed cardinalities above a binary bottom tier, 4 blocks
    ([2] * 9, 8, 0),                        # no mid tier at all (Lm = 0), 2 blocks
    ([2] * 11, 20, 1),                      # no term tables: libm fallback inside the bottom kernel
    ([2, 3, 2, 2, 4, 2, 2, 3, 2, 2, 2], 20, 0),          # exactly 8 binary variables, scattered: digit order != variable order
    ([2, 2, 3, 2, 2, 2, 4, 2, 2, 2, 3, 2, 3], 11, 0),    # 9 binary variables, the last one is not binary; 4 blocks
])
def test_lattice_bottom_tier_equals_hbm_tiers(cards, max_l, no_lut):
    """The shared-memory bottom tier (last 8 binary variables) and the all-HBM lattice build the same
    tables; F differs only by summation order (fixed point vs fp64), so the optimum agrees to 1e-11."""
    n = len(cards)
    net = nw.random_network(n, n + 4, 3, cards, seed=5)
    codes, cards_seen = ranked(net.sample(5000, 7))
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards_seen)
    out = {}
    try:
        L.set_option("lattice_max_l", max_l)
        L.set_option("lattice_no_lut", no_lut)
        for bottom in ("0", "1", "1"):          # the second "1" run replays the cached graph
            L.set_option("lattice_bottom", int(bottom))
            g = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), n - 1, exact_order=0, lattice=True)
            st = L.structure_learning.last_es_stats()
            assert st["lattice"]
            out.setdefault(bottom, []).append((st["best_score"], pdag_to_adj(g, n).tolist()))
    finally:
        for k, v in (("lattice_max_l", 20), ("lattice_no_lut", 0), ("lattice_bottom", 1)):
            L.set_option(k, v)
    (s0, a0), = out["0"]
    (s1, a1), (s2, a2) = out["1"]
    assert s1 == s2 and a1 == a2                      # deterministic, cache hit or not
    assert s1 == pytest.approx(s0, rel=1e-11)
    ods = oracle.Dataset(codes, cards_seen)            # the DAG really has that score (ties may pick another member
    adj = np.array(a1)                                 # of the equivalence class than the fp64-order run)
    total = sum(oracle.bdeu_local_score(ods, v, [int(p) for p in np.nonzero(adj[:, v])[0]], 1.0) for v in range(n))
    assert total == pytest.approx(s1, rel=1e-9)
    # identical structure up to Markov equivalence with the exact-order search and, through it, the oracle
    g_exact = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), n - 1, exact_order=1, lattice=False)
    want = pdag_to_adj(g_exact, n)
    assert np.array_equal(want, oracle.exhaustive_search(ods, 1.0, n - 1))
    assert np.array_equal(sd.pdag2cpdag(adj), sd.pdag2cpdag(want))
    assert np.array_equal(sd.pdag2cpdag(np.array(a0)), sd.pdag2cpdag(want))
