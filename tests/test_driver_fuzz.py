"""Randomised parity of the host drivers (CPU): random networks, small samples (so that CI decisions
and score ties actually vary), both PC modes, several batch sizes -- our C++ drivers over the
checker-backed evaluator against the C restatement, and the C restatement against the compiled
unmodified reference where oracle/_ref is built.  The learned graphs must be identical arc by arc
(enumeration order, first passing separating set, tie-breaks)."""
import numpy as np
import pytest

import oracle
import openbnsl_b200
from helpers import OracleEvaluator, pdag_to_adj, ranked
from openbnsl_b200 import networks as nw

L = openbnsl_b200.openbnsllib
REF = oracle.reference_modules()[0]


def case(seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(4, 10))
    cards = rng.integers(2, 5, size=n)
    net = nw.random_network(n, int(rng.integers(n - 1, 2 * n)), 3, cards, seed)
    rows = int(rng.choice([60, 200, 1000]))
    codes, cards = ranked(net.sample(rows, seed + 1))
    return net, codes, cards


class CountingEvaluator(OracleEvaluator):
    """Also counts the asynchronous rounds of the test seam (bindings.cpp: PyEvaluator::pc_round_submit) and the
    tests a chained round left out."""
    submitted = 0
    skipped = 0


@pytest.mark.parametrize("seed", range(40))
@pytest.mark.parametrize("overlap", [0, 1])
def test_pc_driver_fuzz(seed, overlap):
    # overlap = 1: round r+1 is enumerated AND submitted (chained) while round r is "in flight" (pc(..., overlap=True),
    # pc.cpp) -- same graphs
    net, codes, cards = case(seed)
    ods = oracle.Dataset(codes, cards)
    rng = np.random.default_rng(1000 + seed)
    kind = int(rng.integers(0, 2))
    level = float(rng.choice([0.01, 0.05, 0.2]))
    depth = int(rng.integers(0, 4))
    T = L.citest.ChiSquare(level) if kind == 0 else L.citest.GSquare(level)
    rdf = REF.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards)) if REF is not None else None
    for stable in (True, False):
        want = oracle.pc(ods, kind, level, depth, stable)
        for tb in (0, 1, 5):                                     # default rounds, one test per round, tiny rounds
            ev = CountingEvaluator(ods)
            got = pdag_to_adj(L._testing.pc(net.n, ev, T, depth, stable, None, tb, False, bool(overlap)), net.n)
            assert (got == want).all(), (seed, stable, tb)
            assert (ev.submitted > 0) == bool(overlap and stable)    # the overlapped loop really ran
        if rdf is not None:
            RT = REF.citest.ChiSquare(level) if kind == 0 else REF.citest.GSquare(level)
            assert (pdag_to_adj(REF.structure_learning.pc(rdf, RT, depth, stable), net.n) == want).all(), (seed, stable)


@pytest.mark.parametrize("seed", range(12))
def test_pc_pipelined_rounds_ask_the_same_tests(seed):
    """The pipelined driver submits round r+1 before it knows what round r removed; the evaluator drops those tests
    (chain).  What is finally ASKED -- and accounted in last_pc_stats / recorded -- must be exactly the serial run's."""
    net, codes, cards = case(300 + seed)
    ods = oracle.Dataset(codes, cards)
    T = L.citest.GSquare(0.05)
    seen, skipped_any = [], 0
    for overlap in (False, True):
        ev = CountingEvaluator(ods)
        g = L._testing.pc(net.n, ev, T, 3, True, None, 3, True, overlap)
        st = L.structure_learning.last_pc_stats()
        rec = [np.asarray(a).copy() for a in L.structure_learning.last_pc_tests()]
        tests = sorted((int(x), int(y), tuple(rec[0][int(rec[1][t]):int(rec[1][t + 1])].tolist()))
                       for t, (x, y) in enumerate(zip(rec[2], rec[3])))
        seen.append((pdag_to_adj(g, net.n), int(st["total_tests"]), list(st["tests_per_depth"]), tests))
        skipped_any += ev.skipped
        assert len(tests) == st["total_tests"]
    assert (seen[0][0] == seen[1][0]).all()
    assert seen[0][1:] == seen[1][1:]
    test_pc_pipelined_rounds_ask_the_same_tests.skipped_total = getattr(
        test_pc_pipelined_rounds_ask_the_same_tests, "skipped_total", 0) + skipped_any


def test_pc_pipelined_rounds_did_skip_something():
    assert getattr(test_pc_pipelined_rounds_ask_the_same_tests, "skipped_total", 0) > 0      # the chain path really ran


@pytest.mark.parametrize("seed", range(16))
def test_exhaustive_search_driver_fuzz(seed):
    net, codes, cards = case(100 + seed)
    if net.n > 7:
        codes, cards, n = codes[:, :7], cards[:7], 7
    else:
        n = net.n
    ods = oracle.Dataset(np.ascontiguousarray(codes), cards)
    rng = np.random.default_rng(2000 + seed)
    ess = float(rng.choice([0.5, 1.0, 10.0]))
    rdf = (REF.base.DataframeWrapper(nw.to_reference_frame(np.ascontiguousarray(codes), net.names[:n], cards))
           if REF is not None else None)
    for mp in sorted({0, 1, int(rng.integers(1, n)), n - 1}):
        want = oracle.exhaustive_search(ods, ess, mp)
        got = pdag_to_adj(L._testing.exhaustive_search(n, OracleEvaluator(ods), L.score.BDeu(ess), mp), n)
        assert (got == want).all(), (seed, mp)
        if rdf is not None:
            g = REF.structure_learning.exhaustive_search(rdf, REF.score.BDeu(ess), mp)
            assert (pdag_to_adj(g, n) == want).all(), (seed, mp)


def test_pipelined_driver_survives_an_evaluator_that_raises():
    """An exception in the middle of the pipeline (a CUDA error on the device path) must surface as an exception and
    leave nothing in flight; the next run is unaffected."""
    net, codes, cards = case(7)
    ods = oracle.Dataset(codes, cards)

    class Failing(CountingEvaluator):
        calls = 0

        def ci_pvalues(self, *a):
            Failing.calls += 1
            if Failing.calls == 3:
                raise RuntimeError("boom")
            return super().ci_pvalues(*a)

    T = L.citest.GSquare(0.05)
    with pytest.raises(RuntimeError, match="boom"):
        L._testing.pc(net.n, Failing(ods), T, 3, True, None, 1, False, True)
    want = oracle.pc(ods, 1, 0.05, 3, True)
    got = pdag_to_adj(L._testing.pc(net.n, CountingEvaluator(ods), T, 3, True, None, 1, False, True), net.n)
    assert (got == want).all()
