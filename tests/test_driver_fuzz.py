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
    """Also answers the two-half calls of the test seam (bindings.cpp: PyEvaluator::ci_submit)."""
    submitted = 0


@pytest.mark.parametrize("seed", range(40))
@pytest.mark.parametrize("overlap", [0, 1])
def test_pc_driver_fuzz(seed, overlap):
    # overlap = 1: round r+1 is enumerated while round r is "in flight" (pc(..., overlap=True), pc.cpp) -- same graphs
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
