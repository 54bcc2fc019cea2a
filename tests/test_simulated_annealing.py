"""simulated_annealing: the lockstep multi-chain driver must return
the reference's DAG for the same seed -- the fixtures are outputs of the unmodified reference
(tests/golden/simulated_annealing.json, generator tests/golden/make_golden.py --sa).

CPU part: the C++ driver over the checker's BDeu (test seam).  GPU part: the product path
(structure_learning.simulated_annealing -> obn_bdeu_scores, exact-order mode)."""
import json
import os

import numpy as np
import pytest

import oracle
import openbnsl_b200
from helpers import GOLDEN_DIR, OracleEvaluator, load_golden, pdag_to_adj, ranked
from openbnsl_b200 import networks as nw

L = openbnsl_b200.openbnsllib
with open(os.path.join(GOLDEN_DIR, "simulated_annealing.json")) as f:
    SA = json.load(f)


def _args(e):
    return (e["max_parents"], e["max_iters"], e["init_temp"], e["cooling_rate"], e["is_deterministic"], e["seed"],
            e["num_chains"])


@pytest.mark.parametrize("name", sorted(SA))
def test_driver_matches_golden_on_cpu(name):
    g, net, codes, cards = load_golden(name)
    ods = oracle.Dataset(codes, cards)
    for e in SA[name]:
        ev = OracleEvaluator(ods)
        got = L._testing.simulated_annealing(net.n, ev, L.score.BDeu(e["ess"]), *_args(e))
        assert pdag_to_adj(got, net.n).tolist() == e["adj"], e
        st = L.structure_learning.last_sa_stats()
        assert st["chains"] == e["num_chains"]
        assert st["evaluated"] == ev.n_sc <= st["proposals"] + net.n      # memoised: never more than one score per proposal
        adj = np.array(e["adj"])
        if e["max_parents"]:                # (max_parents = 0 never hits the reference's `size == max_parents` guard)
            assert adj.sum(axis=0).max() <= e["max_parents"]


def test_driver_matches_compiled_reference_live():
    ref, _ = oracle.reference_modules()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    net = nw.hepar2_like()
    codes, cards = ranked(net.sample(1200, 3))
    ods = oracle.Dataset(codes, cards)
    rdf = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
    for det in (False, True):
        want = ref.structure_learning.simulated_annealing(rdf, ref.score.BDeu(2.0), 4, 500, 2.0, 0.995, det, 99, 3)
        got = L._testing.simulated_annealing(net.n, OracleEvaluator(ods), L.score.BDeu(2.0), 4, 500, 2.0, 0.995, det, 99, 3)
        assert (pdag_to_adj(got, net.n) == pdag_to_adj(want, net.n)).all()


@pytest.mark.parametrize("case", range(8))
def test_driver_matches_compiled_reference_random_settings(case):
    """Random (max_parents, temperature schedule, seed, chains, container mode, ess) on a small Alarm sample."""
    ref, _ = oracle.reference_modules()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(100 + case)
    net = nw.alarm()
    codes, cards = ranked(net.sample(int(rng.integers(300, 1500)), int(rng.integers(0, 50))))
    ods = oracle.Dataset(codes, cards)
    rdf = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
    mp = int(rng.integers(1, 6))
    iters = int(rng.integers(50, 400))
    t0 = float(rng.choice([0.1, 1.0, 10.0, 100.0]))
    cr = float(rng.choice([0.9, 0.99, 0.9995, 1.0]))
    det = bool(rng.integers(0, 2))
    seed = int(rng.integers(0, 2 ** 31))
    chains = int(rng.integers(1, 6))
    ess = float(rng.choice([0.5, 1.0, 10.0]))
    want = ref.structure_learning.simulated_annealing(rdf, ref.score.BDeu(ess), mp, iters, t0, cr, det, seed, chains)
    got = L._testing.simulated_annealing(net.n, OracleEvaluator(ods), L.score.BDeu(ess), mp, iters, t0, cr, det, seed, chains)
    assert (pdag_to_adj(got, net.n) == pdag_to_adj(want, net.n)).all(), (mp, iters, t0, cr, det, seed, chains, ess)


def test_errors_like_the_reference():
    g, net, codes, cards = load_golden("asia_10k_s0")
    ev = OracleEvaluator(oracle.Dataset(codes, cards))
    with pytest.raises(ValueError, match="max_parents out of range"):        # simulated_annealing.cpp:109-111
        L._testing.simulated_annealing(net.n, ev, L.score.BDeu(1.0), net.n)
    with pytest.raises(ValueError, match="Unsupported score type"):          # local_score.h:43
        L._testing.simulated_annealing(net.n, ev, L.score.K2(), 2)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SA))
def test_gpu_matches_golden(name):
    g, net, codes, cards = load_golden(name)
    df = L.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
    for e in SA[name]:
        got = L.structure_learning.simulated_annealing(df, L.score.BDeu(e["ess"]), *_args(e))
        assert pdag_to_adj(got, net.n).tolist() == e["adj"], e


@pytest.mark.gpu
def test_gpu_many_chains_equals_checker_and_reference():
    """A larger run (Alarm structure, 20k rows, 16 chains): GPU product path == the same driver over the
    checker's scores, bit for bit (best score included), and == the compiled reference where it exists."""
    net = nw.alarm()
    codes, cards = ranked(net.sample(20_000, 4))
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    args = (3, 400, 1.0, 0.999, False, 5, 16)
    got = L.structure_learning.simulated_annealing(df, L.score.BDeu(1.0), *args)
    st_gpu = L.structure_learning.last_sa_stats()
    chk = L._testing.simulated_annealing(net.n, OracleEvaluator(oracle.Dataset(codes, cards)), L.score.BDeu(1.0), *args)
    st_chk = L.structure_learning.last_sa_stats()
    assert (pdag_to_adj(got, net.n) == pdag_to_adj(chk, net.n)).all()
    assert st_gpu["best_score"] == st_chk["best_score"] and st_gpu["best_chain"] == st_chk["best_chain"]
    assert st_gpu["evaluated"] == st_chk["evaluated"] and st_gpu["batches"] <= 401
    ref, _ = oracle.reference_modules()
    if ref is not None:
        rdf = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
        want = ref.structure_learning.simulated_annealing(rdf, ref.score.BDeu(1.0), *args)
        assert (pdag_to_adj(got, net.n) == pdag_to_adj(want, net.n)).all()
