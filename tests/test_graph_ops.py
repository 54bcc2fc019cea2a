"""CPU tests of the graph post-processing either side of pc():
Meek's rules as the reference states them (src/graph/pdag_with_adjmat.cpp:215-279) and the
structural-distance helpers the reference's tests judge every learned graph with
(helpers/structural_distance.py:43-137).  Golden vectors: tests/golden/graph_kat.json, made by
`tests/golden/make_golden.py --graphs` from the reference's own code; live comparison against
oracle/_ref where it is built."""
import itertools
import json
import os

import numpy as np
import pytest

import oracle
import openbnsl_b200
from helpers import GOLDEN_DIR, pdag_to_adj
from openbnsl_b200 import networks as nw
from openbnsl_b200 import structural_distance as sd

L = openbnsl_b200.openbnsllib

with open(os.path.join(GOLDEN_DIR, "graph_kat.json")) as f:
    KAT = json.load(f)


def to_pdag(adj):
    adj = np.asarray(adj)
    g = L.base.PDAG(len(adj))
    for u, v in zip(*np.nonzero(adj)):
        g.add_edge(int(u), int(v))
    return g


def truth_of(net):
    g = L.base.PDAG(net.n)
    for p, v in net.edges():
        g.add_edge(p, v)
    return g


def test_meek_rules_match_reference_golden():
    fired = 0
    for c in KAT["meek_cpdag"]:
        g = to_pdag(c["adj"])
        assert pdag_to_adj(L.structure_learning.apply_meeks_rules(g), c["n"]).tolist() == c["meek"]
        assert pdag_to_adj(L.structure_learning.apply_meeks_rules(g, apply_r4=True), c["n"]).tolist() == c["meek_r4"]
        assert pdag_to_adj(g, c["n"]).tolist() == c["adj"]          # the argument is not modified
        fired += c["meek"] != c["adj"]
    assert fired >= 10                                               # the fixture exercises the rules


def test_pdag2cpdag_matches_reference_golden():
    for c in KAT["meek_cpdag"]:
        assert sd.pdag2cpdag(np.asarray(c["adj"])).tolist() == c["cpdag"]
        assert sd.PDAG2CPDAG(to_pdag(c["adj"])).tolist() == c["cpdag"]       # PDAG objects and matrices alike


def test_structural_errors_match_reference_golden():
    nonzero = 0
    for c in KAT["structural_errors"]:
        got = sd.structural_errors(np.asarray(c["truth"]), to_pdag(c["pred"]))
        assert got == c["errors"]
        assert got["SHD"] == got["ME"] + got["EE"] + got["DE"] and got["DE"] == got["ED"] + got["MD"] + got["RD"]
        nonzero += got["SHD"] > 0
    assert nonzero >= 10
    with pytest.raises(ValueError):
        sd.structural_errors(np.zeros((3, 3)), np.zeros((4, 4)))


def test_reference_rule1_is_dead_and_textbook_rules_are_not():
    # rule 1 pattern: 0 -> 1 - 2, 0 and 2 not adjacent.  The reference's guards `!is_adjacent(x, z)` and
    # `!creates_unshielded_collider(y, z)` (pdag_with_adjmat.cpp:225-226) contradict each other, so ITS rule 1
    # never fires (pinned by the golden vectors above); the published rule gives 1 -> 2.
    a = np.zeros((3, 3), dtype=np.uint8)
    a[0, 1] = a[1, 2] = a[2, 1] = 1
    assert (pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a)), 3) == a).all()
    out = pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a), textbook=True), 3)
    assert out[1, 2] == 1 and out[2, 1] == 0
    # rule 2: 0 -> 1 -> 2 and 0 - 2  =>  0 -> 2 (both variants)
    a = np.zeros((3, 3), dtype=np.uint8)
    a[0, 1] = a[1, 2] = a[0, 2] = a[2, 0] = 1
    for tb in (False, True):
        out = pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a), textbook=tb), 3)
        assert out[0, 2] == 1 and out[2, 0] == 0
    # rule 3: 0 - 1, 0 - 2, 0 - 3, 2 -> 1 <- 3, 2 and 3 apart  =>  0 -> 1
    a = np.zeros((4, 4), dtype=np.uint8)
    for u, v in ((0, 1), (0, 2), (0, 3)):
        a[u, v] = a[v, u] = 1
    a[2, 1] = a[3, 1] = 1
    out = pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a), textbook=True), 4)
    assert out[0, 1] == 1 and out[1, 0] == 0 and out[0, 2] == 1 and out[2, 0] == 1
    # nothing to orient in an undirected chain; empty and single-vertex graphs pass through
    a = np.zeros((3, 3), dtype=np.uint8)
    a[0, 1] = a[1, 0] = a[1, 2] = a[2, 1] = 1
    for tb in (False, True):
        assert (pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a), textbook=tb), 3) == a).all()
        assert L.structure_learning.apply_meeks_rules(L.base.PDAG(0), textbook=tb).num_vars == 0
        assert L.structure_learning.apply_meeks_rules(L.base.PDAG(1), textbook=tb).num_vars == 1


def test_wide_graph_crosses_word_boundary():
    # 70 vertices: the bit matrix spans two 64-bit words per row; a directed chain start propagates by rule 1
    n = 70
    a = np.zeros((n, n), dtype=np.uint8)
    a[0, 1] = 1
    for v in range(1, n - 1):
        a[v, v + 1] = a[v + 1, v] = 1
    out = pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a), textbook=True), n)
    want = np.zeros((n, n), dtype=np.uint8)
    for v in range(n - 1):
        want[v, v + 1] = 1
    assert (out == want).all()
    # rule 2 of the reference variant across the boundary: 0 -> 66 -> 69, 0 - 69
    a = np.zeros((n, n), dtype=np.uint8)
    a[0, 66] = a[66, 69] = a[0, 69] = a[69, 0] = 1
    out = pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a)), n)
    assert out[0, 69] == 1 and out[69, 0] == 0 and out.sum() == 3


def _is_dag(a):
    a = a.astype(bool).copy()
    alive = np.ones(len(a), dtype=bool)
    for _ in range(len(a)):
        src = [v for v in np.nonzero(alive)[0] if not (a[:, v] & alive).any()]
        if not src:
            return False
        alive[src[0]] = False
    return True


def _brute_force_cpdag(dag):
    """Union of every DAG with the skeleton and unshielded colliders of `dag` (its Markov equivalence class)."""
    n = len(dag)
    pattern = sd.pdag2cpdag(dag)
    skel = [(u, v) for u in range(n) for v in range(u + 1, n) if dag[u, v] or dag[v, u]]
    union = np.zeros((n, n), dtype=np.uint8)
    for bits in itertools.product((0, 1), repeat=len(skel)):
        a = np.zeros((n, n), dtype=np.uint8)
        for (u, v), b in zip(skel, bits):
            a[(u, v) if b else (v, u)] = 1
        if _is_dag(a) and (sd.pdag2cpdag(a) == pattern).all():
            union |= a
    return union


def test_textbook_rules_complete_a_pattern_to_its_cpdag():
    rng = np.random.default_rng(5)
    nontrivial = 0
    for _ in range(120):
        n = int(rng.integers(2, 7))
        perm = rng.permutation(n)
        dag = np.zeros((n, n), dtype=np.uint8)
        for i in range(n):
            for j in range(i + 1, n):
                if rng.random() < 0.45:
                    dag[perm[i], perm[j]] = 1
        pattern = sd.pdag2cpdag(dag)
        got = pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(pattern), textbook=True), n)
        want = _brute_force_cpdag(dag)
        assert (got == want).all(), dag.tolist()
        nontrivial += int((want != pattern).any())
        # idempotent
        assert (pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(got), textbook=True), n) == got).all()
    assert nontrivial >= 8


@pytest.mark.skipif(oracle.reference_modules()[1] is None, reason="oracle/_ref not built")
def test_meek_rules_live_against_compiled_reference():
    har = oracle.reference_modules()[1]
    if not hasattr(har, "meek"):
        pytest.skip("oracle/_ref predates the meek entry point")
    rng = np.random.default_rng(11)
    for _ in range(400):
        n = int(rng.integers(2, 14))
        dens, pdir = rng.uniform(0.1, 0.7), rng.uniform(0.0, 0.9)
        a = np.zeros((n, n), dtype=np.uint8)
        for u in range(n):
            for v in range(u + 1, n):
                if rng.random() < dens:
                    r = rng.random()
                    if r < pdir / 2:
                        a[u, v] = 1
                    elif r < pdir:
                        a[v, u] = 1
                    else:
                        a[u, v] = a[v, u] = 1
        for r4 in (False, True):
            got = pdag_to_adj(L.structure_learning.apply_meeks_rules(to_pdag(a), r4), n)
            assert (got == har.meek(a, r4)).all(), (a.tolist(), r4)


@pytest.mark.parametrize("netf", [nw.asia, nw.alarm])
@pytest.mark.parametrize("stable", [True, False])
def test_pc_oracle_graph_shd_zero_like_the_reference_test(netf, stable):
    # pytest/structure_learning/test_pc.py:11-33: PC with the d-separation oracle, judged by structural_errors
    net = netf()
    truth = truth_of(net)
    df = L.base.DataframeWrapper.from_codes(np.zeros((2, net.n), dtype=np.uint8), net.names, [1] * net.n)
    mc = net.n if net.n <= 8 else 4
    got = L.structure_learning.pc(df, L.citest.OracleGraph(truth), mc, stable)
    err = sd.structural_errors(truth, got)
    assert err["SHD"] == 0, err
    # apply_meek=True is pc() followed by the rules, and only ever orients undirected edges
    meek = L.structure_learning.pc(df, L.citest.OracleGraph(truth), mc, stable, apply_meek=True)
    a, m = pdag_to_adj(got, net.n), pdag_to_adj(meek, net.n)
    assert (m == pdag_to_adj(L.structure_learning.apply_meeks_rules(got), net.n)).all()
    assert ((m | m.T) == (a | a.T)).all() and (m <= a).all()
    # every arc the rules orient agrees with the true DAG (they are sound on a correct pattern)
    t = pdag_to_adj(truth, net.n)
    newly = (a & a.T) & (m ^ m.T) & m
    assert (newly <= t).all()
    assert sd.structural_errors(truth, meek)["SHD"] == 0
    # complete_cpdag=True: the published rules; on Asia the result is the CPDAG of the true DAG (brute force)
    full = pdag_to_adj(L.structure_learning.pc(df, L.citest.OracleGraph(truth), mc, stable, complete_cpdag=True), net.n)
    assert ((full | full.T) == (a | a.T)).all() and (full <= a).all()
    assert (((a & a.T) & (full ^ full.T) & full) <= t).all()
    if net.n <= 8:
        assert (full == _brute_force_cpdag(t)).all()
    else:
        assert (full != a).any()                   # Alarm's pattern has edges only the propagation orients


def test_exhaustive_search_shd_zero_like_the_reference_test():
    # pytest/structure_learning/test_exhaustive_search.py:22-30: 1e5 samples, max_parents = 3, judged by
    # structural_errors against the true graph (here: the Asia structure; the driver runs over the checker seam)
    import oracle as orc
    from helpers import OracleEvaluator, ranked
    net = nw.asia()
    codes, cards = ranked(net.sample(100000, 0))
    ods = orc.Dataset(codes, cards)
    got = L._testing.exhaustive_search(net.n, OracleEvaluator(ods), L.score.BDeu(1.0), 3)
    err = sd.structural_errors(truth_of(net), got)
    assert err["SHD"] == 0, err
