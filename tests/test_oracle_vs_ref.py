"""Live cross-check of the C restatement against the compiled unmodified
reference (oracle/_ref).  Skipped where the reference build is absent."""
import numpy as np
import pytest

import oracle
from helpers import close, pdag_to_adj, ranked
from openbnsl_b200 import networks as nw

ref, har = oracle.reference_modules()
pytestmark = pytest.mark.skipif(ref is None, reason="oracle/_ref not built")


@pytest.fixture(scope="module")
def hepar():
    net = nw.hepar2_like()
    codes, cards = ranked(net.sample(4000, 7))
    return net, codes, cards, oracle.Dataset(codes, cards), ref.base.DataframeWrapper(
        nw.to_reference_frame(codes, net.names, cards))


def test_random_tuples(hepar):
    net, codes, cards, ods, df = hepar
    rng = np.random.default_rng(3)
    for _ in range(40):
        k = int(rng.integers(1, 6))
        vs = sorted(rng.choice(net.n, size=k, replace=False).tolist())
        ct = ref.base.ContingencyTable(vs, df)
        dense = oracle.count_table(ods, vs)
        assert {int(a): int(b) for a, b in enumerate(dense) if b} == dict(ct.counts)
        c = vs[int(rng.integers(0, k))]
        pa = [v for v in vs if v != c]
        assert oracle.bdeu_local_score(ods, c, pa, 2.5) == ref.score.calculate_local_score(c, pa, ct, ref.score.BDeu(2.5))


def test_ci_batch(hepar):
    net, codes, cards, ods, df = hepar
    rng = np.random.default_rng(4)
    xs, ys, zs = [], [], []
    for _ in range(60):
        vs = rng.choice(net.n, size=int(rng.integers(2, 7)), replace=False).tolist()
        xs.append(vs[0]); ys.append(vs[1]); zs.append(sorted(vs[2:]))
    zo = np.cumsum([0] + [len(z) for z in zs]).astype(np.int64)
    zf = np.asarray([v for z in zs for v in z], dtype=np.int64)
    for kind in (0, 1):
        r = har.ci_batch(df, np.asarray(xs, dtype=np.int64), np.asarray(ys, dtype=np.int64), zf, zo, kind, 0.05, 2)
        for i in range(len(xs)):
            stat, dof, p = oracle.citest(ods, xs[i], ys[i], zs[i], kind)
            assert dof == int(r["dof"][i])
            if r["stat"][i] == r["stat"][i]:
                assert close(stat, float(r["stat"][i]), 1e-9, 1e-9)
            assert close(p, float(r["p"][i]), 1e-9, 1e-12)


def test_pc_small_alarm():
    net = nw.alarm()
    codes, cards = ranked(net.sample(2000, 11))
    ods = oracle.Dataset(codes, cards)
    df = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
    for stable in (True, False):
        g = ref.structure_learning.pc(df, ref.citest.GSquare(0.01), 2, stable)
        assert (oracle.pc(ods, 1, 0.01, 2, stable) == pdag_to_adj(g, net.n)).all()


def test_pc_full_config2_alarm_100k():
    """BASELINE config 2 at full size through the STOCK reference entry point (pc.cpp:76-86, serial: about 25 s)
    against the C restatement the GPU tests are checked with (tests/test_gpu_module.py::test_config1_alarm_pc_full_size)."""
    net = nw.alarm()
    codes, cards = ranked(net.sample(100_000, 0))
    ods = oracle.Dataset(codes, cards)
    df = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
    g = ref.structure_learning.pc(df, ref.citest.GSquare(0.05), 4, True)
    assert (oracle.pc(ods, 1, 0.05, 4, True) == pdag_to_adj(g, net.n)).all()


def test_oracle_pc_matches_stock_reference_on_hepar2_like_20k():
    """The fixture of the GPU test (tests/golden/pc_big.json, made by the stock reference) also pins the restatement:
    70 variables, depth <= 3, non-stable order."""
    import json, os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pc_big.json")) as f:
        case = [c for c in json.load(f) if c["name"] == "hepar2_like_20k_nonstable"][0]
    net = getattr(nw, case["network"])()
    codes, cards = ranked(net.sample(case["n_rows"], case["seed"]))
    assert cards == case["cards"]
    adj = oracle.pc(oracle.Dataset(codes, cards), case["kind"], case["level"], case["max_cond_vars"], case["stable"])
    assert [[p, v] for p in range(net.n) for v in range(net.n) if adj[p, v]] == case["arcs"]
