"""Generates tests/golden/*.json from the UNMODIFIED reference compiled in
oracle/_ref (oracle/build_ref.py).  Run in the build container, where
/root/reference exists:   python tests/golden/make_golden.py

The reference's own tests hold no stored vectors for this path (every
expectation is computed live by pgmpy), so the fixtures
are outputs of the reference itself on seeded synthetic data.  The statistic
and dof of citest<>() are read through the Boost-shim recorder
(oracle/ref_harness.cpp); everything else comes from the reference's Python
module.  The enumeration KATs are the gtest known answers
(gtest/utils/next_combination.cpp:19-58, gtest/utils/comb2vec.cpp:7-56).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from openbnsl_b200 import networks as nw  # noqa: E402


def adj_of(g, n):
    return [[int(g.has_edge(p, v)) for v in range(n)] for p in range(n)]


def dataset_cases():
    yield "asia_10k_s0", nw.asia(), 10000, 0
    yield "alarm_3k_s1", nw.alarm(), 3000, 1
    yield "munin_like_2k_s2", nw.munin1_like(), 2000, 2


SA_CASES = [  # max_parents, max_iters, init_temp, cooling_rate, is_deterministic, seed, num_chains
    (3, 800, 1.0, 0.9995, False, 0, 3), (3, 800, 1.0, 0.9995, True, 0, 3), (2, 500, 5.0, 0.99, False, 7, 4),
    (1, 400, 0.5, 0.999, True, 123, 2), (0, 60, 1.0, 0.9995, False, 1, 2)]


def make_sa(ref):
    """simulated_annealing.json: the reference's simulated_annealing() on two of the golden datasets
    (SURVEY.md section 8f rank 1).  Kept in its own file so the other fixtures stay byte-identical."""
    out = {}
    for name, net, n_rows, seed in dataset_cases():
        if net.n > 40:
            continue
        codes = net.sample(n_rows, seed)
        cols, cards = [], []
        for v in range(net.n):
            u, inv = np.unique(codes[:, v], return_inverse=True)
            cols.append(inv.astype(np.uint8)); cards.append(len(u))
        codes = np.stack(cols, axis=1)
        df = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
        runs = []
        for mp, iters, t0, cr, det, sd, chains in SA_CASES:
            g = ref.structure_learning.simulated_annealing(df, ref.score.BDeu(1.0), mp, iters, t0, cr, det, sd, chains)
            runs.append({"max_parents": mp, "max_iters": iters, "init_temp": t0, "cooling_rate": cr,
                         "is_deterministic": det, "seed": sd, "num_chains": chains, "ess": 1.0, "adj": adj_of(g, net.n)})
        out[name] = runs
    with open(os.path.join(HERE, "simulated_annealing.json"), "w") as f:
        json.dump(out, f)
    print("wrote simulated_annealing.json", os.path.getsize(os.path.join(HERE, "simulated_annealing.json")), "bytes")


def random_pdag(rng, n, dens, pdir):
    adj = np.zeros((n, n), dtype=np.uint8)
    for u in range(n):
        for v in range(u + 1, n):
            if rng.random() < dens:
                r = rng.random()
                if r < pdir / 2:
                    adj[u, v] = 1
                elif r < pdir:
                    adj[v, u] = 1
                else:
                    adj[u, v] = adj[v, u] = 1
    return adj


def load_reference_structural_distance():
    """helpers/structural_distance.py of the reference, imported from where it lies.  Its only
    non-stdlib imports are networkx (present) and pgmpy.base.PDAG (absent): pgmpy's PDAG is a
    networkx.DiGraph subclass and the helper uses DiGraph methods only, so a two-line stand-in
    module is enough -- the helper's own code runs unmodified."""
    import importlib.util
    import types
    import networkx as nx
    pg, base = types.ModuleType("pgmpy"), types.ModuleType("pgmpy.base")
    base.PDAG = type("PDAG", (nx.DiGraph,), {})
    pg.base = base
    sys.modules.setdefault("pgmpy", pg)
    sys.modules.setdefault("pgmpy.base", base)
    spec = importlib.util.spec_from_file_location(
        "ref_structural_distance", os.path.join(os.environ.get("OBN_REFERENCE_DIR", "/root/reference"),
                                                "helpers", "structural_distance.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, base.PDAG


def make_graphs(har):
    """graph_kat.json: Meek's rules from the reference's own
    PDAGwithAdjMat::apply_meeks_rules (through oracle/ref_harness.cpp) and PDAG2CPDAG /
    structural_errors from the reference's helpers/structural_distance.py, on seeded random PDAGs
    and on the true DAGs of the example networks."""
    sd, RefPDAG = load_reference_structural_distance()

    def to_ref(adj):
        g = RefPDAG()
        g.add_nodes_from(range(len(adj)))
        g.add_edges_from((int(u), int(v)) for u, v in zip(*np.nonzero(adj)))
        return g

    def from_ref(g, n):
        a = np.zeros((n, n), dtype=np.uint8)
        for u, v in g.edges():
            a[u, v] = 1
        return a

    rng = np.random.default_rng(20260)
    graphs = []
    for i in range(48):
        n = int(rng.integers(2, 11))
        graphs.append(random_pdag(rng, n, rng.uniform(0.15, 0.7), rng.uniform(0.0, 0.9)))
    for net in (nw.asia(), nw.alarm()):
        a = np.zeros((net.n, net.n), dtype=np.uint8)
        for v, ps in enumerate(net.parents):
            for p_ in ps:
                a[p_, v] = 1
        graphs.append(a)
        graphs.append(from_ref(sd.PDAG2CPDAG(to_ref(a)), net.n))     # pattern of the true DAG: Meek has work to do
    cases = []
    for a in graphs:
        n = len(a)
        cases.append({"n": n, "adj": a.tolist(), "meek": har.meek(a, False).tolist(),
                      "meek_r4": har.meek(a, True).tolist(), "cpdag": from_ref(sd.PDAG2CPDAG(to_ref(a)), n).tolist()})
    pairs = []
    for i in range(40):
        n = int(rng.integers(3, 10))
        gt = random_pdag(rng, n, rng.uniform(0.2, 0.6), rng.uniform(0.3, 1.0))
        pr = gt.copy()
        for _ in range(int(rng.integers(0, 6))):                        # perturb: drop / add / flip / undirect
            u, v = rng.choice(n, size=2, replace=False)
            op = int(rng.integers(0, 4))
            if op == 0:
                pr[u, v] = pr[v, u] = 0
            elif op == 1:
                pr[u, v] = 1
            elif op == 2:
                pr[u, v], pr[v, u] = pr[v, u], pr[u, v]
            else:
                if pr[u, v] or pr[v, u]:
                    pr[u, v] = pr[v, u] = 1
        pairs.append({"n": n, "truth": gt.tolist(), "pred": pr.tolist(),
                      "errors": sd.structural_errors(to_ref(gt), to_ref(pr))})
    with open(os.path.join(HERE, "graph_kat.json"), "w") as f:
        json.dump({"meek_cpdag": cases, "structural_errors": pairs}, f)
    print("wrote graph_kat.json", os.path.getsize(os.path.join(HERE, "graph_kat.json")), "bytes")


BIG_PC_CASES = [  # name, network, rows, seed, kind (0 chi2, 1 G2), level, max_cond_vars, stable
    ("hepar2_like_20k", "hepar2_like", 20000, 4, 1, 0.05, 4, True),
    ("hepar2_like_20k_nonstable", "hepar2_like", 20000, 4, 1, 0.05, 3, False),
    ("munin1_like_20k", "munin1_like", 20000, 5, 0, 0.05, 3, True),
    ("munin1_like_20k_g2", "munin1_like", 20000, 5, 1, 0.01, 2, True)]


def make_big_pc(ref):
    """pc_big.json: the STOCK reference pc() (bindings/submodules/structure_learning.cpp:31-37 ->
    src/structure_learning/pc.cpp:76-86) on down-scaled stand-ins of BASELINE configs 4 and 5 -- the
    same networks as the benchmark at a row count the serial reference finishes in minutes.  Depth >= 2
    frontiers with cardinalities up to 21 and every table tier inside a real PC run are compared
    PDAG for PDAG on the GPU (tests/test_gpu_module.py::test_pc_big_networks_match_reference)."""
    import time
    out = []
    for name, netname, n_rows, seed, kind, level, mc, stable in BIG_PC_CASES:
        net = getattr(nw, netname)()
        codes = net.sample(n_rows, seed)
        cols, cards = [], []
        for v in range(net.n):
            u, inv = np.unique(codes[:, v], return_inverse=True)
            cols.append(inv.astype(np.uint8)); cards.append(len(u))
        codes = np.stack(cols, axis=1)
        df = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
        T = ref.citest.ChiSquare(level) if kind == 0 else ref.citest.GSquare(level)
        t0 = time.time()
        g = ref.structure_learning.pc(df, T, mc, stable)
        secs = time.time() - t0
        arcs = [[p, v] for p in range(net.n) for v in range(net.n) if g.has_edge(p, v)]
        out.append({"name": name, "network": netname, "n_rows": n_rows, "seed": seed, "kind": kind, "level": level,
                    "max_cond_vars": mc, "stable": stable, "cards": cards, "arcs": arcs, "reference_seconds": secs,
                    "codes_sha": int(np.frombuffer(codes.tobytes(), dtype=np.uint8).astype(np.uint64).sum())})
        print(name, len(arcs), "arcs", "%.1f s" % secs, flush=True)
    with open(os.path.join(HERE, "pc_big.json"), "w") as f:
        json.dump(out, f)
    print("wrote pc_big.json", os.path.getsize(os.path.join(HERE, "pc_big.json")), "bytes")


def main():
    ref, har = oracle.reference_modules()
    assert ref is not None, "build oracle/_ref first (python oracle/build_ref.py)"
    if "--big-pc" in sys.argv:
        make_big_pc(ref)
        return
    if "--sa" in sys.argv:
        make_sa(ref)
        return
    if "--graphs" in sys.argv:
        make_graphs(har)
        return
    make_sa(ref)
    make_graphs(har)
    for name, net, n_rows, seed in dataset_cases():
        codes = net.sample(n_rows, seed)
        # ranks of observed states, as DataframeWrapper would compute them
        cols, cards = [], []
        for v in range(net.n):
            u, inv = np.unique(codes[:, v], return_inverse=True)
            cols.append(inv.astype(np.uint8)); cards.append(len(u))
        codes = np.stack(cols, axis=1)
        df = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
        assert list(df.num_of_values) == cards
        rng = np.random.default_rng(1000 + seed)
        n = net.n
        out = {"name": name, "n_rows": n_rows, "seed": seed, "cards": cards, "network": name.split("_")[0],
               "codes_sha": int(np.frombuffer(codes.tobytes(), dtype=np.uint8).astype(np.uint64).sum())}
        # ---- contingency tables ------------------------------------------------
        tables = []
        for k in (0, 1, 2, 3, 4, 5):
            for _ in range(3):
                vs = sorted(rng.choice(n, size=k, replace=False).tolist())
                ct = ref.base.ContingencyTable(vs, df)
                tables.append({"var_ids": vs, "counts": {str(a): int(b) for a, b in ct.counts.items()}})
        out["tables"] = tables
        # ---- BDeu local scores ---------------------------------------------------
        scores = []
        for npar in (0, 1, 2, 3, 4, 5):
            for ess in (1.0, 10.0):
                vs = rng.choice(n, size=npar + 1, replace=False).tolist()
                child, pa = vs[0], sorted(vs[1:])
                ct = ref.base.ContingencyTable(sorted(vs), df)
                s = ref.score.calculate_local_score(child, pa, ct, ref.score.BDeu(ess))
                scores.append({"child": child, "parents": pa, "ess": ess, "score": s})
        out["scores"] = scores
        # ---- CI tests (stat / dof / p through the shim recorder) -----------------
        xs, ys, zs = [], [], []
        for nz in (0, 0, 1, 1, 2, 2, 3, 4):
            for _ in range(4):
                vs = rng.choice(n, size=nz + 2, replace=False).tolist()
                xs.append(vs[0]); ys.append(vs[1]); zs.append(sorted(vs[2:]))
        zf = [v for z in zs for v in z]
        zo = np.cumsum([0] + [len(z) for z in zs])
        ci = []
        for kind in (0, 1):
            r = har.ci_batch(df, np.asarray(xs, dtype=np.int64), np.asarray(ys, dtype=np.int64),
                             np.asarray(zf, dtype=np.int64), zo.astype(np.int64), kind, 0.05, 1)
            for i in range(len(xs)):
                ct = ref.base.ContingencyTable(sorted(zs[i] + [xs[i], ys[i]]), df)
                T = ref.citest.ChiSquare(0.05) if kind == 0 else ref.citest.GSquare(0.05)
                dec = bool(ref.citest.citest(xs[i], ys[i], zs[i], ct, T))
                assert dec == bool(r["decision"][i])
                st = float(r["stat"][i])
                ci.append({"x": xs[i], "y": ys[i], "z": zs[i], "kind": kind, "stat": None if st != st else st,
                           "dof": int(r["dof"][i]), "p": float(r["p"][i]), "independent_at_0.05": dec})
        out["ci"] = ci
        # ---- drivers ----------------------------------------------------------------
        if n <= 8:
            out["exhaustive_search"] = [
                {"max_parents": mp, "ess": 1.0,
                 "adj": adj_of(ref.structure_learning.exhaustive_search(df, ref.score.BDeu(1.0), mp), n)}
                for mp in (0, 1, 3, 7)]
        if n <= 40:
            pcs = []
            for kind, level in ((0, 0.01), (1, 0.05)):
                for stable in (True, False):
                    T = ref.citest.ChiSquare(level) if kind == 0 else ref.citest.GSquare(level)
                    mc = 4 if n <= 8 else 2
                    g = ref.structure_learning.pc(df, T, mc, stable)
                    pcs.append({"kind": kind, "level": level, "stable": stable, "max_cond_vars": mc, "adj": adj_of(g, n)})
            out["pc"] = pcs
        with open(os.path.join(HERE, name + ".json"), "w") as f:
            json.dump(out, f)
        print("wrote", name, os.path.getsize(os.path.join(HERE, name + ".json")), "bytes")

    # gtest known answers for the enumeration utilities
    kat = {"next_combmask_counts": [{"n": 5, "k": 3, "count": 10}, {"n": 10, "k": 3, "count": 120},
                                    {"n": 31, "k": 3, "count": 4495}, {"n": 63, "k": 3, "count": 39711},
                                    {"n": 0, "k": 0, "count": 1}, {"n": 31, "k": 0, "count": 1},
                                    {"n": 63, "k": 0, "count": 1}, {"n": 31, "k": 31, "count": 1},
                                    {"n": 63, "k": 63, "count": 1}],
           "combmask2vec": [{"mask": 0, "vec": []}, {"mask": 1, "vec": [0]}, {"mask": 0b1010, "vec": [1, 3]},
                            {"mask": 31, "vec": [0, 1, 2, 3, 4]}, {"mask": (1 << 50) | (1 << 30), "vec": [30, 50]},
                            {"mask": (1 << 64) - 1, "vec": list(range(64))}]}
    with open(os.path.join(HERE, "enumeration_kat.json"), "w") as f:
        json.dump(kat, f)


if __name__ == "__main__":
    main()
