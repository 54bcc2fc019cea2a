"""Shared test helpers.  The OracleEvaluator plugs the CHECKER (oracle/) into the
C++ drivers' test seam so that driver and sharding logic run without a GPU."""
import numpy as np

import oracle


class OracleEvaluator:
    def __init__(self, ods: oracle.Dataset):
        self.ods = ods
        self.n_ci = 0
        self.n_sc = 0

    def ci_pvalues(self, var_ids, offsets, x, y, kind):
        out = np.zeros(len(x), dtype=np.float64)
        for t in range(len(x)):
            vs = [int(v) for v in var_ids[int(offsets[t]):int(offsets[t + 1])]]
            z = [v for v in vs if v != int(x[t]) and v != int(y[t])]
            out[t] = oracle.citest(self.ods, int(x[t]), int(y[t]), z, int(kind))[2]
        self.n_ci += len(x)
        return out

    def subset_scores(self, var_ids, offsets, ess, exact):
        out = np.zeros(len(var_ids), dtype=np.float64)
        for t in range(len(offsets) - 1):
            vs = [int(v) for v in var_ids[int(offsets[t]):int(offsets[t + 1])]]
            for i, c in enumerate(vs):
                out[int(offsets[t]) + i] = oracle.bdeu_local_score(self.ods, c, [v for v in vs if v != c], ess)
        self.n_sc += len(var_ids)
        return out

    def child_scores(self, var_ids, offsets, child, ess, exact):
        out = np.zeros(len(child), dtype=np.float64)
        for t in range(len(child)):
            vs = [int(v) for v in var_ids[int(offsets[t]):int(offsets[t + 1])]]
            out[t] = oracle.bdeu_local_score(self.ods, int(child[t]), [v for v in vs if v != int(child[t])], ess)
        self.n_sc += len(child)
        return out


def pdag_to_adj(g, n):
    return np.array([[1 if g.has_edge(p, v) else 0 for v in range(n)] for p in range(n)], dtype=np.uint8)


# ---------------------------------------------------------------------------
# golden fixtures (tests/golden/*.json, made by tests/golden/make_golden.py)
# ---------------------------------------------------------------------------
import json
import os

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_SETS = ["asia_10k_s0", "alarm_3k_s1", "munin_like_2k_s2"]


def ranked(codes):
    """Re-code every column as ranks of its OBSERVED states (dataframe_wrapper.cpp:36-58)."""
    cols, cards = [], []
    for v in range(codes.shape[1]):
        u, inv = np.unique(codes[:, v], return_inverse=True)
        cols.append(inv.astype(np.uint8))
        cards.append(len(u))
    return np.stack(cols, axis=1), cards


def load_golden(name):
    from openbnsl_b200 import networks as nw
    with open(os.path.join(GOLDEN_DIR, name + ".json")) as f:
        g = json.load(f)
    net = {"asia": nw.asia, "alarm": nw.alarm, "munin": nw.munin1_like}[g["network"]]()
    codes, cards = ranked(net.sample(g["n_rows"], g["seed"]))
    assert cards == g["cards"]
    assert int(np.frombuffer(codes.tobytes(), dtype=np.uint8).astype(np.uint64).sum()) == g["codes_sha"], \
        "synthetic data generator drifted from the one the fixtures were made with"
    return g, net, codes, cards


def close(a, b, rel=1e-9, abs_=0.0):
    return abs(a - b) <= max(rel * max(abs(a), abs(b)), abs_)
