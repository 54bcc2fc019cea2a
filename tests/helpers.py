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


def pdag_to_adj(g, n):
    return np.array([[1 if g.has_edge(p, v) else 0 for v in range(n)] for p in range(n)], dtype=np.uint8)
