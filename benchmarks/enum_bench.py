"""enum_bench.py -- host cost of the PC driver's frontier enumeration, without a GPU: `_testing.pc` over a synthetic
evaluator (hash-based p-values: most pairs independent at depth 0, a quarter of the deeper tests pass) on 186 variables,
serial and pipelined; prints last_pc_stats()["enumerate_ms"] per evaluated test.  CPU only."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import openbnsl_b200
L = openbnsl_b200.openbnsllib

class Ev:
    submitted = 0
    def __init__(self): self.calls = 0
    def ci_pvalues(self, var_ids, offsets, x, y, kind):
        var_ids = np.asarray(var_ids); offsets = np.asarray(offsets).astype(np.int64); x = np.asarray(x).astype(np.int64); y = np.asarray(y).astype(np.int64)
        n = len(x)
        k = np.diff(offsets)
        # hash of the tuple + roles
        h = np.add.reduceat(var_ids.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15), offsets[:-1]) if len(var_ids) else np.zeros(n, np.uint64)
        h = h ^ (np.minimum(x, y).astype(np.uint64) << np.uint64(32)) ^ (np.maximum(x, y).astype(np.uint64) << np.uint64(40))
        h = (h * np.uint64(0xFF51AFD7ED558CCD)) >> np.uint64(40)
        u = (h % np.uint64(10000)).astype(np.float64) / 10000.0
        thr = np.where(k == 2, 0.91, 0.25)          # depth 0: most pairs independent
        return np.where(u < thr, 0.5, 0.001)

T = L.citest.ChiSquare(0.05)
for overlap in (False, True):
    best = None
    for rep in range(5):
        ev = Ev()
        t0 = time.perf_counter()
        g = L._testing.pc(186, ev, T, 3, True, None, 8192, False, overlap)
        dt = time.perf_counter() - t0
        st = L.structure_learning.last_pc_stats()
        if best is None or st["enumerate_ms"] < best[0]:
            best = (st["enumerate_ms"], st["resolve_ms"], st["total_tests"], list(st["tests_per_depth"]), list(st["batches_per_depth"]))
    print("overlap", overlap, "enumerate_ms %.3f resolve_ms %.3f tests %d per depth %s rounds %s -> %.1f ns/test" % (best[0], best[1], best[2], best[3], best[4], 1e6 * best[0] / best[2]))
