#!/usr/bin/env python
"""ab_ci_pool.py -- A/B of the CI-test pooling on one B200: the same pc() call with
OBN_CI_POOL=0 and =1 (and a few table / width caps), wall time per call and CUDA-event time of the counting kernels.

usage: python benchmarks/ab_ci_pool.py [c2] [c4] [c5]      (prints one JSON line per setting)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openbnsl_b200  # noqa: E402
from openbnsl_b200 import networks as nw  # noqa: E402

L = openbnsl_b200.openbnsllib


def frame(net, n_rows, seed):
    codes = net.sample(n_rows, seed)
    cards = []
    for v in range(net.n):
        u = np.unique(codes[:, v])
        if len(u) != int(u.max()) + 1:
            lut = np.zeros(256, dtype=np.uint8)
            lut[u] = np.arange(len(u), dtype=np.uint8)
            codes[:, v] = lut[codes[:, v]]
        cards.append(len(u))
    return L.base.DataframeWrapper.from_codes(codes, net.names, cards)


def adj_of(g, n):
    return np.array([[g.has_edge(a, b) for b in range(n)] for a in range(n)], dtype=np.uint8)


def run(name, net, n_rows, test, depth, reps, settings):
    df = frame(net, n_rows, 0)
    df.to_device()
    L.profile_enable(True)
    base = None
    for pool, cells, max_z in settings:
        os.environ["OBN_CI_POOL"] = str(pool)
        os.environ["OBN_CI_POOL_CELLS"] = str(cells)
        os.environ["OBN_CI_POOL_VARS"] = str(max_z)
        g = L.structure_learning.pc(df, test, depth, True)               # warm-up of this setting
        L.profile_read(True)
        t0 = time.perf_counter()
        for _ in range(reps):
            g = L.structure_learning.pc(df, test, depth, True)
        secs = (time.perf_counter() - t0) / reps
        prof = L.profile_read(True)
        a = adj_of(g, net.n)
        if base is None:
            base = a
        st = L.structure_learning.last_pc_stats()
        print(json.dumps({"config": name, "pool": pool, "pool_cells": cells, "pool_vars": max_z, "ms_per_pc": secs * 1e3,
                          "kernel_ms": prof["count_kernel_ms"] / reps, "launches": prof["count_launches"] / reps,
                          "tests": int(sum(st["tests_per_depth"])), "tests_per_s": sum(st["tests_per_depth"]) / secs,
                          "same_graph_as_first": bool((a == base).all()),
                          "host_prep_ms": prof["host_prep_ms"] / reps, "host_wait_ms": prof["host_wait_ms"] / reps}),
              flush=True)


def main():
    which = [a.lower() for a in sys.argv[1:]] or ["c2"]
    if "c2" in which:
        run("C2_alarm37_100k_g2_d4", nw.alarm(), 100_000, L.citest.GSquare(0.05), 4, 100,
            [(0, 4096, 4), (1, 4096, 4), (1, 4096, 30), (1, 1024, 3), (1, 256, 2)])
    if "c4" in which:
        run("C4_hepar70_1M_g2_d4", nw.hepar2_like(), 1_000_000, L.citest.GSquare(0.05), 4, 5,
            [(0, 4096, 4), (1, 4096, 4), (1, 4096, 30), (1, 1024, 3), (1, 256, 2)])
    if "c5" in which:
        run("C5_munin186_1M_chi2_d3", nw.munin1_like(), 1_000_000, L.citest.ChiSquare(0.05), 3, 3,
            [(0, 4096, 4), (1, 4096, 4), (1, 1024, 3)])


if __name__ == "__main__":
    main()
