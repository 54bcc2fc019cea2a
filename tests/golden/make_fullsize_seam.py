"""Generates tests/golden/pc_fullsize_seam.json: the PDAGs of BASELINE configs 4 and 5 AT FULL SIZE (1M rows),
learned by this repository's host driver (openbnsl_b200/csrc/host/pc.cpp) with every CI decision taken by the
UNMODIFIED reference functions (ContingencyTable<true> + citest<true> through oracle/ref_harness.cpp, OpenMP over the
tests of a round).  The stock reference pc() is serial and would need hours at this size; the driver itself is pinned
against the stock pc() on the down-scaled fixtures (pc_big.json) and on random networks (tests/test_driver_fuzz.py), so
this seam pins the full-size decisions -- the high-cardinality tables and the depth >= 2 frontiers at 1M rows -- that
the GPU run must reproduce arc for arc (tests/test_gpu_large.py::test_full_size_pc_matches_reference_seam).

Run in the build container (needs oracle/_ref; about 10 minutes on 8 cores):
    python tests/golden/make_fullsize_seam.py
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
import oracle  # noqa: E402
import openbnsl_b200  # noqa: E402
from openbnsl_b200 import networks as nw  # noqa: E402


class RefEvaluator:
    """CI batches answered by the unmodified reference functions (oracle/_ref harness)."""

    def __init__(self, har, df, level, threads):
        self.har, self.df, self.level, self.threads, self.tests = har, df, level, threads, 0

    def ci_pvalues(self, var_ids, offsets, x, y, kind):
        n = len(x)
        k = np.diff(offsets.astype(np.int64))
        owner = np.repeat(np.arange(n), k)
        keep = (var_ids != x[owner]) & (var_ids != y[owner])
        zo = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(np.bincount(owner[keep], minlength=n), out=zo[1:])
        r = self.har.ci_batch(self.df, x.astype(np.int64), y.astype(np.int64), var_ids[keep].astype(np.int64), zo,
                              int(kind), self.level, self.threads)
        self.tests += n
        return r["p"]


def main():
    ref, har = oracle.reference_modules()
    assert ref is not None, "build oracle/_ref first (python oracle/build_ref.py)"
    L = openbnsl_b200.openbnsllib          # host driver only: the test seam needs no GPU
    out = []
    for key in ("c4", "c5"):
        c = bench.CONFIGS[key]
        net, codes, cards = bench.make_data(key)
        rdf = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
        T = L.citest.ChiSquare(c["level"]) if c["ci_test"] == "chi2" else L.citest.GSquare(c["level"])
        ev = RefEvaluator(har, rdf, c["level"], os.cpu_count() or 1)
        t0 = time.time()
        g = L._testing.pc(net.n, ev, T, c["max_cond_vars"], True)
        secs = time.time() - t0
        arcs = [[p, v] for p in range(net.n) for v in range(net.n) if g.has_edge(p, v)]
        out.append({"config": key, "workload": c["workload"], "n_rows": c["n_rows"], "cards": cards, "arcs": arcs,
                    "tests": ev.tests, "seconds": secs,
                    "codes_sha": int(np.frombuffer(codes.tobytes(), dtype=np.uint8).astype(np.uint64).sum())})
        print(key, len(arcs), "arcs", ev.tests, "tests", "%.0f s" % secs, flush=True)
    with open(os.path.join(HERE, "pc_fullsize_seam.json"), "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main()
