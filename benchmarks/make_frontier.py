#!/usr/bin/env python
"""make_frontier.py -- record the CI tests one PC run of a bench.py configuration evaluates (needs a GPU).

bench.py's reference arm times the UNMODIFIED reference functions over a uniform sample of these tests
without loading any of this repository's native libraries, so the list is written once by this
script and committed: benchmarks/frontiers/<config>_tests.npz (var_ids, offsets, x, y; the data are
seeded synthetic samples, so the list is reproducible).

usage: python benchmarks/make_frontier.py [c5 c4 c2]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import openbnsl_b200  # noqa: E402

L = openbnsl_b200.openbnsllib
out_dir = os.path.join(ROOT, sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else "benchmarks/frontiers")
os.makedirs(out_dir, exist_ok=True)
for key in [a for a in sys.argv[1:] if a in bench.CONFIGS] or ["c5", "c4", "c2"]:
    c = bench.CONFIGS[key]
    net, codes, cards = bench.make_data(key)
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    test = L.citest.ChiSquare(c["level"]) if c["ci_test"] == "chi2" else L.citest.GSquare(c["level"])
    L.structure_learning.pc(df, test, c["max_cond_vars"], True, record_tests=True)
    var_ids, offsets, x, y = L.structure_learning.last_pc_tests()
    small = np.uint8 if net.n <= 256 else np.uint16
    np.savez_compressed(os.path.join(out_dir, key + "_tests.npz"), var_ids=var_ids.astype(small),
                        offsets=offsets.astype(np.uint32), x=x.astype(small), y=y.astype(small))
    print(key, len(x), "tests", os.path.getsize(os.path.join(out_dir, key + "_tests.npz")), "bytes")
