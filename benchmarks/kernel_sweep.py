#!/usr/bin/env python
"""kernel_sweep.py -- throughput of the counting kernel through the C ABI as a
function of tuple shape and launch geometry.

For each (n_rows, card, k, n_tuples) it evaluates a batch of random CI tests
and prints the algorithmic GB/s of the counting-kernel launches (CUDA events
inside the library, obn_profile) for several forced CTA widths / row splits
(OBN_FORCE_T / OBN_FORCE_SPLITS; 0 = the library's own choice).

usage: python benchmarks/kernel_sweep.py [--rows 100000] [--quick]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openbnsl_b200 import capi  # noqa: E402


def run(ds, ctx, flat, offs, x, y, reps=5, want_p=True):
    import ctypes as C
    n = len(offs) - 1
    stat = np.zeros(n, dtype=np.float64)
    dof = np.zeros(n, dtype=np.uint64)
    pv = np.zeros(n, dtype=np.float64) if want_p else None
    L = capi.load()

    def call():
        capi._check(L.obn_ci_tests(ctx._h, ds._h, capi._ptr(flat, C.c_uint32), capi._ptr(offs, C.c_uint64), n,
                                   capi._ptr(x, C.c_uint32), capi._ptr(y, C.c_uint32), 1,
                                   capi._ptr(stat, C.c_double), capi._ptr(dof, C.c_uint64), capi._ptr(pv, C.c_double)))
    call()
    ctx.profile_read(True)
    for _ in range(reps):
        call()
    p = ctx.profile_read(True)
    return p.algorithmic_bytes / (p.count_kernel_ms * 1e-3) / 1e9, p.count_kernel_ms / reps * 1e3, p.count_launches // reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=100_000)
    ap.add_argument("--vars", type=int, default=40)
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    ctx = capi.Context(0)
    ctx.profile_enable(True)
    rng = np.random.default_rng(0)
    shapes = [(2, 2, 666), (3, 3, 2848), (3, 4, 1400), (3, 6, 400), (4, 4, 1200), (2, 5, 3000)]
    if args.quick:
        shapes = shapes[:3]
    for card, k, n_t in shapes:
        codes = rng.integers(0, card, size=(args.rows, args.vars), dtype=np.uint8)
        codes[:card, :] = np.arange(card, dtype=np.uint8)[:, None]
        ds = ctx.upload(codes, [card] * args.vars)
        tuples = [np.sort(rng.choice(args.vars, size=k, replace=False)) for _ in range(n_t)]
        flat = np.concatenate(tuples).astype(np.uint32)
        offs = np.arange(0, (n_t + 1) * k, k, dtype=np.uint64)
        x = np.array([t[0] for t in tuples], dtype=np.uint32)
        y = np.array([t[-1] for t in tuples], dtype=np.uint32)
        print("rows=%d card=%d k=%d cells=%d tuples=%d" % (args.rows, card, k, card ** k, n_t), flush=True)
        for T, S in [(0, 0), (32, 1), (64, 1), (128, 1), (256, 1)]:
            os.environ["OBN_FORCE_T"] = str(T)
            os.environ["OBN_FORCE_SPLITS"] = str(S)
            gbs, us, nl = run(ds, ctx, flat, offs, x, y)
            gbs2, us2, _ = run(ds, ctx, flat, offs, x, y, want_p=False)
            print("   T=%-4d splits=%-2d  %8.1f GB/s  %8.1f us/batch  launches=%d   | without p-value: %8.1f GB/s %8.1f us" % (T, S, gbs, us, nl, gbs2, us2), flush=True)
        os.environ["OBN_FORCE_T"] = "0"
        os.environ["OBN_FORCE_SPLITS"] = "0"
        ds.close()


if __name__ == "__main__":
    main()
