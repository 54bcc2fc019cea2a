#!/usr/bin/env python
"""run_configs.py -- the five BASELINE.json configurations on one B200.

Prints one JSON line per configuration: learn time, CI tests/s or local
scores/s, and the counting kernel's algorithmic GB/s (CUDA events inside the
library) as a fraction of the measured HBM peak.  C3's full lattice (2^24
tables x 1M rows) is not run to completion on one GPU; its first levels
(max_parents = 3 and 5) are.

usage: python benchmarks/run_configs.py [c1 c2 c3 c4 c5]
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
COMM = None          # set in main() when launched under torchrun (one rank per GPU)
RANK = 0
PC_KW = {}
OPTS = {}
REPS = None          # reps=<n> on the command line overrides the per-config repetition count           # extra keyword arguments of pc() from the command line (tb=<target_batch>, overlap=0/1, shard=<bytes>)


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def frame(net, n_rows, seed):
    t0 = time.perf_counter()
    codes = net.sample(n_rows, seed)
    cards = []
    for v in range(net.n):                     # codes must be ranks of OBSERVED states
        u = np.unique(codes[:, v])
        if len(u) != int(u.max()) + 1:
            lut = np.zeros(256, dtype=np.uint8)
            lut[u] = np.arange(len(u), dtype=np.uint8)
            codes[:, v] = lut[codes[:, v]]
        cards.append(len(u))
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    return df, cards, time.perf_counter() - t0


def report(name, what, units, unit_name, secs, prof, extra=None):
    k_s = prof["count_kernel_ms"] * 1e-3
    gbs = prof["algorithmic_bytes"] / k_s / 1e9 if k_s > 0 else 0.0
    line = {"config": name, "what": what, "learn_time_s": secs, "learn_ms": secs * 1e3, unit_name: units, unit_name + "_per_s": units / secs,
            "kernel_ms": prof["count_kernel_ms"], "algorithmic_GB": prof["algorithmic_bytes"] / 1e9,
            "kernel_algorithmic_GBps": gbs, "hbm_peak_GBps": peak(), "roofline_frac": gbs / peak(),
            "launches": int(prof["count_launches"]), "host_prep_ms": prof["host_prep_ms"],
            "host_wait_ms": prof["host_wait_ms"]}
    if extra:
        line.update(extra)
    if COMM is not None:
        line["n_gpus"] = COMM.world
    if RANK == 0:
        print(json.dumps(line), flush=True)


def run_pc(name, net, n_rows, test, depth, what, reps=1, stable=True):
    reps = REPS or reps
    df, cards, t_gen = frame(net, n_rows, 0)
    L.profile_enable(True)
    df.to_device()
    L.structure_learning.pc(df, test, depth, stable, comm=COMM, **PC_KW)   # warm-up (allocations, first launches, NCCL)
    L.profile_read(True)
    if COMM is not None:
        import torch.distributed as dist
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        g = L.structure_learning.pc(df, test, depth, stable, comm=COMM, **PC_KW)
    secs = (time.perf_counter() - t0) / reps
    prof = L.profile_read(True)
    for key in ("count_kernel_ms", "algorithmic_bytes", "host_prep_ms", "host_wait_ms", "count_launches", "scanned_bytes",
                "increments", "tables", "collectives"):
        prof[key] = prof[key] / reps
    st = L.structure_learning.last_pc_stats()
    n_edges = sum(1 for v in range(net.n) for p in range(net.n) if g.has_edge(p, v))
    report(name, what, st["total_tests"], "ci_tests", secs, prof,
           {"tests_per_depth": [int(x) for x in st["tests_per_depth"]], "n_vars": net.n, "n_rows": n_rows,
            "max_card": int(max(cards)), "pdag_arcs": n_edges, "enumerate_ms": st["enumerate_ms"],
            "evaluate_ms": st["evaluate_ms"], "resolve_ms": st["resolve_ms"], "exposed_ms": st.get("exposed_ms"),
            "collect_wait_ms": st.get("collect_wait_ms"), "data_gen_s": t_gen,
            "batches_per_depth": [int(x) for x in st["batches_per_depth"]], "tables": prof["tables"],
            "scanned_GB": prof["scanned_bytes"] / 1e9, "atoms_ceiling_ms": prof["increments"] / 4.65e12 * 1e3,
            "collectives": prof["collectives"], "pc_kwargs": dict(PC_KW), "options": dict(OPTS),
            "adj_sha": int(sum((i + 1) * (j + 7) for i in range(net.n) for j in range(net.n) if g.has_edge(i, j)))})


def run_es(name, net, n_rows, max_parents, what, exact=2):
    df, cards, t_gen = frame(net, n_rows, 0)
    L.profile_enable(True)
    df.to_device()
    L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, exact_order=exact, comm=COMM)   # warm-up step
    L.profile_read(True)
    if COMM is not None:
        import torch.distributed as dist
        dist.barrier()
    t0 = time.perf_counter()
    L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, exact_order=exact, comm=COMM)
    secs = time.perf_counter() - t0
    prof = L.profile_read(True)
    st = L.structure_learning.last_es_stats()
    report(name, what, st["scores"], "local_scores", secs, prof,
           {"tables": int(st["tables"]), "n_vars": net.n, "n_rows": n_rows, "max_parents": max_parents,
            "exact_order": exact, "data_gen_s": t_gen, "enumerate_ms": st["enumerate_ms"], "score_ms": st["score_ms"],
            "gather_ms": st["gather_ms"], "dp_ms": st["dp_ms"]})


def run_sa(name, net, n_rows, max_parents, iters, chains, what):
    """Multi-chain simulated annealing: lockstep chains, one batch of new local
    scores per step.  Also times the unmodified reference on a bounded sample (2 chains, 100 steps)."""
    df, cards, t_gen = frame(net, n_rows, 0)
    L.profile_enable(True)
    df.to_device()
    L.structure_learning.simulated_annealing(df, L.score.BDeu(1.0), max_parents, 50, 1.0, 0.9995, False, 0, chains, comm=COMM)
    L.profile_read(True)
    t0 = time.perf_counter()
    L.structure_learning.simulated_annealing(df, L.score.BDeu(1.0), max_parents, iters, 1.0, 0.9995, False, 0, chains, comm=COMM)
    secs = time.perf_counter() - t0
    prof = L.profile_read(True)
    st = L.structure_learning.last_sa_stats()
    extra = {"chains": chains, "max_iters": iters, "proposals": int(st["proposals"]), "gpu_scores": int(st["evaluated"]),
             "batches": int(st["batches"]), "evaluate_ms": st["evaluate_ms"], "best_score": st["best_score"],
             "proposals_per_s": st["proposals"] * (COMM.world if COMM is not None else 1) / secs,
             "n_vars": net.n, "n_rows": n_rows, "max_parents": max_parents, "data_gen_s": t_gen}
    if RANK == 0:
        try:
            import oracle
            ref, _ = oracle.reference_modules()
            if ref is not None:
                codes = net.sample(n_rows, 0)
                rdf = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
                t0 = time.perf_counter()
                ref.structure_learning.simulated_annealing(rdf, ref.score.BDeu(1.0), max_parents, 100, 1.0, 0.9995, False, 0, 2)
                dt = time.perf_counter() - t0
                extra["reference_cpu_proposals_per_s_per_chain_thread"] = 100.0 / dt * 2 / min(2, os.cpu_count() or 1)
                extra["reference_sample"] = "2 chains x 100 steps of the unmodified reference (one OpenMP thread per chain)"
        except Exception as e:                      # the checker is optional here
            extra["reference_error"] = str(e)[:200]
    report(name, what, st["proposals"], "sa_proposals", secs, prof, extra)


def main():
    global COMM, RANK, REPS
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch
        import torch.distributed as dist
        from openbnsl_b200.dist import TorchComm, init_native_comm
        local = int(os.environ["LOCAL_RANK"])
        os.environ["OBN_DEVICE"] = str(local)
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        COMM = TorchComm() if "hostcomm" in sys.argv else init_native_comm(L)
        RANK = COMM.rank
    args = [a.lower() for a in sys.argv[1:] if a != "hostcomm"]
    for a in list(args):
        if a.startswith("tb="):
            PC_KW["target_batch"] = int(a[3:]); args.remove(a)
        elif a.startswith("overlap="):
            PC_KW["overlap"] = bool(int(a[8:])); args.remove(a)
        elif a.startswith("opt:"):                      # opt:<name>=<value> -> obn_ctx_set_option
            name, val = a[4:].split("=")
            L.set_option(name, float(val)); OPTS[name] = float(val); args.remove(a)
        elif a.startswith("reps="):
            REPS = int(a[5:]); args.remove(a)
        elif a.startswith("shard="):
            PC_KW["shard_min_bytes"] = float(a[6:]); args.remove(a)
    which = args or ["c1", "c2", "c3", "c4", "c5"]
    if "sa" in which:
        run_sa("SA", nw.alarm(), 100_000, 3, 2000, 64,
               "simulated annealing BDeu, Alarm structure, 37 vars, 100k rows, 64 chains x 2000 steps, max_parents=3")
    if "c1" in which:
        run_es("C1", nw.asia(), 10_000, 7, "exhaustive search BDeu, Asia, 8 vars, 10k rows, full lattice")
        run_es("C1", nw.asia(), 10_000, 3, "exhaustive search BDeu, Asia, 8 vars, 10k rows, max_parents=3")
    if "c2" in which:
        run_pc("C2", nw.alarm(), 100_000, L.citest.GSquare(0.05), 4,
               "PC-stable G2, Alarm structure, 37 vars, 100k rows, depth<=4", reps=20)
    if "c2ns" in which:                   # the original (order-dependent) PC: sequential over pairs by definition, one batch per pair
        run_pc("C2-nonstable", nw.alarm(), 100_000, L.citest.GSquare(0.05), 4,
               "PC (stable=False) G2, Alarm structure, 37 vars, 100k rows, depth<=4", reps=3, stable=False)
    if "c3" in which:
        net = nw.random_binary(24)
        run_es("C3", net, 1_000_000, 3, "exhaustive search BDeu, 24 binary vars, 1M rows, max_parents=3 (12 951 tables)", exact=0)
        run_es("C3", net, 1_000_000, 5, "exhaustive search BDeu, 24 binary vars, 1M rows, max_parents=5 (190 051 tables)", exact=0)
    for w in which:                       # c3mp<k>[x]: C3 with max_parents=k (x = exact-order scores)
        if w.startswith("c3mp"):
            ex = 2 if w.endswith("x") else 0
            mp = int(w[4:].rstrip("x"))
            run_es("C3", nw.random_binary(24), 1_000_000, mp,
                   "exhaustive search BDeu, 24 binary vars, 1M rows, max_parents=%d%s" % (mp, " exact-order" if ex else ""), exact=ex)
    if "c4" in which:
        run_pc("C4", nw.hepar2_like(), 1_000_000, L.citest.GSquare(0.05), 4,
               "PC-stable G2, Hepar2-like, 70 vars, 1M rows, depth<=4", reps=5)
    if "c5" in which:
        run_pc("C5", nw.munin1_like(), 1_000_000, L.citest.ChiSquare(0.05), 3,
               "PC-stable chi2, Munin1-like, 186 vars, cards<=21, 1M rows, depth<=3", reps=5)


if __name__ == "__main__":
    main()
