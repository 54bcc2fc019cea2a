#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 contingency-counting hot path.

Workload (N = 1 and N > 1 alike): BASELINE.json configs[4], the largest single-GPU
configuration and the only one whose sample matrix (186 MB) does not fit the 126 MB L2:
PC-stable with the chi-square test (alpha 0.05, conditioning depth <= 3) on 1 000 000 samples
forward-drawn from a Munin1-like network (186 variables, cardinalities 2..21).
One STEP = one complete `structure_learning.pc` call = every CI test of every depth (contingency
table + statistic + dof + p-value + decision per test), through the C++ driver and the extern-"C"
CUDA layer.

  value     CI tests / s with the sample matrix already resident in HBM
  e2e       the same call starting from HOST memory: dataset upload (H2D of the whole matrix),
            tuple descriptors H2D and the per-edge decisions D2H all inside the timed region
  roofline  the counting kernel, three ways:
              frac        algorithmic bytes (sum over TESTS of (|Z|+2) * N) / CUDA-event
                          time of the kernel launches / measured HBM peak -- the accounting unit; tests
                          that share a table make it exceed what the kernels actually read
              dram_frac   real DRAM bytes (ncu dram__bytes, profiles/r2_traffic.json) / the same time / peak
              atoms_frac  shared-memory increments applied / 4.65e12 per s (ATOMS.POPC.INC ceiling:
                          16 per clk per SM x 148 SMs x 1.965 GHz) / the same time -- the ceiling that binds
              step_frac   algorithmic bytes / whole step time / peak
  cpu_baseline  the UNMODIFIED reference functions (ContingencyTable + citest, oracle/_ref) over a
            uniform sample of the very same tests, OpenMP over tests on all host cores, extrapolated
            (SURVEY.md 8d-iii); the reference's own pc() is serial and cannot finish this configuration

N > 1 (`torchrun`, one rank per GPU): ONE learn task, the CI tests of every round sharded over the
ranks inside the library (obn_pc_round: tables dealt to the ranks, ncclAllReduce(min) of the first
passing candidate per edge on the device buffer) -- "scaling": "strong".  The learned graph is
checked against the unsharded run on the same ranks (`parity_ok`).

`--impl reference` times the reference's CPU implementation alone (rank 0), without loading any of
this repository's native libraries: the unmodified reference functions under OpenMP over a uniform
sample of the workload's tests (list: benchmarks/frontiers/, written by benchmarks/make_frontier.py),
plus the stock `structure_learning.pc` entry point on a row sub-sample it can finish.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ATOMS_PER_S = 16 * 148 * 1.965e9          # shared-memory increments per second at the ATOMS.POPC.INC ceiling

CONFIGS = {
    "c5": {"workload": "pc_stable_chi2_munin1like186_1M_depth3", "net": "munin1_like", "n_rows": 1_000_000,
           "ci_test": "chi2", "level": 0.05, "max_cond_vars": 3},
    "c4": {"workload": "pc_stable_g2_hepar2like70_1M_depth4", "net": "hepar2_like", "n_rows": 1_000_000,
           "ci_test": "G2", "level": 0.05, "max_cond_vars": 4},
    "c2": {"workload": "pc_stable_g2_alarm37_100k_depth4", "net": "alarm", "n_rows": 100_000,
           "ci_test": "G2", "level": 0.05, "max_cond_vars": 4},
}


def config_of(key):
    """The `config` object of the JSON line: identical in both arms."""
    c = CONFIGS[key]
    from openbnsl_b200 import networks as nw
    net = getattr(nw, c["net"])()
    return {"workload": c["workload"], "n_vars": net.n, "n_rows": c["n_rows"], "ci_test": c["ci_test"],
            "level": c["level"], "max_cond_vars": c["max_cond_vars"], "stable": True,
            "l2": "dataset (186 MB) larger than L2 (126 MB) and L2 flushed between steps (512 MiB memset)"
                  if key == "c5" else "flushed between steps (512 MiB memset)"}


def make_data(key, seed=0):
    """Forward-sample the configuration's data; codes are ranks of OBSERVED states (dataframe_wrapper.cpp:48-58)."""
    from openbnsl_b200 import networks as nw
    c = CONFIGS[key]
    net = getattr(nw, c["net"])()
    codes = net.sample(c["n_rows"], seed)
    cards = []
    for v in range(net.n):
        u = np.unique(codes[:, v])
        if len(u) != int(u.max()) + 1:
            lut = np.zeros(256, dtype=np.uint8)
            lut[u] = np.arange(len(u), dtype=np.uint8)
            codes[:, v] = lut[codes[:, v]]
        cards.append(len(u))
    return net, codes, cards


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6),
                              ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


_STDOUT_FD = None


def quiet_stdout():
    """stdout carries ONE JSON line.  Libraries write banners to fd 1 (torch prints "NCCL version ..." when the first
    communicator is created): everything written to fd 1 from here on goes to stderr, the line itself to the real stdout."""
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _STDOUT_FD is None:
        os.write(1, data)
    else:
        os.write(_STDOUT_FD, data)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(key):
    """DRAM bytes read+written by the counting kernels of ONE step, from the committed ncu --set full
    capture of this configuration (profiles/r2_traffic.json; made by benchmarks/ncu_summary.py)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            return json.load(f)[key]
    except Exception:
        return None


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ----------------------------------------------------------------------------
# CPU arm: the reference's own functions over a list of CI tests
# ----------------------------------------------------------------------------
def reference_sample_rate(ref, har, rdf, tests, kind, level, budget_s, cores, rng):
    """Wall-clock rate of the unmodified ContingencyTable<true> + citest<true> over a uniform sample
    of `tests` = (var_ids, offsets, x, y), OpenMP over tests.  Returns (tests, seconds)."""
    var_ids, offsets, x, y = tests
    n = len(x)

    def run(idx):
        zf, zo = [], [0]
        for t in idx:
            vs = var_ids[int(offsets[t]):int(offsets[t + 1])]
            zf.extend(int(v) for v in vs if v != x[t] and v != y[t])
            zo.append(len(zf))
        xa, ya = x[idx].astype(np.int64), y[idx].astype(np.int64)
        za, zoa = np.asarray(zf, dtype=np.int64), np.asarray(zo, dtype=np.int64)
        t0 = time.perf_counter()
        har.ci_batch(rdf, xa, ya, za, zoa, kind, level, cores)
        return time.perf_counter() - t0            # wall clock around the call, not the harness's own figure

    probe = rng.choice(n, size=min(n, 4 * cores), replace=False)
    t_probe = max(run(probe), 1e-6)
    m = int(min(n, max(len(probe), len(probe) * budget_s / t_probe)))
    idx = np.sort(rng.choice(n, size=m, replace=False))
    return m, run(idx)


def load_frontier(key, n_vars):
    """(var_ids, offsets, x, y) of the CI tests one PC run of the configuration evaluates, and where the list came from."""
    path = os.path.join(ROOT, "benchmarks", "frontiers", key + "_tests.npz")
    if os.path.exists(path):
        z = np.load(path)
        return (z["var_ids"].astype(np.uint32), z["offsets"].astype(np.uint64), z["x"].astype(np.uint32),
                z["y"].astype(np.uint32)), "benchmarks/frontiers/%s_tests.npz" % key
    # no recorded frontier: the depth-0 frontier is the same for every data set -- all pairs (pc.cpp:22-31)
    x, y = np.triu_indices(n_vars, 1)
    var_ids = np.stack([x, y], axis=1).reshape(-1).astype(np.uint32)
    return (var_ids, np.arange(0, 2 * len(x) + 1, 2, dtype=np.uint64), x.astype(np.uint32), y.astype(np.uint32)), \
        "depth-0 frontier only (no recorded frontier file)"


def run_reference_arm(args, rank):
    """The reference's CPU implementation of the path on the host cores.  Loads oracle/_ref only."""
    if rank != 0:
        return
    import oracle
    from openbnsl_b200 import networks as nw          # pure-python data generator; no native library
    key = args.workload
    c = CONFIGS[key]
    net, codes, cards = make_data(key)
    ref, har = oracle.reference_modules()
    kind = 0 if c["ci_test"] == "chi2" else 1
    cores = host_cores()
    line = {"impl": "reference", "metric": "CI tests/s", "unit": "CI tests/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
            "scaling": "strong" if args.gpus > 1 else "weak",
            "vs_baseline": None, "dtype": "u8 codes -> int32 counts -> f64 statistics", "data": "synthetic",
            "config": config_of(key)}
    tests, frontier_src = load_frontier(key, net.n)
    n_tests = len(tests[2])
    budget = min(20.0, 150.0 / max(1, args.steps + args.warmup))
    rng = np.random.default_rng(0)
    if ref is not None:
        rdf = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
        done, secs = [], []
        for i in range(args.warmup + args.steps):
            m, s = reference_sample_rate(ref, har, rdf, tests, kind, c["level"], budget, cores, rng)
            if i >= args.warmup:
                done.append(m); secs.append(s)
        kindname = "reference"
        sample = ("uniform sample of %d of the %d CI tests of the PC run per step (%s), "
                  "OpenMP over tests around the unmodified ContingencyTable<true>+citest<true> (oracle/_ref), wall clock "
                  "around the call; the frontier's learn time follows by extrapolation" % (int(np.mean(done)), n_tests, frontier_src))
        # the stock entry point (bindings/submodules/structure_learning.cpp:31-37 -> pc.cpp:76-86: serial, early
        # exit) on a row sub-sample it can finish; wall clock around the call
        sub = min(c["n_rows"], args.stock_rows)
        RT = ref.citest.ChiSquare(c["level"]) if kind == 0 else ref.citest.GSquare(c["level"])
        sdf = rdf if sub == c["n_rows"] else ref.base.DataframeWrapper(
            nw.to_reference_frame(np.ascontiguousarray(codes[:sub]), net.names, cards))
        t0 = time.perf_counter()
        ref.structure_learning.pc(sdf, RT, c["max_cond_vars"], True)
        stock = {"entry": "openbnsllib.structure_learning.pc (unmodified, serial)", "rows": sub,
                 "learn_time_s": time.perf_counter() - t0,
                 "note": "rows sub-sampled: the full %d-row run needs about (frontier tests) / (one-core rate) seconds"
                         % c["n_rows"] if sub != c["n_rows"] else "full size"}
    else:
        ods = oracle.Dataset(codes, cards)
        var_ids, offsets, x, y = tests
        done, secs = [], []
        for i in range(args.warmup + args.steps):
            m, t0 = 0, time.perf_counter()
            for t in rng.permutation(n_tests):
                vs = [int(v) for v in var_ids[int(offsets[t]):int(offsets[t + 1])]]
                oracle.citest(ods, int(x[t]), int(y[t]), [v for v in vs if v != x[t] and v != y[t]], kind)
                m += 1
                if time.perf_counter() - t0 > budget:
                    break
            if i >= args.warmup:
                done.append(m); secs.append(time.perf_counter() - t0)
        kindname, cores, stock = "port", 1, None
        sample = "uniform sample of %d of the %d CI tests per step, scalar C restatement (oracle/)" % (int(np.mean(done)), n_tests)
    val = sum(done) / sum(secs)
    line.update({"value": val, "ms_per_step": 1e3 * sum(secs) / len(secs),
                 "cpu_baseline": {"value": val, "unit": "CI tests/s", "cores": cores, "kind": kindname, "sample": sample},
                 "e2e": {"value": val, "unit": "CI tests/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "frontier_learn_time_s_extrapolated": n_tests / val, "stock_pc": stock, "gpu_launches": 0})
    emit(line)


# ----------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------
def adj_of(g, n):
    return np.array([[g.has_edge(a, b) for b in range(n)] for a in range(n)], dtype=np.uint8)


class Job:
    """One process of the job: torch for the device / process-group plumbing, the library for everything timed."""

    def __init__(self, rank, world, local_rank):
        import torch
        import openbnsl_b200
        self.torch, self.rank, self.world, self.local_rank = torch, rank, world, local_rank
        self.L = openbnsl_b200.openbnsllib        # raises if the CUDA extension is not built
        torch.cuda.set_device(local_rank)
        self.comm = None
        if world > 1:
            import torch.distributed as dist
            from openbnsl_b200.dist import init_native_comm
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            self.comm = init_native_comm(self.L)          # NCCL owned by the library (obn_comm_init)
        self.flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")    # > 126 MB L2
        self.lib_stream = torch.cuda.ExternalStream(self.L.stream_handle(), device=torch.device("cuda", local_rank))
        self.L.profile_enable(True)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, value, op):
        if self.world == 1:
            return value
        import torch.distributed as dist
        t = self.torch.tensor([value], dtype=self.torch.float64, device="cuda")
        dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN, "sum": dist.ReduceOp.SUM}[op])
        return float(t.item())


def time_pc(job, key, steps, warmup, e2e_steps, want_clocks=False, record=False):
    """Times `steps` pc() calls of configuration `key` (sharded over the job's ranks when world > 1)."""
    L, torch = job.L, job.torch
    c = CONFIGS[key]
    net, codes, cards = make_data(key)
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    test = L.citest.ChiSquare(c["level"]) if c["ci_test"] == "chi2" else L.citest.GSquare(c["level"])
    depth = c["max_cond_vars"]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kw = {"comm": job.comm} if job.comm is not None else {}
    event_ms = []

    def step(resident):
        if not resident:
            df.release_device()                 # e2e: the dataset starts in host memory
        job.flush.zero_()                       # evict the previous step from L2
        job.barrier()
        ev0.record(job.lib_stream)
        t0 = time.perf_counter()
        g = L.structure_learning.pc(df, test, depth, True, **kw)
        # pc() is synchronous: it returns after the library has synchronised its stream and the PDAG is in
        # host memory.  The e2e steps also walk the result (the "read of the step's result").
        n_edges = sum(1 for v in range(net.n) for p in range(net.n) if g.has_edge(p, v)) if not resident else 0
        dt = time.perf_counter() - t0
        ev1.record(job.lib_stream)
        torch.cuda.synchronize()
        if resident:
            event_ms.append(ev0.elapsed_time(ev1))
        return job.reduce(dt, "max"), g, n_edges

    df.to_device()
    parity = None
    g_ref = L.structure_learning.pc(df, test, depth, True)            # unsharded, this rank alone (also a warm-up)
    tests_per_step = int(L.structure_learning.last_pc_stats()["total_tests"])
    for _ in range(max(1, warmup)):
        _, g, _ = step(True)
    if job.world > 1:
        parity = bool(job.reduce(float(np.array_equal(adj_of(g, net.n), adj_of(g_ref, net.n))), "min"))
    L.profile_read(True)
    clocks = ClockSampler(job.local_rank) if (want_clocks and job.rank == 0) else None
    if clocks:
        clocks.start()
    job.barrier()
    event_ms.clear()
    times = []
    phases = {"enumerate_ms": 0.0, "evaluate_ms": 0.0, "resolve_ms": 0.0, "exposed_ms": 0.0, "collect_wait_ms": 0.0}
    for _ in range(steps):
        dt, g, _ = step(True)
        times.append(dt)
        st = L.structure_learning.last_pc_stats()
        for k in phases:
            phases[k] += st.get(k, 0.0) / steps
    job.barrier()
    prof = L.profile_read(True)
    st = L.structure_learning.last_pc_stats()
    e2e_times, prof_e2e = [], None
    for i in range(e2e_steps + 1 if e2e_steps else 0):
        dt, _, _ = step(False)
        if i:                                   # the first one warms the upload path
            e2e_times.append(dt)
        else:
            L.profile_read(True)
    if e2e_steps:
        prof_e2e = L.profile_read(True)
    clk = clocks.stop() if clocks else None
    rec = None
    if record:
        L.structure_learning.pc(df, test, depth, True, record_tests=True)
        rec = L.structure_learning.last_pc_tests()
    total_s = sum(times)
    peak, peak_src = measured_peaks()
    kernel_s = prof["count_kernel_ms"] * 1e-3
    # kernel time / launches / bytes are this rank's (rank 0's share of a sharded run); tests are the whole job's
    alg = prof["algorithmic_bytes"]
    achieved = alg / kernel_s / 1e9 if kernel_s > 0 else 0.0
    traffic = ncu_traffic(key) if job.world == 1 else None
    out = {
        "key": key, "net": net, "codes": codes, "cards": cards, "recorded": rec,
        "value": tests_per_step * steps / total_s, "ms_per_step": 1e3 * total_s / steps, "tests_per_step": tests_per_step,
        "ms_per_step_cuda_events": sum(event_ms) / max(1, len(event_ms)),
        "parity_ok": parity, "clocks": clk, "launches": int(prof["total_launches"]),
        "tests_per_depth": [int(v) for v in st["tests_per_depth"]], "rounds_per_depth": [int(v) for v in st["batches_per_depth"]],
        "e2e": None if not e2e_times else {
            "value": tests_per_step * len(e2e_times) / sum(e2e_times), "unit": "CI tests/s",
            "h2d_bytes_per_step": prof_e2e["h2d_bytes"] / len(e2e_times), "d2h_bytes_per_step": prof_e2e["d2h_bytes"] / len(e2e_times),
            "ms_per_step": 1e3 * sum(e2e_times) / len(e2e_times)},
        "roofline": {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "peak_source": peak_src,
            "kernel": "obn::count_kernel (all tiers)", "kernel_ms_per_step": prof["count_kernel_ms"] / steps,
            "kernel_share_of_step": kernel_s / total_s,
            "algorithmic_bytes_per_step": alg / steps,
            "algorithmic_bytes_per_launch": alg / max(1, prof["count_launches"]),
            "step_frac": alg / total_s / 1e9 / peak,
            "tables_per_step": prof["tables"] / steps, "tests_per_table": prof["tuples"] / max(1, prof["tables"]),
            "scanned_bytes_per_step": prof["scanned_bytes"] / steps,
            "scanned_frac": prof["scanned_bytes"] / kernel_s / 1e9 / peak if kernel_s > 0 else None,
            "increments_per_step": prof["increments"] / steps,
            "atoms_frac": prof["increments"] / ATOMS_PER_S / kernel_s if kernel_s > 0 else None,
            "atoms_ceiling_per_s": ATOMS_PER_S,
            "traffic": None if traffic is None else traffic["dram_bytes_per_launch"],
            "dram_bytes_per_step": None if traffic is None else traffic["dram_bytes_per_step"],
            "dram_frac": None if traffic is None or kernel_s <= 0 else traffic["dram_bytes_per_step"] * steps / kernel_s / 1e9 / peak,
            "note": "frac = SURVEY 8d accounting unit (k*N bytes per TEST; tests over one variable set share a table, "
                    "so it exceeds what is read); scanned_frac = k*N bytes per TABLE actually scanned (L2 + HBM); "
                    "dram_frac = real DRAM bytes from the committed ncu capture; atoms_frac = fraction of the "
                    "shared-memory increment ceiling, the resource that binds these kernels"},
        "limiter": dict(phases, kernel_ms=prof["count_kernel_ms"] / steps, host_pack_ms=prof["host_prep_ms"] / steps,
                        collectives_per_step=prof["collectives"] / steps,
                        note="host wall time of the driver phases per step on rank 0: enumerate (frontier), evaluate "
                             "(plan + pack + enqueue + wait), resolve; rounds are pipelined two deep, so only exposed_ms "
                             "(host time with no round in flight: the start and end of every depth) is not hidden behind "
                             "the kernels; collect_wait_ms = host blocked waiting for the device"),
    }
    del df
    return out


def time_es(job, name, net, rows, max_parents, exact, lattice, reps):
    """BASELINE.json's other metric, local scores/s, on an exhaustive-search configuration; sharded over the
    job's ranks when world > 1, with the result checked against the unsharded run (parity gate)."""
    L = job.L
    codes = net.sample(rows, 0)
    cards = [int(codes[:, v].max()) + 1 for v in range(net.n)]
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    df.to_device()
    kw = {"exact_order": exact, "lattice": lattice}
    g1 = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, **kw)    # unsharded (also a warm-up)
    s1 = L.structure_learning.last_es_stats()
    if job.comm is not None:
        kw["comm"] = job.comm
        L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, **kw)      # warm-up of the sharded path
    L.profile_read(True)
    job.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        g = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), max_parents, **kw)
    dt = job.reduce((time.perf_counter() - t0) / reps, "max")
    st = L.structure_learning.last_es_stats()
    prof = L.profile_read(True)
    same = np.array_equal(adj_of(g, net.n), adj_of(g1, net.n)) and st["best_score"] == s1["best_score"]
    parity = bool(job.reduce(float(same), "min")) if job.world > 1 else None
    k_s = prof["count_kernel_ms"] * 1e-3 / reps
    del df
    return {"workload": name, "metric": "local scores/s", "value": st["scores"] / dt, "learn_time_s": dt,
            "local_scores": int(st["scores"]), "tables": int(st["tables"]), "lattice_mode": bool(st["lattice"]),
            "parity_caveat": "lattice (fast) mode: scores within 1e-9 relative, structure checked up to Markov equivalence "
                             "in tests/test_gpu_module.py" if st["lattice"] else "exact-order scores: bit-identical to the reference",
            "n_gpus": job.world, "parity_ok": parity, "kernel_s": k_s,
            "phases_ms": {"score": st["score_ms"], "gather": st["gather_ms"], "dp": st["dp_ms"], "enumerate": st["enumerate_ms"]},
            "collectives": int(prof["collectives"] / reps)}


def run_ours(args, rank, world, local_rank):
    from openbnsl_b200 import networks as nw
    job = Job(rank, world, local_rank)
    L = job.L
    key = args.workload
    head = time_pc(job, key, args.steps, args.warmup, 0 if args.no_e2e else min(args.steps, 5),
                   want_clocks=True, record=(rank == 0 and world == 1 and not args.no_cpu))
    secondary = []
    if not args.no_secondary:
        for k2 in ("c4", "c2"):
            if k2 == key:
                continue
            r = time_pc(job, k2, 10 if k2 == "c4" else 50, 3, 0)
            secondary.append({"workload": CONFIGS[k2]["workload"], "metric": "CI tests/s", "value": r["value"],
                              "ms_per_step": r["ms_per_step"], "tests_per_step": r["tests_per_step"], "n_gpus": world,
                              "parity_ok": r["parity_ok"], "tests_per_depth": r["tests_per_depth"],
                              "roofline": {k: r["roofline"][k] for k in ("frac", "step_frac", "scanned_frac", "atoms_frac", "dram_frac",
                                                                        "kernel_ms_per_step", "kernel_share_of_step", "tests_per_table")},
                              "limiter": r["limiter"]})
        if world == 1:
            secondary.append(time_es(job, "C1_asia8_10k_full_lattice_exact", nw.asia(), 10_000, 7, 1, False, 20))
        secondary.append(time_es(job, "C3_bin24_1M_full_lattice_fast", nw.random_binary(24), 1_000_000, 23, 0, True, 1))

    cpu = None
    if head["recorded"] is not None:
        import oracle
        ref, har = oracle.reference_modules()
        c = CONFIGS[key]
        if ref is not None:
            cores = host_cores()
            rdf = ref.base.DataframeWrapper(nw.to_reference_frame(head["codes"], head["net"].names, head["cards"]))
            m, s = reference_sample_rate(ref, har, rdf, head["recorded"], 0 if c["ci_test"] == "chi2" else 1, c["level"],
                                         args.cpu_budget, cores, np.random.default_rng(0))
            cpu = {"value": m / s, "unit": "CI tests/s", "cores": cores, "kind": "reference",
                   "sample": "uniform sample of %d of the %d CI tests of the timed PC run, OpenMP over tests around the unmodified "
                             "ContingencyTable<true>+citest<true> (oracle/_ref), wall clock; sampled-and-extrapolated "
                             "(SURVEY 8d-iii): the whole frontier would take %.0f s on these cores"
                             % (m, len(head["recorded"][2]), len(head["recorded"][2]) * s / m)}
    if rank != 0:
        return
    conf = config_of(key)
    line = {
        "metric": "CI tests/s", "value": head["value"], "unit": "CI tests/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"],
        "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
        "dtype": "u8 codes -> int32 counts -> f64 statistics", "data": "synthetic",
        "config": conf,
        "detail": {"tests_per_step": head["tests_per_step"], "tests_per_depth": head["tests_per_depth"],
                   "rounds_per_depth": head["rounds_per_depth"], "ms_per_step_cuda_events": head["ms_per_step_cuda_events"],
                   "multi_gpu": "one learn task; every round's tables dealt to the ranks inside obn_pc_round, "
                                "ncclAllReduce(min) of the first passing candidate per edge on the device" if world > 1 else "single GPU",
                   "timing": "host clock around the synchronous pc() call bracketed by barrier + cuda synchronize, max over "
                             "ranks (the step has host work between its device rounds, so this is >= the device span); "
                             "ms_per_step_cuda_events = the same steps between CUDA events on the library's stream (rank 0); "
                             "kernel time by CUDA events around the kernel phases"},
        "parity_ok": head["parity_ok"],
        "e2e": head["e2e"],
        "gpu_launches": head["launches"],
        "roofline": head["roofline"], "limiter": head["limiter"],
        "clocks": head["clocks"], "cpu_baseline": cpu,
        "secondary": secondary,
    }
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c5", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the from-host-memory steps (A/B runs only)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the other configurations")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--stock-rows", type=int, default=20000, help="reference arm: rows of the stock pc() call")
    args = ap.parse_args()
    quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("OBN_DEVICE", str(local_rank))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    args.warmup = max(args.warmup, 3)
    run_ours(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        import openbnsl_b200
        openbnsl_b200.openbnsllib.comm_destroy()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
