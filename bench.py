#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 contingency-counting hot path.

Workload (BASELINE.json configs[1], the configuration the metric is quoted on
at one GPU): PC-stable with the G-square test (alpha 0.05, conditioning depth
<= 4) on 100 000 samples forward-drawn from the 37-node Alarm structure.
One STEP = one complete `structure_learning.pc` call = every CI test of every
depth (contingency table + G2 statistic + dof + p-value per test), through the
C++ driver and the extern-"C" CUDA layer.

  value  CI tests / s with the sample matrix already resident in HBM
  e2e    the same call starting from HOST memory: dataset upload, tuple
         descriptors H2D and p-values D2H all inside the timed region
  roofline  the counting kernel: algorithmic bytes (sum over tests of
         (|Z|+2) * N) / CUDA-event time of its launches
  cpu_baseline  the UNMODIFIED reference functions (ContingencyTable + citest)
         over a bounded sample of the very same tests, OpenMP over tests on all
         host cores (oracle/_ref; the reference has no parallel path of its own)

`--impl reference` times that CPU arm alone.

N>1: one C2 run is 0.8 ms of work -- far below the size at which sharding a PC
round over GPUs pays for its all-reduce ("replicas only"
for this workload) -- so every rank learns from its OWN bootstrap sample of the
same network (seed = rank), with no data-path collective: `value` = CI tests of
all ranks / max-over-ranks time ("scaling": "weak").  The sharded multi-GPU path
(NCCL all-gather of the lattice's F) is what `secondary` C3 measures at N>1.
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

WORKLOAD = "pc_stable_g2_alarm37_100k_depth4"
N_ROWS, SEED, LEVEL, MAX_COND = 100_000, 0, 0.05, 4


def make_data(seed=SEED):
    from openbnsl_b200 import networks as nw
    net = nw.alarm()
    for attempt in range(8):
        codes = net.sample(N_ROWS, seed + 1000 * attempt)
        cards = [int(codes[:, v].max()) + 1 for v in range(net.n)]
        # codes must be ranks of OBSERVED states (dataframe_wrapper.cpp:48-58)
        if all(len(np.unique(codes[:, v])) == cards[v] for v in range(net.n)):
            return net, codes, cards
    raise RuntimeError("no sample with every state observed")


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


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per counting-kernel launch from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as f:
            return float(json.load(f)["count_kernel_dram_bytes_per_launch"])
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
def reference_ci_rate(codes, names, cards, tests, budget_s, threads=0):
    """tests = (var_ids, offsets, x, y). Returns dict(value, cores, kind, sample)."""
    import oracle
    from openbnsl_b200 import networks as nw
    var_ids, offsets, x, y = tests
    n = len(x)
    ref, har = oracle.reference_modules()
    if ref is not None:
        df = ref.base.DataframeWrapper(nw.to_reference_frame(codes, names, cards))
        cores = threads or host_cores()

        def run(idx):
            zf, zo = [], [0]
            for t in idx:
                vs = var_ids[int(offsets[t]):int(offsets[t + 1])]
                zf.extend(int(v) for v in vs if v != x[t] and v != y[t])
                zo.append(len(zf))
            r = har.ci_batch(df, x[idx].astype(np.int64), y[idx].astype(np.int64),
                             np.asarray(zf, dtype=np.int64), np.asarray(zo, dtype=np.int64), 1, LEVEL, cores)
            return r["seconds"]

        rng = np.random.default_rng(0)
        probe = rng.choice(n, size=min(n, 8 * cores), replace=False)
        t_probe = max(run(probe), 1e-6)
        m = int(min(n, max(len(probe), len(probe) * budget_s / t_probe)))
        idx = np.sort(rng.choice(n, size=m, replace=False))
        secs = run(idx)
        return {"value": m / secs, "unit": "CI tests/s", "cores": cores, "kind": "reference",
                "sample": "%d of the %d CI tests of the PC run, uniform sample, OpenMP over tests around the "
                          "unmodified ContingencyTable<true>+citest<true> (oracle/_ref)" % (m, n)}, m, secs
    ods = oracle.Dataset(codes, cards)
    rng = np.random.default_rng(0)
    m, secs = 0, 0.0
    t0 = time.perf_counter()
    for t in rng.permutation(n):
        vs = [int(v) for v in var_ids[int(offsets[t]):int(offsets[t + 1])]]
        oracle.citest(ods, int(x[t]), int(y[t]), [v for v in vs if v != x[t] and v != y[t]], 1)
        m += 1
        secs = time.perf_counter() - t0
        if secs > budget_s:
            break
    return {"value": m / secs, "unit": "CI tests/s", "cores": 1, "kind": "port",
            "sample": "%d of the %d CI tests of the PC run, scalar C restatement (oracle/)" % (m, n)}, m, secs


class _RefEvaluator:
    """CI batches answered by the UNMODIFIED reference functions (oracle/_ref harness)."""

    def __init__(self, har, df, threads):
        self.har, self.df, self.threads, self.seconds, self.tests = har, df, threads, 0.0, 0

    def ci_pvalues(self, var_ids, offsets, x, y, kind):
        n = len(x)
        k = np.diff(offsets.astype(np.int64))
        owner = np.repeat(np.arange(n), k)
        keep = (var_ids != x[owner]) & (var_ids != y[owner])
        zo = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(np.bincount(owner[keep], minlength=n), out=zo[1:])
        r = self.har.ci_batch(self.df, x.astype(np.int64), y.astype(np.int64),
                              var_ids[keep].astype(np.int64), zo, int(kind), LEVEL, self.threads)
        self.seconds += r["seconds"]
        self.tests += n
        return r["p"]


def run_reference_arm(args, rank):
    """The reference's CPU implementation of the path on the host cores."""
    if rank != 0:
        return
    import oracle
    from openbnsl_b200 import networks as nw
    net, codes, cards = make_data()
    ref, har = oracle.reference_modules()
    line = {"impl": "reference", "metric": "CI tests/s", "unit": "CI tests/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64 statistics on int counts", "data": "synthetic",
            "config": {"workload": WORKLOAD, "n_vars": net.n, "n_rows": N_ROWS, "ci_test": "G2",
                       "level": LEVEL, "max_cond_vars": MAX_COND, "stable": True}}
    if ref is None:
        # oracle port, scalar
        ods = oracle.Dataset(codes, cards)
        import openbnsl_b200
        L = openbnsl_b200.openbnsllib
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from helpers import OracleEvaluator
        ev = OracleEvaluator(ods)
        t0 = time.perf_counter()
        L._testing.pc(net.n, ev, L.citest.GSquare(LEVEL), 1, True)    # bounded: depth <= 1
        secs = time.perf_counter() - t0
        val = ev.n_ci / secs
        line.update({"value": val, "ms_per_step": secs * 1e3,
                     "cpu_baseline": {"value": val, "unit": "CI tests/s", "cores": 1, "kind": "port",
                                      "sample": "depth<=1 frontier, %d tests, scalar C restatement" % ev.n_ci}})
    else:
        import openbnsl_b200
        L = openbnsl_b200.openbnsllib     # host driver only: enumerates the frontier, no CUDA on this path
        df = ref.base.DataframeWrapper(nw.to_reference_frame(codes, net.names, cards))
        cores = host_cores()
        ev = _RefEvaluator(har, df, cores)
        L._testing.pc(net.n, ev, L.citest.GSquare(LEVEL), MAX_COND, True, None, 0, True)   # full pass (warm-up)
        full_s, n_tests = ev.seconds, ev.tests
        tests = L.structure_learning.last_pc_tests()
        budget = 150.0 / max(1, args.steps + args.warmup)
        times, done = [], []
        for i in range(args.warmup + args.steps):
            if full_s <= budget:
                ev2 = _RefEvaluator(har, df, cores)
                L._testing.pc(net.n, ev2, L.citest.GSquare(LEVEL), MAX_COND, True)
                s, m = ev2.seconds, ev2.tests
                sample = "the full PC frontier (%d CI tests) per step" % n_tests
            else:
                res, m, s = reference_ci_rate(codes, net.names, cards, tests, budget * 0.8, cores)
                sample = res["sample"]
            if i >= args.warmup:
                times.append(s); done.append(m)
        val = sum(done) / sum(times)
        line.update({"value": val, "ms_per_step": 1e3 * sum(times) / len(times),
                     "cpu_baseline": {"value": val, "unit": "CI tests/s", "cores": cores, "kind": "reference",
                                      "sample": sample + "; OpenMP over tests around the unmodified "
                                      "ContingencyTable<true>+citest<true> (the reference's pc() is serial)"}})
    line["e2e"] = {"value": line["value"], "unit": "CI tests/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    line["gpu_launches"] = 0
    print(json.dumps(line), flush=True)


def secondary_metrics(L, comm, world):
    """BASELINE.json's other metric, local scores/s, on its two exhaustive-search configurations
    (not the headline line: reported beside it).  C1 = Asia, 8 vars, 10k rows, exact-order scores
    (bit-identical to the reference); C3 = 24 binary vars, 1M rows, the FULL 2^24-subset lattice in
    fast mode (marginalisation lattice over the joint table), subsets sharded over the GPUs."""
    import torch
    from openbnsl_b200 import networks as nw
    out = []
    for name, net, rows, mp, exact, reps in (("C1_asia8_10k_full_lattice_exact", nw.asia(), 10_000, 7, 1, 20),
                                            ("C3_bin24_1M_full_lattice_fast", nw.random_binary(24), 1_000_000, 23, 0, 1)):
        codes = net.sample(rows, 0)
        cards = [int(codes[:, v].max()) + 1 for v in range(net.n)]
        df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
        df.to_device()
        kw = {"comm": comm} if (comm is not None and name.startswith("C3")) else {}
        L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), mp, exact_order=exact, **kw)   # warm-up step (same call)
        L.profile_read(True)
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), mp, exact_order=exact, **kw)
        dt = (time.perf_counter() - t0) / reps
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        st = L.structure_learning.last_es_stats()
        prof = L.profile_read(True)
        k_s = prof["count_kernel_ms"] * 1e-3 / reps
        out.append({"workload": name, "metric": "local scores/s", "value": st["scores"] / dt, "learn_time_s": dt,
                    "local_scores": int(st["scores"]), "tables": int(st["tables"]), "lattice_mode": bool(st["lattice"]),
                    "n_gpus": world if kw else 1, "kernel_s": k_s,
                    "algorithmic_GBps_of_kernels": prof["algorithmic_bytes"] / reps / k_s / 1e9 if k_s > 0 else None})
        del df
    return out


def run_ours(args, rank, world, local_rank):
    import torch
    import openbnsl_b200
    L = openbnsl_b200.openbnsllib        # raises if the CUDA extension is not built
    torch.cuda.set_device(local_rank)
    comm = None
    if world > 1:
        import torch.distributed as dist
        from openbnsl_b200.dist import TorchComm
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        comm = TorchComm()
    net, codes, cards = make_data(SEED + rank)          # N>1: independent replicas, one sample per rank
    df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
    test = L.citest.GSquare(LEVEL)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")    # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    # CUDA events on the LIBRARY's stream (torch's current stream never sees the kernels): the span from
    # the step's first device operation being reachable to its last one, next to the host clock
    lib_stream = torch.cuda.ExternalStream(L.stream_handle(), device=torch.device("cuda", local_rank))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    event_ms = []

    def step(resident):
        if not resident:
            df.release_device()                 # e2e: the dataset starts in host memory
        flush.zero_()                           # evict the previous step from L2
        barrier()
        ev0.record(lib_stream)
        t0 = time.perf_counter()
        g = L.structure_learning.pc(df, test, MAX_COND, True)      # no comm: replicas (see the module docstring)
        # pc() is synchronous: it returns after the library has synchronised its stream and the PDAG is
        # in host memory.  The e2e steps also walk the result (the "read of the step's result").
        n_edges = sum(1 for v in range(net.n) for p in range(net.n) if g.has_edge(p, v)) if not resident else 0
        dt = time.perf_counter() - t0
        ev1.record(lib_stream)
        torch.cuda.synchronize()
        if resident:
            event_ms.append(ev0.elapsed_time(ev1))
        nt = L.structure_learning.last_pc_stats()["total_tests"]
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
            t = torch.tensor([nt], dtype=torch.int64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.SUM)                # whole-job units
            nt = int(t.item())
        return dt, nt, n_edges

    L.profile_enable(True)
    df.to_device()
    for _ in range(args.warmup):
        step(True)
    L.profile_read(True)
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    barrier()
    times, tests = [], 0
    event_ms.clear()
    for _ in range(args.steps):
        dt, nt, _ = step(True)
        times.append(dt)
        tests += nt
    barrier()
    prof = L.profile_read(True)
    step_event_ms = sum(event_ms[:args.steps]) / max(1, min(len(event_ms), args.steps))
    # launches / kernel time are per rank; rank 0's numbers describe the roofline
    e2e_times, e2e_tests = [], 0
    for i in range(min(args.steps, 5) + 1):
        dt, nt, _ = step(False)
        if i:                                   # first one is a warm-up of the upload path
            e2e_times.append(dt)
            e2e_tests += nt
        if i == 0:
            L.profile_read(True)
    prof_e2e = L.profile_read(True)
    clk = clocks.stop() if rank == 0 else None

    secondary = None
    if not args.no_secondary:
        secondary = secondary_metrics(L, comm, world)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        L.structure_learning.pc(df, test, MAX_COND, True, record_tests=True)
        rec = L.structure_learning.last_pc_tests()
        cpu, _, _ = reference_ci_rate(codes, net.names, cards, rec, args.cpu_budget)
    if rank != 0:
        return
    total_s = sum(times)
    peak, peak_src = measured_peaks()
    kernel_s = prof["count_kernel_ms"] * 1e-3
    achieved = prof["algorithmic_bytes"] / kernel_s / 1e9 if kernel_s > 0 else 0.0
    n_e2e = max(1, len(e2e_times))
    line = {
        "metric": "CI tests/s", "value": tests / total_s, "unit": "CI tests/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_s / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 codes -> int32 counts -> f64 statistics", "data": "synthetic",
        "config": {"workload": WORKLOAD, "multi_gpu": "independent replicas, one bootstrap sample per rank, no collective", "n_vars": net.n, "n_rows": N_ROWS, "ci_test": "G2", "level": LEVEL,
                   "max_cond_vars": MAX_COND, "stable": True, "tests_per_step": tests // args.steps,
                   "l2": "flushed between steps (512 MiB memset)",
                   "timing": "host clock around the synchronous pc() call bracketed by cuda synchronize, max over ranks "
                             "(the step has host work between its five device rounds, so this is >= the device span); "
                             "ms_per_step_cuda_events = the same steps between CUDA events on the library's stream (rank 0); "
                             "kernel time by CUDA events around the kernel phases",
                   "ms_per_step_cuda_events": step_event_ms},
        "e2e": {"value": e2e_tests / sum(e2e_times) if e2e_times else None, "unit": "CI tests/s",
                "h2d_bytes_per_step": prof_e2e["h2d_bytes"] / n_e2e, "d2h_bytes_per_step": prof_e2e["d2h_bytes"] / n_e2e,
                "ms_per_step": 1e3 * sum(e2e_times) / n_e2e},
        "gpu_launches": int(prof["total_launches"]),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(), "peak_source": peak_src,
                     "traffic_note": "dram bytes read+written per launch (ncu --set full, profiles/r1_traffic.json); the "
                                     "3.7 MB dataset is L2-resident, so DRAM sees it about once per launch while the "
                                     "algorithmic bytes (k*N per test) are served from L2",
                     "algorithmic_bytes_per_launch": prof["algorithmic_bytes"] / max(1, prof["count_launches"]),
                     "kernel": "obn::count_kernel (all tiers)", "kernel_ms_per_step": prof["count_kernel_ms"] / args.steps,
                     "kernel_share_of_step": kernel_s / total_s,
                     "algorithmic_bytes_per_step": prof["algorithmic_bytes"] / args.steps},
        "clocks": clk, "cpu_baseline": cpu,
        "secondary": secondary,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-secondary", action="store_true", help="skip the local-scores/s measurements")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
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
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
