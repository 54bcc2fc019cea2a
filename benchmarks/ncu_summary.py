#!/usr/bin/env python
"""ncu_summary.py -- turn an `ncu --page raw --csv` export of one bench step into the small files kept under
profiles/: a per-launch table (kernel variant, grid, block, duration, DRAM bytes, L2 hit rate, warps active,
L1 data-pipe utilisation, shared-atomic wavefronts) and the DRAM traffic figures bench.py's `roofline` reads.

usage: python benchmarks/ncu_summary.py <raw.csv> <config key> <pc() calls captured> <out prefix>
       e.g.  ... gpurun_out/r2_ncu_c5_raw.csv c5 2 profiles/r2_ncu_c5

The capture holds <pc() calls captured> identical pc() calls (run_configs.py reps=1: a warm-up call and one more);
the launches of the LAST one -- count_kernel and, for tables in global memory, ci_epilogue_kernel -- are one
complete step and are the ones summarised.  Writes <out prefix>_launches.csv and merges
{"<key>": {"dram_bytes_per_step", "dram_bytes_per_launch", ...}} into profiles/r2_traffic.json.
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

COLS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
        ("lts__t_sector_hit_rate.pct", "l2_hit_pct"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
        ("l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed", "l1_data_pipe_pct"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum", "smem_atom_wavefronts"),
        ("smsp__inst_executed_op_shared_atom.sum", "smem_atom_instructions"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
        ("launch__registers_per_thread", "regs"), ("launch__shared_mem_per_block_dynamic", "dyn_smem"),
        ("launch__occupancy_limit_shared_mem", "occ_limit_smem"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_pct")]

SCALE = {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "s": 1.0, "second": 1.0, "nsecond": 1e-9,
         "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    raw, key, n_calls, prefix = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    data = [r for r in data if "count_kernel" in r[col["Kernel Name"]] or "ci_epilogue_kernel" in r[col["Kernel Name"]]]
    assert len(data) % n_calls == 0, (len(data), n_calls)
    data = data[-(len(data) // n_calls):]

    def val(r, name):
        if name not in col or r[col[name]] in ("", "n/a"):
            return None
        v = float(r[col[name]].replace(",", ""))
        return v * SCALE.get(units[col[name]], 1.0)

    out = []
    for r in data:
        m = re.search(r"count_kernel<(.*?)>", r[col["Kernel Name"]])
        line = {"variant_KW_TIER_SCALE4_REP": m.group(1).replace(" ", "") if m else "ci_epilogue", "grid": r[col["Grid Size"]],
                "block": r[col["Block Size"]]}
        for name, short in COLS:
            line[short] = val(r, name)
        out.append(line)
    with open(prefix + "_launches.csv", "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(out[0].keys()))
        w.writeheader()
        w.writerows(out)
    total_t = sum(o["duration"] for o in out)
    dram = sum((o["dram_read"] or 0) + (o["dram_write"] or 0) for o in out)
    wav = sum(o["smem_atom_wavefronts"] or 0 for o in out)
    ins = sum(o["smem_atom_instructions"] or 0 for o in out)
    summary = {"source": os.path.basename(raw), "launches_per_step": len(out), "kernel_time_s_under_ncu": total_t,
               "dram_bytes_per_step": dram, "dram_bytes_per_launch": dram / len(out),
               "dram_GBps_under_ncu": dram / total_t / 1e9, "smem_atom_wavefronts_per_step": wav,
               "smem_atom_instructions_per_step": ins,
               "wavefronts_per_atom_instruction": (wav / ins) if ins else None,
               "note": "per-launch durations under ncu are serialised and cold-cache: the kernels' SHARE of the step "
                       "must agree with bench.py's CUDA-event figure, not the absolute time"}
    path = os.path.join(ROOT, "profiles", "r2_traffic.json")
    try:
        allk = json.load(open(path))
    except Exception:
        allk = {}
    allk[key] = summary
    json.dump(allk, open(path, "w"), indent=1)
    print(json.dumps(summary, indent=1))
    by = {}
    for o in out:
        b = by.setdefault(o["variant_KW_TIER_SCALE4_REP"], [0, 0.0])
        b[0] += 1
        b[1] += o["duration"]
    for k_, (n, t) in sorted(by.items(), key=lambda kv: -kv[1][1]):
        print("%-16s launches %3d  %.3f ms  %.0f %%" % (k_, n, t * 1e3, 100 * t / total_t))


if __name__ == "__main__":
    main()
