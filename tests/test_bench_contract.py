"""The reference arm of bench.py (CPU only): one JSON line on stdout with the contract's keys, the same `config`
object the GPU arm prints, and no native library of this repository in the process (only oracle/_ref)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    import oracle
    if oracle.reference_modules()[0] is None:
        pytest.skip("oracle/_ref is not built")
    code = ("import sys, runpy; sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'c2', '--steps', '1', '--warmup', '0', "
            "'--stock-rows', '300', '--cpu-budget', '1']\n"
            "try:\n    runpy.run_path(%r, run_name='__main__')\nfinally:\n"
            "    libs = sorted({l.split()[-1] for l in open('/proc/self/maps') if l.rstrip().endswith('.so') and %r in l})\n"
            "    sys.stderr.write('LIBS ' + repr(libs) + '\\n')\n") % (os.path.join(ROOT, "bench.py"), ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout[:2000]
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "impl", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["impl"] == "reference" and d["metric"] == "CI tests/s" and d["value"] > 0 and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["e2e"]["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    sys.path.insert(0, ROOT)
    import bench
    assert d["config"] == bench.config_of("c2")                       # identical in both arms
    assert "stock_pc" in d and d["stock_pc"]["learn_time_s"] > 0       # the stock serial entry point was timed too
    libs = eval([l for l in r.stderr.splitlines() if l.startswith("LIBS ")][-1][5:])
    assert libs and all(l.startswith(os.path.join(ROOT, "oracle", "_ref") + os.sep) for l in libs), libs


def test_committed_roofline_is_reproducible_from_the_committed_ncu_capture():
    """profiles/: the DRAM figure bench.py quotes (r2_traffic.json) is the sum over the per-launch table of the ncu
    capture (r2_ncu_c5_launches.csv), and the committed bench line was computed from it with the stated formulas."""
    import csv
    prof = os.path.join(ROOT, "profiles")
    rows = list(csv.DictReader(open(os.path.join(prof, "r2_ncu_c5_launches.csv"))))
    dram = sum(float(r["dram_read"] or 0) + float(r["dram_write"] or 0) for r in rows)
    traffic = json.load(open(os.path.join(prof, "r2_traffic.json")))["c5"]
    assert traffic["launches_per_step"] == len(rows)
    assert abs(traffic["dram_bytes_per_step"] - dram) <= 1e-9 * dram
    line = json.load(open(os.path.join(prof, "r2_bench_n1.json")))
    r = line["roofline"]
    assert line["config"]["workload"] == "pc_stable_chi2_munin1like186_1M_depth3" and line["n_gpus"] == 1
    assert r["dram_bytes_per_step"] == traffic["dram_bytes_per_step"] and r["traffic"] == traffic["dram_bytes_per_launch"]
    kernel_s = r["kernel_ms_per_step"] * 1e-3
    assert abs(r["achieved"] - r["algorithmic_bytes_per_step"] / kernel_s / 1e9) <= 1e-6 * r["achieved"]
    assert abs(r["frac"] - r["achieved"] / r["peak"]) <= 1e-9
    assert abs(r["dram_frac"] - r["dram_bytes_per_step"] / kernel_s / 1e9 / r["peak"]) <= 1e-6
    assert abs(r["atoms_frac"] - r["increments_per_step"] / r["atoms_ceiling_per_s"] / kernel_s) <= 1e-6
    # every launch of the step is one of this repository's kernels, and the variants are the ones DESIGN 3.1 lists
    assert {x["variant_KW_TIER_SCALE4_REP"] for x in rows} <= {"8,0,1,5", "8,0,0,5", "16,0,1,4", "16,0,1,2", "16,0,1,(int)-1",
                                                             "32,0,0,(int)-1", "32,2,0,(int)-1", "32,1,0,(int)-1", "ci_epilogue"}
