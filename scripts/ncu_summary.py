#!/usr/bin/env python
"""Turn ncu outputs brought back in gpurun_out/ into the small, tracked files under profiles/.

  python scripts/ncu_summary.py full  gpurun_out/r01/spmv_stream_full.ncu-rep profiles/r01_spmv_stream_full.csv
  python scripts/ncu_summary.py list  gpurun_out/r01/launches.csv            profiles/r01_launches.md

`full`  keeps the counters the roofline discussion in DESIGN.md cites (one row per metric).
`list`  aggregates the per-launch gpu__time_duration list per kernel: launches, total, mean,
        share of the profiled region.
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

KEEP = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum(\.per_second)?|"
    r"gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"dram__cycles_active\.avg|dram__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"lts__t_sector_hit_rate\.pct|lts__t_sectors_srcunit_tex_op_read(_lookup_hit)?\.sum|"
    r"lts__t_sectors_srcunit_tex_op_write\.sum|lts__t_bytes\.sum|"
    r"l1tex__t_sector_hit_rate\.pct|l1tex__t_(sectors|requests)_pipe_lsu_mem_global_op_(ld|st)\.sum|"
    r"l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_(ld|st)\.ratio|"
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|l1tex__data_pipe_lsu_wavefronts(_mem_shared)?\.sum|"
    r"l1tex__data_pipe_lsu_wavefronts\.avg\.pct_of_peak_sustained_elapsed|"
    r"l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_(hit|miss)\.sum|"
    r"(l1tex|lts)__throughput\.avg\.pct_of_peak_sustained_elapsed|smsp__warps_eligible\.avg\.per_cycle_active|"
    r"launch__(registers_per_thread|grid_size|block_size|waves_per_multiprocessor|"
    r"shared_mem_per_block_dynamic|occupancy_limit_(registers|shared_mem|warps|blocks))|"
    r"sm__warps_active\.avg\.pct_of_peak_sustained_active|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__cycles_active\.avg|"
    r"smsp__average_warps_issue_stalled_[a-z_]+_per_issue_active\.ratio|"
    r"sm__inst_executed_pipe_(lsu|fma|fp64|alu|uniform)\.sum|smsp__inst_executed\.sum|"
    r"sm__cycles_elapsed\.max|gpc__cycles_elapsed\.max|smsp__pcsamp_warps_issue_stalled_[a-z_]+)$")


def full(rep: str, out: str) -> None:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True,
                         check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, unit = rows[0], rows[1]
    with open(out, "w") as f:
        w = csv.writer(f)
        for r in rows[2:]:
            name = r[head.index("Kernel Name")]
            w.writerow(["kernel", "", name])
            w.writerow(["grid/block", "", r[head.index("Grid Size")] + " x " + r[head.index("Block Size")]])
            for i, k in enumerate(head):
                if KEEP.match(k) and r[i] not in ("", "0", "n/a"):
                    w.writerow([k, unit[i], r[i]])
    print("wrote", out)


def launch_list(path: str, out: str) -> None:
    lines = [l for l in open(path) if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    agg = OrderedDict()
    total = 0.0
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        k = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("void ", "").replace("<unnamed>::", "")
        ns = float(r["Metric Value"].replace(",", ""))
        a = agg.setdefault(k, [0, 0.0, r["Grid Size"], r["Block Size"]])
        a[0] += 1
        a[1] += ns
        total += ns
    with open(out, "w") as f:
        f.write(f"ncu --metrics gpu__time_duration.sum --clock-control none (source: {path})\n")
        f.write("cold-cache, serialised launches: compare shares, not absolutes\n\n")
        f.write("| kernel | launches | total us | mean us | share | grid | block |\n|---|---|---|---|---|---|---|\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k[:110]}` | {a[0]} | {a[1] / 1e3:.1f} | {a[1] / a[0] / 1e3:.2f} | "
                    f"{a[1] / total * 100:.1f}% | {a[2]} | {a[3]} |\n")
    print("wrote", out)


if __name__ == "__main__":
    {"full": full, "list": launch_list}[sys.argv[1]](sys.argv[2], sys.argv[3])
