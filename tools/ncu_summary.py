"""Turns an `ncu --set full` report of tools/profile_step.py into
profiles/<tag>_ncu_summary.csv (one row per kernel: the last captured launch)
and profiles/traffic.json (DRAM bytes per launch, read by bench.py).
Usage: ncu_summary.py <report.ncu-rep> <tag>"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

rep, tag = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True,
                     check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
kernels = {}
for r in rows[2:]:
    d = dict(zip(hdr, r))
    name = d["Kernel Name"].split("(")[0].replace("void ", "").replace("<unnamed>::", "").replace("unnamed>::", "").replace(", ", "; ")
    kernels[name.replace("pfb::", "")] = d              # the last launch of each kernel wins
unit = dict(zip(hdr, units))
out = os.path.join(ROOT, "profiles", "%s_ncu_summary.csv" % tag)
with open(out, "w", newline="") as f:
    wr = csv.writer(f)
    wr.writerow(["kernel"] + COLS)
    for name, d in kernels.items():
        wr.writerow([name] + ["%s %s" % (d.get(c, ""), unit.get(c, "")) for c in COLS])


def dram(name_part):
    tot = 0.0
    for name, d in kernels.items():
        if name_part in name:
            for c in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(d[c]) * SCALE[unit[c]]
    return tot


def pct(name_part, col):
    for name, d in kernels.items():
        if name_part in name:
            try:
                return float(d[col])
            except (KeyError, ValueError):
                return None
    return None


rows_json = {"ace_scores_kernel": {
    "issue_active_pct": pct("ace_scores_kernel", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "pipe_fma_cycles_active_pct": pct("ace_scores_kernel", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
    "pipe_xu_pct": pct("ace_scores_kernel", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    "pipe_alu_pct": pct("ace_scores_kernel", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    "warp_instructions": pct("ace_scores_kernel", "smsp__inst_executed.sum"),
    "registers": pct("ace_scores_kernel", "launch__registers_per_thread"),
    "source": "profiles/%s_ncu_summary.csv" % tag}}
with open(os.path.join(ROOT, "profiles", "%s_ncu_rows.json" % tag), "w") as f:
    json.dump(rows_json, f, indent=1)

traffic = {"ace_scores_kernel": dram("ace_scores_kernel"), "ace_all_kernels": dram("ace_"),
           "gaussian": dram("gaussian"),
           "sobel": dram("sobel_kernel"), "canny": dram("canny_"),
           "note": "dram__bytes_read.sum + dram__bytes_write.sum per launch from profiles/%s_ncu_summary.csv "
                   "(ncu --set full, tools/profile_step.py, A4 page; canny = classify + hysteresis). Writes that "
                   "stay in the 126 MB L2 until after the kernel ends are not counted by the counter." % tag}
with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
    json.dump(traffic, f, indent=1)
print(out)
for name in kernels:
    print(" ", name)
