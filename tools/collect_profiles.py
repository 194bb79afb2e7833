#!/usr/bin/env python
"""Turn one measurement set brought back in gpurun_out/<tag>/ into the committed summaries under profiles/<tag>_*.

usage: python tools/collect_profiles.py r1g      (expects bench_*.json, launches_bench_c4.csv, c5_lattice.txt,
       pytest_gpu.log and {paths,skyblock,blocktask}_c4.ncu-rep as written by the measurement script in profiles/README.md)"""
import csv
import json
import os
import shutil
import subprocess
import sys

tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, dst = os.path.join(root, "gpurun_out", tag), os.path.join(root, "profiles")
METRICS = """gpu__time_duration.sum launch__grid_size launch__registers_per_thread launch__shared_mem_per_block_dynamic
sm__warps_active.avg.pct_of_peak_sustained_active smsp__issue_active.avg.pct_of_peak_sustained_active
sm__inst_issued.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active
sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active
smsp__inst_executed.sum smsp__thread_inst_executed_per_inst_executed.ratio smsp__warps_eligible.avg.per_cycle_active
dram__bytes_read.sum dram__bytes_write.sum l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum
l1tex__data_pipe_lsu_wavefronts_mem_shared.sum l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed
sm__cycles_elapsed.max smsp__cycles_active.avg lts__t_sector_hit_rate.pct l1tex__t_sector_hit_rate.pct""".split()

for f in sorted(os.listdir(src)):
    if f.startswith("bench_") and f.endswith(".json") or f.startswith("scale_n") and f.endswith(".json"):
        lines = open(os.path.join(src, f)).read().strip().splitlines()
        if lines:
            open(os.path.join(dst, f"{tag}_{f}"), "w").write(lines[-1] + "\n")
for f in ("launches_bench_c4.csv", "c5_lattice.txt", "pytest_gpu.log", "pytest_gpu_2gpu.log", "pipe_probe.txt", "sweep_probe.txt"):
    if os.path.exists(os.path.join(src, f)):
        shutil.copy(os.path.join(src, f), os.path.join(dst, f"{tag}_{f}"))

kernels, table, traffic = ("skyblock", "blocktask", "paths"), {}, {}
names = []
for k in kernels:
    rep = os.path.join(src, f"{k}_c4.ncu-rep")
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, un, val = rows[0], rows[1], rows[2]
    names.append(val[hdr.index("Kernel Name")].replace(",", ";")[:64])
    for i, h in enumerate(hdr):
        if h in METRICS:
            table.setdefault(h, []).append((val[i] + " " + un[i]).strip())
    byt = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    traffic[f"render_{k}_kernel"] = int(sum(float(val[hdr.index(m)]) * byt[un[hdr.index(m)]]
                                            for m in ("dram__bytes_read.sum", "dram__bytes_write.sum")))
    srcp = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    tmp = os.path.join(src, f"{k}_src.csv")
    open(tmp, "w").write(srcp)
    out = subprocess.run([sys.executable, os.path.join(root, "tools", "ncu_source_summary.py"), tmp, "--kernel", f"render_{k}",
                          "--window", "128", "--top", "12"], capture_output=True, text=True).stdout
    open(os.path.join(dst, f"{tag}_ncu_c4_source_render_{k}.txt"), "w").write(out)
if names:
    with open(os.path.join(dst, f"{tag}_ncu_c4_metrics.csv"), "w") as f:
        f.write("metric," + ",".join(names) + "\n")
        for m in METRICS:
            if m in table:
                f.write(m + "," + ",".join(table[m]) + "\n")
    json.dump({"c4": traffic, "source": f"profiles/{tag}_ncu_c4_metrics.csv (ncu --set full, one launch each): "
               "dram__bytes_read.sum + dram__bytes_write.sum"}, open(os.path.join(dst, "ncu_traffic.json"), "w"), indent=1)
print("collected", tag)
