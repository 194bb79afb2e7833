#!/usr/bin/env python
"""Turn the round-2 measurement sets (gpurun_out/r2p: one GPU, tools/gpu/r2p.sh; gpurun_out/r2s: one 8-GPU box,
tools/gpu/r2s.sh) into the committed summaries under profiles/r2_*.

usage: python tools/collect_r2.py [r2p] [r2s]"""
import csv
import json
import os
import shutil
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
dst = os.path.join(root, "profiles")
METRICS = """gpu__time_duration.sum launch__grid_size launch__block_size launch__registers_per_thread launch__shared_mem_per_block_dynamic
sm__cycles_elapsed.avg sm__cycles_active.avg smsp__cycles_active.avg
sm__warps_active.avg.pct_of_peak_sustained_active smsp__issue_active.avg.pct_of_peak_sustained_active
smsp__inst_executed.sum smsp__thread_inst_executed_per_inst_executed.ratio smsp__warps_eligible.avg.per_cycle_active
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active
smsp__sass_thread_inst_executed_op_fadd_pred_on.sum smsp__sass_thread_inst_executed_op_fmul_pred_on.sum smsp__sass_thread_inst_executed_op_ffma_pred_on.sum
l1tex__data_pipe_lsu_wavefronts_mem_shared.sum l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed
l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum dram__bytes_read.sum dram__bytes_write.sum lts__t_sector_hit_rate.pct
smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio""".split()
CAPS = [("paths_serial_c4", "C4 whole frame, path kernel, serial form (the default for this list length)"),
        ("pool_c4_full", "C4 whole frame, path kernel, pool form forced (RTC_PATHS=pool)"),
        ("pool_c4_n8share", "C4, rank 0's share of an 8-GPU band split, path kernel, pool form (the default for this list length)"),
        ("skyblock_c4", "C4 whole frame, primary kernel (block bounds)"),
        ("blocktask_c4", "C4 whole frame, block-task kernel (remaining samples of the blocks with candidates)"),
        ("paths_stream_c5", "synthetic 1 M triangles, sparse lattice, path kernel, serial form, streamed (TMA tiles)"),
        ("pool_stream_c5", "synthetic 1 M triangles, sparse lattice, path kernel, pool form, streamed (TMA tiles)")]


def last_json(path):
    lines = [l for l in open(path).read().strip().splitlines() if l.startswith("{")]
    return lines[-1] if lines else None


def collect_r2p(src):
    for f in sorted(os.listdir(src)):
        if f.startswith("bench_") and f.endswith(".json"):
            line = last_json(os.path.join(src, f))
            if line:
                open(os.path.join(dst, "r2_" + f), "w").write(line + "\n")
    for f in ("launches_bench_c4.csv", "c5_lattice.txt", "pytest_gpu.log", "pipe_probe.txt", "sweep_probe.txt", "stock_rayt_c1.txt", "e2e_trace_c4.txt"):
        if os.path.exists(os.path.join(src, f)) and os.path.getsize(os.path.join(src, f)) > 0:
            shutil.copy(os.path.join(src, f), os.path.join(dst, "r2_" + f))
    if os.path.exists(os.path.join(src, "fp32_peak.json")):
        shutil.copy(os.path.join(src, "fp32_peak.json"), os.path.join(dst, "fp32_peak.json"))
    table, names, traffic = {}, [], {}
    byt = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for cap, what in CAPS:
        raw = os.path.join(src, cap + ".raw.csv")
        if not os.path.exists(raw):
            continue
        rows = list(csv.reader(open(raw)))
        hdr, un, val = rows[0], rows[1], rows[2]
        names.append(cap)
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                table.setdefault(m, {})[cap] = (val[i] + " " + un[i]).strip()
        tr = sum(float(val[hdr.index(m)]) * byt.get(un[hdr.index(m)], 1) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        traffic[cap] = int(tr)
        for ext in ("source_summary.txt", "lines.txt"):
            p = os.path.join(src, f"{cap}.{ext}")
            if os.path.exists(p):
                shutil.copy(p, os.path.join(dst, f"r2_ncu_{cap}.{ext}"))
    if names:
        with open(os.path.join(dst, "r2_ncu_metrics.csv"), "w") as f:
            f.write("# one `ncu --set full --clock-control none --import-source on` capture per column (tools/gpu/r2p.sh)\n")
            for cap, what in CAPS:
                if cap in names:
                    f.write(f"# {cap}: {what}\n")
            f.write("metric," + ",".join(names) + "\n")
            for m in METRICS:
                if m in table:
                    f.write(m + "," + ",".join(table[m].get(c, "") for c in names) + "\n")
        tj = {"c4": {"path kernel": traffic.get("paths_serial_c4"), "rtc::render_pool_kernel": traffic.get("pool_c4_full"),
                     "render_skyblock_kernel": traffic.get("skyblock_c4"), "render_blocktask_kernel": traffic.get("blocktask_c4")},
              "c5l": {"path kernel": None},
              "source": "profiles/r2_ncu_metrics.csv (ncu --set full, one launch each): dram__bytes_read.sum + dram__bytes_write.sum"}
        json.dump(tj, open(os.path.join(dst, "ncu_traffic.json"), "w"), indent=1)


def collect_r2s(src):
    for f in sorted(os.listdir(src)):
        if f.endswith(".json"):
            line = last_json(os.path.join(src, f))
            if line:
                open(os.path.join(dst, "r2_" + f), "w").write(line + "\n")
    for f in ("pytest_multi_8gpu.log", "gpus.txt"):
        if os.path.exists(os.path.join(src, f)):
            shutil.copy(os.path.join(src, f), os.path.join(dst, "r2_" + f))


for tag in sys.argv[1:] or ["r2p", "r2s"]:
    src = os.path.join(root, "gpurun_out", tag)
    if os.path.isdir(src):
        (collect_r2p if tag == "r2p" else collect_r2s)(src)
        print("collected", tag)
