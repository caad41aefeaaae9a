#!/usr/bin/env python3
"""Summarise an .ncu-rep here (no GPU needed): key raw metrics per kernel launch, per-opcode dynamic
instruction counts and the top stall sites from the source page.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--kernel regex] [--per N] [--top 12]
"""
import argparse
import collections
import csv
import io
import re
import subprocess

RAW = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
       "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
       "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
       "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
       "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
       "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
       "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
       "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
       "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
       "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
       "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
       "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active"]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--kernel", default=None)
    ap.add_argument("--per", type=float, default=None, help="divide instruction counts by this (e.g. frames per launch)")
    ap.add_argument("--top", type=int, default=12)
    a = ap.parse_args()
    rows = list(csv.reader(io.StringIO(ncu(["-i", a.rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    seen = set()
    for r in rows[2:]:
        name = r[ix["Kernel Name"]]
        if a.kernel and not re.search(a.kernel, name):
            continue
        if name in seen:
            continue
        seen.add(name)
        print("==", name)
        for m in RAW:
            if m in ix:
                print(f"   {m:72s} {r[ix[m]]} {units[ix[m]]}")
        print("   warps stalled per issued instruction, by reason:")
        stalls = sorted(((float(r[i].replace(",", "")), h) for h, i in ix.items()
                         if "average_warps_issue_stalled" in h and h.endswith("_per_issue_active.ratio")), reverse=True)
        for v, h in stalls[:10]:
            print(f"      {v:6.3f}  {h.split('stalled_')[1].split('_per_issue')[0]}")
    for name in seen:
        kn = re.sub(r"\(.*", "", name).split("::")[-1].split("<")[0].split()[-1]
        rows = list(csv.reader(io.StringIO(ncu(["-i", a.rep, "--page", "source", "--csv", "--kernel-name",
                                                f"regex:{kn}", "--launch-count", "1"]))))
        hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
        if not hi:
            continue
        hdr = rows[hi[0]]
        ix = {h: i for i, h in enumerate(hdr)}
        data = []
        for r in rows[hi[0] + 1:]:
            if not r or r[0] in ("Kernel Name", "Address"):
                break
            data.append(r)
        byop, stall, tot = collections.Counter(), collections.Counter(), 0
        for r in data:
            m = re.match(r"(@!?U?P\w+\s+)?([A-Z0-9_]+)", r[ix["Source"]].strip())
            op = m.group(2) if m else "?"
            n = int(r[ix["Instructions Executed"]])
            byop[op] += n; tot += n
            stall[op] += int(r[ix["Warp Stall Sampling (All Samples)"]])
        per = a.per or 1.0
        print(f"== {kn}: {tot} warp instructions ({tot / per:.1f} per unit)")
        for op, n in byop.most_common(18):
            print(f"   {op:10s} {n:12d} {n / per:9.1f}  stall samples {stall[op]}")
        print("   top stall sites:")
        for r in sorted(data, key=lambda r: -int(r[ix["Warp Stall Sampling (All Samples)"]]))[:a.top]:
            print("    ", r[ix["Warp Stall Sampling (All Samples)"]], r[ix["Instructions Executed"]], r[ix["Source"]].strip()[:90])


if __name__ == "__main__":
    main()
