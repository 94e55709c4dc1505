#!/usr/bin/env python
"""Digest of `ncu -i X.ncu-rep --page raw --csv` (+ optional `--page source --csv`): the metrics profiles/*.md quote.
usage: ncu_digest.py raw.csv [src.csv]"""
import csv, sys
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_srcunit_tex.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed"]
STALLS = ["long_scoreboard", "barrier", "wait", "no_instruction", "short_scoreboard", "math_pipe_throttle", "not_selected", "branch_resolving",
          "mio_throttle", "lg_throttle", "dispatch_stall", "drain", "membar", "sleeping", "tex_throttle", "imc_miss", "selected"]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("kernel:", d.get("Kernel Name", "?")[:90])
    for k in KEYS:
        if k in d: print(f"  {k:72s} {d[k]} {units[hdr.index(k)]}")
    for s in STALLS:
        k = f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio"
        if k in d: print(f"  stall {s:20s} {float(d[k]):.3f} warps / issue")
if len(sys.argv) > 2:
    rows = list(csv.reader(open(sys.argv[2])))
    hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
    def f(r, k):
        try: return float(r[ix[k]])
        except Exception: return 0.0
    tot = sum(f(r, "# Samples") for r in data)
    cols = [c for c in hdr if c.startswith("stall_") and "(Not Issued)" not in c]
    print(f"source page: {len(data)} SASS instructions, {tot:.0f} samples")
    print("  stall share of all samples:", ", ".join(f"{c[6:]} {100 * sum(f(r, c) for r in data) / tot:.1f}%" for c in cols if sum(f(r, c) for r in data) / tot > 0.01))
    print("  per 256-instruction block: first index, share of samples, no_inst share inside block, warp-instructions executed")
    for b in range(0, len(data), 256):
        blk = data[b:b + 256]
        s = sum(f(r, "# Samples") for r in blk); ni = sum(f(r, "stall_no_inst") for r in blk); ex = sum(f(r, "Instructions Executed") for r in blk)
        print(f"  {b:5d} {100 * s / tot:5.1f}%  no_inst {100 * ni / max(s, 1):4.0f}%  exec {ex / 1e6:7.1f} M")
