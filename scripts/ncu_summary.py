#!/usr/bin/env python
"""Summarise an .ncu-rep (read offline): headline metrics + SASS instruction/stall shares per barrier-delimited phase.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [out.csv]"""
import csv, io, re, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
KEEP = ['dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'smsp__cycles_active.avg', 'sm__cycles_elapsed.max', 'lts__t_bytes.sum', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed']
out = []
for h, u, v in zip(hdr, units, vals):
    if h in KEEP or h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio'):
        out.append((h, u, v))
for o in out:
    print("%-90s %-14s %s" % o)
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
ia, isrc, isamp = h.index("Instructions Executed"), h.index("Source"), h.index("# Samples")
segs, ops, tot, samp = [], {}, 0, 0
for r in rows[2:]:
    if len(r) <= ia:
        continue
    try:
        n, s = int(r[ia]), int(r[isamp])
    except ValueError:
        continue
    t = r[isrc].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    ops[op] = ops.get(op, 0) + n
    tot += n
    samp += s
    if 'BAR.SYNC' in r[isrc]:
        segs.append((tot, samp, ops)); ops, tot, samp = {}, 0, 0
segs.append((tot, samp, ops))
T = sum(s[0] for s in segs) or 1
S = sum(s[1] for s in segs) or 1
print("phase (between BAR.SYNCs): inst M, inst %, stall-sample %, top opcodes (M)")
for i, s in enumerate(segs):
    top = sorted(s[2].items(), key=lambda kv: -kv[1])[:9]
    print("%2d %8.1f %5.1f%% %5.1f%%  %s" % (i, s[0] / 1e6, 100 * s[0] / T, 100 * s[1] / S, ' '.join("%s:%.0f" % (k, v / 1e6) for k, v in top)))
if len(sys.argv) > 2:
    with open(sys.argv[2], "w") as f:
        f.write("metric,unit,value\n")
        for o in out:
            f.write("%s,%s,%s\n" % o)
        f.write("# phase,inst_M,inst_pct,stall_sample_pct\n")
        for i, s in enumerate(segs):
            f.write("phase%d,%.1f,%.1f,%.1f\n" % (i, s[0] / 1e6, 100 * s[0] / T, 100 * s[1] / S))
