#!/usr/bin/env python3
"""Summarise an .ncu-rep: per-launch key metrics (raw page) and, per kernel, the SASS opcode mix and
the hottest source lines (source page; needs -lineinfo and --import-source on).
usage: tools/ncu_summary.py report.ncu-rep [kernel_regex]"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
kre = sys.argv[2] if len(sys.argv) > 2 else None
WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H, U = rows[0], rows[1]
ki = H.index('Kernel Name')
seen = set()
for r in rows[2:]:
    name = r[ki].split('(')[0]
    if kre and not re.search(kre, name):
        continue
    print('===', name)
    for w in WANT:
        if w in H:
            i = H.index(w)
            print(f"   {w:72s} {r[i]:>18s} {U[i]}")
    if name in seen:
        continue
    seen.add(name)
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + name.replace('void ', '').split('::')[-1].split('<')[0]],
                         capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    hdr = [i for i, x in enumerate(srows) if x and x[0] == 'Address']
    if not hdr:
        continue
    SH = srows[hdr[0]]
    ia, ie, isamp = SH.index('Source'), SH.index('Instructions Executed'), SH.index('# Samples')
    end = hdr[1] - 1 if len(hdr) > 1 else len(srows)
    data = [(x[ia].strip(), int(x[ie]), int(x[isamp])) for x in srows[hdr[0] + 1:end] if len(x) > ie and x[ie].isdigit()]
    tot = sum(d[1] for d in data) or 1
    mix = collections.Counter(); samp = collections.Counter()
    for s, e, sa in data:
        t = s.split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        mix[op] += e; samp[op] += sa
    print(f"   SASS: {len(data)} instructions, {tot} warp-level executions; opcode mix:")
    print('   ' + ', '.join(f"{op} {100 * e / tot:.1f}%" for op, e in mix.most_common(14)))
    tsamp = sum(samp.values()) or 1
    print('   stall samples by opcode: ' + ', '.join(f"{op} {100 * e / tsamp:.1f}%" for op, e in samp.most_common(8)))
