#!/usr/bin/env python3
"""Per-source-line share of executed warp instructions, stall samples and shared-memory wavefronts of one kernel
in an .ncu-rep captured with --import-source on.   usage: tools/lineprof.py report.ncu-rep kernel_regex source.cu [min_pct]"""
import csv, collections, sys, subprocess
rep, kern, srcfile = sys.argv[1:4]
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.7
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdrs = [i for i, r in enumerate(rows) if r and r[0] == 'Line No']
h = hdrs[0]; H = rows[h]
cols = {'ex': 'Instructions Executed', 's': '# Samples', 'wf': 'L1 Wavefronts Shared', 'wfi': 'L1 Wavefronts Shared Ideal'}
ix = {k: H.index(v) for k, v in cols.items() if v in H}
end = hdrs[1] - 2 if len(hdrs) > 1 else len(rows)
per = {}; tot = collections.Counter()
for r in rows[h + 1:end]:
    if len(r) <= max(ix.values()) or not r[0].isdigit():
        continue
    d = per.setdefault(int(r[0]), collections.Counter())
    for k, i in ix.items():
        try: v = int(r[i])
        except ValueError: v = 0
        d[k] += v; tot[k] += v
print(dict(tot))
src = open(srcfile).read().split('\n')
for ln, d in sorted(per.items()):
    if d['ex'] * 100 / tot['ex'] > thr or d['s'] * 100 / max(1, tot['s']) > thr or d['wf'] * 100 / max(1, tot['wf']) > thr:
        print(f"{ln:4d} ex{d['ex'] * 100 / tot['ex']:5.1f}% st{d['s'] * 100 / max(1, tot['s']):5.1f}% wf{d['wf'] * 100 / max(1, tot['wf']):5.1f}% (ideal{d['wfi'] * 100 / max(1, tot['wf']):5.1f}%)  {src[ln - 1][:110] if ln <= len(src) else ''}")
