#!/usr/bin/env python3
"""Turns gpurun_out/<launch list csv> and <.ncu-rep> into the small CSV summaries kept under profiles/.
usage: tools/profiles_export.py TAG launches.csv report.ncu-rep "free-text description" """
import collections, csv, io, subprocess, sys
tag, launches, rep, desc = sys.argv[1:5]
rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == 'ID'][0]
H = rows[hdr]; ki = H.index('Kernel Name'); vi = H.index('Metric Value'); ui = H.index('Metric Unit')
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    v = float(r[vi].replace(',', '')); u = r[ui]
    v = v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
    a = agg.setdefault(r[ki].split('(')[0].replace('void ', ''), [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
with open(f'profiles/{tag}_launch_shares.csv', 'w') as f:
    f.write(f"# {desc}\n# ncu --metrics gpu__time_duration.sum --clock-control none -c 400 over `python bench.py --steps 2 --warmup 3 --no-cpu-baseline`\n")
    f.write("# per-launch times under ncu are cold-cache and serialised: compare SHARES with bench.py's stage_ms_per_step, not absolutes\n")
    f.write("kernel,launches,total_us,avg_us,share_pct\n")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k},{n},{t:.1f},{t / n:.1f},{100 * t / tot:.1f}\n")
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); H, U = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size']
idx = [H.index(w) for w in want]; ki = H.index('Kernel Name')
with open(f'profiles/{tag}_ncu_full_one_step.csv', 'w') as f:
    f.write(f"# {desc}\n# ncu --set full --clock-control none --import-source on -k regex:k_ --launch-skip 39 --launch-count 13 (all launches of one step)\n")
    f.write("kernel," + ",".join(f"{w} [{U[i]}]" for w, i in zip(want, idx)) + "\n")
    for r in rows[2:]:
        f.write(r[ki].split('(')[0].replace('void ', '') + "," + ",".join(r[i].replace(',', '') for i in idx) + "\n")
print(open(f'profiles/{tag}_launch_shares.csv').read())
