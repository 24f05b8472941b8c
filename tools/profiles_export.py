#!/usr/bin/env python3
"""Turns gpurun_out/<launch list csv> and <.ncu-rep> into the small CSV summaries kept under profiles/.
usage: tools/profiles_export.py TAG launches.csv report.ncu-rep "free-text description" [stereo pairs per step]
Also rewrites profiles/traffic.json (per-stage DRAM bytes and issue utilisation that bench.py quotes)."""
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
    f.write(f"# {desc}\n# ncu --metrics gpu__time_duration.sum --clock-control none -c 400 over `python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary --no-verify`\n")
    f.write("# per-launch times under ncu are cold-cache and serialised: compare SHARES with bench.py's stage_ms_per_step, not absolutes\n")
    f.write("kernel,launches,total_us,avg_us,share_pct\n")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"\"{k}\",{n},{t:.1f},{t / n:.1f},{100 * t / tot:.1f}\n")      # quoted: template arguments carry commas
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
    f.write(f"# {desc}\n# ncu --set full --clock-control none --import-source on -k regex:k_ --launch-skip 45 --launch-count 15 (all launches of one step)\n")
    f.write("kernel," + ",".join(f"{w} [{U[i]}]" for w, i in zip(want, idx)) + "\n")
    for r in rows[2:]:
        f.write('"' + r[ki].split('(')[0].replace('void ', '') + '",' + ",".join(r[i].replace(',', '') for i in idx) + "\n")
print(open(f'profiles/{tag}_launch_shares.csv').read())

# ---- profiles/traffic.json: per-stage DRAM bytes per step and issue / pipe utilisation, read by bench.py ----
import json
STAGE = [('k_border_level0', 'pyramid'), ('k_resize', 'pyramid'), ('k_fast_cells', 'fast'), ('k_octree', 'quadtree'), ('k_blur', 'blur'), ('k_describe', 'describe'),
         ('k_stereo_rows', 'stereo_match'), ('k_stereo_match', 'stereo_match'), ('k_stereo_median', 'stereo_median')]
def col(name): return H.index(name)
dram = collections.OrderedDict(); acc = {}
for r in rows[2:]:
    kn = r[ki]
    st = next((s for k, s in STAGE if k in kn), None)
    if st is None:
        continue
    f = lambda n: float(r[col(n)].replace(',', ''))
    mul = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    dram[st] = dram.get(st, 0) + int(f('dram__bytes_read.sum') * mul[U[col('dram__bytes_read.sum')]] + f('dram__bytes_write.sum') * mul[U[col('dram__bytes_write.sum')]])
    t = f('gpu__time_duration.sum')
    a = acc.setdefault(st, [0.0, 0.0, 0.0, 0.0])
    a[0] += t; a[1] += t * f('smsp__issue_active.avg.pct_of_peak_sustained_active')
    a[2] += t * f('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'); a[3] += t * f('l1tex__throughput.avg.pct_of_peak_sustained_active')
pairs = int(sys.argv[5]) if len(sys.argv) > 5 else 256
src = f"profiles/{tag}_ncu_full_one_step.csv"
out = {"source": src + " (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, summed over the launches of a stage)",
       "stereo_pairs_per_step": pairs, "workload": "KITTI", "dram_bytes_per_step": dram,
       "issue": {"source": src + " (smsp__issue_active, sm__inst_executed_pipe_alu, l1tex__throughput; % of peak sustained, time-weighted over the launches of a stage, ncu --set full of one step)"}}
for st, a in acc.items():
    if st != 'stereo_median':
        out["issue"][st] = {"issue_slot_pct": round(a[1] / a[0], 1), "alu_pipe_pct": round(a[2] / a[0], 1), "l1tex_pct": round(a[3] / a[0], 1)}
json.dump(out, open('profiles/traffic.json', 'w'), indent=1)
print(json.dumps(out["dram_bytes_per_step"]))
