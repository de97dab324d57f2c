#!/bin/bash
# per-kernel launch list of the config-4 training step (cold-cache, serialised: shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_c4.csv python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c4.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_c4.log | cut -c1-200
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/launches_c4.csv')) if len(r) > 5]
hdr = rows[0]; ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
rows = rows[1:]
# last two steps = last 2*L launches; take the tail half of all launches after warmup: use last 2 steps' worth by counting 'prep_kernel'
idx = [i for i, r in enumerate(rows) if 'prep_kernel' in r[ki]]
start = idx[-2] if len(idx) >= 2 else 0
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[start:]:
    k = r[ki].split('(')[0]
    agg[k][0] += 1; agg[k][1] += float(r[vi].replace(',', '')) / 1e3
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:40s} n={v[0]:4d} {v[1]:10.1f} us {100*v[1]/tot:5.1f}%")
print("total us", tot)
PY
