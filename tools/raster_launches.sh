#!/bin/bash
# ncu launch list (time + DRAM bytes) of the raster bench; prints per-kernel averages
set -u
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 70 --csv --log-file gpurun_out/raster_launches.csv python bench.py --mode raster --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/raster_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows = list(csv.reader(open('gpurun_out/raster_launches.csv')))
hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
h = rows[hi]; kn = h.index('Kernel Name'); mn = h.index('Metric Name'); mv = h.index('Metric Value')
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= mv or 'smplb' not in r[kn]: continue
    agg.setdefault(r[kn].split('(')[0], {}).setdefault(r[mn], []).append(float(r[mv].replace(',', '')))
for k, d in agg.items():
    print('%-40s n=%d  %.1f us  rd %.0f MB  wr %.0f MB' % (k, len(d['gpu__time_duration.sum']), sum(d['gpu__time_duration.sum']) / len(d['gpu__time_duration.sum']) / 1e3,
          sum(d['dram__bytes_read.sum']) / len(d['dram__bytes_read.sum']) / 1e6, sum(d['dram__bytes_write.sum']) / len(d['dram__bytes_write.sum']) / 1e6))
PY
