#!/bin/bash
# in-pipeline DRAM traffic per kernel: one ncu pass (no replay), caches left as the step leaves them
OUT=gpurun_out/${1:-dram}; mkdir -p $OUT
CMD="python bench.py --steps 6 --warmup 3 --skip-cpu-baseline"
timeout 300 $CMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --cache-control none --clock-control none -k regex:'bf16_|adam_polyak|loss_kernel|replay_gather' -s 30 -c 12 --csv --log-file $OUT/dram.csv $CMD > $OUT/ncu.log 2>&1
echo rc=$?
python - $OUT/dram.csv <<'PY'
import csv,sys,collections
lines=[l for l in open(sys.argv[1]) if l.startswith('"')]
rd=csv.reader(lines); hdr=next(rd)
ki,mi,vi,ui,ii=hdr.index("Kernel Name"),hdr.index("Metric Name"),hdr.index("Metric Value"),hdr.index("Metric Unit"),hdr.index("ID")
rows=collections.OrderedDict()
for r in rd:
    rows.setdefault((r[ii],r[ki].split("(")[0]),{})[r[mi]]=(r[vi],r[ui])
for (i,k),m in rows.items():
    print(i,k[:40],{a.split("__")[-1]:" ".join(b) for a,b in m.items()})
PY
