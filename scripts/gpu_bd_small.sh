#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
timeout 1700 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
F="--skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 --steps 300 --warmup 20"
for i in 1 2 3; do
  timeout 300 python bench.py --agents-per-gpu 8 $F 2>/dev/null | tail -1 > gpurun_out/bds_8ag_persist_$i.json
  IQLB_BD_PERSIST=0 timeout 300 python bench.py --agents-per-gpu 8 $F 2>/dev/null | tail -1 > gpurun_out/bds_8ag_old_$i.json
done
for n in 1 16 32; do
  timeout 300 python bench.py --agents-per-gpu $n $F 2>/dev/null | tail -1 > gpurun_out/bds_${n}ag_persist.json
  IQLB_BD_PERSIST=0 timeout 300 python bench.py --agents-per-gpu $n $F 2>/dev/null | tail -1 > gpurun_out/bds_${n}ag_old.json
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bds_*.json')):
    try:
        d=json.loads(open(f).read()); print(f, round(d['value'],1), round(d['ms_per_step']*1e3,2), [(k['kernel'][:12], round(k['ms_per_step']*1e3,1)) for k in d['kernels'] if 'bwd' in k['kernel']])
    except Exception as e: print(f, 'ERR', e)
PY
