#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
F="--skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 --steps 200 --warmup 20"
for z in 0 16 8 24; do
  IQLB_DBG_Z=$z timeout 300 python bench.py $F 2>/dev/null | tail -1 > gpurun_out/bwx_64_z$z.json
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bwx_*.json')):
    try:
        d=json.loads(open(f).read()); print(f, round(d['ms_per_step']*1e3,2), [(k['kernel'][:12], round(k['ms_per_step']*1e3,1)) for k in d['kernels'] if 'bwd' in k['kernel']])
    except Exception as e: print(f, 'ERR', e)
PY
