#!/bin/bash
# whole GPU suite, then the two bench workloads (kernel experiments: no CPU legs)
mkdir -p gpurun_out; cd /root/repo
timeout 1700 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 300 python bench.py --workload cfg4 --steps 300 --warmup 20 --skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 2>/dev/null | tail -1 > gpurun_out/suite_cfg4.json
timeout 300 python bench.py --steps 300 --warmup 20 --skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 2>/dev/null | tail -1 > gpurun_out/suite_cfg2.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/suite_*.json')):
    try:
        d=json.loads(open(f).read()); print(f, round(d['value'],1), round(d['ms_per_step']*1e3,2), [(k['kernel'], round(k['ms_per_step']*1e3,1)) for k in d['kernels']])
    except Exception as e: print(f, 'ERR', e)
PY
