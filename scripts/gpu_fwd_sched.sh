#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
timeout 900 python -m pytest tests/test_gpu_bf16_step.py tests/test_gpu_iql_step.py tests/test_gpu_pipeline.py -x -q -m gpu 2>&1 | tail -3
F="--skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 --steps 300 --warmup 20"
for i in 1 2; do timeout 300 python bench.py $F 2>/dev/null | tail -1 > gpurun_out/fs_cfg2_$i.json; done
timeout 300 python bench.py --workload cfg4 $F 2>/dev/null | tail -1 > gpurun_out/fs_cfg4.json
timeout 300 python bench.py --agents-per-gpu 8 $F 2>/dev/null | tail -1 > gpurun_out/fs_8ag.json
timeout 300 python bench.py --agents-per-gpu 32 $F 2>/dev/null | tail -1 > gpurun_out/fs_32ag.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/fs_*.json')):
    try:
        d=json.loads(open(f).read()); print(f, round(d['value'],1), round(d['ms_per_step']*1e3,2), [(k['kernel'][:12], round(k['ms_per_step']*1e3,1)) for k in d['kernels']])
    except Exception as e: print(f, 'ERR', e)
PY
