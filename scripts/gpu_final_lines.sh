#!/bin/bash
# bench lines of the secondary workloads at the final tree (kernel-only legs: no CPU baseline, no relabel block)
mkdir -p gpurun_out/final; cd /root/repo
F="--skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 --steps 300 --warmup 20"
timeout 300 python bench.py --workload cfg4 $F 2>/dev/null | tail -1 > gpurun_out/final/bench_bf16_cfg4_pen_dropout_n1.json
timeout 300 python bench.py --agents-per-gpu 8 $F 2>/dev/null | tail -1 > gpurun_out/final/bench_bf16_8agents_n1.json
timeout 300 python bench.py --precision fp32 $F 2>/dev/null | tail -1 > gpurun_out/final/bench_fp32_n1.json
timeout 600 python scripts/pipeline_cfg3_cfg4.py > gpurun_out/final/pipelines.jsonl 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/final/bench_*.json')):
    try:
        d=json.loads(open(f).read()); print(f, round(d['value'],1), round(d['ms_per_step']*1e3,2), round(d['e2e']['value'],1), [(k['kernel'][:12], round(k['ms_per_step']*1e3,1)) for k in d['kernels']])
    except Exception as e: print(f, 'ERR', e)
PY
cut -c1-400 gpurun_out/final/pipelines.jsonl
