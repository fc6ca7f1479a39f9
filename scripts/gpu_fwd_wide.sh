#!/bin/bash
# wide persistent forward kernel (KXB = 2, action_dim <= 24, actor dropout): parity, then A/B against the non-persistent kernel
# (historical: IQLB_FWD_PERSIST=0 selected the one-tile-per-CTA forward kernel, removed after this A/B - profiles/r02f_fwd_wide_ab.jsonl)
mkdir -p gpurun_out; cd /root/repo
timeout 900 python -m pytest tests/test_gpu_bf16_step.py tests/test_gpu_iql_step.py -x -q -m gpu 2>&1 | tail -5
for i in 1 2; do
  timeout 300 python bench.py --workload cfg4 --steps 300 --warmup 20 --skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 2>/dev/null | tail -1 > gpurun_out/fw_cfg4_persist_$i.json
  IQLB_FWD_PERSIST=0 timeout 300 python bench.py --workload cfg4 --steps 300 --warmup 20 --skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 2>/dev/null | tail -1 > gpurun_out/fw_cfg4_nonpersist_$i.json
done
timeout 300 python bench.py --steps 300 --warmup 20 --skip-reference-cuda --skip-relabel --skip-cpu-baseline --skip-cfg1 2>/dev/null | tail -1 > gpurun_out/fw_cfg2.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/fw_*.json')):
    try:
        d=json.loads(open(f).read()); print(f, d['value'], d['ms_per_step'], [(k.get('name'),k.get('us')) for k in d.get('kernels',[])] if isinstance(d.get('kernels'),list) else d.get('kernels'))
    except Exception as e: print(f, 'ERR', e)
PY
