#!/bin/bash
# A/B of the weight-gradient kernel schedules on one box: scripts/gpu_bw_ab.sh TAG
set -u
TAG=${1:-bw}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_bf16_step.py tests/test_gpu_iql_step.py tests/test_gpu_checkpoint.py -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest.log | cut -c1-300
run() { name=$1; shift; timeout 300 env "$@" python bench.py --skip-cpu-baseline --skip-reference-cuda --skip-relabel --steps 200 --warmup 10 ${BENCH_ARGS:-} > $OUT/bench_$name.json 2> $OUT/bench_$name.err;
  python - $OUT/bench_$name.json $name <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), " ".join(f"{k['kernel']}={k['ms_per_step']*1e3:.1f}" for k in d["kernels"]))
except Exception as e: print(sys.argv[2], "FAILED", e)
PY
}
run new IQLB_X=0
run new2 IQLB_X=0
timeout 200 python scripts/timeline.py > $OUT/timeline.txt 2>&1; grep -A12 "bwd_weight CTAs" $OUT/timeline.txt | tail -8
