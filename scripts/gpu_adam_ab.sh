#!/bin/bash
# A/B of optimizer-kernel launch shapes on one box (same tree): scripts/gpu_adam_ab.sh TAG
set -u
TAG=${1:-adam}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_bf16_step.py tests/test_gpu_replay.py tests/test_gpu_iql_step.py -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/pytest.log
run() { name=$1; shift; timeout 300 env "$@" python bench.py --skip-cpu-baseline --skip-reference-cuda --skip-relabel --steps 200 --warmup 10 > $OUT/bench_$name.json 2> $OUT/bench_$name.err; 
  python - $OUT/bench_$name.json $name <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), " ".join(f"{k['kernel']}={k['ms_per_step']*1e3:.1f}" for k in d["kernels"]))
except Exception as e: print(sys.argv[2], "FAILED", e)
PY
}
run base IQLB_X=0
run nolook IQLB_PRESAMPLE=0
run persist0 IQLB_ADAM_PERSIST=0
run persist0_nolook IQLB_ADAM_PERSIST=0 IQLB_PRESAMPLE=0
run base2 IQLB_X=0
