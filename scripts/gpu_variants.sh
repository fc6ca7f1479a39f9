#!/bin/bash
# quick A/B of step variants on one box: tests first, then short benches
set -u
TAG=${1:-var}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest_gpu.log
run() { name=$1; shift; timeout 300 env "$@" python bench.py --skip-cpu-baseline --steps 200 --warmup 10 > $OUT/bench_$name.json 2> $OUT/bench_$name.err; echo "$name rc=$?" | tee -a $OUT/rc.txt;
  python - $OUT/bench_$name.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print(sys.argv[1], "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]))
    print("   ", " ".join(f"{k['kernel']}={k['ms_per_step']*1e3:.1f}" for k in d["kernels"]))
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
}
run base IQLB_X=0
run groups2 IQLB_GROUPS=2
