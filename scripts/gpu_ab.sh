#!/bin/bash
# A/B of environment-selected variants: scripts/gpu_ab.sh tag "NAME=ENV=VAL ..." (each item name:env assignments comma separated)
set -u
TAG=$1; shift; OUT=gpurun_out/$TAG; mkdir -p $OUT
if [ "${TESTS:-1}" = "1" ]; then timeout 600 python -m pytest tests/test_gpu_bf16_step.py tests/test_gpu_iql_step.py -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/pytest.log; fi
for item in "$@"; do
  name=${item%%:*}; envs=${item#*:}; envs=${envs//,/ }
  timeout 300 env $envs python bench.py --skip-cpu-baseline --steps 200 --warmup 10 > $OUT/bench_$name.json 2> $OUT/bench_$name.err; rc=$?
  python - $OUT/bench_$name.json $name $rc <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print(sys.argv[2], "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), "|", " ".join(f"{k['kernel']}={k['ms_per_step']*1e3:.1f}" for k in d["kernels"]))
except Exception as e: print(sys.argv[2], "FAILED rc", sys.argv[3], e)
PY
done
