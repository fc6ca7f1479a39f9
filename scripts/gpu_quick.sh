#!/bin/bash
# bf16 step tests + timeline + one short bench
set -u
TAG=${1:-q}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_bf16_step.py tests/test_gpu_iql_step.py tests/test_gpu_replay.py -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest.log
timeout 200 python scripts/timeline.py > $OUT/timeline.txt 2>&1; cat $OUT/timeline.txt | tail -${TL_LINES:-30}
timeout 300 python bench.py --skip-cpu-baseline --steps 200 --warmup 10 ${BENCH_ARGS:-} > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
python - $OUT/bench.json <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
print("value", round(d["value"]), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]))
print("   ", " ".join(f"{k['kernel']}={k['ms_per_step']*1e3:.1f}" for k in d["kernels"]))
PY
