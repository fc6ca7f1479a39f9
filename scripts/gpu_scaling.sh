#!/bin/bash
# strong (64 agents total, BASELINE configs[4]) and weak (64 agents per GPU) scaling on one 8-GPU box
OUT=gpurun_out/${TAG:-scaling_r02}; mkdir -p $OUT
for mode in strong weak; do
  for n in ${NS:-1 2 4 8}; do
    timeout 600 python bench.py --gpus $n --steps 200 --warmup 10 --scaling $mode --agents-total 64 --skip-cpu-baseline --skip-reference-cuda --skip-relabel \
      > $OUT/bench_${mode}_n$n.json 2> $OUT/bench_${mode}_n$n.err
    echo "$mode n=$n rc=$?"
    python - $OUT/bench_${mode}_n$n.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print("   value", round(d["value"]), "ms/step", round(d["ms_per_step"],4), "agents/gpu", d["config"]["agents_per_gpu"], "e2e", round(d["e2e"]["value"]))
except Exception as e: print("   no line:", e)
PY
  done
done
