#!/bin/bash
# group timeline + kernel times of a 32-agent chain (run on the GPU box)
for ctas in 2 4 8; do
  echo "== timeline G=2 ADAM_CTAS=$ctas"
  IQLB_GROUPS=2 IQLB_GROUP_TL=1 IQLB_ADAM_CTAS=$ctas python bench.py --steps 40 --warmup 10 --skip-cpu-baseline --skip-reference-cuda 2>&1 >/dev/null | grep "group timeline" | head -4
done
for n in 32 64; do
  echo "== kernel times, $n agents, one group"
  IQLB_GROUPS=1 python bench.py --steps 50 --warmup 10 --agents-per-gpu $n --skip-cpu-baseline --skip-reference-cuda 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step']*1e3, [(k['kernel'], round(k['ms_per_step']*1e3,1)) for k in d['kernels']])"
done
