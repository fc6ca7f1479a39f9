#!/bin/bash
# A/B of the relabel micro-benchmark by environment: scripts/gpu_relabel_ab.sh TAG name:ENV=V[,ENV2=V] ...
set -u
TAG=$1; shift; OUT=gpurun_out/$TAG; mkdir -p $OUT
export RELABEL_CASES=${RELABEL_CASES:-antmaze_q_mlp,antmaze_ens10_cvar,pen_human_ens981} RELABEL_PRECISIONS=split_bf16
for item in "$@"; do
  name=${item%%:*}; envs=${item#*:}; envs=${envs//,/ }
  timeout 300 env $envs python scripts/bench_relabel.py > $OUT/$name.json 2> $OUT/$name.err
  python - $OUT/$name.json $name <<'PY'
import json,sys
out=[]
for l in open(sys.argv[1]):
    d=json.loads(l); out.append(f"{d['case']}={d['split_bf16']['ms']:.3f}ms")
print(sys.argv[2], " ".join(out))
PY
done
