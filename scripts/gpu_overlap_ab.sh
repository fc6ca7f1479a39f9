#!/bin/bash
# A/B of the optimizer-overlap schedule (run on the GPU box): one bench line per setting into gpurun_out/overlap_ab.jsonl
out=gpurun_out/overlap_ab.jsonl
: > $out
run() {
  echo "== $*" >&2
  env "$@" python bench.py --steps 100 --warmup 10 --skip-cpu-baseline --skip-reference-cuda 2>>gpurun_out/overlap_ab.err | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'env':'$*','value':d['value'],'ms':d['ms_per_step'],'e2e':d['e2e']['value'],'e2e_block':d['e2e']['blocking_api_value']}))" >> $out
}
run IQLB_GROUPS=1
run IQLB_GROUPS=2 IQLB_OVERLAP_ARG=0
run IQLB_GROUPS=2
run IQLB_GROUPS=2 IQLB_ADAM_CTAS=1
run IQLB_GROUPS=2 IQLB_ADAM_CTAS=3
run IQLB_GROUPS=2 IQLB_ADAM_CTAS=4
run IQLB_GROUPS=2 IQLB_L2HINT=0
run IQLB_GROUPS=4
run IQLB_GROUPS=3
cat $out
