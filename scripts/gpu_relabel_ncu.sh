#!/bin/bash
# launch list + one --set full capture of the tensor-core relabel: scripts/gpu_relabel_ncu.sh TAG [cases]
set -u
TAG=$1; OUT=gpurun_out/$TAG; mkdir -p $OUT
export RELABEL_CASES=${2:-antmaze_q_mlp,antmaze_ens10_cvar} RELABEL_PRECISIONS=split_bf16
CMD="python scripts/bench_relabel.py"
timeout 300 $CMD > $OUT/relabel.json 2> $OUT/relabel.err; rc=$?; echo "bench rc=$rc"; cat $OUT/relabel.json
[ $rc -eq 0 ] || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"relabel_mlp_tc|mlp_tc_prep|cvar_select" -c 60 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu.log 2>&1; echo "ncu launches rc=$?"
timeout 300 $CMD > $OUT/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"relabel_mlp_tc" -s 5 -c 1 -f -o $OUT/prof_step $CMD > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
