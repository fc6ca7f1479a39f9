#!/bin/bash
# Round-2 ncu captures (run on the GPU box, one GPU).  Every profiled command first exits 0 without ncu.
OUT=gpurun_out/prof_r02; mkdir -p $OUT
CMD="python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-reference-cuda --skip-relabel"
K='regex:bf16_|adam_polyak|loss_kernel|replay_gather'
timeout 300 $CMD > $OUT/plain.json 2> $OUT/plain.err || { echo "plain bench failed"; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "$K" -s 18 -c 6 -o $OUT/prof_step -f $CMD > $OUT/ncu_full.log 2>&1; echo "full rc=$?"
# the generic tensor-core GEMM and the PT relabel
timeout 300 python scripts/bench_pt.py 4096 1 > $OUT/pt_plain.log 2>&1 || { echo "pt plain failed"; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'pt_|gemm' -c 300 --csv --log-file $OUT/pt_launches.csv python scripts/bench_pt.py 4096 1 > $OUT/pt_ncu.log 2>&1; echo "pt launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gemm_tc' -s 5 -c 1 -o $OUT/prof_gemm_tc -f python scripts/bench_pt.py 4096 1 > $OUT/pt_ncu_full.log 2>&1; echo "gemm_tc full rc=$?"
ls -la $OUT
