#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (both arms), ncu launch list + full capture of the step kernels.
# Usage (under gpurun): bash scripts/gpu_check.sh [tag]
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench_n1.json 2> $OUT/bench_n1.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" | tee -a $OUT/rc.txt
if [ "${NCU:-1}" = "1" ]; then
CMD="python bench.py --steps 3 --warmup 3 --skip-cpu-baseline --skip-reference-cuda --skip-relabel"
timeout 300 $CMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launch.log 2>&1
echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
timeout 300 $CMD > $OUT/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'bf16_|adam_polyak|loss_kernel|replay_gather|pack_x' -s 18 -c 6 -f -o $OUT/prof_step $CMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $OUT/rc.txt
fi
tail -3 $OUT/pytest_gpu.log; tail -2 $OUT/smoke.log; cat $OUT/bench_n1.json | cut -c1-600
