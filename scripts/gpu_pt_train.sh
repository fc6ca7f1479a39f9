#!/bin/bash
# PT trainer: parity tests, launch list of one step (ncu), micro-benchmark.  scripts/gpu_pt_train.sh TAG
set -u
TAG=${1:-pt}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_pt_training.py tests/test_gpu_mr_training.py -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest.log | cut -c1-200
PT_CASES=${PT_CASES:-antmaze_E256_L1_B64,antmaze_E256_L1_B256,pen_E128_L2_B64} PT_CPU=${PT_CPU:-0} timeout 300 python scripts/bench_pt_train.py > $OUT/bench.jsonl 2>/dev/null; cat $OUT/bench.jsonl
if [ "${NCU:-1}" = "1" ]; then
PT_CASES=antmaze_E256_L1_B64 PT_CPU=0 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s ${NCU_SKIP:-1400} -c 300 --csv --log-file $OUT/launches.csv python scripts/bench_pt_train.py > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
python scripts/launch_shares.py $OUT/launches.csv | head -30
fi
