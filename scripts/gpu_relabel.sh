#!/bin/bash
# relabel parity tests + micro-benchmark of the MLP relabel (fp32 vs split-bf16): scripts/gpu_relabel.sh TAG
set -u
TAG=$1; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_relabel.py -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest.log
timeout 600 python scripts/bench_relabel.py > $OUT/relabel.json 2> $OUT/relabel.err; echo "bench rc=$?"; cat $OUT/relabel.json; tail -3 $OUT/relabel.err
