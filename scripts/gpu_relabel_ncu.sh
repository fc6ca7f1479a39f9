#!/bin/bash
# launch list (durations) of the tensor-core relabel cases: scripts/gpu_relabel_ncu.sh TAG [cases]
set -u
TAG=$1; OUT=gpurun_out/$TAG; mkdir -p $OUT
export RELABEL_CASES=${2:-antmaze_q_mlp,antmaze_ens10_cvar} RELABEL_PRECISIONS=split_bf16
timeout 300 python scripts/bench_relabel.py > $OUT/relabel.json 2> $OUT/relabel.err; rc=$?; echo "bench rc=$rc"; cat $OUT/relabel.json
[ $rc -eq 0 ] || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_uniform.sum --clock-control none -k regex:"relabel_mlp_tc|mlp_tc_prep|cvar_select" -c 60 --csv --log-file $OUT/launches.csv python scripts/bench_relabel.py > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
python - $OUT/launches.csv <<'PY'
import csv,sys,collections
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); mi=hdr.index("Metric Name"); vi=hdr.index("Metric Value"); ii=hdr.index("ID")
d=collections.OrderedDict()
for r in rows[1:]:
    d.setdefault((r[ii],r[ki][:40]),{})[r[mi]]=r[vi]
for k,v in list(d.items())[:40]: print(k, v)
PY
