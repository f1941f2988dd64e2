#!/bin/bash
# Profiling recipe run on the GPU box (gpurun): tests, plain bench, ncu launch list, full captures.
# Usage: bash profiles/capture.sh <tag>     (outputs land in gpurun_out/<tag>_*)
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" >> $OUT/${TAG}_pytest.log 2>&1
timeout 600 python bench.py --steps 200 --warmup 10 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?" >> $OUT/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err
if [ "${NCU:-1}" = "1" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 > $OUT/${TAG}_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'igemm_fwd_kernel|wgrad_kernel' -s 20 -c 8 \
  -f -o $OUT/${TAG}_gemm_full python bench.py --steps 2 --warmup 3 > $OUT/${TAG}_ncu_full.log 2>&1
fi
timeout 300 python profiles/layer_bench.py > $OUT/${TAG}_layers.txt 2>&1
timeout 300 python profiles/trace_step.py $OUT/${TAG}_trace.txt > /dev/null 2>&1
LMSFC_SEQUENTIAL=1 timeout 300 python profiles/trace_step.py $OUT/${TAG}_trace_seq.txt > /dev/null 2>&1
timeout 300 python profiles/timeline.py > $OUT/${TAG}_timeline.txt 2>&1
tail -3 $OUT/${TAG}_pytest.log; cat $OUT/${TAG}_bench.json; tail -2 $OUT/${TAG}_bench.err
