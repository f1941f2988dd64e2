#!/bin/bash
# Round-2 evidence pass on one B200 (gpurun): bench line, ncu launch list of the same command, ncu --set full of the
# roofline kernels (profiles/ncu_kernels.py), CUPTI trace of one replay, per-layer table.   bash profiles/r2_capture.sh
set -u
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/r2_smi.txt 2>&1
timeout 600 python bench.py --steps 200 --warmup 10 > $OUT/r2_bench.json 2> $OUT/r2_bench.err; echo "bench rc=$?" >> $OUT/r2_bench.err
timeout 600 python profiles/ncu_kernels.py > $OUT/r2_ncu_kernels_plain.log 2>&1; echo "plain rc=$?" >> $OUT/r2_ncu_kernels_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $OUT/r2_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-baselines > $OUT/r2_ncu_list.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k regex:'igemm_fwd_kernel|wgrad_kernel|adam_stream_kernel|adam_kernel|sumsq_kernel|mse_multi_kernel|gc_forward_kernel|gc_backward_kernel|eb_forward_kernel|eb_backward_kernel' \
  -f -o $OUT/r2_kernels_full python profiles/ncu_kernels.py > $OUT/r2_ncu_full.log 2>&1
timeout 300 python profiles/layer_bench.py > $OUT/r2_layers.txt 2>&1
timeout 300 python profiles/trace_step.py $OUT/r2_trace_step_n1.txt > /dev/null 2>&1
ls -la $OUT; tail -2 $OUT/r2_bench.err; tail -3 $OUT/r2_ncu_full.log; tail -2 $OUT/r2_ncu_list.log
