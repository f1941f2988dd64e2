#!/bin/bash
# 2-GPU pass: DP parity test (kept log), bench at N=2 with both exchanges
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_dp.py -q -x -s > $OUT/r2_dp2_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/r2_dp2_pytest.log
tail -5 $OUT/r2_dp2_pytest.log
for mode in p2p nccl; do
  LMSFC_DP=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus 2 --steps 200 --warmup 10 > $OUT/r2_dp2_bench_$mode.json 2> $OUT/r2_dp2_bench_$mode.err
  echo "$mode rc=$?"; cut -c1-260 $OUT/r2_dp2_bench_$mode.json; tail -3 $OUT/r2_dp2_bench_$mode.err
done
