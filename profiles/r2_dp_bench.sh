#!/bin/bash
# usage: r2_dp_bench.sh N "modes" [trace]
N=${1:-2}; MODES=${2:-"p2p nccl"}; OUT=gpurun_out; mkdir -p $OUT
for mode in $MODES; do
  LMSFC_VERBOSE=1 LMSFC_DP=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus $N --steps 200 --warmup 10 > $OUT/r2_dp${N}_bench_$mode.json 2> $OUT/r2_dp${N}_bench_$mode.err
  echo "$mode rc=$?"; grep "lmsfc exchange" $OUT/r2_dp${N}_bench_$mode.json $OUT/r2_dp${N}_bench_$mode.err | head -2; python -c "
import json,sys
d=json.loads([l for l in open('$OUT/r2_dp${N}_bench_$mode.json') if l.startswith('{')][-1]); print('$mode', 'N=$N', round(d['ms_per_step'],4), 'ms/step  e2e', round(d['e2e']['ms_per_step'],4), ' value', round(d['value']), 'replicas', d['replicas_bit_identical'])" || tail -5 $OUT/r2_dp${N}_bench_$mode.err
done
if [ "${3:-}" = "trace" ]; then
  LMSFC_DP=p2p timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 profiles/trace_step.py $OUT/r2_dp${N}_trace_p2p.txt > /dev/null 2>&1
fi
