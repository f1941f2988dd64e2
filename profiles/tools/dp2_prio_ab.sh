# 2-GPU A/B: exchange stream at default vs high priority
for prio in 0 1 0 1; do
  LMSFC_XCHG_PRIO=$prio timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus 2 --steps 300 --warmup 20 --no-baselines 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1])
print('XCHG_PRIO=$prio', 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['ms_per_step'], 4), 'bit-identical', d['replicas_bit_identical'])"
done
