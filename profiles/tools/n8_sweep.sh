run() { tag=$1; shift; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) bench.py --gpus 8 --steps 60 --warmup 5 2> gpurun_out/n8_$tag.err | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$tag', round(d['value']), round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']))
"; }
run b16 LMSFC_BUCKET_MB=16
run b64 LMSFC_BUCKET_MB=64
run ctas8 NCCL_MAX_CTAS=8
run ring NCCL_ALGO=Ring
