for nb in 2 3 2 3 4; do
LMSFC_BENCH_BUFFERS=$nb timeout 300 python bench.py --no-baselines --steps 300 --warmup 20 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('NB=$nb ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['ms_per_step'], 4), d['loss_check']['ok'])"
done
