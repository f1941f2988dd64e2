run() { timeout 300 python bench.py --no-baselines --steps 300 --warmup 20 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['ms_per_step'], 4))"; }
for i in 1 2; do
LMSFC_FUSE=2 run FUSE=2
LMSFC_FUSE=6 run FUSE=6
LMSFC_FUSE=7 run FUSE=7
done
