# bench.py --no-baselines a few times: ms/step of the current tree (A/B against a number from the same box)
for i in 1 2 3; do
  timeout 300 python bench.py --no-baselines --steps 300 --warmup 20 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('run $i ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['ms_per_step'], 4), 'sumsq GB/s', round(d['kernels']['sumsq_kernel']['achieved']))"
done
