for v in 0 148 96 64 48 32; do
  if [ "$v" = "0" ]; then unset LMSFC_ADAM_STREAM; else export LMSFC_ADAM_STREAM=$v; fi
  timeout 300 python bench.py --no-baselines --steps 300 --warmup 20 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ADAM_STREAM=$v', 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['ms_per_step'], 4), 'loss', d['final_loss'])"
done
