# run-to-run spread of bench.py's final loss (same seeds): is the trajectory deterministic, and does it move with the optimiser's speed?
for v in 0 0 0 32 32 148; do
  if [ "$v" = "0" ]; then unset LMSFC_ADAM_STREAM; else export LMSFC_ADAM_STREAM=$v; fi
  timeout 300 python bench.py --no-baselines --steps 300 --warmup 20 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ADAM_STREAM=$v', 'ms/step', round(d['ms_per_step'], 4), 'first', d['first_loss'], 'final', d['final_loss'])"
done
