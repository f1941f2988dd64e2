# grid of adam_stream_kernel (LMSFC_ADAM_STREAM = number of CTAs; chunks are dealt round-robin): more, shorter-lived CTAs
# let the forward convolutions' CTAs take an SM sooner
run() { timeout 300 python bench.py --no-baselines --steps 300 --warmup 20 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['ms_per_step'], 4))"; }
for v in 148 296 592 1184 148 592; do LMSFC_ADAM_STREAM=$v run ADAM_STREAM=$v; done
