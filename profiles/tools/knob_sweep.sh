# one bench line per knob setting on the same box: LMSFC_N_WGRAD (weight-gradient streams), LMSFC_FUSE (GDN fusion modes)
run() { timeout 300 python bench.py --no-baselines --steps 300 --warmup 20 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['ms_per_step'], 4))"; }
run default
LMSFC_N_WGRAD=2 run N_WGRAD=2
LMSFC_N_WGRAD=4 run N_WGRAD=4
LMSFC_FUSE=0 run FUSE=0
LMSFC_FUSE=3 run FUSE=3
LMSFC_FUSE=6 run FUSE=6
run default
