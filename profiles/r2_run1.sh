#!/bin/bash
# round 2, first GPU pass: the whole -m gpu suite, smoke, both bench arms, one long (sustained-clock) bench line
set -u
OUT=gpurun_out; TAG=${1:-r2a}
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x -s > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" >> $OUT/${TAG}_pytest.log 2>&1
timeout 600 python bench.py --steps 200 --warmup 10 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?" >> $OUT/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err
timeout 600 python bench.py --steps 5000 --warmup 10 --no-baselines > $OUT/${TAG}_bench_long.json 2>> $OUT/${TAG}_bench.err
grep -E "passed|failed|error|rc=" $OUT/${TAG}_pytest.log | tail -5; cat $OUT/${TAG}_bench.json | cut -c1-1500; tail -3 $OUT/${TAG}_bench.err
