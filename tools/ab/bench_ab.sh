#!/bin/bash
# power-capped A/B: the full device-resident bench loop (300 steps) alternately on tools/ab/A.so and B.so
keep=$(mktemp); cp hivemind_b200/libhm_engine.so $keep
for rep in 1 2 3; do for v in A B; do cp tools/ab/$v.so hivemind_b200/libhm_engine.so; echo -n "$v: "; timeout 200 python bench.py --steps 300 --warmup 10 --no-cpu-baseline --selfplay-seconds 0 --no-crossboard 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['roofline']['avg_launch_ms'], d['clocks']['sm_mhz'])"; done; done
cp $keep hivemind_b200/libhm_engine.so
