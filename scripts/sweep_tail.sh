#!/bin/bash
# Tuning aid: the tail bands of the marching kernels' launch plan (solver.cu: march_plan) at 8192 x 8192.
cd "$(dirname "$0")/.."
run() { python bench.py --no-cpu-baseline --no-e2e --steps 5 "${ARGS[@]}" 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['value']/1e9,2), round(d['roofline']['k1_ms'],4), round(d['roofline']['k2_ms'],4))"; }
ARGS=("$@")
run default
PB200_TAIL_WAVES=1.0 run waves1.0
PB200_TAIL_WAVES=2.5 run waves2.5
PB200_TAIL_WAVES=4 run waves4
PB200_TAIL=0 run tail_off
