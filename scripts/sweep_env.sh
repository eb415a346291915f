#!/bin/bash
# Tuning aid: bench.py under several values of one environment variable:  scripts/sweep_env.sh VAR v1 v2 ...
cd "$(dirname "$0")/.."
var=$1; shift
for v in "$@"; do
  env $var=$v python bench.py --no-cpu-baseline --steps 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$var=$v', 'Gpt-it/s', round(d['value']/1e9,2), 'k1_ms', round(d['roofline']['k1_ms'],4), 'k2_ms', round(d['roofline']['k2_ms'],4))"
done
