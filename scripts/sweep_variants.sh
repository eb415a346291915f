#!/bin/bash
# Tuning aid: time K1/K2 for every libpoisson_b200_<variant>.so built with `make variant` (see the package Makefile).
cd "$(dirname "$0")/.."
for lib in poisson-ellipse-openmp-mpi-cuda-new_b200/libpoisson_b200_*.so; do
  v=$(basename "$lib" .so); v=${v#libpoisson_b200_}
  PB200_LIB=$PWD/$lib python bench.py --no-cpu-baseline --steps 3 "$@" 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', 'Gpt-it/s', round(d['value']/1e9,2), 'k1_ms', round(d['roofline']['k1_ms'],4), 'k2_ms', round(d['roofline']['k2_ms'],4))"
done
