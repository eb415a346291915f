#!/bin/bash
# Tuning aid: per-iteration time of a fixed-iteration run for several band heights:  scripts/sweep_rows.sh M N rows...
cd "$(dirname "$0")/../poisson-ellipse-openmp-mpi-cuda-new_b200"
M=$1; N=$2; shift 2
for r in "$@"; do
  if [ "$r" = "auto" ]; then unset PB200_ROWS; else export PB200_ROWS=$r; fi
  ./poisson_b200 $M $N --fixed-iters 300 --json - 2>&1 | grep t_loop | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$M x $N rows=$r', 'us/it', round(1e6*d['t_loop_s']/d['iterations'],1))"
done
