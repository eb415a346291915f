#!/bin/bash
# Tuning aid: per-iteration time of small solves through the resident loop kernels (default) and the two marching kernels.
cd "$(dirname "$0")/../poisson-ellipse-openmp-mpi-cuda-new_b200"
for g in "$@"; do
  for mode in 0 1; do PB200_LOOP=$mode timeout 120 ./poisson_b200 $g --json - 2>&1 | grep t_loop | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$g loop=$mode', d['iterations'], round(1e6*d['t_loop_s']/d['iterations'],1), 'us/it', d['kernel_launches'], 'launches')"; done
done
