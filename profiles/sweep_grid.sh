#!/bin/bash
# Tuning aid (run on the GPU box): rebuild casim_sensors.cu with k_grid's compile-time knobs and time the `full` workload.
# Usage: bash profiles/sweep_grid.sh "W,MINB,U ..."    Leaves the default build in place at the end.
cd gym-exploration-2d_b200/casim/csrc
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC"
for v in $1; do
  IFS=, read W B U <<< "$v"
  $NV -DGRID_WARPS=$W -DGRID_MINB=$B -DGRID_U=$U -c casim_sensors.cu -o casim_sensors.o 2>/dev/null && \
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libcasim.so *.o || { echo "$v build failed"; continue; }
  echo -n "W=$W MINB=$B U=$U "
  (cd ../../.. && python bench.py --workload full --no-cpu-baseline --e2e-steps 3 --steps 1000 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('step_us', round(d['ms_per_step']*1000,2), 'k_grid_us', round(r['duration_us'],2), 'frac', round(r['frac'],3))")
done
touch casim_sensors.cu; make -s >/dev/null 2>&1; echo rebuilt default
