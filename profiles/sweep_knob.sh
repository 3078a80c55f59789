#!/bin/bash
# Tuning aid (run on the GPU box): rebuild ONE kernel source with compile-time knobs and time a bench workload.
# Usage: bash profiles/sweep_knob.sh <file.cu> <workload> "<-Dflags set 1>" "<-Dflags set 2>" ...   (restores the default build at the end)
SRC=$1; WL=$2; shift 2
cd gym-exploration-2d_b200/casim/csrc
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC"
for v in "$@"; do
  $NV $v -c $SRC -o ${SRC%.cu}.o 2>/dev/null && \
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libcasim.so *.o || { echo "$v build failed"; continue; }
  echo -n "[$v] "
  (cd ../../.. && python bench.py --workload $WL --no-cpu-baseline --e2e-steps 3 --steps 500 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('step_us', round(d['ms_per_step']*1000,2), r['kernel'], round(r['duration_us'],2), 'frac', round(r['frac'],4))")
done
touch $SRC; make -s >/dev/null 2>&1; echo rebuilt default
