# k_explore CTA size sweep (run on the GPU box): rebuilds casim_explore.o per setting, C4 bench, kernel time alone.
for T in 96 128 192 256; do
  rm -f gym-exploration-2d_b200/casim/csrc/casim_explore.o
  make -s -C gym-exploration-2d_b200/casim/csrc EXTRA="-DX_THREADS=$T" > /dev/null 2>&1
  python bench.py --workload c4 --no-cpu-baseline --e2e-steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('X_THREADS=$T', round(d['ms_per_step']*1e3,1), {k:round(v['duration_us'],1) for k,v in d['roofline']['kernels'].items()})"
done
rm -f gym-exploration-2d_b200/casim/csrc/casim_explore.o
make -s -C gym-exploration-2d_b200/casim/csrc > /dev/null 2>&1
