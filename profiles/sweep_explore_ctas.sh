# k_explore register cap sweep at 128 threads (run on the GPU box): __launch_bounds__(128, M)
for M in ${MS:-5 6 7 8}; do
  rm -f gym-exploration-2d_b200/casim/csrc/casim_explore.o
  make -s -C gym-exploration-2d_b200/casim/csrc EXTRA="-DX_THREADS=128 -DX_MIN_CTAS=$M" > /dev/null 2>&1
  grep -A2 "k_exploreILb0" gym-exploration-2d_b200/casim/csrc/casim_explore.ptxas.log | grep -o "Used [0-9]* registers\|[0-9]* bytes spill stores" | tr '\n' ' '
  python bench.py --workload c4 --no-cpu-baseline --e2e-steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('X_MIN_CTAS=$M', round(d['ms_per_step']*1e3,1), {k:round(v['duration_us'],1) for k,v in d['roofline']['kernels'].items()})"
done
rm -f gym-exploration-2d_b200/casim/csrc/casim_explore.o
make -s -C gym-exploration-2d_b200/casim/csrc > /dev/null 2>&1
