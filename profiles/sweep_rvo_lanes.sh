for L in 4 8 16; do
  rm -f gym-exploration-2d_b200/casim/csrc/casim_rvo.o
  make -s -C gym-exploration-2d_b200/casim/csrc EXTRA="-DR4_LPA=$L" > /dev/null 2>&1
  python bench.py --workload rvo --no-cpu-baseline --e2e-steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('LPA=$L', {k:round(v['duration_us'],1) for k,v in d['roofline']['kernels'].items()})"
done
