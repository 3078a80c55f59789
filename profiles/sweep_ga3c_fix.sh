# fix-up tile size sweep of the exact GA3C mode (run on the GPU box): agents per CTA = 8 * G_FIX_AG
for G in ${GS:-1 2}; do
  rm -f gym-exploration-2d_b200/casim/csrc/casim_ga3c.o
  make -s -C gym-exploration-2d_b200/casim/csrc EXTRA="-DG_FIX_AG=$G" > /dev/null 2>&1
  python bench.py --workload c5 --no-cpu-baseline --e2e-steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('G_FIX_AG=$G', round(d['ms_per_step']*1e3,1), {k:round(v['duration_us'],1) for k,v in d['roofline']['kernels'].items()})"
done
rm -f gym-exploration-2d_b200/casim/csrc/casim_ga3c.o
make -s -C gym-exploration-2d_b200/casim/csrc > /dev/null 2>&1
