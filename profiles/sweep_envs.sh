#!/bin/bash
# E-sweep of the step kernel: config C2 (k_core alone) at E = 4096 ... 131072 environments per GPU,
# one step per launch and 16 steps per launch (casim_step_n); one JSON line per point.
#   bash profiles/sweep_envs.sh <tag>  ->  gpurun_out/<tag>_sweep_envs.jsonl
TAG=${1:-r2}
OUT=gpurun_out/${TAG}_sweep_envs.jsonl
: > $OUT
for E in 4096 16384 65536 131072; do
  for SPL in 1 16; do
    python bench.py --workload c2 --envs $E --steps 160 --warmup 16 --steps-per-launch $SPL --no-cpu-baseline --e2e-steps 5 \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(json.dumps({'workload':'c2','envs':$E,'steps_per_launch':$SPL,'us_per_step':d['ms_per_step']*1e3,'agent_steps_per_s':d['value'],'frac':r['whole_step']['frac_of_hbm_peak'],'frac_must_move':r['whole_step']['frac_must_move']}))" >> $OUT
  done
done
cat $OUT
