#!/bin/bash
# Round profiling pass (run on the GPU box through gpurun): for every bench workload
#   1. the bench itself, un-profiled (its JSON line is the only source of bench values),
#   2. the ncu launch list (gpu__time_duration.sum, --clock-control none) of a short run of the same command,
#   3. one `ncu --set full` capture of the workload's dominant kernel (c2: the STEP instantiation k_core<0, A>, not the
#      RESET one that runs once per ring batch at set-up).
# Outputs land in gpurun_out/; profiles/summarize_ncu.py + profiles/make_traffic.py turn them into the tracked
# summaries and profiles/traffic.json.   Usage: bash profiles/capture.sh <tag> [workloads...]
TAG=${1:-r2}; shift
WLS=${@:-"c2 c3 full c4 c5 rvo"}
declare -A KERN=( [c2]="k_core<\(int\)0" [c3]="k_laser" [full]="k_grid_tma" [c4]="k_explore" [c5]="k_ga3c_tc" [rvo]="k_rvo4" )
mkdir -p gpurun_out
for wl in $WLS; do
  python bench.py --workload $wl > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err; echo "bench $wl rc=$?"
  ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv \
      python bench.py --workload $wl --steps 60 --warmup 3 --no-graph --no-cpu-baseline --e2e-steps 3 > gpurun_out/${TAG}_ncu_$wl.log 2>&1; echo "launches $wl rc=$?"
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:${KERN[$wl]}" -s 10 -c 3 -f -o gpurun_out/${TAG}_full_$wl \
      python bench.py --workload $wl --steps 40 --warmup 3 --no-graph --no-cpu-baseline --e2e-steps 3 > gpurun_out/${TAG}_ncufull_$wl.log 2>&1; echo "full $wl rc=$?"
  python profiles/summarize_ncu.py gpurun_out/${TAG}_full_$wl.ncu-rep > gpurun_out/${TAG}_ncu_${wl}_summary.txt 2>&1
done
python profiles/make_traffic.py $TAG > gpurun_out/${TAG}_traffic.log 2>&1; cp profiles/traffic.json gpurun_out/${TAG}_traffic.json
# gpurun brings back at most 64 MiB: keep the raw reports of the two headline workloads only
for wl in $WLS; do case $wl in c2|full) ;; *) rm -f gpurun_out/${TAG}_full_$wl.ncu-rep;; esac; done
du -sh gpurun_out
