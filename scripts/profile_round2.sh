#!/bin/bash
# Runs ON THE GPU BOX (gpurun).  Round 2: (1) launch list of one eager cfg2 bench run (per-launch times, serialised and
# cold-cache: only the SHARES are comparable with bench.py's CUDA-event numbers); (2) ONE `ncu --set full` pass over a
# profiled region that contains every kernel of the path once (scripts/profile_step.py), exported as raw CSV; (3) a
# source-level capture of the two tcgen05 edge kernels.  Outputs: gpurun_out/prof2/ (summaries are copied to profiles/
# by scripts/summarise_round2.py).  Numbers printed under ncu are never bench values.
set -u
O=gpurun_out/prof2; mkdir -p $O
B="python bench.py --eager --steps 2 --warmup 3 --sampling-structures 0"
$B > $O/plain_cfg2.json 2> $O/plain_cfg2.err || { echo "plain bench failed"; tail -5 $O/plain_cfg2.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_cfg2.csv $B > $O/ncu_launch_cfg2.log 2>&1
python scripts/profile_step.py cfg2 > $O/plain_step.log 2>&1 || { echo "plain profile_step failed"; tail -5 $O/plain_step.log; exit 1; }
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o $O/all_kernels python scripts/profile_step.py cfg2 > $O/ncu_all.log 2>&1
ncu -i $O/all_kernels.ncu-rep --page raw --csv --print-units base > $O/all_kernels_raw.csv 2>/dev/null
for K in tc_edge_bwd tc_edge_fwd; do
  ncu -i $O/all_kernels.ncu-rep --kernel-name "regex:^$K\$" --page source --csv > $O/${K}_source.csv 2>/dev/null
  ncu -i $O/all_kernels.ncu-rep --kernel-name "regex:^$K\$" --page details > $O/${K}_details.txt 2>/dev/null
done
# reduce ON THE BOX: gpurun returns at most 64 MiB, the report and the per-instruction CSVs are far larger
mkdir -p gpurun_out/profiles_r2
python scripts/summarise_round2.py $O gpurun_out/profiles_r2 > $O/summarise.log 2>&1
rm -f $O/all_kernels.ncu-rep $O/*_source.csv $O/launches_cfg2.csv
du -sh gpurun_out; ls -la gpurun_out/profiles_r2
