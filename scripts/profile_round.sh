#!/bin/bash
# Runs ON THE GPU BOX (gpurun): launch lists and one `ncu --set full` capture per hot kernel, cfg2 and cfg4.
# Outputs go to gpurun_out/prof/ (text summaries are copied to profiles/ by hand).  Numbers printed under ncu are never
# bench values; the bench lines come from separate runs.
set -u
O=gpurun_out/prof; mkdir -p $O
B="python bench.py --steps 2 --warmup 3"
$B > $O/plain_cfg2.json 2> $O/plain_cfg2.err || { echo "plain cfg2 bench failed"; exit 1; }
if [ "${ONLY128:-0}" != 1 ]; then
$B --workload cfg4 > $O/plain_cfg4.json 2> $O/plain_cfg4.err || { echo "plain cfg4 bench failed"; exit 1; }
for W in cfg2 cfg4; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_$W.csv $B --workload $W > $O/ncu_launch_$W.log 2>&1
done
fi
cap() {  # name regex workload skip
  ncu --set full --clock-control none --import-source on --kernel-name "regex:$2" --launch-skip $4 --launch-count 1 -f -o $O/$1 $B --workload $3 > $O/ncu_$1.log 2>&1
  ncu -i $O/$1.ncu-rep --page details > $O/$1_details.txt 2>&1
  ncu -i $O/$1.ncu-rep --page source --csv > $O/$1_source.csv 2>/dev/null
  ncu -i $O/$1.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
h=rows[0]; v=rows[-1]
want=['dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','lts__t_bytes.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed','dram__throughput.avg.pct_of_peak_sustained_elapsed','smsp__inst_executed.sum','sm__cycles_elapsed.max']
for k,(a,b,c) in enumerate(zip(h,rows[1],v)):
    if any(a.startswith(w) for w in want): print(a,b,c)
" > $O/$1_raw.txt
}
cap tc_edge_bwd128 '^tc_edge_bwd$' cfg2 14   # ncu matches the base name (no template arguments)
cap tc_edge_fwd128 '^tc_edge_fwd$' cfg2 14
if [ "${ONLY128:-0}" = 1 ]; then ls -la $O; exit 0; fi
cap tc_edge_fwd256 'tc_edge_fwd256' cfg4 12
cap tc_edge_bwd256 'tc_edge_bwd256' cfg4 12
cap tc_wgrad256 'tc_wgrad256' cfg4 12
ls -la $O | head -40
