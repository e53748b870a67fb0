#!/bin/bash
# ONE `ncu --set full` pass over the tile-GEMM kernels of the parameters.yaml step (8 graphs of 96 atoms): the forward of the
# 4th step (30 GEMMs) and the backward of its last layer (16 GEMMs).  Reduced on the box; the .ncu-rep stays there.
O=gpurun_out/wide; mkdir -p $O
timeout 300 python bench_yaml.py --batch 8 --steps 1 --skip-cpu > $O/plain.json 2> $O/plain.err || exit 1
timeout 800 ncu --set full --clock-control none --import-source on --kernel-name regex:tc_wide_gemm --launch-skip 330 --launch-count 46 \
    -f -o $O/wide python bench_yaml.py --batch 8 --steps 1 --skip-cpu > $O/ncu.log 2>&1
ncu -i $O/wide.ncu-rep --page raw --csv --print-units base > $O/wide_raw.csv 2>/dev/null
python scripts/summarise_wide.py $O/wide_raw.csv $O/r2_ncu_wide_gemm.csv
# the heaviest forward kernel (ROWDOT, x2: 128 x 256 tiles, K = 1024) and the heaviest backward one: details pages
ncu -i $O/wide.ncu-rep --kernel-name "regex:tc_wide_gemm" --launch-skip 3 --launch-count 1 --page details > $O/r2_ncu_wide_x2_fwd_details.txt 2>/dev/null
rm -f $O/wide.ncu-rep $O/wide_raw.csv
cat $O/r2_ncu_wide_gemm.csv
