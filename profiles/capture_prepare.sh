#!/bin/bash
# usage (GPU box): profiles/capture_prepare.sh <tag>   -- one `ncu --set full` launch of prepare_kernel on the stick240 batch
tag=$1
ncu --set full --clock-control none --import-source on -k regex:prepare_kernel -s 6 -c 1 -f -o /tmp/prep_$tag \
    python profiles/exp_kernel.py stick240 3 > gpurun_out/ncu_prep_$tag.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/prep_$tag.ncu-rep --page raw --csv > gpurun_out/prep_${tag}_raw.csv
ncu -i /tmp/prep_$tag.ncu-rep --page source --csv --print-source sass > gpurun_out/prep_${tag}_sass.csv
