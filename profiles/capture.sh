#!/bin/bash
# usage (on the GPU box, under gpurun): profiles/capture.sh <tag> [bench args...]
# plain run first (must exit 0), then one `ncu --set full` launch of the specialised kernel; exports the raw / sass pages as CSV
# into gpurun_out/ (the .ncu-rep itself only when it is small enough to travel).
tag=$1; shift
python bench.py --no-cpu-baseline --no-workloads --sustain 0 --steps 2 --warmup 3 "$@" > gpurun_out/plain_$tag.log 2>&1 || { echo "plain run failed"; exit 1; }
OMO_JIT_DUMP=gpurun_out/jit_$tag ncu --set full --clock-control none --import-source on -k regex:omo_jit_kernel -s 3 -c 1 -f -o /tmp/prof_$tag \
    python bench.py --no-cpu-baseline --no-workloads --sustain 0 --steps 2 --warmup 3 "$@" > gpurun_out/ncu_$tag.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/prof_$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv
ncu -i /tmp/prof_$tag.ncu-rep --page source --csv --print-source sass > gpurun_out/${tag}_sass.csv
S=$(stat -c %s /tmp/prof_$tag.ncu-rep); if [ "$S" -lt 20000000 ]; then cp /tmp/prof_$tag.ncu-rep gpurun_out/; fi
ls -la gpurun_out | grep $tag
