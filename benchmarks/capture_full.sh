#!/bin/bash
cd "$(dirname "$0")/.."
CMD="python bench.py --steps 2 --warmup 1 --history 12500000 --no-cpu-baseline --e2e-steps 0"
$CMD > gpurun_out/r1h_plain2.json 2> gpurun_out/r1h_plain2.err || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"sample_kernel|set_dedup" -s 6 -c 3 -f -o gpurun_out/prof_r1h $CMD > gpurun_out/r1h_ncu2.log 2>&1
echo "full capture rc=$?"
