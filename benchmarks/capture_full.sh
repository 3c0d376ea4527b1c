#!/bin/bash
# ncu --set full of the step's kernels (suggest, dedup passes) in one short bench run
cd "$(dirname "$0")/.."
TAG=${1:-r2}
CMD="python bench.py --steps 2 --warmup 1 --history 12500000 --no-cpu-baseline --e2e-steps 0 --sustain 0 --no-extra"
$CMD > gpurun_out/${TAG}_plain2.json 2> gpurun_out/${TAG}_plain2.err || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"sample_kernel|set_dedup" -s 4 -c 3 -f -o gpurun_out/prof_${TAG} $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "full capture rc=$?"
