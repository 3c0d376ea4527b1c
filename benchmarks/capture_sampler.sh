#!/bin/bash
# ncu evidence for the sampler variants (run under gpurun; profiles only after the plain run exited 0)
cd "$(dirname "$0")/.."
TAG=${1:-r2}
CMD="python benchmarks/prof_sampler.py 1"
$CMD > gpurun_out/${TAG}_sampler_plain.json 2> gpurun_out/${TAG}_sampler_plain.err || { echo "plain run failed"; exit 1; }
cat gpurun_out/${TAG}_sampler_plain.json
ncu --set full --clock-control none -k regex:"sample_kernel" -c 7 --import-source on -f -o gpurun_out/prof_${TAG}_sampler $CMD > gpurun_out/${TAG}_sampler_ncu.log 2>&1
echo "full capture rc=$?"
