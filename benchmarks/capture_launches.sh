#!/bin/bash
# ncu launch list of one short bench run (run under gpurun; profiles only after the plain run exited 0)
cd "$(dirname "$0")/.."
TAG=${1:-r2}
CMD="python bench.py --steps 2 --warmup 1 --history 12500000 --no-cpu-baseline --e2e-steps 0 --sustain 0"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "launch list rc=$?"
