#!/bin/bash
# ncu evidence for the one-kernel suggest build (run under gpurun, after the plain run exited 0)
cd "$(dirname "$0")/.."
CMD="python bench.py --steps 2 --warmup 1 --history 12500000 --no-cpu-baseline --e2e-steps 0"
$CMD > gpurun_out/r1h_plain.json 2> gpurun_out/r1h_plain.err || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r1h.csv $CMD > gpurun_out/r1h_ncu1.log 2>&1
echo "launch list rc=$?"
