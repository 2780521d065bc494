#!/bin/bash
mkdir -p gpurun_out/c2list
ncu --clock-control none --metrics gpu__time_duration.sum --csv --log-file gpurun_out/c2list/launches.csv python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/c2list/log.txt 2>&1
python tools/launch_summary.py gpurun_out/c2list/launches.csv 3 | head -45
