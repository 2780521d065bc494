#!/bin/bash
# round 3, third call: (1) compute-sanitizer attempt (evidence either way), (2) interleaved A/B: HEAD / chunk-acc off / r1 slots /
# norms skipped / the round-1 tree
mkdir -p gpurun_out/r3c
for tool in memcheck racecheck; do
  echo "\$ compute-sanitizer --tool $tool python tools/sanitize_run.py" > gpurun_out/r3c/sanitizer_$tool.log
  timeout 600 compute-sanitizer --tool $tool python tools/sanitize_run.py >> gpurun_out/r3c/sanitizer_$tool.log 2>&1
  echo "exit code $?" >> gpurun_out/r3c/sanitizer_$tool.log
  tail -5 gpurun_out/r3c/sanitizer_$tool.log
done
python tools/ab_interleaved.py 8 gpurun_out/r3c/ab.json > gpurun_out/r3c/ab.log 2>&1; echo "ab exit $?"; tail -8 gpurun_out/r3c/ab.log
