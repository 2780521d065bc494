#!/bin/bash
# round 3, first call: the no-allocation entry point alone in a fresh process, then the whole GPU suite (it runs last there),
# then a default bench line as the round's starting point
mkdir -p gpurun_out/r3a
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r3a/noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -5 gpurun_out/r3a/noalloc_alone.log
python -m pytest tests -x -q -m gpu --durations=15 > gpurun_out/r3a/gpu_tests.log 2>&1; echo "gpu tests exit $?"; tail -25 gpurun_out/r3a/gpu_tests.log
python bench.py --no-extra --steps 20 --warmup 3 > gpurun_out/r3a/bench_noextra.json 2> gpurun_out/r3a/bench_noextra.err; echo "bench exit $?"; tail -c 1500 gpurun_out/r3a/bench_noextra.json
