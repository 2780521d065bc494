#!/bin/bash
# round 3, second call: the ADVICE fixes (stream detach on ctx destroy, emit buffers, error paths, BF16 message) under the whole GPU suite
mkdir -p gpurun_out/r3b
python -m pytest tests -x -q -m gpu > gpurun_out/r3b/gpu_tests.log 2>&1; echo "gpu tests exit $?"; tail -8 gpurun_out/r3b/gpu_tests.log
