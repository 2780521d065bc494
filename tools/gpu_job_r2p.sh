#!/bin/bash
mkdir -p gpurun_out/r2p
python -m pytest tests/test_gpu_robustness.py tests/test_gpu_streaming.py tests/test_batch_tts.py -x -q -m gpu > gpurun_out/r2p/tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r2p/tests.log
