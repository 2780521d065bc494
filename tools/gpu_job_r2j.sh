#!/bin/bash
mkdir -p gpurun_out/r2j
python -m pytest tests/test_gpu_parity.py -q -m gpu -s 2>&1 | grep -v "^miocodec\|^ggml\|loaded" > gpurun_out/r2j/parity_s.log
grep -i "dB" gpurun_out/r2j/parity_s.log | head -60; tail -2 gpurun_out/r2j/parity_s.log
