#!/bin/bash
mkdir -p gpurun_out/r2f
python -m pytest tests/test_gpu_variants.py tests/test_batch_tts.py -q -m gpu -s > gpurun_out/r2f/variants.log 2>&1
echo "variants exit $?" >> gpurun_out/r2f/variants.log
grep -v "^miocodec\|^ggml\|loaded" gpurun_out/r2f/variants.log | grep -i "dB\|passed\|failed\|exit" | tail -40
