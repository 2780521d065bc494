#!/bin/bash
# round 2, job e: parametric upsampler + stored matrix types, then the whole GPU suite and a bench line
mkdir -p gpurun_out/r2e
python -m pytest tests/test_gpu_variants.py -x -q -m gpu -s > gpurun_out/r2e/variants.log 2>&1
echo "variants exit $?" >> gpurun_out/r2e/variants.log
python -m pytest tests -q -m gpu --deselect tests/test_gpu_variants.py > gpurun_out/r2e/suite.log 2>&1
echo "suite exit $?" >> gpurun_out/r2e/suite.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2e/bench.json 2> gpurun_out/r2e/bench.err
tail -5 gpurun_out/r2e/variants.log; tail -5 gpurun_out/r2e/suite.log; cat gpurun_out/r2e/bench.json
