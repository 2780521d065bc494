mkdir -p gpurun_out
python tools/debug_transparency.py > gpurun_out/r2b_transparency.log 2>&1
python tests/accuracy_study.py gemm > gpurun_out/r2b_gemm.log 2>&1
python tests/accuracy_study.py stages > gpurun_out/r2b_stages.log 2>&1
python tests/accuracy_study.py e2e --n 32 > gpurun_out/r2b_e2e.log 2>&1
python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2b_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r2b_tests.log
python bench.py > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
MIO_CHUNK_ACC=off python bench.py > gpurun_out/r2b_bench_unchunked.json 2> gpurun_out/r2b_bench_unchunked.err
MIO_CHUNK_ACC=all python bench.py > gpurun_out/r2b_bench_chunkall.json 2> gpurun_out/r2b_bench_chunkall.err
tail -4 gpurun_out/r2b_tests.log; cat gpurun_out/r2b_e2e.log
