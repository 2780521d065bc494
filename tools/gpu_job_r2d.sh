mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2d_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r2d_tests.log
python __graft_entry__.py smoke > gpurun_out/r2d_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r2d_smoke.log
python tests/accuracy_study.py e2e --n 32 > gpurun_out/r2d_e2e.log 2>&1
python tests/accuracy_study.py stages > gpurun_out/r2d_stages.log 2>&1
python bench.py > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err
grep -E "passed|failed" gpurun_out/r2d_tests.log | tail -3; grep FAILED gpurun_out/r2d_tests.log; tail -3 gpurun_out/r2d_smoke.log; cat gpurun_out/r2d_e2e.log
