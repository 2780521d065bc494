mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r2a_env.log 2>&1
lscpu | head -20 >> gpurun_out/r2a_env.log
python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2a_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r2a_tests.log
python tests/accuracy_study.py gemm > gpurun_out/r2a_gemm.log 2>&1
python tests/accuracy_study.py stages > gpurun_out/r2a_stages.log 2>&1
python tests/accuracy_study.py e2e --n 32 > gpurun_out/r2a_e2e.log 2>&1
python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/r2a_sanitizer_$tool.log python tools/sanitize_run.py > gpurun_out/r2a_sanitizer_$tool.out 2>&1; echo "exit $?" >> gpurun_out/r2a_sanitizer_$tool.out
done
tail -5 gpurun_out/r2a_tests.log; tail -3 gpurun_out/r2a_e2e.log
