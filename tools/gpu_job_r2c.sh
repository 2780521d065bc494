mkdir -p gpurun_out
python tools/chunk_sweep.py 6 > gpurun_out/r2c_chunk_sweep.log 2>&1
cat gpurun_out/r2c_chunk_sweep.log
