#!/bin/bash
# round 3, 2-GPU call: the two-devices-in-one-process test (skipped on 1-GPU boxes) and the N = 2 bench (P2P gather with the explicit stream ordering)
mkdir -p gpurun_out/r3f
nvidia-smi -L > gpurun_out/r3f/gpus.txt
python -m pytest tests/test_gpu_robustness.py -x -q -m gpu -k "two_devices or two_contexts" -rA > gpurun_out/r3f/two_devices.log 2>&1; echo "two devices exit $?"; tail -6 gpurun_out/r3f/two_devices.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r3f/bench_n2.json 2> gpurun_out/r3f/bench_n2.err; echo "bench n2 exit $?"; tail -c 900 gpurun_out/r3f/bench_n2.json
