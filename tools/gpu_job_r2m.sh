#!/bin/bash
# N=2: default bench line with the c5 block (NCCL gather to rank 0, shared pinned buffer)
mkdir -p gpurun_out/r2i
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2i/bench_n8.json 2> gpurun_out/r2i/bench_n8.err
echo "exit $?" >> gpurun_out/r2i/bench_n8.err
tail -15 gpurun_out/r2i/bench_n8.err
python - <<'PY'
import json
l = [x for x in open('gpurun_out/r2i/bench_n8.json') if x.startswith('{')]
d = json.loads(l[-1])
for k in ('value', 'ms_per_step', 'e2e', 'c5'):
    print(k, json.dumps(d.get(k))[:2500])
PY
