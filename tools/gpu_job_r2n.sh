#!/bin/bash
mkdir -p gpurun_out/r2n
python -m pytest tests/test_gpu_parity.py tests/test_gpu_stages.py tests/test_gpu_variants.py -x -q -m gpu > gpurun_out/r2n/tests.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/r2n/tests.log
python bench.py --workload c2 --steps 20 --no-cpu-baseline > gpurun_out/r2n/c2.json 2> gpurun_out/r2n/c2.err
python -c "
import json
d = json.loads([x for x in open('gpurun_out/r2n/c2.json') if x.startswith('{')][-1])
print('c2 ms', d['ms_per_step'], d['latency'])
for k, v in d['stages'].items(): print(k, round(v['ms_per_step'], 4), v['launches_per_step'])
"
