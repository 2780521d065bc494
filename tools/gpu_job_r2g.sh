#!/bin/bash
mkdir -p gpurun_out/r2g
( time python bench.py ) > gpurun_out/r2g/bench_n1.json 2> gpurun_out/r2g/bench_n1.err
echo "exit $?" >> gpurun_out/r2g/bench_n1.err
tail -5 gpurun_out/r2g/bench_n1.err
python - <<'PY'
import json
l = [x for x in open('gpurun_out/r2g/bench_n1.json') if x.startswith('{')]
d = json.loads(l[-1])
for k in ('value', 'ms_per_step', 'e2e', 'c5', 'c2', 'c4', 'cpu_baseline'):
    print(k, json.dumps(d.get(k))[:1500])
PY
