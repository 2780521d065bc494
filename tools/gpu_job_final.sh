#!/bin/bash
# what the driver runs at round end: GPU tests, smoke, bench (N=1) -- on the current tree
mkdir -p gpurun_out/final_r3
python -m pytest tests -x -q -m gpu > gpurun_out/final_r3/gpu_tests.log 2>&1; echo "gpu tests exit $?"; tail -3 gpurun_out/final_r3/gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_r3/smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/final_r3/smoke.log
python bench.py > gpurun_out/final_r3/bench.json 2> gpurun_out/final_r3/bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_r3/bench_ref.json 2> gpurun_out/final_r3/bench_ref.err; echo "ref exit $?"
python - <<'PY'
import json
d = json.loads([x for x in open('gpurun_out/final_r3/bench.json') if x.startswith('{')][-1])
r = json.loads([x for x in open('gpurun_out/final_r3/bench_ref.json') if x.startswith('{')][-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'pcm16', d['e2e_pcm16']['value'], 'frac', d['roofline']['frac'], d['clocks'])
print('ref', r['value'], r['cpu_baseline']['cores'], 'same config', r['config'] == d['config'])
PY
