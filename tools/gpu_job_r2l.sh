#!/bin/bash
mkdir -p gpurun_out/r2l
python -m pytest tests/test_gpu_streaming.py tests/test_batch_tts.py tests/test_gpu_dropin.py tests/test_gpu_robustness.py -x -q -m gpu > gpurun_out/r2l/stream_tests.log 2>&1
tail -3 gpurun_out/r2l/stream_tests.log
python bench.py > gpurun_out/r2l/bench_n1.json 2> gpurun_out/r2l/bench_n1.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads([x for x in open('gpurun_out/r2l/bench_n1.json') if x.startswith('{')][-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'e2e_pcm16', d['e2e_pcm16']['value'], d['clocks'])
print('c4', d['c4']['value'], d['c4']['seconds_per_session'], d['c4']['decoded_audio_s_per_s'], d['c4']['roofline']['frac'])
print('c2', d['c2']['device_ms'], d['c2']['latency'])
print('c5', d['c5']['value'], d['c5']['e2e'])
PY
