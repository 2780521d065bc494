#!/bin/bash
# PDL A/B (C2 latency, C3 throughput), parity suite with PDL on, then the full default bench line
mkdir -p gpurun_out/r2h
for pdl in 0 1; do
  MIO_PDL=$pdl python bench.py --workload c2 --steps 20 --no-cpu-baseline > gpurun_out/r2h/c2_pdl$pdl.json 2> gpurun_out/r2h/c2_pdl$pdl.err
  MIO_PDL=$pdl python bench.py --no-extra --no-cpu-baseline > gpurun_out/r2h/c3_pdl$pdl.json 2> gpurun_out/r2h/c3_pdl$pdl.err
done
python - <<'PY'
import json
for w in ('c2', 'c3'):
    for p in (0, 1):
        try:
            d = json.loads([x for x in open(f'gpurun_out/r2h/{w}_pdl{p}.json') if x.startswith('{')][-1])
            print(w, 'pdl', p, 'ms/step %.4f' % d['ms_per_step'], 'e2e ms %.3f' % d['e2e']['ms_per_step'], d.get('latency'), d['clocks'])
        except Exception as e:
            print(w, p, 'FAILED', e)
PY
MIO_PDL=1 python -m pytest tests/test_gpu_parity.py tests/test_gpu_stages.py tests/test_gpu_streaming.py -x -q -m gpu > gpurun_out/r2h/parity.log 2>&1
tail -3 gpurun_out/r2h/parity.log
bash tools/gpu_job_r2g.sh
