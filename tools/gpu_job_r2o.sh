#!/bin/bash
mkdir -p gpurun_out/r2o
python -m pytest tests/test_gpu_parity.py tests/test_gpu_stages.py tests/test_gpu_streaming.py tests/test_gpu_variants.py -x -q -m gpu > gpurun_out/r2o/tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r2o/tests.log
for f in 0 1; do
  MIO_FUSE_TAIL=$f python bench.py --no-extra --no-cpu-baseline > gpurun_out/r2o/c3_fuse$f.json 2> gpurun_out/r2o/c3_fuse$f.err
done
python - <<'PY'
import json
for f in (0, 1):
    d = json.loads([x for x in open(f'gpurun_out/r2o/c3_fuse{f}.json') if x.startswith('{')][-1])
    st = d['stages']
    print('fuse', f, 'ms/step %.3f' % d['ms_per_step'], 'launches', d['gpu_launches'], 'split', st.get('split', {}).get('ms_per_step'), 'gemm_f16', st['gemm_f16']['ms_per_step'], d['clocks'])
PY
