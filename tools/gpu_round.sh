#!/bin/bash
# One GPU-box session: the GPU test suite, then the bench (plain), logs under gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider ${PYTEST_ARGS} > gpurun_out/${TAG:-r02}_tests.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/${TAG:-r02}_tests.log
tail -5 gpurun_out/${TAG:-r02}_tests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG:-r02}_bench_n1.json 2> gpurun_out/${TAG:-r02}_bench_n1.err
echo "bench rc=$?"
python - <<'PY'
import json, os
tag = os.environ.get('TAG', 'r02')
try:
    d = json.loads(open('gpurun_out/%s_bench_n1.json' % tag).read().strip().splitlines()[-1])
    print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'])
    print('roofline kernel', d['roofline']['kernel'], 'frac', d['roofline']['frac'], 'step frac', d['roofline']['step']['frac'])
    for k, v in d['roofline']['kernels'].items():
        print('  %-16s %.4f ms/step  frac %s' % (k, v['ms_per_step'], v.get('frac')))
except Exception as e:
    print('no bench line:', e)
PY
