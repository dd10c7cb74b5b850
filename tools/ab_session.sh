#!/bin/bash
# GPU test suite, then bench.py (N=1, no CPU legs) under the environment settings given as arguments (one run each)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/ab_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/ab_tests.log
for kv in "$@"; do
  env $kv timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab.json 2> gpurun_out/ab.err
  python -c "
import json,sys; d=json.loads(open('gpurun_out/ab.json').read().strip().splitlines()[-1]); k=d['roofline']['kernels']; print(sys.argv[1], 'ms_per_step', round(d['ms_per_step'],4), {n: round(v['ms_per_step'],4) for n,v in k.items()})" "$kv" || tail -2 gpurun_out/ab.err
done
