#!/bin/bash
# bench.py (N=1, no CPU legs) with alternative builds of the library: KGS_LIB_PATH
for so in "$@"; do
  KGS_LIB_PATH=$PWD/kube_gpu_stats_b200/$so timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/var_$so.json 2> gpurun_out/var_$so.err
  python -c "
import json; d=json.loads(open('gpurun_out/var_$so.json').read().strip().splitlines()[-1]); k=d['roofline']['kernels']; print('$so ms_per_step', round(d['ms_per_step'],4), {n: round(v['ms_per_step'],4) for n,v in k.items()})" || tail -2 gpurun_out/var_$so.err
done
