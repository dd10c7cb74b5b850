#!/bin/bash
# bench.py (N=1, no CPU legs) under environment knobs of the library
run() { env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/knob.json 2> gpurun_out/knob.err
  python -c "
import json,sys; d=json.loads(open('gpurun_out/knob.json').read().strip().splitlines()[-1]); k=d['roofline']['kernels']; print(' '.join(sys.argv[1:]), 'ms_per_step', round(d['ms_per_step'],4), {n: round(v['ms_per_step'],4) for n,v in k.items()})" "$@" || tail -2 gpurun_out/knob.err; }
run KGS_ACCUM_WAVES=1
run KGS_ACCUM_WAVES=2
run KGS_ACCUM_WAVES=3
run KGS_ACCUM_WAVES=4
run KGS_ACCUM_TMA=1
run KGS_LIB_PATH=$PWD/kube_gpu_stats_b200/libkgs_nt64.so
run KGS_LIB_PATH=$PWD/kube_gpu_stats_b200/libkgs_nt256.so
