#!/bin/bash
# bench.py at N GPUs, launched as the driver launches it; result line -> gpurun_out/r02_bench_n<N>.json
N=${N:-2}; T=${TAG:-r02}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_n${N}.json 2> gpurun_out/${T}_bench_n${N}.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open('gpurun_out/${T}_bench_n${N}.json').read().strip().splitlines()[-1])
print('n_gpus', d['n_gpus'], 'ms_per_step', round(d['ms_per_step'],4), 'value', round(d['value']/1e9,1), 'e2e', round(d['e2e']['value']/1e9,2), 'copy-only GB/s', round(d['e2e']['copy_only']['aggregate_h2d_gbs'],1), d['clocks'])
PY
