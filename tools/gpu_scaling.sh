#!/bin/bash
# One N-GPU session: bench.py at N GPUs (as the driver launches it), then BASELINE configs[3] at full size
# (10 000 nodes x 8 GPUs x 24 h = 6.91 G samples split over the N GPUs: STRONG scaling) with --check, and with
# SLIDING=1 also configs[4] (5-min windows every 10 s, p50/p99).  Text results go to gpurun_out/ as r02_*_n<N>.*
N=${N:-2}; T=${TAG:-r02}
mkdir -p gpurun_out
run() { timeout ${TMO:-900} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
if [ "${MULTIRANK_TEST:-0}" = "1" ]; then
  timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/${T}_multirank_n${N}.log 2>&1; echo "multirank rc=$?"; tail -3 gpurun_out/${T}_multirank_n${N}.log
fi
run 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_n${N}.json 2> gpurun_out/${T}_bench_n${N}.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/${T}_bench_n${N}.json').read().strip().splitlines()[-1])
    print('n_gpus', d['n_gpus'], 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['copy_only']['aggregate_h2d_gbs'], d['config']['exchange_bytes_per_cell'])
except Exception as e:
    print('no bench line:', e)
PY
run 29542 tools/run_cfg4.py --nodes $((10000 / N)) --base-nodes 10 --steps 3 --check > gpurun_out/${T}_cfg4_n${N}.log 2>&1; echo "cfg4 rc=$?"; grep -v "^W\|OMP\|\*\*\*" gpurun_out/${T}_cfg4_n${N}.log | tail -6
if [ "${SLIDING:-0}" = "1" ]; then
  run 29543 tools/run_cfg4.py --nodes $((10000 / N)) --base-nodes 10 --steps 3 --sliding --sliding-windows 8611 --check > gpurun_out/${T}_cfg5_n${N}.log 2>&1; echo "cfg5 rc=$?"; grep -v "^W\|OMP\|\*\*\*" gpurun_out/${T}_cfg5_n${N}.log | tail -6
fi
