#!/bin/bash
# Multi-GPU session: the multi-rank parity test, then the bench at N GPUs (plain torchrun launch as the driver does).
N=${N:-2}; TAG=${TAG:-r02}
mkdir -p gpurun_out
if [ "${SKIP_TEST:-0}" != "1" ]; then
timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/${TAG}_multirank_n${N}.log 2>&1
echo "multirank rc=$?"; tail -15 gpurun_out/${TAG}_multirank_n${N}.log
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 ${BENCH_ARGS} > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err
echo "bench rc=$?"; tail -5 gpurun_out/${TAG}_bench_n${N}.err
python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/${TAG}_bench_n${N}.json').read().strip().splitlines()[-1])
    print('n_gpus', d['n_gpus'], 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'copy_only', d['e2e']['copy_only'])
    print(d['config']['collective'], d['config']['exchange_bytes_per_cell'])
except Exception as e:
    print('no bench line:', e)
PY
