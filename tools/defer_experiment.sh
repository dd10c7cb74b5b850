#!/bin/bash
# Exchange copies queued at once (KGS_PUSH_DEFER=0) against behind the next accumulate kernel (1): ms per step at N GPUs
N=${N:-4}
timeout 400 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_window_stats.py -k "multirank or shards or packed" -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/defer_test.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/defer_test.log
for d in 0 1; do
  KGS_PUSH_DEFER=$d timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2957$d bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/defer_${N}_$d.json 2> gpurun_out/defer_${N}_$d.err
  python -c "
import json; d=json.loads(open('gpurun_out/defer_${N}_$d.json').read().strip().splitlines()[-1]); print('N=$N defer $d ms_per_step', round(d['ms_per_step'],4), 'value', round(d['value']/1e9,1))" || tail -3 gpurun_out/defer_${N}_$d.err
done
