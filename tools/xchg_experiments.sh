#!/bin/bash
# Where does the step time at N > 1 go?  bench.py at N GPUs with parts of the exchange switched off (KGS_XCHG_DEBUG).
N=${N:-2}
for dbg in 0 1 2 4 7; do
  KGS_XCHG_DEBUG=$dbg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$dbg bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/xd_$dbg.json 2> gpurun_out/xd_$dbg.err
  python -c "
import json; d=json.loads(open('gpurun_out/xd_$dbg.json').read().strip().splitlines()[-1]); print('dbg $dbg ms_per_step', round(d['ms_per_step'],4))"
done
