#!/bin/bash
# What does the single multicast push cost per step against the unicast pushes?  (bytes checked by tests/test_gpu_multirank.py with KGS_TEST_MULTICAST=1)
N=${N:-2}
if [ "${WITH_TEST:-0}" = "1" ]; then
KGS_TEST_MULTICAST=1 timeout 300 python -m pytest tests/test_gpu_multirank.py -m gpu -q -s --tb=short -p no:cacheprovider > gpurun_out/mc_test.log 2>&1; echo "multicast test rc=$?"; grep -E "transports|passed|failed|Error|error" gpurun_out/mc_test.log | head -8
fi
for mc in 0 1; do
  KGS_EXCHANGE_MULTICAST=$mc timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$mc bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/mc_${N}_$mc.json 2> gpurun_out/mc_${N}_$mc.err
  python -c "
import json; d=json.loads(open('gpurun_out/mc_${N}_$mc.json').read().strip().splitlines()[-1]); print('N=$N multicast $mc ms_per_step', round(d['ms_per_step'],4), d['config']['collective'][:60])" || tail -3 gpurun_out/mc_${N}_$mc.err
done
