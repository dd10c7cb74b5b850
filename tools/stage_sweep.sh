#!/bin/bash
# Host threads copying a pageable answer into the pinned staging pieces: ms per 543 MB answer
for t in 1 2 4 8 12; do
  KGS_STAGE_THREADS=$t timeout 200 python tools/decode_bench.py --reps 1 --cpu-series 8 > gpurun_out/stage_$t.json 2>gpurun_out/stage_$t.err
  python -c "
import json; d=json.loads(open('gpurun_out/stage_$t.json').read().strip().splitlines()[-1]); print('threads $t pageable ms', d['gpu_ms_from_pageable_bytes'], 'pinned ms', d['gpu_ms_per_answer'])" || tail -2 gpurun_out/stage_$t.err
done
