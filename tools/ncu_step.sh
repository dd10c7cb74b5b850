#!/bin/bash
# ncu --set full on the kernels of one bench step (configs[1] resident), after a plain run of the same command.
TAG=${TAG:-r02}
mkdir -p gpurun_out
python tools/profile_step.py --steps 3 ${PROF_ARGS} > gpurun_out/${TAG}_prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:${KREGEX:-k_accum|k_finalize}" --launch-skip ${SKIP:-4} -c ${COUNT:-4} -f \
    -o gpurun_out/${TAG}_full python tools/profile_step.py --steps 3 ${PROF_ARGS} > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/${TAG}_ncu_full.log; cat gpurun_out/${TAG}_prof_plain.log
