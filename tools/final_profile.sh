#!/bin/bash
# Round-2 evidence session on ONE GPU: test suite, smoke, bench (both arms), launch list, ncu --set full of the step kernels.
# Small text results go to profiles/ (tracked), the .ncu-rep to gpurun_out/ (summarised here afterwards).
set -x
mkdir -p gpurun_out profiles
T=${TAG:-r02}
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_tests.log; tail -4 gpurun_out/${T}_tests.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"; tail -c 400 gpurun_out/${T}_bench_n1.json
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; tail -c 300 gpurun_out/${T}_bench_reference.json
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --min-timed-ms 1 > gpurun_out/${T}_plain_short.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --min-timed-ms 1 > gpurun_out/${T}_ncu_bench.log 2>&1; tail -c 200 gpurun_out/${T}_ncu_bench.log
timeout 200 python tools/profile_step.py --steps 3 > gpurun_out/${T}_prof_plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k "regex:k_accum|k_finalize" --launch-skip 4 -c 4 -f -o gpurun_out/${T}_full python tools/profile_step.py --steps 3 > gpurun_out/${T}_ncu_full.log 2>&1; tail -2 gpurun_out/${T}_ncu_full.log; cat gpurun_out/${T}_prof_plain.log
