set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 500 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 300 gpurun_out/bench_n1.json
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 400 gpurun_out/bench_reference.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; tail -c 200 gpurun_out/ncu_bench.log
timeout 300 ncu --set full --clock-control none --import-source on -k "regex:k_accum|k_finalize_cells" --launch-skip 2 -c 2 -f -o gpurun_out/r01_full python tools/profile_step.py --steps 3 > gpurun_out/ncu_full.log 2>&1; tail -2 gpurun_out/ncu_full.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/decode_launches.csv python tools/decode_bench.py --reps 1 --cpu-series 8 > gpurun_out/ncu_decode.log 2>&1; tail -c 150 gpurun_out/ncu_decode.log
KGS_PARSE_PIECE_KB=2000000 timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_mx_ --launch-skip 14 -c 7 -f -o gpurun_out/r01_decode python tools/decode_bench.py --series 2000 --reps 1 --cpu-series 4 > gpurun_out/ncu_decode_full.log 2>&1; tail -1 gpurun_out/ncu_decode_full.log | cut -c1-120
