#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_full.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>&1; tail -c 600 gpurun_out/bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --no-extra --steps 2 --warmup 1 > gpurun_out/ncu_bench.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:ntt_fwd_p --launch-skip 4 -c 2 -o gpurun_out/full_bench_ntt -f python bench.py --no-extra --steps 2 --warmup 1 > gpurun_out/ncu_bench_full.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:ntt_inv_p --launch-skip 4 -c 2 -o gpurun_out/full_bench_intt -f python bench.py --no-extra --steps 2 --warmup 1 >> gpurun_out/ncu_bench_full.log 2>&1
ls -la gpurun_out | grep full_bench
