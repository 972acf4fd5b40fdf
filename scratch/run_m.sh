#!/bin/bash
# full GPU suite + smoke + both bench arms (state check after container re-creation)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/tests_m.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/tests_m.log
tail -5 gpurun_out/tests_m.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3 | tee gpurun_out/smoke_m.log
timeout 600 python bench.py > gpurun_out/bench_m.json 2> gpurun_out/bench_m.err; echo "bench rc=$?"; cat gpurun_out/bench_m.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_m.json 2> gpurun_out/bench_ref_m.err; cat gpurun_out/bench_ref_m.json
