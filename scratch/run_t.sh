#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/tests_t.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/tests_t.log
tail -4 gpurun_out/tests_t.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; echo "bench rc=$?"; python - <<'PY'
import json
l=json.load(open('gpurun_out/bench_t.json'))
print({k:l[k] for k in ('value','ms_per_step','gpu_launches')}, l['roofline']['frac'], l['e2e']['value'])
print(json.dumps(l['extra'].get('resnet20_layers'))[:900])
print(json.dumps(l['extra'].get('bootstrap'))[:600])
PY
tail -3 gpurun_out/bench_t.err
