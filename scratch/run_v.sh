#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_bootstrap_gpu.py -x -q -m gpu -k "setup_matches or batched" 2>&1 | tail -25
timeout 600 python - <<'PY' 2>&1 | tail -5
import sys, json
sys.path.insert(0, '.')
import torch, bench
import phantom_fhe_boot_b200 as fhe
print(json.dumps(bench.extras_keyswitch_chain(torch, fhe)))
PY
