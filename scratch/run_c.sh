#!/bin/bash
mkdir -p gpurun_out
for v in "" _tpb128; do
  export B200FHE_LIB=$PWD/phantom-fhe-boot_b200/libb200fhe$v.so
  echo "== lib$v"
  timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_ref_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
  timeout 300 python tools/op_bench.py 2>&1 | tail -1 | tee gpurun_out/op_bench$v.json
  timeout 300 python tools/boot_bench.py --mode synthetic --iters 5 2>&1 | tail -1 | tee gpurun_out/boot_syn$v.json
  timeout 300 python bench.py --no-extra --steps 10 2>&1 | tail -1 | tee gpurun_out/bench$v.json
done
