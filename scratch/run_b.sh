#!/bin/bash
mkdir -p gpurun_out
cap() { # name regex skip
  timeout 400 ncu --profile-from-start off --set full --import-source on --clock-control none --kernel-name-base demangled -k "regex:$2" --launch-skip $3 -c 1 -o gpurun_out/full_$1 -f python tools/boot_bench.py --mode synthetic --iters 1 --profile > gpurun_out/ncu_full_$1.log 2>&1
  echo "$1 rc=$?"; tail -2 gpurun_out/ncu_full_$1.log
}
cap p2_1 'ntt_fwd_p2<.int.1>' 30
cap p2_0 'ntt_fwd_p2<.int.0>' 10
cap fast_rot 'fast_rot_tma_kernel<.bool.1, .bool.0>' 40
cap bconv10 'bconv_kernel<.int.10>' 30
cap p1 'ntt_fwd_p1<.int.4>' 60
ls -la gpurun_out | grep ncu-rep
