#!/bin/bash
# tests + warm op timings + synthetic bootstrap + launch list + set-full captures
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/tests.log
tail -3 gpurun_out/tests.log
timeout 300 python tools/op_bench.py > gpurun_out/op_bench.json 2> gpurun_out/op_bench.err; cat gpurun_out/op_bench.json
timeout 300 python tools/boot_bench.py --mode synthetic --iters 5 > gpurun_out/boot_syn.json 2> gpurun_out/boot_syn.err; cat gpurun_out/boot_syn.json
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/boot_bench.py --mode synthetic --iters 1 --profile > gpurun_out/ncu_list.log 2>&1
for k in "ntt_fwd_p2<1>" "fast_rot_tma_kernel<1, 0>" "bconv_kernel<10>" "ntt_fwd_p1<4>" "inner_prod_kernel"; do
  tag=$(echo "$k" | tr -c 'a-zA-Z0-9' '_')
  timeout 400 ncu --profile-from-start off --set full --import-source on --clock-control none -k "regex:^$(echo "$k" | sed 's/[<>,]/./g')" --launch-skip 20 -c 1 -o gpurun_out/full_$tag -f python tools/boot_bench.py --mode synthetic --iters 1 --profile > gpurun_out/ncu_full_$tag.log 2>&1
  echo "$k rc=$?"
done
ls -la gpurun_out
