#!/bin/bash
# ncu --set full of one search step's non-scan kernels (MLP Linear layers on tcgen05, softmax/top-n,
# inversion, merge) on the default bench workload, and of the multi-level navigation kernels.
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_small.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on \
  -k 'regex:linear_tc_kernel|softmax_topn|invert_|merge_topk|split_tf32' -s 60 -c 16 -o gpurun_out/small_kernels \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_small.log 2>&1
timeout 300 python bench.py --workload ml --n 200000 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/plain_ml.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on \
  -k 'regex:navigate_tree|softmax_segments|linear_tc_kernel' -s 30 -c 8 -o gpurun_out/nav_kernels \
  python bench.py --workload ml --n 200000 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/ncu_nav.log 2>&1
echo done
