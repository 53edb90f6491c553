#!/bin/bash
# round-2 GPU pass: energy ablations of the scan kernel, then the ncu evidence (launch list, scan --set full,
# fused navigation kernel --set full)
mkdir -p gpurun_out
timeout 900 python tools/scan_ablate.py --ablate 0,2,17,19,32,34,0 --impls 0 --steps 150 --out gpurun_out/scan_ablate.jsonl > gpurun_out/scan_ablate.log 2>&1; echo "rc=$?" >> gpurun_out/scan_ablate.log
KREGEX='regex:linear_|softmax_topn|invert_|scan_topk|merge_|cast_f16|nav_fused'
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/plain_r02.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREGEX" -c 400 --csv --log-file gpurun_out/launches_r02.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/ncu_list_r02.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_topk_tc -s 4 -c 1 -o gpurun_out/scan_r02 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/ncu_full_r02.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nav_fused -s 2 -c 1 -o gpurun_out/nav_r02 python tools/nav_lab.py --nq 10000 --once fused > gpurun_out/ncu_nav_r02.log 2>&1
echo done
