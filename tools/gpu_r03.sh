#!/bin/bash
# round-3 evidence pass on ONE GPU: the default invocations of both bench arms (the headline line now carries
# configs A, C, the budget / batch sweeps, the 2-level workload and the cold call), ncu launch list and
# ncu --set full of the scan at the same commit
TAG=${1:-r03}
mkdir -p gpurun_out
timeout 1500 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?" >> gpurun_out/bench_$TAG.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "rc=$?" >> gpurun_out/bench_ref_$TAG.err
if [ "$2" = "ncu" ]; then
KREGEX='regex:linear_|softmax_topn|invert_|scan_topk|merge_|cast_f16|nav_fused'
timeout 600 python bench.py --steps 2 --warmup 3 --only-headline > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREGEX" -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --only-headline > gpurun_out/ncu_list_$TAG.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_topk_tc -s 4 -c 1 -o gpurun_out/scan_$TAG python bench.py --steps 2 --warmup 3 --only-headline > gpurun_out/ncu_full_$TAG.log 2>&1
fi
tail -c 600 gpurun_out/bench_$TAG.err
echo done
