#!/bin/bash
# round-2 evidence pass on ONE GPU: parity tests, the default bench (headline + appended 100M point), the reference
# arm on configs B and A, our arm on config A, the budget / batch sweeps, ncu launch list, ncu --set full of the scan
# and of the one-launch navigation kernel
TAG=${1:-r02}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?" >> gpurun_out/bench_$TAG.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_b_$TAG.json 2> gpurun_out/bench_ref_b_$TAG.err; echo "rc=$?" >> gpurun_out/bench_ref_b_$TAG.err
timeout 600 python bench.py --workload a --steps 20 --warmup 5 > gpurun_out/bench_a_$TAG.json 2> gpurun_out/bench_a_$TAG.err; echo "rc=$?" >> gpurun_out/bench_a_$TAG.err
timeout 600 python bench.py --workload a --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_a_$TAG.json 2> gpurun_out/bench_ref_a_$TAG.err; echo "rc=$?" >> gpurun_out/bench_ref_a_$TAG.err
timeout 600 python tools/sweeps.py > gpurun_out/sweeps_$TAG.jsonl 2> gpurun_out/sweeps_$TAG.err
KREGEX='regex:linear_|softmax_topn|invert_|scan_topk|merge_|cast_f16|nav_fused'
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREGEX" -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/ncu_list_$TAG.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_topk_tc -s 4 -c 1 -o gpurun_out/scan_$TAG python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/ncu_full_$TAG.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nav_fused -s 1 -c 1 -o gpurun_out/nav_$TAG python tools/nav_lab.py --nq 10000 --once fused > gpurun_out/ncu_nav_$TAG.log 2>&1
echo done
