#!/bin/bash
# round-2 GPU pass 1: parity tests, scan lab (impl 0 vs pair kernel), navigation lab, bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/scan_lab.py --steps 12 --rounds 6 --configs "${LAB_CONFIGS:-one_cta:;n128:LMI_SCAN_IMPL=3;pair:LMI_SCAN_IMPL=2}" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
timeout 300 python tools/nav_lab.py > gpurun_out/nav_lab.log 2>&1; echo "nav rc=$?" >> gpurun_out/nav_lab.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-workload-c > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench rc=$?" >> gpurun_out/bench_b.err
echo done
