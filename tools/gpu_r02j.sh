#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r02j.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r02j.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-workload-c --no-cpu-baseline > gpurun_out/bench_r02j.json 2> gpurun_out/bench_r02j.err; echo "rc=$?" >> gpurun_out/bench_r02j.err
timeout 600 python tools/sweeps.py --probes 4 --batches 1,10,100,1000 > gpurun_out/sweeps_r02j.jsonl 2> gpurun_out/sweeps_r02j.err
