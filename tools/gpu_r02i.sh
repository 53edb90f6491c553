#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r02i.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r02i.log
timeout 600 python bench.py --workload ml > gpurun_out/bench_ml_r02.json 2> gpurun_out/bench_ml_r02.err; echo "rc=$?" >> gpurun_out/bench_ml_r02.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r02.log 2>&1; echo "rc=$?" >> gpurun_out/smoke_r02.log
