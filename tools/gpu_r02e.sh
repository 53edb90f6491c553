#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or medium or cluster_pair or invert" > gpurun_out/pytest_lab.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_lab.log
timeout 600 python tools/scan_lab.py --steps 12 --rounds 6 --configs "n64:;n128_split:LMI_SCAN_IMPL=4;n128:LMI_SCAN_IMPL=3" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
timeout 600 python tools/scan_ablate.py --ablate 0,8 --impls 0,4 --steps 150 --out gpurun_out/scan_ablate_z.jsonl > gpurun_out/scan_ablate_z.log 2>&1
