#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/scan_lab.py --steps 12 --rounds 6 --configs "gate:;nogate:LMI_SCAN_SHARE_GATE=0;n128_gate:LMI_SCAN_IMPL=3" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
LMI_SCAN_SHARE_GATE=1 timeout 600 python tools/scan_ablate.py --ablate 0,2,8 --impls 0 --steps 150 --out gpurun_out/scan_ablate_gate.jsonl > gpurun_out/scan_ablate_gate.log 2>&1
