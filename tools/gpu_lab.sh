#!/bin/bash
# quick GPU pass: the scan-impl parity tests, then the interleaved scan lab
# usage: tools/gpu_lab.sh "<scan_lab --configs string>" [pytest -k expression]
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "${2:-golden or medium or cluster_pair}" > gpurun_out/pytest_lab.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_lab.log
timeout 600 python tools/scan_lab.py --steps 12 --rounds 6 --configs "$1" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
