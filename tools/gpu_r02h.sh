#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python tools/scan_lab.py --steps 12 --rounds 14 --configs "base:;dense:LMI_SCAN_FLAGS=16;r32k:LMI_RANGE_ROWS=32768;r32k_dense:LMI_RANGE_ROWS=32768,LMI_SCAN_FLAGS=16;r49k_dense:LMI_RANGE_ROWS=49152,LMI_SCAN_FLAGS=16" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
