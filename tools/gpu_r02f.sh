#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/scan_lab.py --steps 12 --rounds 6 --configs "n64:;n64_pf4:LMI_SCAN_FLAGS=8;n64_pf2:LMI_SCAN_FLAGS=4;split:LMI_SCAN_IMPL=4;split_pf2:LMI_SCAN_IMPL=4,LMI_SCAN_FLAGS=4;split_pf3:LMI_SCAN_IMPL=4,LMI_SCAN_FLAGS=6" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
