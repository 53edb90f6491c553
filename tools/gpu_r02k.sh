#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/scan_lab.py --steps 12 --rounds 12 --configs "base:;sib_start:LMI_SCAN_FLAGS=2" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
LMI_SCAN_FLAGS=2 timeout 600 ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:scan_topk_tc -s 4 -c 2 --csv --log-file gpurun_out/sib_dram.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/ncu_sib.log 2>&1
