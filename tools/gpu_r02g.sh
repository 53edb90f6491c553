#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "shared_gate or cluster_pair" > gpurun_out/pytest_lab.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_lab.log
timeout 900 python tools/scan_lab.py --steps 12 --rounds 6 --configs "base:;dense_refresh:LMI_SCAN_FLAGS=16;r16k:LMI_RANGE_ROWS=16384;r32k:LMI_RANGE_ROWS=32768;tail4k:LMI_TAIL_ROWS=4096;r16k_dense:LMI_RANGE_ROWS=16384,LMI_SCAN_FLAGS=16" --out gpurun_out/scan_lab.jsonl > gpurun_out/scan_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/scan_lab.log
LMI_SCAN_IMPL=2 timeout 600 ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,l1tex__m_xbar2l1tex_read_bytes.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:scan_topk_pair -s 4 -c 2 --csv --log-file gpurun_out/pair_dram.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-workload-c > gpurun_out/ncu_pair.log 2>&1
