#!/bin/bash
# The library with the scan kernel's profiling switches compiled IN (-DLMI_PROFILE: LMI_SCAN_ABLATE, the
# MMA-warp cycle counters).  Results are WRONG when an ablation is set; the product library never has them.
# tools/scan_ablate.py points the package loader at it (the product package never loads it).
cd "$(dirname "$0")/../learnedmetricindex_b200/csrc" || exit 1
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -shared -DLMI_PROFILE \
  -o ../../tools/liblmi_b200_profile.so lmi_mlp.cu lmi_mlp_tc.cu lmi_invert.cu lmi_scan.cu lmi_merge.cu lmi_layout.cu lmi_navigate.cu lmi_nav_fused.cu -lcuda
