#!/bin/bash
cd /root/repo
timeout 120 python tools/ae_k1_once.py > gpurun_out/r2_t54_plain.log 2>&1 || exit 1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel" -s 2 -c 2 -o gpurun_out/r2_ae_unit -f python tools/ae_k1_once.py > gpurun_out/r2_t54_ncu.log 2>&1
