#!/bin/bash
cd /root/repo
ECHO_B200_LIB=debug timeout 300 python tools/gemm_trace.py --case ae_k7,ae_k1,ae_k1_bf16 > gpurun_out/r2_t50.log 2>&1
