#!/bin/bash
cd /root/repo
for lib in base ""; do
  echo "== ECHO_B200_LIB=$lib" >> gpurun_out/r2_t55.log
  ECHO_B200_LIB=$lib timeout 300 python tools/ae_unit_probe.py 2>&1 | grep "k7\|k1 full\|k1 no snake\|k1 bf16" >> gpurun_out/r2_t55.log
done
