#!/bin/bash
cd /root/repo
timeout 300 python tools/ae_unit_probe.py > gpurun_out/r2_t49.log 2>&1
