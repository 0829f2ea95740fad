#!/bin/bash
mkdir -p gpurun_out
for w in 1 2 4 8; do timeout 120 python tools/kbench.py --cases car4k,car800 --world $w 2>&1 | tail -1; done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests13.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/gpu_tests13.log
