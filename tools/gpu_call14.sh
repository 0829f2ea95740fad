#!/bin/bash
mkdir -p gpurun_out
for r in 0 1 2 3 4 5 6 7; do timeout 120 python tools/kbench.py --cases car4k --world 8 --rank $r 2>&1 | tail -1; done
for r in 0 1; do timeout 120 python tools/kbench.py --cases car4k --world 2 --rank $r 2>&1 | tail -1; done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests14.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/gpu_tests14.log
