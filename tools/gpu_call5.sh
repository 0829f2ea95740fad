#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -s > gpurun_out/gpu_tests5.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/gpu_tests5.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench5_n1.json 2> gpurun_out/bench5_n1.err; echo "bench rc=$?"
tail -3 gpurun_out/bench5_n1.err
