#!/bin/bash
# call 2: pair walk — parity first, then variant timings
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_upstream.py -m gpu -x -q > gpurun_out/gpu_tests2.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/gpu_tests2.log
echo "--- in-tree (call, 8)"
timeout 300 python tools/kbench.py --cases car4k,car800,monkey800,soup4k 2>&1 | tail -1 | tee gpurun_out/kb_base.json
for v in w2_in_8 w2_in_6 w2_in_5 w2_in_4 w2_in_6_s8; do
  echo "--- $v"
  RT_B200_LIB=build_var/$v.so timeout 300 python tools/kbench.py --cases car4k,car800,monkey800,soup4k 2>&1 | tail -1 | tee gpurun_out/kb_$v.json
done
