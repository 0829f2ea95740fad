#!/bin/bash
mkdir -p gpurun_out
for v in w3_ln_8 w3_ln_7 w3_ln_6 w2_in_7 w2_in_6_st; do
  echo "--- $v"
  RT_B200_LIB=build_var/$v.so timeout 300 python tools/kbench.py --cases car4k,car800,monkey800,soup4k 2>&1 | tail -1 | tee gpurun_out/kb_$v.json
done
