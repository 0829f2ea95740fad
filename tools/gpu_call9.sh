#!/bin/bash
mkdir -p gpurun_out
echo "--- default"; timeout 200 python tools/kbench.py --cases car4k,car800,soup4k 2>&1 | tail -1
echo "--- default + RT_L2_WINDOW=1"; RT_L2_WINDOW=1 timeout 200 python tools/kbench.py --cases car4k,car800,soup4k 2>&1 | tail -1
for v in top_256 top_1024 mo_opq mo_lr1 mo_opqlr1 mo_s8; do
  echo "--- $v"; RT_B200_LIB=build_var/$v.so timeout 200 python tools/kbench.py --cases car4k,car800,soup4k 2>&1 | tail -1
done
