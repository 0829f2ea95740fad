#!/bin/bash
for v in lm4 lm6 lm3st2 lm3st4 lm3st8 lm4st6; do
  echo "--- $v"; RT_B200_LIB=build_var/$v.so timeout 200 python tools/kbench.py --cases car4k,car800,monkey800,soup4k 2>&1 | tail -1
  echo "--- w8 $v"; RT_B200_LIB=build_var/$v.so timeout 200 python tools/kbench.py --cases car4k --world 8 2>&1 | tail -1
done
