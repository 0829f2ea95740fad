#!/bin/bash
echo "--- default"; timeout 200 python tools/kbench.py --cases car4k,car800,monkey800,soup4k 2>&1 | tail -1
for v in on l2 l2r1 l2r3 l2r4 l2on l2onr1 l2onr3; do
  echo "--- $v"; RT_B200_LIB=build_var/$v.so timeout 200 python tools/kbench.py --cases car4k,car800,monkey800,soup4k 2>&1 | tail -1
done
