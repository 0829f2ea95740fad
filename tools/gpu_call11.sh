#!/bin/bash
echo "--- default"; timeout 200 python tools/kbench.py --cases car4k,car800,monkey800 2>&1 | tail -1
for v in sl2 sl3 sl6 sl8 fl4 fl6 fl12 fl4sl2; do
  echo "--- $v"; RT_B200_LIB=build_var/ly_$v.so timeout 200 python tools/kbench.py --cases car4k,car800,monkey800 2>&1 | tail -1
done
