#!/bin/bash
for w in 1 2 4 8; do
echo "=== world $w"
timeout 120 python tools/kbench.py --cases car4k,car800 --world $w 2>&1 | tail -1
for v in hf4 hf6 hf8 hf6k4 hf4k3 k3 hs32 hs16; do
  RT_B200_LIB=build_var/sp_$v.so timeout 120 python tools/kbench.py --cases car4k,car800 --world $w 2>&1 | tail -1
done
done
