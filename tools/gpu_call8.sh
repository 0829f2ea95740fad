#!/bin/bash
for r in 0 1 2 3 4 5 6 7; do timeout 120 python tools/kbench.py --cases car4k --world 8 --rank $r 2>&1 | tail -1; done
for r in 0 1 2 3; do timeout 120 python tools/kbench.py --cases car4k --world 4 --rank $r 2>&1 | tail -1; done
