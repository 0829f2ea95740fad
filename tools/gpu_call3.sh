#!/bin/bash
mkdir -p gpurun_out
for v in w2_in_6 w2_in_8; do
export RT_B200_LIB=build_var/$v.so
python tools/ncu_target.py --width 3840 --height 2160 > gpurun_out/target_$v.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -f -o gpurun_out/prof_$v python tools/ncu_target.py --width 3840 --height 2160 > gpurun_out/ncu_$v.log 2>&1
echo "ncu $v rc=$?"
done
