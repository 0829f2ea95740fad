#!/bin/bash
# evidence of a round on one B200 (run under gpurun): ncu --set full captures of the render kernel at 4K, 800x600 and as rank 0 of 8
# (read here with tools/ncu_summary.py, tools/ncu_traffic.py, tools/sass_lines.py), the launch list, the GPU test suite
mkdir -p gpurun_out
for cfg in "4k:--width 3840 --height 2160" "800:--width 800 --height 600" "4k_r0of8:--width 3840 --height 2160 --world 8 --rank 0"; do
  name=${cfg%%:*}; args=${cfg#*:}
  timeout 120 python tools/ncu_target.py $args > gpurun_out/target_$name.log 2>&1 &&
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -f -o gpurun_out/head_$name python tools/ncu_target.py $args > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
done
timeout 120 python tools/ncu_target.py --width 3840 --height 2160 --frames 3 > /dev/null 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_head.csv python tools/ncu_target.py --width 3840 --height 2160 --frames 3 > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q -s > gpurun_out/gpu_tests_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/gpu_tests_final.log
