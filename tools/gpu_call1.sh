#!/bin/bash
# round 2, call 1: baseline evidence on HEAD — L1/L2 read peaks, the GPU suite incl. the quoted-size parity tests, bench N=1,
# ncu --set full of the HEAD render kernel at 4K, at 800x600 and as rank 0 of 8.
set -x
mkdir -p gpurun_out
tools/l2_probe > gpurun_out/l2_peak.json 2> gpurun_out/l2_probe.err
python -m pytest tests -m gpu -x -q -s > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/gpu_tests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
for cfg in "4k:--width 3840 --height 2160" "800:--width 800 --height 600" "4k_r0of8:--width 3840 --height 2160 --world 8 --rank 0"; do
  name=${cfg%%:*}; args=${cfg#*:}
  python tools/ncu_target.py $args > gpurun_out/target_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -f -o gpurun_out/prof_$name python tools/ncu_target.py $args > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
done
ls -la gpurun_out
