#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_build.py tests/test_gpu_quoted_sizes.py -m gpu -x -q -s -k "k9 or animated or device_layout" > gpurun_out/gpu_tests16.log 2>&1; echo "pytest rc=$?"; grep -E "C5|passed|failed|Error" gpurun_out/gpu_tests16.log | tail -12
timeout 300 python - <<'PY'
import sys, json, torch
sys.path.insert(0, '.')
import bench, rtload
rt = rtload.load()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    print(json.dumps(bench.time_animated(torch, rt, s)))
PY
