#!/usr/bin/env python3
"""Generates tests/golden/camera_traces.json from the REFERENCE'S OWN Camera class compiled in place
(oracle/_ref/libref_camera.so, built by oracle/build_ref.py from /root/reference/src/camera.hpp): for a few seeded input
traces (LookAt, keyboard, mouse, scroll, shift — what processInput / mouse_callback feed the camera, main.cpp:509-551)
the 80 serializeCamera bytes (main.cpp:806-816) after every event, as hex.  Run where /root/reference is present:
    python tests/golden/make_camera_golden.py"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from test_camera import RefCamera, make_trace, play   # noqa: E402

out = {"about": "serializeCamera bytes of the reference Camera class (g++ build, GLM 0.9.8.5 vendored upstream) after every event of make_trace(seed)",
       "traces": []}
for seed in range(6):
    start, aspect, events = make_trace(seed, 60)
    frames = play(RefCamera(start, aspect), events)
    out["traces"].append({"seed": seed, "n": len(events), "flat_hex": [f.tobytes().hex() for f in frames]})
with open(os.path.join(HERE, "camera_traces.json"), "w") as f:
    json.dump(out, f)
print("wrote camera_traces.json:", sum(t["n"] for t in out["traces"]), "events")
