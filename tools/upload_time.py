"""Host + device latency of rtc_scene_upload for the room scene (13 shapes): the part of every end-to-end step that is not the
render call.  usage: python tools/upload_time.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import rtc_b200 as rtc
r = rtc.Renderer([0])
flat = rtc.scenes.reflect_refract_scene().flatten()
for _ in range(5): r.upload_flat(*flat)
ts = []
for _ in range(200):
    a = time.perf_counter(); r.upload_flat(*flat); ts.append(time.perf_counter() - a)
ts.sort()
print("rtc_scene_upload (13 shapes): min %.1f us median %.1f us p90 %.1f us" % (ts[0] * 1e6, ts[100] * 1e6, ts[180] * 1e6))
