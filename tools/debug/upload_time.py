import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mitsuba_renderer_b200 import api, scenes
s = scenes.path_scene()
ctx = api.Context(0)
for k in range(3):
    t0 = time.perf_counter(); ctx.upload_scene(s, build=False); t1 = time.perf_counter(); ctx.build_accel(); t2 = time.perf_counter()
    print(f"upload (host copies) {1e3*(t1-t0):.1f} ms, build_accel {1e3*(t2-t1):.1f} ms (GPU part {ctx.accel_info()['buildMs']:.1f} ms)")
ctx.close()
