import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mitsuba_renderer_b200 import api, scenes
s = scenes.path_scene()
ctx = api.Context(0)
for k in range(3):
    t0 = time.perf_counter(); ctx.upload_scene(s); dt = time.perf_counter() - t0
    print(f"upload + build {dt*1e3:.1f} ms (GPU build {ctx.accel_info()['buildMs']:.1f} ms)")
ctx.close()
