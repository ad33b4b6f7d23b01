import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
from mitsuba_renderer_b200 import api, scenes
for name, s in (("c2", scenes.microbench_scene()), ("c4", scenes.path_scene())):
    s.spp = 4
    c = api.Context(0); c.upload_scene(s)
    c.set_counters(True)
    rays = np.concatenate([scenes.coherent_rays(s, 512), scenes.incoherent_rays(s, 500000, 5)]).astype(np.float32)
    c.trace(rays); c.trace(scenes.shadow_segments(rays, s), api.TRACE_ANY)
    a = c.counters()["maxStackEntries"]
    if name == "c4":
        c.render(api.Context.params_from_scene(s, seed=1))
    b = c.counters()["maxStackEntries"]
    print(name, "depth", c.accel_info()["depth"], "deepest stack: ray queries", a, "after render", b, flush=True)
    c.close()
