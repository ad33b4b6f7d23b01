#!/usr/bin/env python
"""Queue lengths per bounce of a C4 render at a given spp (the wavefront loop keeps them on the device): renders with
maxDepth = 1, 2, ... and differences the ray statistics. Used to attribute an ncu capture of "the k-th closest-hit launch" to
a ray count (profiles/r02_kernel_traffic.json).   python tools/bounce_counts.py [spp] [bounces]"""
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mitsuba_renderer_b200 import api, scenes  # noqa: E402


def main():
    spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    bounces = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    s = scenes.path_scene(); s.spp = spp
    ctx = api.Context(0)
    ctx.upload_scene(s)
    prev_n = prev_s = 0
    out = []
    for d in range(1, bounces + 2):
        sd = copy.copy(s); sd.max_depth = d
        _, st = ctx.render(api.Context.params_from_scene(sd, seed=1))
        out.append({"max_depth": d, "closest_rays_of_bounce": d - 1, "closest_rays": st["normalRays"] - prev_n,
                    "shadow_rays_of_bounce": d - 2, "shadow_rays": st["shadowRays"] - prev_s})
        prev_n, prev_s = st["normalRays"], st["shadowRays"]
    print(json.dumps({"workload": "C4", "spp": spp, "seed": 1, "per_bounce": out}))
    ctx.close()


if __name__ == "__main__":
    main()
