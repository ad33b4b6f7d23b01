"""A small but structurally complete case for debugging runs (a bounds-checking build, printf tracing, a memory checker where one
is available): a 300-cell terrain with spheres (~ 200k triangles, so the SAH build runs many levels and the traversal kernels use
their shared leaf phase and both stack tiers), closest + any-hit queries and a short `path` render with roulette, then the
instancing / cylinder / bitmap-texture scenes.

    python tools/debug/sanitize_case.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np  # noqa: E402

from mitsuba_renderer_b200 import api, scenes  # noqa: E402


def main():
    s = scenes.path_scene(terrain_cells=300, n_spheres=60, width=96, height=54, spp=4)
    ctx = api.Context(0)
    ctx.upload_scene(s)
    rays = np.concatenate([scenes.coherent_rays(s, 128), scenes.incoherent_rays(s, 20000, 3)]).astype(np.float32)
    h = ctx.trace(rays)
    o = ctx.trace(scenes.shadow_segments(rays, s), api.TRACE_ANY)
    film, stats = ctx.render(api.Context.params_from_scene(s, seed=1))
    print("terrain:", s.triangle_count, "triangles, hits", int((h["shape"] != 0xFFFFFFFF).sum()), "occluded", int(o.sum()),
          "rays", stats["normalRays"] + stats["shadowRays"], "film mean", float(film[..., :3].mean()))
    ctx.close()
    for make in (scenes.instances_test_scene, scenes.cylinders_test_scene, scenes.bitmap_test_scene):
        t = make(48, 48, 2)
        c = api.Context(0)
        c.upload_scene(t)
        f, st = c.render(api.Context.params_from_scene(t, seed=1))
        print(make.__name__, "film mean", float(f[..., :3].mean()))
        c.close()


if __name__ == "__main__":
    main()
