"""Diagnostic: closest/any consistency on the C2 mesh under each builder; prints mismatch counts and a few offenders."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from mitsuba_renderer_b200 import api, scenes
from tests import oracle_util

scene = scenes.microbench_scene()
rays_sets = {"incoherent": scenes.incoherent_rays(scene, 1 << 20, 2), "coherent": scenes.coherent_rays(scene, 1024)}
sess = oracle_util.Session(scene) if oracle_util.available() else None
for builder in ("sah", "lbvh"):
    os.environ["MTSGPU_BUILDER"] = builder
    ctx = api.Context(0)
    ctx.upload_scene(scene)
    print(builder, ctx.accel_info())
    for label, rays in rays_sets.items():
        h1 = ctx.trace(rays)
        hit = h1["shape"] != 0xFFFFFFFF
        seg = rays.copy(); seg[:, 7] = np.where(hit, h1["t"] * np.float32(1 + 1e-5), rays[:, 7])
        occ = ctx.trace(seg, api.TRACE_ANY).astype(bool)
        bad = np.nonzero(occ != hit)[0]
        print(f"  {label}: hits {hit.sum()}  any!=closest {len(bad)}  (occluded-but-miss {np.count_nonzero(occ & ~hit)}, hit-but-unoccluded {np.count_nonzero(~occ & hit)})")
        for i in bad[:5]:
            print("     ray", i, rays[i], "closest", h1[i], "occ", occ[i])
        if sess is not None:
            ref, _ = sess.trace(rays, "closest", os.cpu_count() or 1)
            rhit = ref["shape"] != 0xFFFFFFFF
            same = (h1["shape"] == ref["shape"]) & (h1["prim"] == ref["prim"])
            print(f"     vs reference: hit/miss differs {np.count_nonzero(rhit != hit)}, id differs {np.count_nonzero(~same)}")
            refany, _ = sess.trace(seg, "any", os.cpu_count() or 1)
            print(f"     reference any on the same segments: differs from gpu any {np.count_nonzero(refany.astype(bool) != occ)}, from hit {np.count_nonzero(refany.astype(bool) != hit)}")
            for i in np.nonzero(~same)[:5]:
                print("     id mismatch ray", i, rays[i], "gpu", h1[i], "ref", ref[i])
    ctx.close()
