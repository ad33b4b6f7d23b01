#!/usr/bin/env python
"""Traversal statistics on a render scene: primary camera rays and synthetic diffuse-bounce rays from the
primary hits, traced through the ray-query API with counters. Usage: python tools/trace_stats.py [c3|c4] [res]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from mitsuba_renderer_b200 import api, scenes  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c4"
    res = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
    scene = scenes.path_scene() if which == "c4" else scenes.direct_scene()
    ctx = api.Context(0, os.environ.get("MTSGPU_LIB"))
    ctx.upload_scene(scene)
    print("accel", ctx.accel_info())
    prim = scenes.coherent_rays(scene, res)
    hits = ctx.trace(prim)
    ok = hits["shape"] != 0xFFFFFFFF
    p = prim[ok, 0:3] + prim[ok, 4:7] * hits["t"][ok, None]
    rng = np.random.Generator(np.random.PCG64(1))
    n = len(p)
    u = rng.random((n, 2))
    r, phi = np.sqrt(u[:, 0]), 2 * np.pi * u[:, 1]
    d = np.stack([r * np.cos(phi), np.sqrt(1 - u[:, 0]), r * np.sin(phi)], 1)     # cosine hemisphere about +y
    sec = np.empty((n, 8), np.float32)
    sec[:, 0:3] = p + np.array([0, 1e-3, 0]); sec[:, 3] = 1e-4; sec[:, 4:7] = d; sec[:, 7] = np.inf
    lights = np.array([[0.7, 1.6, 0.0], [-0.7, 1.6, 0.0], [0.0, 1.6, 0.7], [0.0, 1.6, -0.7]])
    tgt = lights[rng.integers(0, 4, n)] + (rng.random((n, 3)) - 0.5) * [0.5, 0, 0.5]
    sh = np.empty((n, 8), np.float32)
    sh[:, 0:3] = p; sh[:, 3] = 1e-3; sh[:, 4:7] = tgt - p; sh[:, 7] = 1 - 1e-3
    for label, rays, mode in (("primary/closest", prim, 0), ("bounce/closest", sec, 0), ("shadow/any", sh, 1)):
        d_rays = torch.from_numpy(np.ascontiguousarray(rays)).cuda()
        d_out = torch.empty(len(rays) * (1 if mode else 5), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        ms = min(ctx.trace_device(d_rays.data_ptr(), len(rays), mode, d_out.data_ptr()) for _ in range(4))
        ctx.set_counters(True); ctx.trace_device(d_rays.data_ptr(), len(rays), mode, d_out.data_ptr()); c = ctx.counters(); ctx.set_counters(False)
        print(f"{label:18s} rays {len(rays):9d}  {ms:8.3f} ms  {len(rays)/ms*1e-3:8.1f} Mrays/s  nodes/ray {c['nodeVisits']/len(rays):6.1f}  tris/ray {c['triTests']/len(rays):6.1f}")


if __name__ == "__main__":
    main()
