#!/usr/bin/env python
"""Measures the BASELINE.md table rows on the current box: each render config on the GPU at full spec
(mtsgpu_render_device, one GPU) and on the host cores with the compiled reference (oracle/_ref/timed, all
cores) at a reduced spp that is scaled linearly (stated). Prints one JSON object per config.
    python tools/baseline_table.py [c1 c3 c4 c5] [--gpu-spp N]"""
import copy
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from mitsuba_renderer_b200 import api, scenes  # noqa: E402
from tests import oracle_util  # noqa: E402

CONFIGS = {
    "c1": (lambda: scenes.cornell_box(512, 512, 64), 64, "C1 Cornell 512x512, path maxDepth 8, 64 spp, box"),
    "c3": (lambda: scenes.direct_scene(), 4, "C3 direct 1+1, 1M triangles, 1920x1080, 256 spp"),
    "c4": (lambda: scenes.path_scene(), 2, "C4 path maxDepth 16 + RR, 10M triangles, 1920x1080, 1024 spp"),
    "c5": (lambda: scenes.path_scene(width=3840, height=2160, spp=4096), 0, "C5 4K, 4096 spp, 10M triangles"),
}


def main():
    names = [a for a in sys.argv[1:] if a in CONFIGS] or ["c1", "c3", "c4"]
    gpu_spp = int(sys.argv[sys.argv.index("--gpu-spp") + 1]) if "--gpu-spp" in sys.argv else 0
    threads = os.cpu_count() or 1
    for name in names:
        make, cpu_spp, label = CONFIGS[name]
        scene = make()
        full_spp = scene.spp
        out = {"config": label, "triangles": scene.triangle_count, "host_cores": threads}
        # ---- GPU ----
        g = copy.copy(scene)
        if gpu_spp:
            g.spp = gpu_spp
        ctx = api.Context(0)
        t0 = time.perf_counter(); ctx.upload_scene(g); out["gpu_upload_build_s"] = time.perf_counter() - t0
        film = torch.zeros(g.width * g.height * 5, dtype=torch.float32, device="cuda")
        warm = api.Context.params_from_scene(g, 0, 1, seed=1)
        ctx.render_device(warm, film.data_ptr())
        torch.cuda.synchronize(); t0 = time.perf_counter()
        st = ctx.render_device(api.Context.params_from_scene(g, seed=1), film.data_ptr())
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        rays = st["normalRays"] + st["shadowRays"]
        scale = full_spp / g.spp
        out.update({"gpu_spp": g.spp, "gpu_s": dt, "gpu_s_at_full_spp": dt * scale, "gpu_samples_per_s": g.width * g.height * g.spp / dt,
                    "gpu_mrays_per_s": rays / dt * 1e-6, "gpu_trace_closest_ms": st["traceClosestMs"], "gpu_trace_any_ms": st["traceAnyMs"],
                    "gpu_other_ms": st["shadeMs"], "gpu_bvh_build_ms": ctx.accel_info()["buildMs"]})
        ctx.close(); del film; torch.cuda.empty_cache()
        # ---- CPU reference ----
        if cpu_spp and oracle_util.available("timed"):
            c = copy.copy(scene); c.spp = cpu_spp
            _, info = oracle_util.Session(c, "timed").render(threads)
            crays = oracle_util.parse_count(info["normal_rays"]) + oracle_util.parse_count(info["shadow_rays"])
            out.update({"cpu_spp": cpu_spp, "cpu_s": info["render_s"], "cpu_s_at_full_spp": info["render_s"] * full_spp / cpu_spp,
                        "cpu_samples_per_s": info["samples_per_s"], "cpu_mrays_per_s": crays / info["render_s"] * 1e-6,
                        "cpu_kd_build_s": info["kd_build_s"], "speedup_samples_per_s": out["gpu_samples_per_s"] / info["samples_per_s"]})
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
