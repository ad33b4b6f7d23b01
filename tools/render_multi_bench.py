#!/usr/bin/env python
"""One process, several GPUs: the C4 (or C5) render through mtsgpu_render_multi — one host thread per GPU, sample ranges
split, films summed on GPU 0 by a kernel that reads the peers' accumulators over NVLink (csrc/multi.cu). This is the launch
shape of the plugin (`<integer name="gpus" value="N"/>`); bench.py measures the one-process-per-GPU shape.

    python tools/render_multi_bench.py [--gpus N] [--workload c4|c5] [--steps K] [--spp S]
Prints one JSON line."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mitsuba_renderer_b200 import api, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--workload", default="c4")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--spp", type=int, default=0)
    a = ap.parse_args()
    s = scenes.path_scene() if a.workload == "c4" else scenes.path_scene(width=3840, height=2160, spp=4096)
    if a.spp:
        s.spp = a.spp
    ctxs = [api.Context(g) for g in range(a.gpus)]
    for c in ctxs:
        c.upload_scene(s)
    p = api.Context.params_from_scene(s, seed=1)
    api.render_multi(ctxs, p)                      # warm-up (allocations, peer mappings)
    times = []
    for _ in range(a.steps):
        t0 = time.perf_counter()
        film, st = api.render_multi(ctxs, p)
        times.append(time.perf_counter() - t0)
    rays = st["normalRays"] + st["shadowRays"]
    print(json.dumps({"what": "mtsgpu_render_multi (one process, one host thread per GPU, peer-load film reduce; film to host included)",
                      "workload": a.workload, "spp": s.spp, "n_gpus": a.gpus, "seconds_per_step": times, "best_s": min(times),
                      "mrays_per_s": rays / min(times) * 1e-6, "samples_per_s": s.width * s.height * s.spp / min(times),
                      "film_mean": float(film[..., :3].sum() / max(film[..., 4].sum(), 1e-20)), "stats": st}))
    for c in ctxs:
        c.close()


if __name__ == "__main__":
    main()
