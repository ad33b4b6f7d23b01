"""Render one config several times under different MTSGPU_* environment settings (experiment driver).

    python tools/render_sweep.py --render c4 --spp 64 --set MTSGPU_SORT=0 --set MTSGPU_SORT=15,6,9 ...

Each --set is a comma-free list of VAR=value pairs joined by '+', e.g. MTSGPU_SORT=15,6,9+MTSGPU_SORT_MIN=65536.
Prints one JSON line per setting: seconds, closest/any/other ms, film checksum.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from mitsuba_renderer_b200 import api, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--render", default="c4")
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--set", action="append", default=[])
    ap.add_argument("--repeat", type=int, default=2)
    ap.add_argument("--lib", default=None, help="alternative libmtscuda build to load")
    args = ap.parse_args()
    if args.render == "c1":
        rs = scenes.cornell_box(512, 512, 64)
    elif args.render == "c3":
        rs = scenes.direct_scene()
    elif args.render == "mat":
        rs = scenes.material_test_scene(512, 512, 64)
    else:
        rs = scenes.path_scene()
    rs.spp = args.spp
    ctx = api.Context(0, args.lib)
    ctx.upload_scene(rs)
    film = torch.zeros(rs.width * rs.height * 5, dtype=torch.float32, device="cuda")
    params = api.Context.params_from_scene(rs, 0, rs.spp, seed=1)
    for setting in args.set or [""]:
        pairs = [kv.split("=", 1) for kv in setting.split("+") if kv]
        for k, v in pairs:
            os.environ[k] = v
        best = None
        for _ in range(args.repeat):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            st = ctx.render_device(params, film.data_ptr())
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if best is None or dt < best[0]:
                best = (dt, st)
        dt, st = best
        print(json.dumps({"setting": setting, "seconds": round(dt, 4), "closest_ms": round(st["traceClosestMs"], 1),
                          "any_ms": round(st["traceAnyMs"], 1), "other_ms": round(st["shadeMs"], 1),
                          "rays": st["normalRays"] + st["shadowRays"],
                          "film_sum": float(film.double().sum().item())}), flush=True)
        for k, _ in pairs:
            os.environ.pop(k, None)
    ctx.close()


if __name__ == "__main__":
    main()
