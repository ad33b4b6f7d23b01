#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): ray-intersection throughput in Mrays/s on config C2 — a
synthetic 1M-triangle mesh, 16M coherent camera rays + 16M incoherent (kdbench-protocol) rays, each
traced as closest-hit and as shadow any-hit — plus samples/s of a path-traced render, beside the
reference's own CPU code on the host cores.

One "step" = one pass of all four ray batches (2 x 16M closest-hit + 2 x 16M any-hit = 67M queries)
through the CUDA library with the rays resident in HBM. `e2e` repeats it through the host-buffer C ABI
call (mtsgpu_trace) with pinned host memory, H2D and D2H copies inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rays R]
Multi-GPU: one process per GPU under torchrun; every rank traces its own ray batches (weak scaling,
no data-path collective) and the render leg shards the sample range and reduces the film with one
NCCL reduce.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ray-intersection throughput, closest-hit + shadow any-hit (config C2)"
UNIT = "Mrays/s"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=3)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def build_ray_sets(scene, n_rays, seed):
    from mitsuba_renderer_b200 import scenes
    res = int(round(np.sqrt(n_rays)))
    coh = scenes.coherent_rays(scene, res)
    inc = scenes.incoherent_rays(scene, n_rays, seed)
    return [("coherent/closest", coh, 0), ("coherent/any", scenes.shadow_segments(coh, scene, seed + 1), 1),
            ("incoherent/closest", inc, 0), ("incoherent/any", scenes.shadow_segments(inc, scene, seed + 2), 1)]


def cpu_reference_trace(scene, n_rays, threads, passes):
    """The reference's own kd-tree (oracle/_ref/timed: its release flags) on a bounded sample of the same
    workload. Returns (Mrays/s over all four batches, description, per-pass seconds)."""
    from tests import oracle_util
    variant = "timed" if oracle_util.available("timed") else "parity"
    if not oracle_util.available(variant):
        return None, "oracle/_ref not built", []
    sess = oracle_util.Session(scene, variant)
    sets = build_ray_sets(scene, n_rays, 2)
    per_pass = np.zeros(passes)
    queries = 0
    for label, rays, mode in sets:
        _, info = sess.trace(rays, "any" if mode else "closest", threads, passes)
        per_pass += np.array(info["trace_s_all"])
        queries += len(rays)
    desc = f"{queries} queries/pass ({len(sets[0][1])} coherent + {n_rays} incoherent rays, closest + any-hit) on the 1M-triangle C2 mesh, " \
           f"{threads} threads, reference kd-tree ({variant} build), kd-tree build excluded"
    return queries, desc, per_pass


def run_reference(args, rank, world):
    if rank != 0:
        return
    from mitsuba_renderer_b200 import scenes
    threads = os.cpu_count() or 1
    scene = scenes.microbench_scene()
    n = args.cpu_rays or (1 << 20 if threads >= 16 else 1 << 18)
    queries, desc, per_pass = cpu_reference_trace(scene, n, threads, args.warmup + args.steps)
    if queries is None:
        print(json.dumps({"impl": "reference", "unavailable": desc}))
        return
    timed = per_pass[args.warmup:]
    value = queries * len(timed) / timed.sum() * 1e-6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": float(timed.mean() * 1e3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C2: 1M-triangle displaced grid, coherent + incoherent rays, closest-hit + any-hit", "sample": desc},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference", "sample": desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from mitsuba_renderer_b200 import api, scenes
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    scene = scenes.microbench_scene()
    ctx = api.Context(local_rank)
    ctx.upload_scene(scene)
    accel = ctx.accel_info()
    sets = build_ray_sets(scene, args.rays, 2 + 100 * rank)
    dev = []
    for label, rays, mode in sets:
        d_rays = torch.from_numpy(rays).cuda()
        d_out = torch.empty(len(rays) * (1 if mode else 5), dtype=torch.int32, device="cuda")
        dev.append((label, d_rays, d_out, mode, len(rays)))
    torch.cuda.synchronize()
    queries_per_step = sum(d[4] for d in dev)

    def step():
        return [ctx.trace_device(d_rays.data_ptr(), n, mode, d_out.data_ptr()) for _, d_rays, d_out, mode, n in dev]

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local_rank); sampler.start()
    torch.cuda.synchronize(); barrier()
    t0 = time.perf_counter()
    kernel_ms = np.zeros(len(dev))
    for _ in range(args.steps):
        kernel_ms += np.array(step())
    torch.cuda.synchronize(); barrier()
    elapsed = max_over_ranks(time.perf_counter() - t0)
    clocks = sampler.summary()
    kernel_ms /= args.steps
    value = queries_per_step * world * args.steps / elapsed * 1e-6

    # ---- roofline of the dominant kernel: algorithmic bytes = rays x (queue + nodes x 128 + tris x 48) ----
    dom = int(np.argmax(kernel_ms))
    label, d_rays, d_out, mode, n = dev[dom]
    ctx.set_counters(True)
    ctx.trace_device(d_rays.data_ptr(), n, mode, d_out.data_ptr())
    cnt = ctx.counters()
    ctx.set_counters(False)
    nodes_per_ray, tris_per_ray = cnt["nodeVisits"] / n, cnt["triTests"] / n
    queue_bytes = 36 if mode else 48
    bytes_per_ray = queue_bytes + nodes_per_ray * 128 + tris_per_ray * 48
    peak, peak_src = measured_peaks()
    achieved = n * bytes_per_ray / (kernel_ms[dom] * 1e-3) * 1e-9
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r01_kernel_traffic.json")
    kname = "k_trace_any" if mode else "k_trace_closest"
    if os.path.exists(tpath):
        rec = json.load(open(tpath)).get(kname)
        if rec and rec.get("batch") == label:       # per launch, scaled to this run's ray count
            traffic, traffic_src = rec["traffic"] * n / rec["rays"], rec["source"]
    roofline = {"bound": "hbm", "kernel": kname, "batch": label, "achieved": achieved, "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "bytes_per_ray": bytes_per_ray, "nodes_per_ray": nodes_per_ray, "tris_per_ray": tris_per_ray,
                "kernel_ms": float(kernel_ms[dom]), "mrays_per_s": n / kernel_ms[dom] * 1e-3,
                "per_batch_ms": {d[0]: float(ms) for d, ms in zip(dev, kernel_ms)},
                "per_batch_mrays_per_s": {d[0]: d[4] / ms * 1e-3 for d, ms in zip(dev, kernel_ms)}}

    # ---- e2e: the host-buffer C ABI call, pinned memory, copies inside the timed region ----
    pinned = []
    for label, rays, mode in sets:
        t = torch.empty(rays.shape, dtype=torch.float32, pin_memory=True)
        t.numpy()[...] = rays
        o = torch.empty(len(rays) * (1 if mode else 5), dtype=torch.int32, pin_memory=True)
        pinned.append((t, o, mode, len(rays)))
    h2d = sum(p[3] * 32 for p in pinned); d2h = sum(p[3] * (4 if p[2] else 20) for p in pinned)

    def e2e_step():
        for t, o, mode, n in pinned:
            ctx._check(ctx.lib.mtsgpu_trace(ctx.h, t.data_ptr(), n, mode, o.data_ptr()))
    e2e_step()
    e2e_steps = max(1, min(args.steps, 3))
    torch.cuda.synchronize(); barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize(); barrier()
    e2e_elapsed = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": queries_per_step * world * e2e_steps / e2e_elapsed * 1e-6, "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "steps": e2e_steps}
    del pinned, dev
    torch.cuda.empty_cache()

    # ---- render leg: config C1 (Cornell box, path maxDepth 8, 64 spp, 512x512, box filter) -----------
    render = None
    if not args.no_render:
        from mitsuba_renderer_b200 import multigpu
        if args.render == "c1":
            rs, rname = scenes.cornell_box(512, 512, 64), "C1: Cornell box 512x512, path maxDepth 8, 64 spp, box filter"
        elif args.render == "c3":
            rs, rname = scenes.direct_scene(), "C3: direct 1+1, 1M triangles, 3 BSDFs + mirror, 8 area lights, 1920x1080, 256 spp"
        elif args.render == "c4":
            rs, rname = scenes.path_scene(), "C4: path maxDepth 16 + RR, 10M triangles (roughmetal terrain + roughmetal/dielectric spheres), 1920x1080, 1024 spp"
        else:
            rs, rname = scenes.path_scene(width=3840, height=2160, spp=4096), "C5: as C4 at 3840x2160, 4096 spp"
        if args.render_spp:
            rs.spp = args.render_spp
            rname += f" [spp overridden to {args.render_spp}]"
        rctx = api.Context(local_rank)
        rctx.upload_scene(rs)
        raccel = rctx.accel_info()
        film = torch.zeros(rs.width * rs.height * 5, dtype=torch.float32, device="cuda")
        lo, hi = multigpu.sample_range(rs.spp, rank, world)             # sample-range shard of this rank
        params = api.Context.params_from_scene(rs, lo, hi, seed=1)
        warm = api.Context.params_from_scene(rs, lo, min(hi, lo + 1), seed=1)
        rctx.render_device(warm, film.data_ptr())                       # warm-up: 1 spp
        torch.cuda.synchronize(); barrier()
        t0 = time.perf_counter()
        st = rctx.render_device(params, film.data_ptr())
        torch.cuda.synchronize()
        t_render = time.perf_counter() - t0
        if world > 1:
            multigpu.reduce_film(film, dst=0)                   # the one collective: film accumulators over NVLink
            torch.cuda.synchronize()
        t_reduce = time.perf_counter() - t0 - t_render
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        rays = torch.tensor([float(st["normalRays"]), float(st["shadowRays"])], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(rays)
        render = {"workload": rname, "triangles": rs.triangle_count, "seconds": dt, "scaling": "strong",
                  "samples_per_s": rs.width * rs.height * rs.spp / dt, "mrays_per_s": float(rays.sum().item()) / dt * 1e-6,
                  "normal_rays": float(rays[0].item()), "shadow_rays": float(rays[1].item()),
                  "kernel_launches_rank0": st["kernelLaunches"], "trace_closest_ms_rank0": st["traceClosestMs"],
                  "trace_any_ms_rank0": st["traceAnyMs"], "other_ms_rank0": st["shadeMs"], "render_s_rank0": t_render,
                  "film_reduce_s_rank0": t_reduce, "film_reduce": "ncclReduce(sum) of W*H*5 floats" if world > 1 else "none (1 GPU)",
                  "bvh_build_ms": raccel["buildMs"], "bvh_nodes": raccel["nodes"]}
        rctx.close()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        n_cpu = args.cpu_rays or (1 << 20 if threads >= 16 else 1 << 18)
        q, desc, per_pass = cpu_reference_trace(scene, n_cpu, threads, 2)
        if q is not None:
            cpu_baseline = {"value": q / per_pass.min() * 1e-6, "unit": UNIT, "cores": threads, "kind": "reference", "sample": desc}
        else:
            cpu_baseline = {"value": None, "unit": UNIT, "cores": threads, "kind": "reference", "sample": desc}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": elapsed / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": "C2: 1M-triangle displaced grid, coherent + incoherent rays, closest-hit + any-hit",
                           "triangles": scene.triangle_count, "rays_per_batch": args.rays, "batches": 4, "queries_per_step_per_gpu": queries_per_step,
                           "l2_policy": "inputs larger than L2 (each ray batch is >= 512 MB at 16M rays)",
                           "bvh_nodes": accel["nodes"], "bvh_node_bytes": accel["nodeBytes"], "triaccel_bytes": accel["triBytes"], "bvh_build_ms": accel["buildMs"],
                           "parallelism": f"{world} x independent ray batches" if world > 1 else "1 GPU"},
                "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "clocks": clocks,
                "gpu_launches": 4 * args.steps, "render": render}
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rays", type=int, default=1 << 24, help="rays per batch (coherent batch is the nearest square)")
    ap.add_argument("--cpu-rays", type=int, default=0, help="rays per batch of the bounded CPU sample (0 = auto)")
    ap.add_argument("--no-render", action="store_true")
    ap.add_argument("--render", default="c4", choices=["c1", "c3", "c4", "c5"], help="config of the render leg")
    ap.add_argument("--render-spp", type=int, default=0, help="override the render leg's samples per pixel")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); local_rank = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
