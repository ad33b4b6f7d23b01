#!/usr/bin/env python
"""Per-ray traversal work, reference kd-tree vs this library's 4-wide BVH, on the same rays: the reference's numbers
come from its own instrumented traversal (rayIntersectHavranCollectStatistics, sahkdtree3.h:310-410, through the oracle harness),
ours from the counter build of the trace kernels. Prints one JSON line per ray set.   python tools/kd_vs_bvh_stats.py [c2|c4]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
from mitsuba_renderer_b200 import api, scenes  # noqa: E402
from tests import oracle_util  # noqa: E402
from tests.test_c4_scene import bounce_like_rays  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c4"
    scene = scenes.path_scene() if which == "c4" else scenes.microbench_scene()
    ctx = api.Context(0)
    ctx.upload_scene(scene)
    sets = {"camera": scenes.coherent_rays(scene, 512), "incoherent (kdbench protocol)": scenes.incoherent_rays(scene, 262144, 5)}
    if which == "c4":
        sets["bounce-like (origin on the surface, mint = Epsilon)"] = bounce_like_rays(scene, ctx, 262144, 6)
    sess = oracle_util.Session(scene)
    threads = os.cpu_count() or 1
    for label, rays in sets.items():
        counts, info = sess.kd_stats(rays, threads)
        ctx.set_counters(True)
        ctx.trace(rays)
        c = ctx.counters()
        ctx.set_counters(False)
        n = len(rays)
        print(json.dumps({"scene": which, "rays": label, "n": n,
                          "kd_nodes_traversed_per_ray": float(counts[:, 0].mean()), "kd_primitives_intersected_per_ray": float(counts[:, 1].mean()),
                          "kd_bytes_per_ray": float(counts[:, 0].mean() * 8 + counts[:, 1].mean() * (4 + 48)),
                          "bvh4_node_visits_per_ray": c["nodeVisits"] / n, "bvh4_triangle_tests_per_ray": c["triTests"] / n,
                          "bvh4_bytes_per_ray": c["nodeVisits"] / n * 64 + c["triTests"] / n * 48}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
