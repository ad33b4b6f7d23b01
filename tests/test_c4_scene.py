"""BASELINE config C4's own scene (10,000,008 triangles: rough-metal terrain + rough-metal / glass spheres + 4 area lights)
under the default GPU run: hit ids on the 10M-triangle tree against the reference's kd-tree for three kinds of rays, and a
deterministic (halton) `path` render on a small film against the reference's film. One scene file and one GPU context
serve the whole module (generating and writing the scene and building the reference's kd-tree take most of the time)."""
import copy

import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes
from mitsuba_renderer_b200.api import Context, TRACE_ANY
from tests.test_trace_parity import check_closest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c4(oracle):
    scene = scenes.path_scene(width=48, height=27, spp=8)
    scene.sampler_type = api.SAMPLER_HALTON
    assert scene.triangle_count == 10_000_008
    ctx = api.Context(0)
    ctx.upload_scene(scene)
    yield scene, ctx, oracle.Session(scene)
    ctx.close()


def bounce_like_rays(scene, ctx, n, seed):
    """Rays as the integrator makes them after a hit: origin ON the surface (the hit point of a camera ray, computed the way
    the shading kernel does, from the barycentrics), mint = Epsilon, maxt = inf, direction cosine-distributed about +y
    (the terrain's mean normal) — plus a share pointing down into the surface so that self-hits are exercised."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cam = scenes.coherent_rays(scene, int(np.ceil(np.sqrt(1.3 * n))))
    hits = ctx.trace(cam)
    ok = np.nonzero(hits["shape"] != 0xFFFFFFFF)[0][:n]
    p = cam[ok, 0:3] + cam[ok, 4:7] * hits["t"][ok, None]
    u = rng.random((len(ok), 2))
    r, phi = np.sqrt(u[:, 0]), 2 * np.pi * u[:, 1]
    d = np.stack([r * np.cos(phi), np.sqrt(1 - u[:, 0]), r * np.sin(phi)], 1)
    d[::7, 1] *= -1.0
    rays = np.empty((len(ok), 8), np.float32)
    rays[:, 0:3] = p; rays[:, 3] = 1e-4; rays[:, 4:7] = d; rays[:, 7] = np.inf
    return rays


def test_hit_ids_on_the_10m_triangle_tree(c4):
    """>= 1M rays of three kinds — camera rays, kdbench-protocol incoherent rays (kdbench.cpp:221-231) and bounce-like rays
    that start on the surface with mint = Epsilon — against ShapeKDTree::rayIntersect (skdtree.cpp:108-132,180-199):
    ids exact up to exact-t ties, t / u / v bit-identical, any-hit arrays equal."""
    import os
    scene, ctx, sess = c4
    info = ctx.accel_info()
    assert info["builder"] == 2 and info["depth"] * 3 <= 96, info       # binned-SAH tree; worst-case stack demand fits
    cam = scenes.coherent_rays(scene, 640)
    inc = scenes.incoherent_rays(scene, 400_000, 5)
    bnc = bounce_like_rays(scene, ctx, 300_000, 6)
    rays = np.concatenate([cam, inc, bnc]).astype(np.float32)
    assert len(rays) >= 1_000_000
    threads = os.cpu_count() or 1
    ref, rinfo = sess.trace(rays, "closest", threads)
    gpu = ctx.trace(rays)
    ties = check_closest(gpu, ref, "C4 closest")
    hit = ref["shape"] != 0xFFFFFFFF
    parts = {"camera": slice(0, len(cam)), "incoherent": slice(len(cam), len(cam) + len(inc)), "bounce": slice(len(cam) + len(inc), len(rays))}
    for k, sl in parts.items():
        assert 0.02 < hit[sl].mean() <= 1.0, (k, hit[sl].mean())
    ctx.set_counters(True)              # counter build: deepest traversal stack any of these rays needs (limit: 96 entries)
    ctx.trace(rays[::8]); ctx.trace(scenes.shadow_segments(rays[::8], scene), TRACE_ANY)
    deepest = ctx.counters()["maxStackEntries"]
    ctx.set_counters(False)
    assert 0 < deepest <= 3 * info["depth"] <= 96, (deepest, info)
    seg = scenes.shadow_segments(rays, scene)
    assert np.array_equal(sess.trace(seg, "any", threads)[0], ctx.trace(seg, TRACE_ANY))
    print(f"C4 tree: {len(rays)} rays, hit rates { {k: round(float(hit[sl].mean()), 3) for k, sl in parts.items()} }, exact-t ties {ties}, "
          f"wide-tree depth {info['depth']}, deepest stack {deepest} entries, kd build {rinfo['kd_build_s']:.1f} s")


def test_halton_render_on_the_c4_scene(c4):
    """`path` maxDepth 16 + roulette on a 48 x 27 film, 8 spp, deterministic sampler: the film against the reference's."""
    scene, ctx, sess = c4
    for max_depth, share in ((2, 0.999), (16, 0.99)):
        s = copy.copy(scene); s.max_depth = max_depth
        cpu = type(sess)(s).render()[0]
        gpu, stats = ctx.render(Context.params_from_scene(s, seed=1))
        assert np.array_equal(gpu[..., 4], cpu[..., 4]) and np.count_nonzero(gpu[..., 3] != cpu[..., 3]) <= 4
        w = np.maximum(cpu[..., 4:5], 1e-20)
        c, g = cpu[..., :3] / w, gpu[..., :3] / w
        err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
        print(f"C4 scene, maxDepth {max_depth}: within 2e-3: {(err < 2e-3).mean():.4f}, median {np.median(err):.2e}, means {c.mean():.6f} {g.mean():.6f}")
        assert (err < 2e-3).mean() >= share and np.median(err) < 1e-6 and abs(g.mean() - c.mean()) < 2e-3 * c.mean()
