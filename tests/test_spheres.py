"""Analytic spheres (src/shapes/sphere.cpp) next to triangle meshes in one acceleration structure:
ray queries, intersection records and deterministic renders against the compiled reference."""
import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile
from mitsuba_renderer_b200.api import Context, TRACE_CLOSEST, TRACE_ANY
from tests import golden_cases as gc
from tests.test_trace_parity import check_closest
from tests.test_shading_parity import close

pytestmark = pytest.mark.gpu
SPHERES = [11, 12, 13, 14, 15]          # shape indices of the spheres in scenes.spheres_test_scene


def test_ray_queries_and_intersection_records(oracle, gpu_ctx_factory):
    scene = scenes.spheres_test_scene(32, 32, 1)
    rays = gc.trace_rays(scene, n=6000)
    sess = oracle.Session(scene)
    ref, ref_full, _ = sess.trace(rays, "closest_full")
    ctx = gpu_ctx_factory(scene)
    hits = ctx.trace(rays, TRACE_CLOSEST)
    on_sphere = np.isin(ref["shape"], SPHERES)
    ref["u"][on_sphere] = 0; ref["v"][on_sphere] = 0       # the kd-tree's intersection cache leaves u, v unset for non-triangles
    check_closest(hits, ref, "spheres")
    assert on_sphere.sum() > 500 and (hits["prim"][on_sphere] == 0xFFFFFFFF).all()       # KNoTriangleFlag
    assert np.array_equal(hits["t"].view(np.uint32), ref["t"].view(np.uint32))          # sphere t included: bit-identical
    seg = scenes.shadow_segments(rays, scene)
    assert np.array_equal(ctx.trace(seg, TRACE_ANY), sess.trace(seg, "any")[0])
    assert np.array_equal(ctx.trace(rays, TRACE_ANY), sess.trace(rays, "any")[0])
    # Sphere::fillIntersectionRecord: p, normals, wi in the (dpdu, dpdv, n) frame
    _, full = ctx.intersect_full(rays)
    same = (hits["shape"] == ref["shape"]) & (ref["shape"] != 0xFFFFFFFF)
    ok = close(full[same], ref_full[same], 2e-5, 2e-6)
    assert ok.all(), f"{np.count_nonzero(~ok.all(axis=1))} records differ"
    # its.uv of the sphere parametrisation and the checkerboard on the textured ball
    uv_ref = sess.uv(rays)
    uv = ctx.intersect_uv(rays)
    assert close(uv[same, :2], uv_ref[same, :2], 2e-5, 2e-6).all()
    ball = same & (ref["shape"] == 13)
    assert ball.sum() > 50 and (~close(uv[ball, 2:5], uv_ref[ball, 2:5], 1e-6, 1e-7).all(axis=1)).sum() <= 2


@pytest.mark.parametrize("mode", [scenefile.MODE_PATH, scenefile.MODE_DIRECT])
def test_halton_render_with_spheres_matches_reference(oracle, gpu_ctx_factory, mode):
    scene = scenes.spheres_test_scene(64, 64, 16, mode, max_depth=7)
    scene.rr_depth = 3
    scene.sampler_type = api.SAMPLER_HALTON
    cpu = oracle.Session(scene).render()[0]
    gpu, _ = gpu_ctx_factory(scene).render(Context.params_from_scene(scene, seed=1))
    assert np.array_equal(gpu[..., 3:], cpu[..., 3:])
    w = np.maximum(cpu[..., 4:5], 1e-20)
    c, g = cpu[..., :3] / w, gpu[..., :3] / w
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    print(f"mode {mode}: within 2e-3: {(err < 2e-3).mean():.4f}, median {np.median(err):.2e}, means {c.mean():.6f} {g.mean():.6f}")
    assert (err < 2e-3).mean() > 0.995 and np.median(err) < 1e-6


def test_sphere_luminaire_sampling(oracle, gpu_ctx_factory):
    """AreaLuminaire::sample / pdf on a sphere (area.cpp:71-86): Sphere::sampleSolidAngle / pdfSolidAngle
    (sphere.cpp:191-247) — the cone of directions seen from outside, uniform area sampling from inside or on the
    surface — next to the mesh luminaires, through Scene::sampleLuminaire / pdfLuminaire."""
    scene = scenes.sphere_lights_test_scene(16, 16, 1)
    ctx = gpu_ctx_factory(scene)
    rng = np.random.Generator(np.random.PCG64(6))
    lo, hi = scenes.scene_bounds(scene)
    q = np.empty((30000, 5), np.float32)
    q[:, :3] = lo + (hi - lo) * rng.random((len(q), 3))
    q[:, 3:] = rng.random((len(q), 2))
    for k, sh in enumerate(s for s in scene.shapes if s.sphere is not None and s.emitter >= 0):
        (c, r, _) = sh.sphere
        d = rng.normal(size=(3000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
        rows = slice(6000 * k, 6000 * k + 3000)
        q[rows, :3] = np.asarray(c) + d * r                                    # on the surface: the "inside" branch
        rows = slice(6000 * k + 3000, 6000 * k + 6000)
        q[rows, :3] = np.asarray(c) + d * r * rng.random((3000, 1))            # inside the ball
    ref = oracle.Session(scene).emitter(q)
    gpu = ctx.emitter_sample(q)
    assert np.array_equal(ref[:, 13], gpu[:, 13]), f"selected luminaire / success flag differs in {(ref[:, 13] != gpu[:, 13]).sum()} rows"
    n_mesh = sum(1 for s in scene.shapes if s.sphere is None and s.emitter >= 0)
    assert set(np.unique(ref[:, 13])) >= {-1.0, float(n_mesh), float(n_mesh + 1)}
    ok_rows = ref[:, 13] >= 0                       # a failed cone sample leaves sRec unset in the reference
    on_sphere = ref[:, 13] >= n_mesh
    assert on_sphere.sum() > 5000
    # A cone sample reaches the sphere through Sphere::rayIntersect: t is a root of the fp32 quadratic, whose
    # discriminant B^2 - 4AC cancels by (dist / r)^2, more so towards the silhouette. One last-bit difference in the
    # sampled direction (sin / cos of the cone angle) therefore moves sRec.p by about eps * dist^2 / (r * cos(theta_l))
    # and sRec.n by 1/r of that; measured worst case 1.1e-5 at dist = 13 r. The tolerance follows that bound per row.
    radius = np.ones(len(q)); centre = np.zeros((len(q), 3))
    for k, sh in enumerate(s for s in scene.shapes if s.sphere is not None and s.emitter >= 0):
        rows = ref[:, 13] == n_mesh + k
        centre[rows] = sh.sphere[0]; radius[rows] = sh.sphere[1]
    dist = np.linalg.norm(q[:, :3] - centre, axis=1)
    cos_l = np.maximum(np.abs(np.sum(ref[:, 3:6] * ref[:, 6:9], axis=1)), 1e-3)
    tol_p = np.where(on_sphere, 2e-6 + 4e-7 * dist ** 2 / (radius * cos_l), 2e-6)
    for c in range(15):
        abs_tol = tol_p if c < 3 else (np.maximum(tol_p / radius, 1e-5) if c < 6 else (1e-5 if c < 9 else 2e-6))
        ok = (np.abs(gpu[:, c] - ref[:, c]) <= 2e-5 * np.abs(ref[:, c]) + abs_tol) | ~ok_rows
        assert ok.all(), f"column {c}: {np.count_nonzero(~ok)} mismatches, e.g. gpu={gpu[~ok][0, c]} ref={ref[~ok][0, c]}"
    assert np.median(np.abs(gpu[on_sphere, :3] - ref[on_sphere, :3])) < 1e-6       # and typically nothing moves at all


@pytest.mark.parametrize("mode", [scenefile.MODE_PATH, scenefile.MODE_DIRECT])
def test_halton_render_with_sphere_luminaires_matches_reference(oracle, gpu_ctx_factory, mode):
    scene = scenes.sphere_lights_test_scene(64, 64, 16, mode, max_depth=7)
    scene.rr_depth = 3
    scene.sampler_type = api.SAMPLER_HALTON
    cpu = oracle.Session(scene).render()[0]
    gpu, _ = gpu_ctx_factory(scene).render(Context.params_from_scene(scene, seed=1))
    assert np.array_equal(gpu[..., 3:], cpu[..., 3:])
    w = np.maximum(cpu[..., 4:5], 1e-20)
    c, g = cpu[..., :3] / w, gpu[..., :3] / w
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    print(f"mode {mode}: within 2e-3: {(err < 2e-3).mean():.4f}, median {np.median(err):.2e}, means {c.mean():.6f} {g.mean():.6f}")
    assert (err < 2e-3).mean() > 0.995 and np.median(err) < 1e-6
