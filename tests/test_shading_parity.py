"""GPU parity of the per-function stages through the C ABI test hooks vs the compiled reference:
BSDF f/pdf/sample (1e-5, north_star), luminaire sampling, camera rays, intersection record, film splat."""
import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile

pytestmark = pytest.mark.gpu
REL, ABS = 1e-5, 1e-7      # tolerance stated by north_star: 1e-5 relative (absolute floor 1e-7)


def close(a, b, rel=REL, abs_=ABS):
    return np.abs(a - b) <= abs_ + rel * np.abs(b)


def bsdf_materials():
    return [scenefile.lambertian((0.5, 0.3, 0.7)),
            scenefile.microfacet((0.5, 0.5, 0.5), (0.5, 0.5, 0.5), 0.5, 0.5, 0.1),      # data/tests/test_bsdf.xml
            scenefile.microfacet((0.2, 0.4, 0.1), (1, 0.9, 0.8), 1.0, 1.0, 0.3, 1.33, 1.0),
            scenefile.mirror((0.8, 0.7, 0.9)),
            scenefile.roughmetal((1, 1, 1), 0.1),                                         # data/tests/test_bsdf.xml
            scenefile.roughmetal((0.9, 0.8, 0.7), 0.4, (0.2, 0.92, 1.1), (3.9, 2.45, 2.14)),
            scenefile.dielectric(), scenefile.dielectric(1.33, 1.0, (0.9, 1, 1), (1, 0.8, 0.9))]


def bsdf_queries(n_mat, n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    q = np.zeros((n, 9), np.float32)
    q[:, 0] = rng.integers(0, n_mat, n)

    def dirs(k):
        d = rng.normal(size=(k, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
        return d
    wi, wo = dirs(n), dirs(n)
    flip = rng.random(n) < 0.7
    wi[flip, 2] = np.abs(wi[flip, 2]); wo[flip, 2] = np.abs(wo[flip, 2])
    q[:, 1:4] = wi; q[:, 4:7] = wo
    q[:, 7:9] = rng.random((n, 2))
    # renormalise in fp32 so both sides see the same unit vectors
    for c in (slice(1, 4), slice(4, 7)):
        q[:, c] /= np.linalg.norm(q[:, c].astype(np.float64), axis=1, keepdims=True).astype(np.float32)
    return q


def test_bsdf_f_pdf_sample(oracle, gpu_ctx_factory):
    scene = scenes.cornell_box(16, 16, 1)
    scene.materials = bsdf_materials()
    for s in scene.shapes:
        s.material = 0
    ctx = gpu_ctx_factory(scene)
    q = bsdf_queries(len(scene.materials), 40000, 21)
    ref = oracle.Session(scene).bsdf(q)
    gpu = ctx.bsdf_eval(q)
    names = ["f.r", "f.g", "f.b", "pdf", "wo.x", "wo.y", "wo.z", "sf.r", "sf.g", "sf.b", "spdf", "type", "s2.r", "s2.g", "s2.b"]
    # which samples succeeded must agree exactly
    assert np.array_equal(ref[:, 11], gpu[:, 11]), "sampled lobe type differs"
    for c, nm in enumerate(names):
        rel, abs_ = REL, ABS
        if nm.startswith("wo"):
            abs_ = 1e-5       # sampled directions are unit vectors: 1e-5 of the vector length, not of each component
        if nm.startswith("sf") or nm == "spdf":
            # f and pdf re-evaluated AT the sampled direction: the direction itself agrees to ~1e-7 (libm vs CUDA
            # sinf/cosf/atanf/logf differ by an ulp), and a Beckmann lobe amplifies that by
            # |d ln D / d theta| = 2 tan(theta) / (alpha^2 cos^2(theta)) ~ 1e2 at alphaB = 0.1. The quantity the
            # integrators consume, f/pdf (columns s2.* and the ratio checked below), is held to 1e-5.
            rel = 2e-4
        ok = close(gpu[:, c], ref[:, c], rel, abs_)
        assert ok.all(), f"{nm}: {np.count_nonzero(~ok)} of {len(q)} outside {rel}; worst gpu={gpu[~ok][0, c]} ref={ref[~ok][0, c]} q={q[~ok][0]}"
    got = gpu[:, 11] != 0
    for c in (7, 8, 9):
        rg = gpu[got, c] / gpu[got, 10]; rr = ref[got, c] / ref[got, 10]
        assert close(rg, rr, REL, 1e-6).all(), f"sampled f/pdf ratio column {c}: max rel {np.max(np.abs(rg - rr) / np.maximum(np.abs(rr), 1e-6))}"


def test_bsdf_f_pdf_at_the_references_sampled_directions(oracle, gpu_ctx_factory):
    """north_star's 1e-5 bar where it is hardest, stated in its own terms: take the directions the REFERENCE sampled (sharp
    Beckmann / GGX / Phong lobes included) and evaluate f and pdf there on both sides as plain (wi, wo) queries. Unlike the
    re-evaluation at each side's own sampled direction (2e-4 above: an ulp of difference in the direction is amplified by
    the lobe), both sides see bit-identical inputs here, so f / pdf must hold 1e-5."""
    scene = scenes.cornell_box(16, 16, 1)
    scene.materials = bsdf_materials()
    for s in scene.shapes:
        s.material = 0
    ctx = gpu_ctx_factory(scene)
    sess = oracle.Session(scene)
    q = bsdf_queries(len(scene.materials), 40000, 33)
    ref = sess.bsdf(q)
    ok = (ref[:, 11] != 0) & ((ref[:, 11].astype(np.int64) & 0x000C) == 0)      # a sample was drawn, and not from a delta lobe
    assert ok.sum() > 10000
    q2 = q[ok].copy()
    q2[:, 4:7] = ref[ok, 4:7]                                                   # wo := the reference's sampled direction
    r2, g2 = sess.bsdf(q2), ctx.bsdf_eval(q2)
    for c, nm in enumerate(["f.r", "f.g", "f.b", "pdf"]):
        good = close(g2[:, c], r2[:, c], REL, ABS)
        assert good.all(), f"{nm} at the reference's sampled wo: {np.count_nonzero(~good)} of {len(q2)} outside {REL}; worst gpu={g2[~good][0, c]} ref={r2[~good][0, c]}"
    assert (r2[:, 3] > 0).mean() > 0.9                                          # these are directions inside the lobes


def test_emitter_sampling(oracle, gpu_ctx_factory):
    scene = scenes.material_test_scene(16, 16, 1)
    ctx = gpu_ctx_factory(scene)
    rng = np.random.Generator(np.random.PCG64(4))
    lo, hi = scenes.scene_bounds(scene)
    q = np.empty((20000, 5), np.float32)
    q[:, :3] = lo + (hi - lo) * rng.random((len(q), 3))
    q[:, 3:] = rng.random((len(q), 2))
    ref = oracle.Session(scene).emitter(q)
    gpu = ctx.emitter_sample(q)
    assert np.array_equal(ref[:, 13], gpu[:, 13]), "selected luminaire / success flag differs"
    assert set(np.unique(ref[:, 13])) >= {0.0, 1.0}
    for c in range(15):
        ok = close(gpu[:, c], ref[:, c], 1e-5, 1e-6)
        assert ok.all(), f"column {c}: {np.count_nonzero(~ok)} mismatches, e.g. gpu={gpu[~ok][0, c]} ref={ref[~ok][0, c]}"


def test_emitter_sampling_all_luminaire_types(oracle, gpu_ctx_factory):
    """point.cpp:58-66, spot.cpp:85-125, directional.cpp:83-90, constant.cpp:72-90 through Scene::sampleLuminaire /
    Scene::pdfLuminaire, next to the two area lights."""
    scene = scenes.lights_test_scene(16, 16, 1)
    ctx = gpu_ctx_factory(scene)
    rng = np.random.Generator(np.random.PCG64(5))
    lo, hi = scenes.scene_bounds(scene)
    q = np.empty((30000, 5), np.float32)
    q[:, :3] = lo + (hi - lo) * rng.random((len(q), 3))
    q[:, 3:] = rng.random((len(q), 2))
    ref = oracle.Session(scene).emitter(q)
    gpu = ctx.emitter_sample(q)
    assert np.array_equal(ref[:, 13], gpu[:, 13]), "selected luminaire / success flag differs"
    assert set(np.unique(ref[:, 13])) >= {0.0, 1.0, 2.0, 3.0, 4.0, 5.0}
    has_n = ref[:, 13] >= 3                      # the delta luminaires leave sRec.n unset in the reference
    for c in range(15):
        rows = has_n if c in (3, 4, 5) else np.ones(len(q), bool)
        ok = close(gpu[rows, c], ref[rows, c], 2e-5, 1e-6)
        assert ok.all(), f"column {c}: {np.count_nonzero(~ok)} mismatches, e.g. gpu={gpu[rows][~ok][0, c]} ref={ref[rows][~ok][0, c]}"
    spot = ref[:, 13] == 1
    inside = spot & (ref[:, 10] > 0)
    assert inside.any() and (spot & (ref[:, 10] == 0)).any()          # both sides of the cone were exercised


@pytest.mark.parametrize("aperture", [0.0, 0.05])
def test_camera_rays(oracle, gpu_ctx_factory, aperture):
    scene = scenes.cornell_box(200, 120, 1)
    scene.aperture_radius = aperture
    scene.focus_depth = 9.0
    ctx = gpu_ctx_factory(scene)
    rng = np.random.Generator(np.random.PCG64(8))
    q = rng.random((20000, 4)).astype(np.float32)
    q[:, 0] *= scene.width; q[:, 1] *= scene.height
    ref = oracle.Session(scene).camera(q)
    gpu = ctx.camera_rays(q)
    assert close(gpu, ref, 1e-5, 1e-6).all(), f"max abs diff {np.abs(gpu - ref).max()}"


def test_intersection_record(oracle, gpu_ctx_factory):
    """fillIntersectionRecord<true> (skdtree.h:352-432): barycentric p, geometric + shading normal, wi."""
    scene = scenes.microbench_scene(cells=48, pad_cells=5)
    ctx = gpu_ctx_factory(scene)
    rays = np.vstack([scenes.incoherent_rays(scene, 20000, 3), scenes.coherent_rays(scene, 128)])
    hits_ref, full_ref, _ = oracle.Session(scene).trace(rays, "closest_full")
    hits, full = ctx.intersect_full(rays)
    same = (hits["shape"] == hits_ref["shape"]) & (hits["prim"] == hits_ref["prim"]) & (hits["shape"] != 0xFFFFFFFF)
    assert same.sum() > 1000
    assert close(full[same], full_ref[same], 1e-5, 1e-6).all(), f"max abs diff {np.abs(full[same] - full_ref[same]).max()}"


@pytest.mark.parametrize("filter_type,crop", [(scenefile.FILTER_BOX, None), (scenefile.FILTER_GAUSSIAN, None),
                                              (scenefile.FILTER_GAUSSIAN, (5, 9, 50, 37))])
def test_film_splat(oracle, gpu_ctx_factory, filter_type, crop):
    """ImageBlock::putSample + Film::putImageBlock (imageblock.h:80-138): footprint, table lookup,
    borders, crop window; invalid (NaN / negative) spectra are dropped."""
    scene = scenes.cornell_box(70, 53, 1)
    scene.filter_type = filter_type
    scene.crop = crop
    sess = oracle.Session(scene)
    cx, cy, cw, ch = scene.crop_rect
    rng = np.random.Generator(np.random.PCG64(12))
    n = 30000
    q = np.empty((n, 6), np.float32)
    q[:, 0] = cx + rng.random(n) * cw; q[:, 1] = cy + rng.random(n) * ch
    q[:, 0] = np.minimum(q[:, 0], np.float32(cx + cw) - np.float32(1e-3)); q[:, 1] = np.minimum(q[:, 1], np.float32(cy + ch) - np.float32(1e-3))
    q[:, 2:5] = rng.random((n, 3)) * 3; q[:, 5] = rng.integers(0, 2, n)
    q[::97, 2] = np.nan; q[::89, 3] = -1.0
    q[:50, 0] = np.floor(q[:50, 0]) + 0.5      # exactly on pixel centres
    ref, table, info = sess.splat(q)
    ctx = gpu_ctx_factory()
    ctx.upload_scene(scene, build=True)
    ctx.set_film(scene, table, info["filter_size"][0])       # the reference's own TabulatedFilter values
    gpu = ctx.film_splat(q)
    assert np.abs(gpu - ref).max() <= 2e-4 * max(1.0, np.abs(ref).max()), np.abs(gpu - ref).max()
    assert np.isclose(gpu[..., 4].sum(), ref[..., 4].sum(), rtol=1e-5)
    # the Python-side filter tables equal the reference's TabulatedFilter
    from mitsuba_renderer_b200.api import filter_table
    mine, size = filter_table(filter_type)
    assert size == info["filter_size"][0]
    assert np.allclose(mine, table, rtol=2e-6, atol=1e-9)


def test_orthographic_camera_rays(oracle, gpu_ctx_factory):
    """OrthographicCamera::generateRay (orthographic.cpp:104-119). The reference's class shadows m_nearClip / m_farClip
    with uninitialised members of its own (orthographic.cpp:176), so its rays carry maxt = 0 and its renders are empty;
    origins and directions are compared with it, the interval with the intended [0, farClip - nearClip]."""
    scene = scenes.cornell_box(200, 120, 1)
    scene.orthographic = True
    m = np.array(scene.camera_to_world, np.float32).copy(); m[:3, 0] *= 2.5; m[:3, 1] *= 2.5     # x / y scale: the image plane's extent
    scene.camera_to_world = m
    ctx = gpu_ctx_factory(scene)
    rng = np.random.Generator(np.random.PCG64(8))
    q = rng.random((20000, 4)).astype(np.float32)
    q[:, 0] *= scene.width; q[:, 1] *= scene.height
    ref = oracle.Session(scene).camera(q)
    gpu = ctx.camera_rays(q)
    cols = [0, 1, 3, 4, 5, 6]                                                           # o.x, o.y, mint = 0, d: bit-identical
    assert np.array_equal(gpu[:, cols].view(np.uint32), ref[:, cols].view(np.uint32))
    # the same uninitialised near plane (0 instead of nearClip) puts the reference's origins on the camera plane itself
    assert np.allclose(gpu[:, 2] - ref[:, 2], scene.near_clip, rtol=0, atol=1e-6)
    assert (ref[:, 7] == 0).all()                                                        # the reference's empty interval
    assert np.allclose(gpu[:, 7], np.float32(scene.far_clip) - np.float32(scene.near_clip), rtol=0, atol=0)
    # and the camera sees the box: a render is not empty
    film, _ = ctx.render(api.Context.params_from_scene(scene, seed=1))
    assert (film[..., 3] > 0).mean() > 0.3
