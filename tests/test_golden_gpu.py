"""GPU parity against the committed golden vectors (tests/golden/*.npz, produced by the compiled reference
with tests/golden/make_golden.py). These run on any B200 box, also where neither /root/reference nor
oracle/_ref exist; the live-oracle tests in test_*_parity.py cover larger randomised inputs."""
import os

import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes
from mitsuba_renderer_b200.api import TRACE_CLOSEST, TRACE_ANY, Context
from tests import golden_cases as gc
from tests.test_trace_parity import check_closest
from tests.test_shading_parity import close

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


@pytest.mark.parametrize("name,make", [("trace_materials", gc.scene_materials), ("trace_grid", gc.scene_grid)])
def test_ray_queries_match_golden(name, make, gpu_ctx_factory):
    scene = make()
    g = gold(name)
    rays = gc.trace_rays(scene)
    assert gc.digest(rays) == str(g["rays_digest"])
    ctx = gpu_ctx_factory(scene)
    ref = g["hits"].view(api.HIT_DTYPE).reshape(-1)
    check_closest(ctx.trace(rays, TRACE_CLOSEST), ref, name)
    assert np.array_equal(ctx.trace(scenes.shadow_segments(rays, scene), TRACE_ANY), g["any_segments"])
    assert np.array_equal(ctx.trace(rays, TRACE_ANY), g["any_rays"])
    hits, full = ctx.intersect_full(rays)
    same = (hits["shape"] == ref["shape"]) & (hits["prim"] == ref["prim"]) & (ref["shape"] != 0xFFFFFFFF)
    assert close(full[same], g["full"][same], 1e-5, 1e-6).all()


def test_texture_coordinates_and_textured_reflectance_match_golden(gpu_ctx_factory):
    """its.uv (skdtree.h:408-415) and Texture2D / Checkerboard::getValue at it, seen through
    BSDF::getDiffuseReflectance(its) of the lambertian, microfacet and phong materials."""
    scene = gc.scene_textured()
    g = gold("uv_textured")
    rays = gc.trace_rays(scene)
    assert gc.digest(rays) == str(g["rays_digest"])
    ctx = gpu_ctx_factory(scene)
    hits = ctx.trace(rays, TRACE_CLOSEST)
    out, ref = ctx.intersect_uv(rays), g["out"]
    hit = hits["shape"] != 0xFFFFFFFF
    assert hit.sum() > 5000 and close(out[hit, :2], ref[hit, :2], 1e-5, 1e-6).all()
    assert np.array_equal(out[~hit], np.zeros_like(out[~hit]))
    want = gc.diffuse_reflectance_from_slots(scene, hits, out[:, 2:5])
    # a uv within rounding of a checker edge may land on the other colour: allow a handful
    bad = ~close(want, ref[:, 2:5], 1e-6, 1e-7).all(axis=1)
    assert bad.sum() <= 3, bad.sum()
    # shading frames: the meshes under the anisotropic Ward materials use tangents computed from the texture
    # coordinates (trimesh.cpp:547-662, skdtree.h:385-398); `wi` is expressed in that frame
    ref_hits = g["hits"].view(api.HIT_DTYPE).reshape(-1)
    check_closest(hits, ref_hits, "textured")
    full = ctx.intersect_full(rays)[1]
    same = (hits["shape"] == ref_hits["shape"]) & (hits["prim"] == ref_hits["prim"]) & hit
    assert close(full[same], g["full"][same], 1e-5, 2e-6).all()
    on_ward = same & np.isin(hits["shape"], [4, 6])
    assert on_ward.sum() > 300
    textured = np.array([h["shape"] != 0xFFFFFFFF and scene.shapes[int(h["shape"])].material in (1, 2, 4, 9) for h in hits])
    assert len(np.unique(ref[textured, 2:5].round(4), axis=0)) >= 4        # both colours of several textures were hit


def test_deterministic_halton_render_matches_golden_film(gpu_ctx_factory):
    """Whole-integrator parity without statistics: with the halton sampler (halton.cpp:68-84) both sides trace the same
    paths in every pixel; the golden film is the reference's `path` integrator on the textured all-BSDF scene."""
    scene = gc.scene_halton()
    ref = gold("render_halton")["film"]
    gpu, _ = gpu_ctx_factory(scene).render(Context.params_from_scene(scene, seed=0))
    assert np.array_equal(gpu[..., 3:], ref[..., 3:])                  # alpha and weight accumulators: identical
    w = np.maximum(ref[..., 4:5], 1e-20)
    c, g = ref[..., :3] / w, gpu[..., :3] / w
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    assert (err < 2e-3).mean() > 0.995 and np.median(err) < 2e-6, ((err < 2e-3).mean(), np.median(err))
    assert abs(g.mean() - c.mean()) < 5e-4 * c.mean()


def test_triaccel_matches_golden(gpu_ctx_factory):
    g = gold("triaccel_materials")
    ref = g["records"].view(api.TRIACCEL_DTYPE)
    ctx = gpu_ctx_factory(gc.scene_materials())
    rec, box = ctx.triaccel(len(ref))
    assert np.array_equal(box.view(np.uint32), g["box"].view(np.uint32))
    for f in ("n_u", "n_v", "n_d", "a_u", "a_v", "b_nu", "b_nv", "c_nu", "c_nv"):
        assert np.array_equal(rec[f].view(np.uint32), ref[f].view(np.uint32)), f
    assert np.array_equal(rec["k"] & 3, ref["k"] & 3)


def test_bsdf_matches_golden(gpu_ctx_factory):
    bs = gc.bsdf_scene()
    q = gc.bsdf_queries(len(bs.materials), 4000, 21)
    g = gold("bsdf")
    assert gc.digest(q) == str(g["queries_digest"])
    out = gpu_ctx_factory(bs).bsdf_eval(q)
    ref = g["out"]
    assert np.array_equal(out[:, 11], ref[:, 11])
    assert close(out[:, :4], ref[:, :4]).all()                 # f, pdf at given directions: 1e-5
    assert close(out[:, 4:7], ref[:, 4:7], 1e-5, 1e-5).all()   # sampled direction: 1e-5 of unit length
    assert close(out[:, 7:11], ref[:, 7:11], 2e-4, 1e-7).all() and close(out[:, 12:15], ref[:, 12:15], 2e-4, 1e-7).all()
    bit_exact = np.mean(out[:, 4:7].view(np.uint32) == ref[:, 4:7].view(np.uint32))
    print(f"sampled directions bit-identical: {bit_exact:.4f}")
    assert bit_exact > 0.95


def test_emitter_and_camera_match_golden(gpu_ctx_factory):
    ms = gc.scene_materials()
    out = gpu_ctx_factory(ms).emitter_sample(gc.emitter_queries(ms, 4000))
    ref = gold("emitter")["out"]
    assert np.array_equal(out[:, 13], ref[:, 13]) and close(out[:, :15], ref[:, :15], 1e-5, 1e-6).all()
    # point / spot / directional / constant luminaires next to the area lights; the reference leaves sRec.n of the
    # delta luminaires unset, so those three columns are compared for the others only
    ls = gc.scene_lights()
    q = gc.emitter_queries(ls, 6000)
    ref = gold("emitter_lights")
    assert gc.digest(q) == str(ref["queries_digest"])
    out, ref = gpu_ctx_factory(ls).emitter_sample(q), ref["out"]
    assert np.array_equal(out[:, 13], ref[:, 13]) and set(np.unique(ref[:, 13])) >= {0.0, 1.0, 2.0, 3.0, 4.0, 5.0}
    cols = [c for c in range(15) if c not in (3, 4, 5)]
    assert close(out[:, cols], ref[:, cols], 2e-5, 1e-6).all()
    has_n = ref[:, 13] >= 3
    assert close(out[has_n, 3:6], ref[has_n, 3:6], 1e-5, 1e-6).all()
    for ap, name in ((0.0, "camera_pinhole"), (0.05, "camera_thinlens")):
        cs = gc.camera_scene(ap)
        rays = gpu_ctx_factory(cs).camera_rays(gc.camera_samples(cs, 4000))
        ref_rays = gold(name)["rays"]
        assert close(rays, ref_rays, 1e-5, 1e-6).all()
        if ap == 0.0:      # rasterToCamera is restated in fp32 (scenefile.raster_to_camera): pinhole rays are bit-identical
            assert np.array_equal(rays.view(np.uint32), ref_rays.view(np.uint32))


@pytest.mark.parametrize("name", list(gc.SPLAT_CASES))
def test_film_splat_matches_golden(name, gpu_ctx_factory):
    ft, crop = gc.SPLAT_CASES[name]
    ss = gc.splat_scene(ft, crop)
    g = gold("splat_" + name)
    q = gc.splat_samples(ss, 8000)
    assert gc.digest(q) == str(g["samples_digest"])
    ctx = gpu_ctx_factory()
    ctx.set_film(ss, g["table"], float(g["filter_size"]))
    film = ctx.film_splat(q)
    assert np.abs(film - g["film"]).max() <= 2e-4 * max(1.0, np.abs(g["film"]).max())


def test_render_matches_golden_reference_image(gpu_ctx_factory):
    rs = gc.render_scene()
    g = gold("render_cornell48")
    dev = lambda f: f[..., :3] / np.maximum(f[..., 4:5], 1e-20)
    ref = dev(g["ref_4096"])
    rel = lambda a: float(np.mean((a - ref) ** 2 / (ref ** 2 + 1e-4)))
    ctx = gpu_ctx_factory(rs)
    gpu, _ = ctx.render(Context.params_from_scene(rs, seed=1))
    assert rel(dev(gpu)) <= 1.25 * rel(dev(g["cpu_64"])) + 1e-4
    import copy
    hi = copy.copy(rs); hi.spp = 4096
    gpu_hi, _ = ctx.render(Context.params_from_scene(hi, seed=2))
    assert abs(dev(gpu_hi).mean() - ref.mean()) <= 0.01 * ref.mean()
    assert rel(dev(gpu_hi)) < 0.2 * rel(dev(g["cpu_64"]))


def test_filtered_bitmap_lookups_match_golden(gpu_ctx_factory):
    """ldrtexture.cpp / mipmap.cpp lookups (EWA, trilinear, unfiltered; repeat, clamp, black) with the uv partials of
    intersection.cpp:5-75 at camera-ray hits, thin lens on, differentials narrowed for 4 samples per pixel."""
    scene = gc.scene_bitmaps()
    g = gold("texlookup_bitmaps")
    q = gc.lens_camera_samples(scene, 6000, 31)
    assert gc.digest(q) == str(g["samples_digest"])
    ctx = gpu_ctx_factory(scene)
    got, ref = ctx.texture_lookup(q, scene.spp), g["out"]
    hits = ctx.trace(ctx.camera_rays(q), TRACE_CLOSEST)
    hit = hits["shape"] != 0xFFFFFFFF
    assert hit.sum() > 5000 and close(got[hit, :2], ref[hit, :2], 2e-5, 2e-6).all()
    scale = np.abs(ref[:, 2:6]).max(axis=1, keepdims=True)
    assert (scale > 0).mean() > 0.5
    assert (np.abs(got[:, 2:6] - ref[:, 2:6]) <= 1e-4 * (scale + 1e-3)).all()
    want = gc.diffuse_reflectance_from_slots(scene, hits, got[:, 6:9])
    err = np.abs(want - ref[:, 6:9]).max(axis=1)
    assert (err > 1e-5).sum() <= 6 and np.median(err) < 1e-7, ((err > 1e-5).sum(), err.max())


def test_envmap_sampling_matches_golden(gpu_ctx_factory):
    """EnvMapLuminaire::sample / pdf (envmap.cpp:134-199) on the rotated 64 x 32 map of the envmap test scene."""
    es = gc.scene_envmap()
    q = gc.emitter_queries(es, 6000)
    ref = gold("emitter_envmap")
    assert gc.digest(q) == str(ref["queries_digest"])
    out, ref = gpu_ctx_factory(es).emitter_sample(q), ref["out"]
    assert np.array_equal(out[:, 13], ref[:, 13]) and (ref[:, 13] == 3).sum() > 1000
    cols = [c for c in range(15) if c not in (3, 4, 5)]
    ok = ref[:, 13] >= 0                                    # a point outside the beam leaves the reference's record unset
    assert close(out[ok][:, cols], ref[ok][:, cols], 2e-5, 2e-6).all()
    has_n = (ref[:, 13] == 3) | (ref[:, 13] >= 5)           # not the delta luminaires (the collimated beam is number 4)
    assert close(out[has_n, 3:6], ref[has_n, 3:6], 2e-5, 2e-6).all()
    assert (ref[:, 13] == 4).sum() > 20


@pytest.mark.parametrize("name,make", [("trace_spheres", gc.scene_spheres), ("trace_instances", gc.scene_instances)])
def test_ray_queries_on_non_triangle_primitives_match_golden(name, make, gpu_ctx_factory):
    """Analytic spheres (sphere.cpp:92-131) and instances of shape groups (instance.cpp:72-98) next to the meshes."""
    scene = make()
    g = gold(name)
    rays = gc.trace_rays(scene)
    assert gc.digest(rays) == str(g["rays_digest"])
    ctx = gpu_ctx_factory(scene)
    ref = g["hits"].view(api.HIT_DTYPE).reshape(-1)
    hits = ctx.trace(rays, TRACE_CLOSEST)
    got = hits.copy()
    nontri = got["prim"] == 0xFFFFFFFF
    assert nontri.sum() > 500
    got["u"][nontri] = 0; got["v"][nontri] = 0
    check_closest(got, ref, name)
    assert np.array_equal(hits["t"].view(np.uint32), ref["t"].view(np.uint32))
    assert np.array_equal(ctx.trace(scenes.shadow_segments(rays, scene), TRACE_ANY), g["any_segments"])
    assert np.array_equal(ctx.trace(rays, TRACE_ANY), g["any_rays"])
    _, full = ctx.intersect_full(rays)
    same = (hits["shape"] == ref["shape"]) & (ref["shape"] != 0xFFFFFFFF)
    assert close(full[same], g["full"][same], 2e-5, 2e-6).all()
