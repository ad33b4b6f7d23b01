"""Bitmap textures: src/textures/ldrtexture.cpp over the MIPMap of src/librender/mipmap.cpp, filtered
with the uv partials that Intersection::computePartials (src/librender/intersection.cpp:5-75) derives from the camera's
ray differentials. Checked against the compiled reference — its own ldrtexture plugin reading a BMP file — per lookup
and through deterministic renders."""
import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile
from mitsuba_renderer_b200.api import Context, TRACE_CLOSEST
from tests import golden_cases as gc

pytestmark = pytest.mark.gpu


def camera_samples(scene, n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    q = np.empty((n, 4), np.float32)
    q[:, 0] = rng.random(n) * scene.width; q[:, 1] = rng.random(n) * scene.height
    q[:, 2:] = rng.random((n, 2))
    return q


@pytest.mark.parametrize("aperture,spp", [(0.0, 1), (0.08, 16)])
def test_filtered_lookups_match_reference(oracle, gpu_ctx_factory, aperture, spp):
    scene = scenes.bitmap_test_scene(64, 64, spp)          # the differentials are narrowed by 1 / sqrt(spp)
    scene.aperture_radius = aperture; scene.focus_depth = 6.0
    q = camera_samples(scene, 40000, 11)
    ref = oracle.Session(scene).texlookup(q)
    ctx = gpu_ctx_factory(scene)
    got = ctx.texture_lookup(q, spp)
    hits = ctx.trace(ctx.camera_rays(q), TRACE_CLOSEST)
    hit = hits["shape"] != 0xFFFFFFFF
    mats = np.array([scene.shapes[int(s)].material if s != 0xFFFFFFFF else -1 for s in hits["shape"]])
    tex = np.array([scene.materials[m].tex[0] if m >= 0 else 0 for m in mats])
    bitmap = np.array([t > 0 and scene.textures[t - 1].type == scenefile.TEX_BITMAP for t in tex])
    assert bitmap.sum() > 20000
    # its.uv
    assert np.allclose(got[hit, :2], ref[hit, :2], rtol=2e-5, atol=2e-6)
    # du/dx .. dv/dy: differences of two intersection points a pixel apart, so the last bits of the offset rays show:
    # 1e-4 of the largest partial of the row
    scale = np.abs(ref[:, 2:6]).max(axis=1, keepdims=True)
    perr = np.abs(got[:, 2:6] - ref[:, 2:6]) / (scale + 1e-3)
    print(f"aperture {aperture}: partials set on {np.mean(scale[:, 0] > 0):.3f} of the samples, max rel err {perr.max():.2e}")
    assert (scale[bitmap, 0] > 0).mean() > 0.95 and (scale[~bitmap, 0] == 0).all()
    assert np.array_equal((got[:, 2:6] == 0).all(axis=1), (ref[:, 2:6] == 0).all(axis=1))
    assert perr.max() < 1e-4
    # the filtered texture value, as the BSDF's diffuse reflectance. The filters are continuous in the partials except
    # where a weight-table index or a mip level flips (mipmap.cpp:273-286, 344-347): a handful of rows may differ.
    want = gc.diffuse_reflectance_from_slots(scene, hits, got[:, 6:9])
    err = np.abs(want - ref[:, 6:9]).max(axis=1)
    for kind in (scenefile.MIP_EWA, scenefile.MIP_TRILINEAR, scenefile.MIP_NONE):
        rows = bitmap & np.array([t > 0 and scene.textures[t - 1].p[0] == kind for t in tex])
        print(f"  filter {kind}: {rows.sum()} lookups, max err {err[rows].max():.2e}, outside 1e-5: {(err[rows] > 1e-5).sum()}")
        assert rows.sum() > 1000
        assert (err[rows] > 1e-5).mean() < 2e-3 and np.median(err[rows]) < 1e-7
    assert err[~bitmap].max() < 1e-6


@pytest.mark.parametrize("mode", [scenefile.MODE_PATH, scenefile.MODE_DIRECT])
def test_halton_render_with_bitmap_textures_matches_reference(oracle, gpu_ctx_factory, mode):
    scene = scenes.bitmap_test_scene(64, 64, 16, mode, max_depth=7)
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


def test_float_pixels_equal_byte_pixels_and_errors(gpu_ctx_factory):
    """MTSGPU_BITMAP_F32 (the EXRTexture route, exrtexture.cpp:52-60) against the byte route with gamma 1, whose table
    is i / 255 exactly; and the refusals."""
    scene = scenes.bitmap_test_scene(32, 32, 1)
    for t in scene.textures:
        if t.pixels is not None:
            t.gamma = 1.0
    q = camera_samples(scene, 5000, 3)
    a = gpu_ctx_factory(scene).texture_lookup(q)
    ctx = Context(0)
    ctx.upload_scene(scene, build=False)
    for i, t in enumerate(scene.textures):
        if t.pixels is None or t.pixels.dtype != np.uint8:        # the EXR texture is float already
            continue
        px = (t.pixels.astype(np.float32) / np.float32(255)).astype(np.float32)
        rgba = np.concatenate([px, np.ones(px.shape[:2] + (1,), np.float32)], axis=2)
        ctx._check(ctx.lib.mtsgpu_upload_bitmap(ctx.h, i, px.shape[1], px.shape[0], 4, ctx._ptr(np.ascontiguousarray(rgba)), api.BITMAP_F32, 0.0))
    ctx.build_accel()
    b = ctx.texture_lookup(q)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    with pytest.raises(api.MtsGpuError, match="not a bitmap texture"):
        ctx._check(ctx.lib.mtsgpu_upload_bitmap(ctx.h, 4, 4, 4, 3, ctx._ptr(np.zeros(48, np.uint8)), api.BITMAP_U8, -1.0))
    with pytest.raises(api.MtsGpuError, match="not supported"):
        ctx._check(ctx.lib.mtsgpu_upload_bitmap(ctx.h, 0, 4, 4, 5, ctx._ptr(np.zeros(80, np.uint8)), api.BITMAP_U8, -1.0))
    ctx.close()
    ctx = Context(0)
    missing = scenes.bitmap_test_scene(16, 16, 1)
    missing.textures[1].pixels = None
    with pytest.raises(api.MtsGpuError, match="pixels were not uploaded"):
        ctx.upload_scene(missing)
    ctx.close()
