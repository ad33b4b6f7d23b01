"""Low-discrepancy sampler (src/samplers/ldsampler.cpp:104-196).

The reference draws its scrambles and shuffles from a per-thread MT19937 stream, so sample values cannot be
compared one to one. What is checked instead: (a) the structural guarantees of the construction — every pixel's
point set for a low-discrepancy draw is a scrambled (0,2)-net visited as a permutation — through the test hook
mtsgpu_sampler_values, and (b) statistical image parity against the reference rendering with its own ldsampler."""
import copy

import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile
from mitsuba_renderer_b200.api import Context

pytestmark = pytest.mark.gpu


def ld_params(scene, spp, depth=3, seed=1):
    s = copy.copy(scene); s.spp = spp; s.sampler_type = api.SAMPLER_LD; s.sampler_depth = depth
    return Context.params_from_scene(s, seed=seed)


def is_02_net(pts):
    """Every elementary interval of area 1/N (N = 2^k points) holds exactly one point."""
    n = len(pts)
    k = n.bit_length() - 1
    for a in range(k + 1):
        cx = np.minimum((pts[:, 0].astype(np.float64) * (1 << a)).astype(np.int64), (1 << a) - 1)
        cy = np.minimum((pts[:, 1].astype(np.float64) * (1 << (k - a))).astype(np.int64), (1 << (k - a)) - 1)
        if len(np.unique(cx * (1 << (k - a)) + cy)) != n:
            return False
    return True


def test_low_discrepancy_draws_are_scrambled_02_nets(gpu_ctx_factory):
    scene = scenes.cornell_box(32, 32, 64)
    ctx = gpu_ctx_factory(scene)
    p = ld_params(scene, 64)
    seen = []
    for pixel in (0, 5, 17 * 32 + 9):
        for draw in range(3):                                   # depth = 3: draws 0..2 are low-discrepancy
            v = ctx.sampler_values(p, pixel, draw)
            assert v.shape == (64, 2) and (v >= 0).all() and (v <= 1).all()
            assert is_02_net(v), (pixel, draw)
            seen.append(v)
        v = ctx.sampler_values(p, pixel, 3)                     # past the depth: independent samples
        assert not is_02_net(v)
    # different pixels / draws use different scrambles and visiting orders
    for i in range(len(seen)):
        for j in range(i + 1, len(seen)):
            assert not np.array_equal(seen[i], seen[j])
            assert not np.array_equal(np.sort(seen[i][:, 0]), np.sort(seen[j][:, 0]))
    # the visiting order is not the generation order: consecutive samples do not walk the van der Corput sequence
    v = seen[0]
    assert np.abs(np.diff(v[:, 0])).min() < 0.4 and len(np.unique(v[:, 0])) == 64
    # a sample-range shard sees the same values as the whole job (keys are per pixel, not per batch)
    whole = ctx.sampler_values(p, 5, 1)
    p2 = ld_params(scene, 64); p2.sppBegin, p2.sppEnd = 16, 32
    assert np.array_equal(ctx.sampler_values(p2, 5, 1), whole)
    # independent sampler: no structure, and the count need not be a power of two
    pi = Context.params_from_scene(scene, seed=1); pi.sppTotal = 48; pi.sppEnd = 48
    assert ctx.sampler_values(pi, 0, 0).shape == (48, 2)
    bad = ld_params(scene, 48)
    with pytest.raises(api.MtsGpuError, match="power of two"):  # ldsampler.cpp:61-65 rounds up with a warning; the ABI refuses
        ctx.render(bad)


def test_stratified_draws_cover_every_stratum_once(gpu_ctx_factory):
    """stratified.cpp:166-176: the R x R cells of the unit square, one jittered sample in each, in a shuffled order."""
    scene = scenes.cornell_box(32, 32, 36)
    ctx = gpu_ctx_factory(scene)
    s = copy.copy(scene); s.spp = 36; s.sampler_type = api.SAMPLER_STRATIFIED; s.sampler_depth = 2
    p = Context.params_from_scene(s, seed=4)
    orders = []
    for pixel in (0, 7, 300):
        for draw in range(2):
            v = ctx.sampler_values(p, pixel, draw)
            cells = (np.floor(v[:, 0] * 6).astype(int) % 6) + 6 * np.floor(v[:, 1] * 6).astype(int)
            assert sorted(cells) == list(range(36)), (pixel, draw)
            frac = v * 6 - np.floor(v * 6)
            assert frac.std() > 0.2                              # jittered, not cell centres
            orders.append(tuple(cells))
        v = ctx.sampler_values(p, pixel, 2)                      # past the depth: independent
        cells = np.floor(v[:, 0] * 6).astype(int) + 6 * np.floor(v[:, 1] * 6).astype(int)
        assert len(set(cells)) < 36
    assert len(set(orders)) == len(orders)                       # every pixel / draw has its own visiting order
    assert all(o != tuple(range(36)) for o in orders)
    bad = Context.params_from_scene(s, seed=4); bad.sppTotal = bad.sppEnd = 40
    with pytest.raises(api.MtsGpuError, match="perfect square"):
        ctx.render(bad)


def develop(film):
    w = film[..., 4:5]
    return np.where(w > 0, film[..., :3] / np.maximum(w, 1e-20), 0.0)


def rel_mse(a, b, eps=1e-4):
    return float(np.mean((a - b) ** 2 / (b ** 2 + eps)))


@pytest.mark.parametrize("mode,sampler", [(scenefile.MODE_DIRECT, api.SAMPLER_LD), (scenefile.MODE_PATH, api.SAMPLER_LD),
                                          (scenefile.MODE_DIRECT, api.SAMPLER_STRATIFIED)])
def test_ldsampler_render_matches_reference_statistics(oracle, gpu_ctx_factory, mode, sampler):
    """Same scene, sampler = ldsampler on both sides. The error against a converged image must match the CPU's at
    equal spp, and both must beat the independent sampler on direct lighting (that is the point of the sampler)."""
    spp = 16
    scene = scenes.material_test_scene(64, 64, spp, mode, max_depth=5)
    scene.sampler_type, scene.sampler_depth = sampler, 3
    ref_scene = copy.copy(scene); ref_scene.spp = 1024
    ref = develop(oracle.Session(ref_scene).render()[0])
    cpu = develop(oracle.Session(scene).render()[0])
    ctx = gpu_ctx_factory(scene)
    gpu = develop(ctx.render(Context.params_from_scene(scene, seed=1))[0])
    gpu_hi = develop(ctx.render(Context.params_from_scene(ref_scene, seed=2))[0])
    ind = copy.copy(scene); ind.sampler_type = api.SAMPLER_INDEPENDENT
    gpu_ind = develop(ctx.render(Context.params_from_scene(ind, seed=1))[0])
    # relMSE at 16 spp is heavy-tailed (measured over 12 GPU seeds / 8 CPU runs of the direct case: 0.020-0.063 and
    # 0.018-0.088, medians 0.024 on both sides): medians over several runs
    e_cpu = float(np.median([rel_mse(cpu, ref)] + [rel_mse(develop(oracle.Session(scene).render()[0]), ref) for _ in range(14)]))
    e_gpu = float(np.median([rel_mse(gpu, ref)] + [rel_mse(develop(ctx.render(Context.params_from_scene(scene, seed=k))[0]), ref) for k in range(3, 13)]))
    e_ind = float(np.median([rel_mse(gpu_ind, ref)] + [rel_mse(develop(ctx.render(Context.params_from_scene(ind, seed=k))[0]), ref) for k in (3, 5)]))
    print(f"mode {mode}: relMSE cpu-ld {e_cpu:.5f} gpu-ld {e_gpu:.5f} gpu-independent {e_ind:.5f}; "
          f"mean ref {ref.mean():.5f} gpu@1024 {gpu_hi.mean():.5f}")
    # (the CPU side's median still moves by +-15 % between runs — its streams depend on thread scheduling — while the GPU's is
    # a fixed set of seeds: measured over five runs of the stratified case 0.030-0.041 against a steady 0.041)
    assert e_gpu <= 1.4 * e_cpu + 1e-4
    assert abs(gpu_hi.mean() - ref.mean()) <= 0.01 * ref.mean()
    for ch in range(3):
        assert abs(gpu_hi[..., ch].mean() - ref[..., ch].mean()) <= 0.015 * ref[..., ch].mean() + 1e-4
    if mode == scenefile.MODE_DIRECT:
        assert e_gpu < 0.9 * e_ind


@pytest.mark.parametrize("mode,make", [(scenefile.MODE_PATH, scenes.textured_test_scene), (scenefile.MODE_DIRECT, scenes.textured_test_scene),
                                       (scenefile.MODE_PATH, scenes.lights_test_scene), (scenefile.MODE_DIRECT, scenes.lights_test_scene)])
def test_halton_render_matches_reference_pixel_by_pixel(oracle, gpu_ctx_factory, mode, make):
    """The halton sampler is deterministic (halton.cpp:68-84: value d of sample i is radicalInverse(prime d, i), the
    same for every pixel), so here the whole integrator can be compared with the reference without statistics: every
    pixel traces the same paths on both sides. A path whose direction differs in the last bit can still land on the
    other side of an edge, so a small share of pixels may differ; the rest must agree to fp32 accuracy. (With the camera
    transform restated in fp32 the camera rays are bit-identical, and so, it turns out, is nearly everything after.)"""
    scene = make(64, 64, 16, mode, max_depth=7)         # all BSDFs + textures + anisotropy, or all luminaire types + background
    scene.rr_depth = 3                                  # roulette values shift the primes of everything after them
    scene.sampler_type = api.SAMPLER_HALTON
    cpu = oracle.Session(scene).render()[0]
    ctx = gpu_ctx_factory(scene)
    gpu, stats = ctx.render(Context.params_from_scene(scene, seed=1))
    assert np.array_equal(gpu[..., 4], cpu[..., 4]) and np.array_equal(gpu[..., 3], cpu[..., 3])     # weights, alpha
    c, g = develop(cpu), develop(gpu)
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    close_px = err < 2e-3
    print(f"mode {mode}: pixels within 2e-3: {close_px.mean():.4f}, median rel err {np.median(err):.2e}, "
          f"mean image {c.mean():.6f} vs {g.mean():.6f}; {stats}")
    assert close_px.mean() > 0.999                      # measured: 1.0000
    assert np.median(err) < 1e-6                        # measured: 0 (more than half of the pixels are bit-identical)
    assert abs(g.mean() - c.mean()) < 1e-4 * c.mean()
    # another seed changes nothing: the sampler has no random state
    gpu2, _ = ctx.render(Context.params_from_scene(scene, seed=99))
    assert np.array_equal(gpu2, gpu)


@pytest.mark.parametrize("mode", [scenefile.MODE_PATH, scenefile.MODE_DIRECT])
def test_hammersley_render_matches_reference_pixel_by_pixel(oracle, gpu_ctx_factory, mode):
    """hammersley.cpp:72-99: as halton, but the first value of sample i is i / N and the primes start one value later.
    Thin lens on, so that first value is the lens sample's second coordinate."""
    scene = scenes.textured_test_scene(64, 64, 16, mode, max_depth=7)
    scene.rr_depth = 3
    scene.aperture_radius, scene.focus_depth = 0.05, 7.0
    scene.sampler_type = api.SAMPLER_HAMMERSLEY
    cpu = oracle.Session(scene).render()[0]
    gpu, _ = gpu_ctx_factory(scene).render(Context.params_from_scene(scene, seed=1))
    assert np.array_equal(gpu[..., 3:], cpu[..., 3:])
    c, g = develop(cpu), develop(gpu)
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    print(f"mode {mode}: pixels within 2e-3: {(err < 2e-3).mean():.4f}, median rel err {np.median(err):.2e}")
    assert (err < 2e-3).mean() > 0.999 and np.median(err) < 1e-6
    halton = copy.copy(scene); halton.sampler_type = api.SAMPLER_HALTON
    assert not np.array_equal(gpu_ctx_factory(halton).render(Context.params_from_scene(halton, seed=1))[0], gpu)


def _halton_variants():
    def thin_lens():
        s = scenes.cornell_box(64, 64, 16, max_depth=6)
        s.aperture_radius, s.focus_depth = 0.08, 9.0        # lens sample first: every later value moves two primes up
        s.filter_type = scenefile.FILTER_GAUSSIAN           # splats reach the neighbouring pixels
        s.crop = (5, 9, 50, 40)
        return s

    def strict_grid():
        s = scenes.direct_scene(cells=40, width=64, height=48, spp=16)
        s.mode, s.max_depth, s.rr_depth, s.strict_normals = scenefile.MODE_PATH, 5, 2, True
        return s

    def c3_geometry():
        s = scenes.direct_scene(width=96, height=54, spp=16)   # BASELINE config C3's 1M-triangle scene and integrator, small film
        return s
    def shutter():
        # a shutter interval: renderBlock draws a time sample between the lens and the pixel sample (integrator.cpp:159-160,
        # camera.h:64), which moves every later halton value one prime up
        s = scenes.textured_test_scene(64, 64, 16, scenefile.MODE_PATH, max_depth=7)
        s.rr_depth = 3
        s.aperture_radius, s.focus_depth = 0.05, 7.0
        s.shutter_open, s.shutter_close = 0.25, 0.75
        return s

    def shutter_direct():
        s = scenes.lights_test_scene(64, 64, 16, scenefile.MODE_DIRECT, max_depth=5)
        s.shutter_open, s.shutter_close = 0.0, 1.0
        return s

    def uniform_luminaire_choice():
        # Scene property importanceSampleLuminaires = false (scene.cpp:51,320-330,381-394) with unequal sampling weights
        s = scenes.lights_test_scene(64, 64, 16, scenefile.MODE_PATH, max_depth=6)
        assert len({e.sampling_weight for e in s.emitters}) > 1
        s.importance_sample_luminaires = False
        return s

    def black_at_depth_zero():
        s = scenes.cornell_box(32, 32, 4, max_depth=0)      # path.cpp:60: the loop never runs; only alpha is set
        return s
    return [("thin lens + gaussian + crop", thin_lens), ("strict normals on smooth terrain", strict_grid),
            ("C3 scene (1M triangles), direct", c3_geometry), ("shutter interval (time sample), path", shutter),
            ("shutter interval, direct", shutter_direct), ("importanceSampleLuminaires = false", uniform_luminaire_choice),
            ("maxDepth = 0", black_at_depth_zero)]


@pytest.mark.parametrize("name,make", _halton_variants())
def test_halton_render_variants_match_reference(oracle, gpu_ctx_factory, name, make):
    """More of the integrator's switches under the deterministic sampler: lens samples, a wide reconstruction filter
    with a crop window, strictNormals. (Unbounded depth is left out on purpose: under this sampler sample 0 draws the
    value 0 for every roulette decision and never dies, in the reference as here.)"""
    scene = make()
    scene.sampler_type = api.SAMPLER_HALTON
    cpu = oracle.Session(scene).render()[0]
    gpu, _ = gpu_ctx_factory(scene).render(Context.params_from_scene(scene, seed=1))
    assert np.allclose(gpu[..., 3:], cpu[..., 3:], rtol=1e-5, atol=1e-5)
    c, g = develop(cpu), develop(gpu)
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    print(f"{name}: pixels within 2e-3: {(err < 2e-3).mean():.4f}, median {np.median(err):.2e}, means {c.mean():.6f} {g.mean():.6f}")
    assert (err < 2e-3).mean() > 0.99 and np.median(err) < 5e-6
    assert abs(g.mean() - c.mean()) <= 1e-3 * c.mean()
    if "importanceSampleLuminaires" in name:         # and the flag does change the image
        on = copy.copy(scene); on.importance_sample_luminaires = True
        assert not np.array_equal(gpu_ctx_factory(on).render(Context.params_from_scene(on, seed=1))[0], gpu)
    if "shutter" in name:
        off = copy.copy(scene); off.shutter_open = off.shutter_close = 0.0
        assert not np.array_equal(gpu_ctx_factory(off).render(Context.params_from_scene(off, seed=1))[0], gpu)


def test_halton_prime_table_limit_is_reported(gpu_ctx_factory):
    """halton.cpp:74-84: the reference stops with "Lookup depth exceeds the prime number table size!" once a path has drawn
    1000 values. A mirror box with unbounded depth and no roulette gets there; the call must fail with that message
    (MTSGPU_ERR_LIMIT), not clamp the prime index silently."""
    # two large mirrors facing each other, the camera between them: every camera ray bounces several hundred times
    s = scenefile.SceneDesc()
    s.materials = [scenefile.mirror((1.0, 1.0, 1.0))]
    for z, toward in ((0.0, (0, 0, 1)), (1.0, (0, 0, 0))):
        p, t, n = scenes.quad((-100, -100, z), (100, -100, z), (100, 100, z), (-100, 100, z), toward=toward)
        s.shapes.append(scenefile.Shape(p, t, n, material=0))
    s.camera_to_world = scenefile.look_at((0.0, 0.0, 0.5), (0.0, 0.0, 1.0), (0, 1, 0))
    s.fov, s.width, s.height, s.spp = 20.0, 16, 16, 2
    s.mode, s.max_depth, s.rr_depth = scenefile.MODE_PATH, -1, 100000
    s.filter_type = scenefile.FILTER_BOX
    s.sampler_type = api.SAMPLER_HALTON
    ctx = gpu_ctx_factory(s)
    with pytest.raises(api.MtsGpuError, match="prime number table"):
        ctx.render(Context.params_from_scene(s, seed=1))
    s2 = copy.copy(s); s2.max_depth = 200           # 2 + 4 * 199 values: inside the table
    ctx.render(Context.params_from_scene(s2, seed=1))
