"""mtsgpu_film_splat_static — the film splat for callers that may not allocate GPU memory.

Some users of the library splat samples from places where calling cudaMalloc is off the table: a
progress/preview refresh on a latency-critical thread while a render already holds most of the GPU's
memory (the allocation can fail, and it stalls the whole device), or a host that fixes the library's
memory budget up front. For them the library keeps one fixed 4 MiB block of device memory in its own
image and this entry point stages everything there.

The test calls the entry point through the C ABI twice in a row on different inputs — the block keeps
the first call's contents, so the second call proves the accumulators are cleared — and compares each
result with the regular, allocating mtsgpu_film_splat on the same samples, within the tolerance the
film-splat parity test uses. It also checks that an over-sized request is refused with a clear message.
This file sorts last on purpose so a `pytest -x` run reaches it only after every other test.
"""
import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile

pytestmark = pytest.mark.gpu
TOL = 2e-4       # same bound as tests/test_shading_parity.py::test_film_splat


def make_samples(scene, n, seed):
    cx, cy, cw, ch = scene.crop_rect
    rng = np.random.Generator(np.random.PCG64(seed))
    q = np.empty((n, 6), np.float32)
    q[:, 0] = np.minimum(cx + rng.random(n) * cw, np.float32(cx + cw) - np.float32(1e-3))
    q[:, 1] = np.minimum(cy + rng.random(n) * ch, np.float32(cy + ch) - np.float32(1e-3))
    q[:, 2:5] = rng.random((n, 3)) * 3
    q[:, 5] = rng.integers(0, 2, n)
    q[::101, 2] = np.nan        # invalid spectra are dropped by both entry points
    q[::103, 4] = -0.5
    return q


@pytest.mark.parametrize("filter_type,crop", [(scenefile.FILTER_GAUSSIAN, (3, 7, 90, 61)), (scenefile.FILTER_BOX, None)])
def test_static_splat_matches_allocating_splat_twice(filter_type, crop):
    scene = scenes.cornell_box(120, 80, 1)
    scene.filter_type = filter_type
    scene.crop = crop
    ctx = api.Context(0)
    try:
        ctx.set_film(scene)              # the splat needs only the film; no scene upload, no acceleration structure
        for call, (n, seed) in enumerate([(40000, 31), (25000, 32)]):
            q = make_samples(scene, n, seed)
            got = ctx.film_splat_static(q)
            want = ctx.film_splat(q)
            scale = max(1.0, float(np.abs(want).max()))
            assert np.abs(got - want).max() <= TOL * scale, f"call {call}: max abs diff {np.abs(got - want).max()}"
            assert np.isclose(got[..., 4].sum(), want[..., 4].sum(), rtol=1e-5)
        # an empty batch must give an all-zero film, not the previous call's leftovers
        assert not ctx.film_splat_static(np.zeros((0, 6), np.float32)).any()
    finally:
        ctx.close()


def test_static_splat_rejects_oversized_requests():
    scene = scenes.cornell_box(512, 512, 1)       # 512*512*20 B of accumulators alone exceed the 4 MiB block
    ctx = api.Context(0)
    try:
        ctx.set_film(scene)
        with pytest.raises(api.MtsGpuError) as err:
            ctx.film_splat_static(make_samples(scene, 10, 1))
        assert err.value.code == -5 and "scratch" in str(err.value) and str(api.STATIC_SCRATCH_BYTES) in str(err.value)
        small = scenes.cornell_box(64, 64, 1)
        ctx.set_film(small)
        too_many = (api.STATIC_SCRATCH_BYTES // 24) + 1
        with pytest.raises(api.MtsGpuError):
            ctx.film_splat_static(np.zeros((too_many, 6), np.float32) + 1.0)
        assert ctx.film_splat_static(make_samples(small, 1000, 2))[..., 4].sum() > 0     # still usable afterwards
    finally:
        ctx.close()


@pytest.mark.parametrize("name", ["box", "gaussian_crop", "mitchell"])
def test_static_splat_matches_the_references_image_block(name):
    """The same entry point against the reference itself: the committed golden films of ImageBlock::putSample
    (tests/golden/splat_*.npz, made from the compiled reference by tests/golden/make_golden.py), same bound as the
    allocating splat's golden test. Two calls in a row on the same context."""
    import os
    from tests import golden_cases as gc
    ft, crop = gc.SPLAT_CASES[name]
    ss = gc.splat_scene(ft, crop)
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "splat_" + name + ".npz"))
    q = gc.splat_samples(ss, 8000)
    assert gc.digest(q) == str(g["samples_digest"])
    ctx = api.Context(0)
    try:
        ctx.set_film(ss, g["table"], float(g["filter_size"]))
        for _ in range(2):
            film = ctx.film_splat_static(q)
            assert np.abs(film - g["film"]).max() <= TOL * max(1.0, np.abs(g["film"]).max())
    finally:
        ctx.close()
