"""Environment-map luminaire (src/luminaires/envmap.cpp): importance sampling through the coarse MIP
level, Le for rays that leave the scene, pdf for MIS — against the compiled reference's own envmap plugin reading the
OpenEXR file this library wrote."""
import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile
from mitsuba_renderer_b200.api import Context
from tests.test_shading_parity import close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rotate", [False, True])
def test_envmap_sampling_matches_reference(oracle, gpu_ctx_factory, rotate):
    """EnvMapLuminaire::sample / pdf (envmap.cpp:134-199) through Scene::sampleLuminaire / pdfLuminaire."""
    scene = scenes.envmap_test_scene(16, 16, 1, rotate=rotate)
    ctx = gpu_ctx_factory(scene)
    rng = np.random.Generator(np.random.PCG64(9))
    lo, hi = scenes.scene_bounds(scene)
    q = np.empty((40000, 5), np.float32)
    q[:, :3] = lo + (hi - lo) * rng.random((len(q), 3))
    q[:, 3:] = rng.random((len(q), 2))
    ref = oracle.Session(scene).emitter(q)
    gpu = ctx.emitter_sample(q)
    assert np.array_equal(ref[:, 13], gpu[:, 13]), "selected luminaire / success flag differs"
    env = ref[:, 13] == 3
    assert env.sum() > 8000
    has_n = (ref[:, 13] == 3) | (ref[:, 13] >= 5)          # the delta luminaires (0..2 and the collimated beam, 4) leave sRec.n unset
    assert (ref[:, 13] == 4).sum() > 100                     # points inside the beam (collimated.cpp:63-75); outside it the sample fails
    for c in range(15):
        rows = has_n if c in (3, 4, 5) else ref[:, 13] >= 0
        ok = close(gpu[rows, c], ref[rows, c], 2e-5, 2e-6 if c < 9 else 1e-6)
        assert ok.all(), f"column {c}: {np.count_nonzero(~ok)} mismatches, e.g. gpu={gpu[rows][~ok][0, c]} ref={ref[rows][~ok][0, c]}"
    # importance sampling: the density follows the map (a uniform one would be 1 / 4 pi everywhere)
    pdf = ref[env, 14]
    print(f"rotate {rotate}: {env.sum()} environment samples, pdf {pdf.min():.4f} .. {pdf.max():.2f}")
    assert pdf.max() > 20 * pdf.min()


@pytest.mark.parametrize("mode", [scenefile.MODE_PATH, scenefile.MODE_DIRECT])
def test_halton_render_with_envmap_matches_reference(oracle, gpu_ctx_factory, mode):
    scene = scenes.envmap_test_scene(64, 64, 16, mode, max_depth=7)
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
    assert (cpu[..., 3] == 0).any()               # some camera rays leave the scene and see the map directly


def test_envmap_refusals():
    scene = scenes.envmap_test_scene(16, 16, 1)
    scenefile.add_light(scene, scenefile.constant_light((0.1, 0.1, 0.1)))
    ctx = Context(0)
    with pytest.raises(api.MtsGpuError, match="at most one background"):
        ctx.upload_scene(scene)
    ctx.close()
    scene = scenes.envmap_test_scene(16, 16, 1)
    for e in scene.emitters:
        e.pixels = None
    ctx = Context(0)
    with pytest.raises(api.MtsGpuError, match="environment map was not uploaded"):
        ctx.upload_scene(scene)
    ctx.close()
