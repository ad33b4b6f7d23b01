"""Drop-in test of plugins/gpupath.so: the REAL Mitsuba PluginManager (compiled reference) dlopens the
plugin, Scene::render calls GPUPathIntegrator::render, the image comes back through Film::putImageBlock.
The scene is built by the harness exactly as SceneHandler would from
<integrator type="gpupath">; nothing else in the scene changes."""
import os

import numpy as np
import pytest

from mitsuba_renderer_b200 import scenes, scenefile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN_DIR = os.path.join(ROOT, "mitsuba-renderer_b200")
PLUGIN = os.path.join(PLUGIN_DIR, "plugins", "gpupath.so")


def develop(film):
    return film[..., :3] / np.maximum(film[..., 4:5], 1e-20)


def need_plugin():
    if not os.path.exists(PLUGIN):
        pytest.skip("plugins/gpupath.so is not built (needs the reference headers)")


def test_plugin_fails_loudly_without_a_gpu(oracle):
    """No CPU fallback: without a CUDA device the job is cancelled through the reference's own error path
    (Log(EError) -> exception -> RenderJob marks the job cancelled, renderjob.cpp:126-130)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    need_plugin()
    s = scenes.cornell_box(16, 16, 1)
    film, info = oracle.Session(s).render(1, "gpupath", env={"MTS_PLUGIN_PATH": PLUGIN_DIR})
    assert info["integrator"] == "gpupath" and not film.any()


@pytest.mark.gpu
@pytest.mark.parametrize("mode,lights", [(scenefile.MODE_PATH, False), (scenefile.MODE_DIRECT, False), (scenefile.MODE_PATH, True),
                                         (scenefile.MODE_PATH, "textures")])
def test_plugin_renders_through_mitsuba(oracle, mode, lights):
    need_plugin()
    # `lights`: point / spot / directional / constant luminaires, decoded by the plugin from Luminaire::serialize
    make = scenes.textured_test_scene if lights == "textures" else scenes.lights_test_scene if lights else scenes.material_test_scene
    s = make(64, 64, 256, mode, max_depth=8)       # "textures": checkerboards decoded from the BSDFs' serialised children
    if lights is True:
        s.sampler_type, s.sampler_depth = 1, 4          # ldsampler: the plugin reads `depth` from LowDiscrepancySampler::serialize
    s.filter_type = scenefile.FILTER_GAUSSIAN if mode == scenefile.MODE_DIRECT else scenefile.FILTER_BOX
    sess = oracle.Session(s)
    gpu, info = sess.render(2, "gpupath", env={"MTS_PLUGIN_PATH": PLUGIN_DIR})
    cpu, _ = sess.render()
    assert gpu[..., 4].sum() > 0, "the plugin produced no image (did the job get cancelled?)"
    assert np.isclose(gpu[..., 4].sum(), cpu[..., 4].sum(), rtol=1e-3)
    assert np.isclose(gpu[..., 3].sum(), cpu[..., 3].sum(), rtol=1e-2)
    g, c = develop(gpu), develop(cpu)
    assert abs(g.mean() - c.mean()) <= 0.02 * c.mean()
    for ch in range(3):
        assert abs(g[..., ch].mean() - c[..., ch].mean()) <= 0.03 * c[..., ch].mean() + 1e-4
    rel = float(np.mean((g - c) ** 2 / (c ** 2 + 1e-4)))
    print(f"mode {mode}: relMSE gpu-plugin vs cpu at 256 spp = {rel:.5f}; {info}")
    assert rel < 0.2


@pytest.mark.gpu
@pytest.mark.parametrize("make", [scenes.textured_test_scene, scenes.lights_test_scene, scenes.sphere_lights_test_scene, scenes.bitmap_test_scene, scenes.envmap_test_scene, scenes.instances_test_scene, scenes.masks_test_scene, scenes.cylinders_test_scene])
def test_plugin_with_the_halton_sampler_reproduces_the_cpu_image(oracle, make):
    """The deterministic case through the whole drop-in path: Mitsuba parses nothing differently, the plugin decodes
    BSDFs, textures, luminaires, sampler and camera from the live objects (rasterToCamera rebuilt with the reference's
    own Transform operations), renders on the GPU and hands the film back through ImageBlock / Film::putImageBlock —
    and the result has to be the CPU `path` integrator's image, pixel by pixel."""
    need_plugin()
    s = make(64, 64, 16, scenefile.MODE_PATH, max_depth=7)
    s.rr_depth = 3
    s.sampler_type = 3          # halton
    sess = oracle.Session(s)
    gpu, info = sess.render(2, "gpupath", env={"MTS_PLUGIN_PATH": PLUGIN_DIR})
    cpu, _ = sess.render()
    assert info["integrator"] == "gpupath" and gpu[..., 4].sum() > 0
    assert np.array_equal(gpu[..., 3:], cpu[..., 3:])
    g, c = develop(gpu), develop(cpu)
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    print(f"plugin vs cpu: within 2e-3: {(err < 2e-3).mean():.4f}, median {np.median(err):.2e}, means {c.mean():.6f} {g.mean():.6f}")
    assert (err < 2e-3).mean() > 0.999 and np.median(err) < 1e-6


@pytest.mark.gpu
def test_plugin_reads_the_scene_level_switches(oracle):
    """importanceSampleLuminaires = false with unequal sampling weights (scene.cpp:51; the plugin infers it, the member
    has no getter) and a shutter interval (the time sample of renderBlock): still the CPU image, pixel by pixel."""
    need_plugin()
    s = scenes.lights_test_scene(64, 64, 16, scenefile.MODE_PATH, max_depth=6)
    s.rr_depth = 3
    s.sampler_type = 3
    s.importance_sample_luminaires = False
    s.shutter_open, s.shutter_close = 0.0, 0.5
    sess = oracle.Session(s)
    gpu, info = sess.render(2, "gpupath", env={"MTS_PLUGIN_PATH": PLUGIN_DIR})
    cpu, _ = sess.render()
    assert info["integrator"] == "gpupath" and gpu[..., 4].sum() > 0
    assert np.array_equal(gpu[..., 3:], cpu[..., 3:])
    g, c = develop(gpu), develop(cpu)
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    assert (err < 2e-3).mean() > 0.999 and np.median(err) < 1e-6
