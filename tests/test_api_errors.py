"""The C ABI refuses what it does not implement — loudly, with the reference's wording where the reference has an
error for the same situation — instead of approximating it."""
import copy

import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile
from mitsuba_renderer_b200.api import Context, MtsGpuError

pytestmark = pytest.mark.gpu


def fresh(scene, build=True):
    ctx = Context(0)
    ctx.upload_scene(scene, build=build)
    return ctx


def test_unsupported_scene_contents_are_refused():
    base = scenes.material_test_scene(16, 16, 1)
    cases = []
    s = copy.deepcopy(base); s.materials[0] = scenefile.Material(42, np.zeros(16, np.float32)); cases.append((s, "unsupported BSDF type"))
    s = copy.deepcopy(base); s.materials[7] = scenefile.Material(scenefile.MAT_DIELECTRIC | scenefile.BSDF_TWOSIDED, s.materials[7].p)
    cases.append((s, "only BRDF child instances"))                      # twosided.cpp:72-74
    s = copy.deepcopy(base); s.textures = [scenefile.checkerboard()]; s.materials[5] = scenefile.textured(s.materials[5], 0, 0)
    cases.append((s, "cannot be textured"))                             # mirror keeps a plain Spectrum (mirror.cpp:32-34)
    s = copy.deepcopy(base); s.materials[0] = scenefile.textured(s.materials[0], 0, 3); cases.append((s, "was not uploaded"))
    s = copy.deepcopy(base); s.materials.append(scenefile.ward(alpha_x=0.05, alpha_y=0.3)); s.shapes[3].material = len(s.materials) - 1
    cases.append((s, "texture coordinates are required"))               # trimesh.cpp:549-556
    s = copy.deepcopy(base); s.emitters[1] = scenefile.point_light((1, 1, 1)); cases.append((s, "not an area luminaire"))
    s = copy.deepcopy(base); s.shapes[5].emitter = -1; cases.append((s, "not bound to any shape"))
    s = copy.deepcopy(base); scenefile.add_light(s, scenefile.constant_light()); scenefile.add_light(s, scenefile.constant_light())
    cases.append((s, "at most one background"))                         # scene.cpp:608-611
    s = copy.deepcopy(base); e = scenefile.spot_light((0, 5, 0), (0, 0, 0), (0, 0, 1), cutoff_angle=10.0, beam_width=20.0)
    scenefile.add_light(s, e); cases.append((s, "cutoffAngle must be >= beamWidth"))   # spot.cpp:42
    s = copy.deepcopy(base); s.shapes[2].triangles = s.shapes[2].triangles.copy(); s.shapes[2].triangles[0, 0] = 999
    cases.append((s, "references vertex"))
    for scene, message in cases:
        ctx = Context(0)
        with pytest.raises(MtsGpuError, match=message):
            ctx.upload_scene(scene)
        ctx.close()


def test_calls_out_of_order_and_bad_render_parameters():
    scene = scenes.cornell_box(16, 16, 4)
    ctx = fresh(scene, build=False)
    rays = scenes.incoherent_rays(scene, 16, 1)
    with pytest.raises(MtsGpuError, match="build_accel first"):
        ctx.trace(rays, api.TRACE_CLOSEST)
    with pytest.raises(MtsGpuError, match="build_accel first"):
        ctx.render(Context.params_from_scene(scene))
    ctx.build_accel()
    with pytest.raises(MtsGpuError, match="unknown mode"):
        ctx.trace(rays, 7)
    assert len(ctx.trace(np.zeros((0, 8), np.float32), api.TRACE_CLOSEST)) == 0          # an empty batch is not an error
    p = Context.params_from_scene(scene); p.mode = 5
    with pytest.raises(MtsGpuError, match="unknown mode"):
        ctx.render(p)
    p = Context.params_from_scene(scene); p.sppBegin, p.sppEnd = 3, 2
    with pytest.raises(MtsGpuError, match="bad sample range"):
        ctx.render(p)
    p = Context.params_from_scene(scene); p.samplerType = 9
    with pytest.raises(MtsGpuError, match="unknown sampler"):
        ctx.render(p)
    d = copy.copy(scene); d.mode = scenefile.MODE_DIRECT; d.luminaire_samples = d.bsdf_samples = 0
    with pytest.raises(MtsGpuError, match="must be > 0"):                # direct.cpp:38 Assert
        ctx.render(Context.params_from_scene(d))
    d.luminaire_samples, d.bsdf_samples, d.sampler_type = 2, 1, api.SAMPLER_HALTON
    with pytest.raises(MtsGpuError, match="not supported by QMC samplers"):   # halton.cpp:99-105
        ctx.render(Context.params_from_scene(d))
    bad = copy.copy(scene); bad.crop = (10, 10, 10, 10)                  # film.cpp:40-45
    with pytest.raises(MtsGpuError, match="Invalid crop window"):
        ctx.set_film(bad)
    # an empty sample range renders nothing and says so
    p = Context.params_from_scene(scene); p.sppBegin = p.sppEnd = 2
    film, stats = ctx.render(p)
    assert not film.any() and stats["cameraRays"] == 0
    ctx.close()
