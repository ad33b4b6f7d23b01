"""Seeded inputs of the golden fixtures under tests/golden/. Shared by the generator
(tests/golden/make_golden.py, which runs the compiled reference) and by the tests that replay them."""
import numpy as np

from mitsuba_renderer_b200 import scenes, scenefile


def scene_materials():
    return scenes.material_test_scene(32, 32, 1)


def scene_lights():
    return scenes.lights_test_scene(32, 32, 1)


def scene_textured():
    return scenes.textured_test_scene(32, 32, 1)


def scene_halton():
    s = scenes.textured_test_scene(48, 48, 16, scenefile.MODE_PATH, max_depth=7)
    s.rr_depth = 3
    s.sampler_type = 3          # halton
    return s


def scene_grid():
    return scenes.microbench_scene(cells=64, pad_cells=9)


def scene_smoke():
    return scenes.cornell_box(64, 64, 16)


def trace_rays(scene, n=3000, seed=11):
    """Six ray families: kdbench incoherent, pinhole coherent, interior origins, axis-parallel, box-missing,
    finite intervals. Returns one [N,8] array."""
    rng = np.random.Generator(np.random.PCG64(seed))
    inc = scenes.incoherent_rays(scene, n, seed)
    coh = scenes.coherent_rays(scene, int(np.sqrt(n)))
    lo, hi = scenes.scene_bounds(scene)
    o = lo + (hi - lo) * rng.random((n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    inside = np.empty((n, 8), np.float32)
    inside[:, :3] = o; inside[:, 3] = 1e-4; inside[:, 4:7] = d; inside[:, 7] = np.inf
    axis = inside[: n // 4].copy()
    ax = rng.integers(0, 3, len(axis))
    axis[:, 4:7] = 0
    axis[np.arange(len(axis)), 4 + ax] = rng.choice([-1.0, 1.0], len(axis))
    miss = inc[: n // 4].copy(); miss[:, :3] += (hi - lo).max() * 10
    finite = inc[: n // 4].copy()
    finite[:, 7] = rng.random(len(finite)).astype(np.float32) * np.float32(np.linalg.norm(hi - lo))
    finite[:, 3] = 1e-3
    return np.ascontiguousarray(np.vstack([inc, coh, inside, axis, miss, finite]), np.float32)


def bsdf_scene():
    s = scenes.cornell_box(16, 16, 1)
    s.materials = [scenefile.lambertian((0.5, 0.3, 0.7)),
                   scenefile.microfacet((0.5, 0.5, 0.5), (0.5, 0.5, 0.5), 0.5, 0.5, 0.1),      # data/tests/test_bsdf.xml
                   scenefile.microfacet((0.2, 0.4, 0.1), (1, 0.9, 0.8), 1.0, 1.0, 0.3, 1.33, 1.0),
                   scenefile.mirror((0.8, 0.7, 0.9)),
                   scenefile.roughmetal((1, 1, 1), 0.1),                                         # data/tests/test_bsdf.xml
                   scenefile.roughmetal((0.9, 0.8, 0.7), 0.4, (0.2, 0.92, 1.1), (3.9, 2.45, 2.14)),
                   scenefile.dielectric(), scenefile.dielectric(1.33, 1.0, (0.9, 1, 1), (1, 0.8, 0.9)),
                   scenefile.twosided(scenefile.lambertian((0.2, 0.6, 0.4))),                     # twosided.cpp adapter
                   scenefile.twosided(scenefile.roughmetal((1, 1, 1), 0.2)),
                   scenefile.twosided(scenefile.microfacet((0.3, 0.3, 0.3), (1, 1, 1), 1.0, 1.0, 0.15)),
                   scenefile.phong((0.5, 0.4, 0.3), (0.2, 0.3, 0.4), 1.0, 1.0, 10.0),                # phong.cpp defaults
                   scenefile.twosided(scenefile.phong((0.3, 0.3, 0.3), (0.6, 0.6, 0.6), 0.9, 0.8, 60.0, 0.7)),
                   scenefile.roughglass((0.9, 1.0, 0.8), (0.8, 0.9, 1.0), 0.1, 1.5046, 1.0, "beckmann"),  # roughglass.cpp defaults
                   scenefile.roughglass((1, 1, 1), (1, 1, 1), 0.3, 1.33, 1.0, "ggx"),
                   scenefile.roughglass((1, 1, 1), (0.9, 0.9, 0.9), 0.25, 1.0, 1.5, "phong"),
                   scenefile.ward((0.5, 0.5, 0.5), (0.2, 0.2, 0.2), 1.0, 1.0, 0.1, 0.1, "balanced"),     # ward.cpp defaults
                   scenefile.ward((0.3, 0.2, 0.1), (0.5, 0.6, 0.7), 0.8, 0.9, 0.05, 0.3, "ward"),
                   scenefile.twosided(scenefile.ward((0.1, 0.2, 0.3), (0.7, 0.6, 0.5), 1.0, 1.0, 0.25, 0.08, "ward-duer", 0.6)),
                   scenefile.difftrans((0.6, 0.7, 0.4)),                                           # difftrans.cpp
                   scenefile.masked(scenefile.lambertian((0.7, 0.6, 0.5)), (0.3, 0.6, 0.9)),        # mask.cpp
                   scenefile.masked(scenefile.twosided(scenefile.roughmetal((1, 1, 1), 0.2)), (0.8, 0.8, 0.8))]
    for sh in s.shapes:
        sh.material = 0
    return s


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
    for c in (slice(1, 4), slice(4, 7)):
        q[:, c] /= np.linalg.norm(q[:, c].astype(np.float64), axis=1, keepdims=True).astype(np.float32)
    return q


def emitter_queries(scene, n, seed=4):
    rng = np.random.Generator(np.random.PCG64(seed))
    lo, hi = scenes.scene_bounds(scene)
    q = np.empty((n, 5), np.float32)
    q[:, :3] = lo + (hi - lo) * rng.random((n, 3))
    q[:, 3:] = rng.random((n, 2))
    return q


def camera_scene(aperture):
    s = scenes.cornell_box(200, 120, 1)
    s.aperture_radius = aperture
    s.focus_depth = 9.0
    return s


def camera_samples(scene, n, seed=8):
    q = np.random.Generator(np.random.PCG64(seed)).random((n, 4)).astype(np.float32)
    q[:, 0] *= scene.width; q[:, 1] *= scene.height
    return q


def splat_scene(filter_type, crop):
    s = scenes.cornell_box(70, 53, 1)
    s.filter_type = filter_type
    s.crop = crop
    return s


def splat_samples(scene, n, seed=12):
    cx, cy, cw, ch = scene.crop_rect
    rng = np.random.Generator(np.random.PCG64(seed))
    q = np.empty((n, 6), np.float32)
    q[:, 0] = np.minimum(cx + rng.random(n) * cw, np.float32(cx + cw) - np.float32(1e-3))
    q[:, 1] = np.minimum(cy + rng.random(n) * ch, np.float32(cy + ch) - np.float32(1e-3))
    q[:, 2:5] = rng.random((n, 3)) * 3; q[:, 5] = rng.integers(0, 2, n)
    q[::97, 2] = np.nan; q[::89, 3] = -1.0
    q[:50, 0] = np.floor(q[:50, 0]) + 0.5
    return q


SPLAT_CASES = {"box": (scenefile.FILTER_BOX, None), "gaussian": (scenefile.FILTER_GAUSSIAN, None),
               "gaussian_crop": (scenefile.FILTER_GAUSSIAN, (5, 9, 50, 37)),
               "mitchell": (scenefile.FILTER_MITCHELL, None), "catmullrom": (scenefile.FILTER_CATMULLROM, (3, 2, 40, 45)),
               "wsinc": (scenefile.FILTER_WSINC, None)}          # negative lobes: weights below zero


def render_scene():
    return scenes.cornell_box(48, 48, 64, max_depth=8)


def digest(*arrays):
    """Cheap content hash to notice when a generator drifts away from the stored fixtures."""
    import hashlib
    h = hashlib.sha1()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()[:16]


# ---- `.serialized` meshes -----------------------------------------------------------------------------------
SERIALIZED_FIXTURE = "/root/reference/data/blender/mitsuba/matpreview/matpreview.serialized"   # the reference's own file
SERIALIZED_TO_WORLD = np.array([[0, -2, 0, 1], [1.5, 0, 0, 2], [0, 0, -1, 3], [0, 0, 0, 1]], np.float32)   # det < 0: winding swaps


def serialized_meshes():
    """Two small meshes for the writer / reader / computeNormals cases: an icosphere without normals (so the
    loader has to compute them) and the Cornell short block with normals and texture coordinates."""
    v, f = scenes.icosphere(2)
    ico = {"positions": (v * 0.7 + np.array([1.9, 3.2, 1.6])).astype(np.float32), "triangles": f.astype(np.uint32)}
    sh = scenes.cornell_box(16, 16, 1).shapes[3]
    uv = (sh.positions[:, :2] * np.float32(0.25)).astype(np.float32)
    block = {"positions": sh.positions, "triangles": sh.triangles, "normals": sh.normals, "texcoords": uv}
    return [ico, block]


def mesh_digest(m):
    parts = [np.ascontiguousarray(m["positions"], np.float32), np.ascontiguousarray(m["triangles"], np.uint32)]
    for k in ("normals", "texcoords"):
        if m.get(k) is not None:
            parts.append(np.ascontiguousarray(m[k], np.float32))
    return digest(*parts)


# ---- film development ------------------------------------------------------------------------------------------
def film_accumulators(h=9, w=13, seed=31):
    """Accumulators {sum rgb, sum alpha, sum weight} with the awkward cases: zero weight, tiny and huge values,
    a negative lobe (gaussian filters never produce one, tabulated ones with negative lobes would)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    a = np.empty((h, w, 5), np.float32)
    wt = rng.integers(1, 64, (h, w)).astype(np.float32) * np.float32(0.37)
    a[..., :3] = rng.random((h, w, 3)).astype(np.float32) * wt[..., None] * np.float32(2.5)
    a[..., 3] = rng.random((h, w)).astype(np.float32) * wt
    a[..., 4] = wt
    a[0, 0] = 0                                   # untouched pixel: weight 0
    a[1, 2, :3] = (1e-7, 3e-6, 2e-8)
    a[2, 3, :3] = (7e4, 1.2e5, 3e3); a[2, 3, 4] = 1.0      # beyond the half range
    a[3, 1, :3] = (-0.5, 0.25, -1e-3)
    a[4, 4, 4] = 0.0                              # colour without weight
    return a


def diffuse_reflectance_from_slots(scene, hits, slots):
    """What BSDF::getDiffuseReflectance(its) returns, from the hit's material and its (texture-resolved) slot 0:
    lambertian: reflectance (lambertian.cpp:91-93); microfacet / phong / ward: diffuseReflectance * kd; zero for the rest."""
    out = np.zeros((len(hits), 3), np.float32)
    for i, h in enumerate(hits):
        if h["shape"] == 0xFFFFFFFF:
            continue
        mid = scene.shapes[int(h["shape"])].material
        if mid < 0:
            continue
        m = scene.materials[mid]
        t = m.type & 0xFF
        if m.type & 0x100:          # twosided forwards to the nested BSDF
            pass
        if t == 0:
            out[i] = slots[i]
        elif t in (1, 5, 7):
            out[i] = (slots[i] * np.float32(m.p[6])).astype(np.float32)
    return out


def scene_bitmaps():
    s = scenes.bitmap_test_scene(64, 64, 4)
    s.aperture_radius = 0.05; s.focus_depth = 6.0
    return s


def lens_camera_samples(scene, n, seed):
    """rows {raster x, raster y, lens x, lens y}"""
    rng = np.random.Generator(np.random.PCG64(seed))
    q = np.empty((n, 4), np.float32)
    q[:, 0] = rng.random(n) * scene.width; q[:, 1] = rng.random(n) * scene.height
    q[:, 2:] = rng.random((n, 2))
    return q


def scene_envmap():
    return scenes.envmap_test_scene(16, 16, 1)


def scene_spheres():
    return scenes.sphere_lights_test_scene(32, 32, 1)


def scene_instances():
    return scenes.instances_test_scene(32, 32, 1)
