"""Deterministic synthetic scenes and ray sets for the BASELINE.json configs.

Everything here is seeded and produced with numpy only; the same `SceneDesc` goes to the CUDA
library (through the C ABI) and, via `scenefile.write_scene`, to the compiled reference.
"""
from __future__ import annotations

import numpy as np

from .scenefile import (SceneDesc, Shape, Emitter, lambertian, microfacet, mirror, roughmetal,
                        dielectric, twosided, phong, roughglass, look_at, point_light, spot_light, directional_light,
                        constant_light, envmap_light, collimated_light, add_light, checkerboard, textured, ward, sphere_shape, instance_shape, cylinder_shape, bitmap_texture, exr_texture, grid_texture, masked, difftrans, MIP_EWA, MIP_TRILINEAR, MIP_NONE,
                        WRAP_REPEAT, WRAP_BLACK, WRAP_WHITE, WRAP_CLAMP, MODE_PATH, MODE_DIRECT, FILTER_BOX)

F32 = np.float32


# --------------------------------------------------------------------------------------------
# mesh helpers
# --------------------------------------------------------------------------------------------
def quad(p0, p1, p2, p3, toward=None, away_from=None):
    """Two triangles (p0,p1,p2),(p0,p2,p3) with explicit per-vertex normals = face normal.
    The winding is flipped so that the normal points toward `toward` / away from `away_from`."""
    P = np.array([p0, p1, p2, p3], np.float64)
    n = np.cross(P[1] - P[0], P[2] - P[0])
    n /= np.linalg.norm(n)
    c = P.mean(0)
    flip = False
    if toward is not None:
        flip = np.dot(np.asarray(toward, np.float64) - c, n) < 0
    if away_from is not None:
        flip = np.dot(np.asarray(away_from, np.float64) - c, n) > 0
    if flip:
        P = P[::-1].copy()
        n = -n
    tris = np.array([[0, 1, 2], [0, 2, 3]], np.uint32)
    return P.astype(F32), tris, np.tile(n.astype(F32), (4, 1))


def merge(parts):
    """Concatenate (positions, triangles, normals) tuples into one mesh."""
    pos, tri, nrm, base = [], [], [], 0
    for p, t, n in parts:
        pos.append(p); nrm.append(n); tri.append(t + np.uint32(base)); base += len(p)
    return np.concatenate(pos).astype(F32), np.concatenate(tri).astype(np.uint32), np.concatenate(nrm).astype(F32)


def box_faces(corners_bottom, height_pts, center):
    """Five faces (top + 4 sides) of a block given its 4 top corners (Cornell data layout)."""
    top = [np.asarray(p, np.float64) for p in height_pts]
    bot = [np.array([p[0], 0.0, p[2]]) for p in top]
    parts = [quad(top[0], top[1], top[2], top[3], away_from=center)]
    for i in range(4):
        j = (i + 1) % 4
        parts.append(quad(top[i], top[j], bot[j], bot[i], away_from=center))
    return parts


def value_noise(x, y, seed, octaves=5, base_freq=4.0):
    """Seeded lattice value noise with smoothstep interpolation, summed over octaves."""
    out = np.zeros_like(x, dtype=np.float64)
    amp, freq = 1.0, base_freq
    for o in range(octaves):
        n = int(freq) + 2
        lattice = np.random.Generator(np.random.PCG64(seed * 1000 + o)).random((n, n))
        fx, fy = x * freq, y * freq
        ix, iy = np.floor(fx).astype(np.int64), np.floor(fy).astype(np.int64)
        tx, ty = fx - ix, fy - iy
        tx, ty = tx * tx * (3 - 2 * tx), ty * ty * (3 - 2 * ty)
        ix0, iy0 = ix % n, iy % n
        ix1, iy1 = (ix + 1) % n, (iy + 1) % n
        v = (lattice[iy0, ix0] * (1 - tx) + lattice[iy0, ix1] * tx) * (1 - ty) + \
            (lattice[iy1, ix0] * (1 - tx) + lattice[iy1, ix1] * tx) * ty
        out += amp * (v - 0.5)
        amp *= 0.5; freq *= 2.0
    return out


def displaced_grid(cells_x, cells_z, seed=1, extent=1.0, amplitude=0.25, x0=None, z0=None, y0=0.0):
    """(cells_x x cells_z)-cell heightfield over [-extent, extent]^2 in the xz-plane, y = noise.
    Two triangles per cell, smooth vertex normals from central differences. Normals face +y."""
    nx, nz = cells_x + 1, cells_z + 1
    xs = np.linspace(-extent if x0 is None else x0[0], extent if x0 is None else x0[1], nx)
    zs = np.linspace(-extent if z0 is None else z0[0], extent if z0 is None else z0[1], nz)
    X, Z = np.meshgrid(xs, zs, indexing="xy")     # [nz, nx]
    u, v = (X + extent) / (2 * extent), (Z + extent) / (2 * extent)
    Y = y0 + amplitude * value_noise(u, v, seed)
    pos = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(F32)
    # normals via finite differences of the (float32-rounded) heights
    Yf = pos[:, 1].reshape(nz, nx).astype(np.float64)
    dYdx = np.gradient(Yf, xs, axis=1) if nx > 2 else np.zeros_like(Yf)
    dYdz = np.gradient(Yf, zs, axis=0) if nz > 2 else np.zeros_like(Yf)
    N = np.stack([-dYdx, np.ones_like(Yf), -dYdz], -1)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    nrm = N.reshape(-1, 3).astype(F32)
    i, j = np.meshgrid(np.arange(cells_x), np.arange(cells_z), indexing="xy")
    v00 = (j * nx + i).ravel(); v10 = v00 + 1; v01 = v00 + nx; v11 = v01 + 1
    # winding chosen so cross(p1-p0, p2-p0) has +y
    t0 = np.stack([v00, v01, v11], -1); t1 = np.stack([v00, v11, v10], -1)
    tri = np.empty((2 * len(v00), 3), np.uint32)
    tri[0::2] = t0; tri[1::2] = t1
    return pos, tri, nrm


def icosphere(subdiv=2):
    """Unit icosphere: 20 * 4**subdiv triangles (subdiv=2 -> 320), outward normals = positions."""
    t = (1.0 + 5 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5],
                  [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(subdiv):
        verts = list(map(tuple, v)); cache = {}; nf = []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = (np.array(verts[a]) + np.array(verts[b])) / 2.0
                m /= np.linalg.norm(m)
                verts.append(tuple(m)); cache[key] = len(verts) - 1
            return cache[key]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        v = np.array(verts, np.float64); f = np.array(nf, np.int64)
    return v, f


def textured_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """The all-BSDF scene with texture coordinates on every mesh and checkerboard textures in the reflectance slots of
    the lambertian, microfacet, roughmetal, phong and roughglass materials (different scales and offsets)."""
    s = material_test_scene(width, height, spp, mode, max_depth)
    for sh in s.shapes:
        P = sh.positions.astype(F32)
        sh.texcoords = np.stack([P[:, 0] * F32(0.7) + P[:, 1] * F32(0.3), P[:, 2] * F32(0.6) - P[:, 1] * F32(0.2)], 1).astype(F32)
    s.textures = [checkerboard((0.75, 0.7, 0.6), (0.15, 0.2, 0.3), 2.0, 2.0),
                  checkerboard((0.9, 0.3, 0.2), (0.2, 0.8, 0.3), 3.0, 1.5, 0.25, -0.4),
                  checkerboard((1.0, 1.0, 1.0), (0.5, 0.6, 0.9), 6.0, 6.0, 0.1, 0.1)]
    m = s.materials
    m[1] = textured(m[1], 0, 1)        # red wall: lambertian
    m[2] = textured(m[2], 0, 0)        # green wall: lambertian
    m[4] = textured(m[4], 0, 0)        # walls: microfacet, diffuse slot
    m[6] = textured(m[6], 0, 2)        # short block: roughmetal specularReflectance
    m[9] = textured(textured(m[9], 0, 1), 1, 2)    # panel: phong, both slots
    m[10] = textured(m[10], 1, 2)      # ball: roughglass specularTransmittance
    # anisotropic Ward on the smooth sphere (tangent frames from the texture coordinates) and on the tall block
    m.append(textured(ward((0.25, 0.2, 0.15), (0.6, 0.6, 0.6), 1.0, 1.0, 0.05, 0.25, "balanced"), 0, 1))
    m.append(ward((0.2, 0.3, 0.4), (0.5, 0.5, 0.5), 1.0, 1.0, 0.3, 0.06, "ward"))
    s.shapes[6].material = len(m) - 2
    s.shapes[4].material = len(m) - 1
    return s


def spheres_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """The textured all-BSDF scene plus analytic spheres (src/shapes/sphere.cpp): a glass ball, a mirror ball, a
    checkerboard-textured lambertian ball (uv from the sphere's own parametrisation), an anisotropic Ward ball (its
    dpdu frame) and an inverted sphere the camera looks into."""
    s = textured_test_scene(width, height, spp, mode, max_depth)
    m = s.materials
    m.append(textured(lambertian((0.5, 0.5, 0.5)), 0, 2))
    tex_ball = len(m) - 1
    s.shapes.append(sphere_shape((4.4, 0.6, 1.0), 0.6, material=7))              # dielectric
    s.shapes.append(sphere_shape((0.9, 0.45, 0.6), 0.45, material=5))             # mirror
    s.shapes.append(sphere_shape((2.9, 0.5, 0.4), 0.5, material=tex_ball))
    s.shapes.append(sphere_shape((1.6, 4.6, 4.4), 0.7, material=len(m) - 3))      # anisotropic ward (index 11)
    s.shapes.append(sphere_shape((4.6, 4.4, 4.6), 0.8, material=0, inverted=True))
    return s


def _test_image(width, height, seed):
    """A deterministic 8-bit RGB image with structure at every scale: stripes, a gradient and per-texel noise."""
    rng = np.random.Generator(np.random.PCG64(seed))
    y, x = np.mgrid[0:height, 0:width]
    img = np.empty((height, width, 3), np.float64)
    img[..., 0] = 0.5 + 0.5 * np.sin(x * 2 * np.pi * 3 / width + seed)
    img[..., 1] = (x + y) / float(width + height)
    img[..., 2] = ((x // 2 + y // 2) % 2) * 0.8 + 0.1
    img = 0.75 * img + 0.25 * rng.random((height, width, 3))
    return np.clip(img * 255.0, 0, 255).astype(np.uint8)


def bitmap_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """The sphere scene with bitmap textures (src/textures/ldrtexture.cpp over src/librender/mipmap.cpp) in place of the
    checkerboards: an EWA-filtered 64 x 32 image on the walls (seen at grazing angles), a trilinear 20 x 24 image (not a
    power of two: Lanczos up-sampling) with `clamp` wrapping and gamma 2.2, an unfiltered 16 x 8 image, an EWA image
    with `black` wrapping inside the inverted sphere, an HDR image through `exrtexture` on an analytic sphere, and one
    remaining checkerboard."""
    s = spheres_test_scene(width, height, spp, mode, max_depth)
    t = s.textures
    green_wall = t[0]
    t[0] = bitmap_texture(_test_image(64, 32, 1), uscale=2.0, vscale=2.0)
    t[1] = bitmap_texture(_test_image(20, 24, 2), gamma=2.2, filter_type=MIP_TRILINEAR, wrap_mode=WRAP_CLAMP,
                          uscale=0.3, vscale=0.25, uoffset=0.05, voffset=-0.1)
    t[2] = bitmap_texture(_test_image(16, 8, 3), filter_type=MIP_NONE, uscale=6.0, vscale=6.0, uoffset=0.1, voffset=0.1)
    t.append(bitmap_texture(_test_image(32, 32, 4), wrap_mode=WRAP_BLACK, max_anisotropy=4.0, uscale=1.5, vscale=1.2))
    t.append(green_wall)
    s.materials[0] = textured(s.materials[0], 0, 3)
    s.materials[2] = textured(s.materials[2], 0, 4)
    # an HDR image through exrtexture (src/textures/exrtexture.cpp: linear, EWA, repeat) on the textured analytic sphere
    t.append(exr_texture((_test_image(24, 16, 5).astype(F32) / F32(255)) ** F32(2.0) * F32(1.6), uscale=2.0, vscale=1.0))
    s.materials[13] = textured(s.materials[13], 0, 5)
    return s


def _sky_image(width, height, seed):
    """A small latitude-longitude HDR map: a bright sun disc, a sky gradient and dim noisy ground."""
    rng = np.random.Generator(np.random.PCG64(seed))
    v, u = np.mgrid[0:height, 0:width]
    theta = (v + 0.5) / height * np.pi
    phi = (u + 0.5) / width * 2 * np.pi
    img = np.empty((height, width, 3), np.float64)
    up = np.clip(np.cos(theta), 0, 1)
    img[..., 0] = 0.2 + 0.3 * up; img[..., 1] = 0.3 + 0.4 * up; img[..., 2] = 0.5 + 0.9 * up
    ground = np.cos(theta) < 0
    img[ground] = np.array([0.12, 0.1, 0.08]) * (0.5 + rng.random((int(ground.sum()), 1)))
    sun = (theta - 0.9) ** 2 + (phi - 4.2) ** 2 < 0.05
    img[sun] = [40.0, 36.0, 30.0]
    return (img * (0.9 + 0.2 * rng.random((height, width, 1)))).astype(F32)


def envmap_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6, rotate=True) -> SceneDesc:
    """The open all-luminaires scene with an environment map (src/luminaires/envmap.cpp) as its background instead of the
    constant one: a 64 x 32 latitude-longitude map, intensityScale 1.5, rotated about two axes."""
    s = lights_test_scene(width, height, spp, mode, max_depth, background=False)
    m = None
    if rotate:
        a, b = F32(0.6), F32(-0.35)
        ry = np.array([[np.cos(a), 0, np.sin(a), 0], [0, 1, 0, 0], [-np.sin(a), 0, np.cos(a), 0], [0, 0, 0, 1]], F32)
        rx = np.array([[1, 0, 0, 0], [0, np.cos(b), -np.sin(b), 0], [0, np.sin(b), np.cos(b), 0], [0, 0, 0, 1]], F32)
        m = (ry @ rx).astype(F32)
    add_light(s, envmap_light(_sky_image(64, 32, 7), 1.5, m, sampling_weight=2.0))
    # a collimated beam (src/luminaires/collimated.cpp) slanting down onto the floor
    add_light(s, collimated_light((4.6, 5.0, 0.2), (3.0, 0.0, 2.6), (0, 1, 0), 0.9, (14.0, 12.0, 9.0), sampling_weight=1.25))
    return s


def _xform(scale=(1, 1, 1), rot_y=0.0, rot_x=0.0, translate=(0, 0, 0)):
    a, b = F32(rot_y), F32(rot_x)
    ry = np.array([[np.cos(a), 0, np.sin(a), 0], [0, 1, 0, 0], [-np.sin(a), 0, np.cos(a), 0], [0, 0, 0, 1]], F32)
    rx = np.array([[1, 0, 0, 0], [0, np.cos(b), -np.sin(b), 0], [0, np.sin(b), np.cos(b), 0], [0, 0, 0, 1]], F32)
    sc = np.diag(np.array(list(scale) + [1], F32)).astype(F32)
    m = (ry @ rx @ sc).astype(F32)
    m[:3, 3] = translate
    return m


def instances_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """The textured all-BSDF scene plus two shape groups (src/shapes/shapegroup.cpp) shown through five instances
    (src/shapes/instance.cpp): a group of two meshes — a smooth icosphere with a checkerboard and a rough-metal box —
    instanced three times (translated; rotated and uniformly scaled; NON-uniformly scaled, which makes the transformed
    shading frame non-orthogonal), and a glass icosphere instanced twice, once overlapping a wall."""
    s = textured_test_scene(width, height, spp, mode, max_depth)
    ball = icosphere(2)
    uv = np.stack([np.arctan2(ball[0][:, 2], ball[0][:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(ball[0][:, 1], -1, 1)) / np.pi], 1).astype(F32)
    ball = (ball[0].astype(F32), ball[1].astype(np.uint32), ball[0].astype(F32))      # unit sphere: normals = positions
    g0_ball = Shape((ball[0] * F32(0.4)).astype(F32), ball[1], ball[2], material=1, texcoords=uv)
    bx = merge(box_faces(None, [(-0.3, 0.5, -0.3), (-0.3, 0.5, 0.3), (0.3, 0.5, 0.3), (0.3, 0.5, -0.3)], np.array([0.0, 0.75, 0.0])))
    g0_box = Shape((bx[0] + np.array([0.0, -1.0, 0.0], F32)).astype(F32), bx[1], bx[2], material=6,
                   texcoords=(bx[0][:, [0, 2]] * F32(0.8)).astype(F32))
    g1_ball = Shape((ball[0] * F32(0.35)).astype(F32), ball[1], ball[2], material=7)
    s.groups = [[g0_ball, g0_box], [g1_ball]]

    def inst(g, m):
        pts = np.concatenate([sh.positions for sh in s.groups[g]])
        lo, hi = pts.min(0), pts.max(0)
        corners = np.array([[x, y, z, 1] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])], np.float64)
        w = (np.asarray(m, np.float64) @ corners.T).T[:, :3]
        return instance_shape(g, m, [w.min(0), w.max(0)])
    s.shapes.append(inst(0, _xform(translate=(1.2, 1.6, 2.2))))
    s.shapes.append(inst(0, _xform((1.4, 1.4, 1.4), 0.7, 0.3, (3.9, 2.0, 1.8))))
    s.shapes.append(inst(0, _xform((0.8, 1.7, 1.1), -0.5, 0.0, (2.6, 3.6, 3.4))))
    s.shapes.append(inst(1, _xform((1.0, 1.0, 1.0), 0.0, 0.0, (2.7, 0.9, 1.0))))
    s.shapes.append(inst(1, _xform((1.3, 0.9, 1.3), 1.1, -0.4, (5.3, 2.8, 3.0))))
    return s


def masks_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """The textured scene with the wrappers and transmitters of src/bsdfs: the phong panel behind a checkerboard opacity
    mask (mask.cpp), the rough-metal block behind a constant one, the tall block a diffuse transmitter (difftrans.cpp)."""
    s = textured_test_scene(width, height, spp, mode, max_depth)
    s.textures.append(checkerboard((1.0, 1.0, 1.0), (0.05, 0.05, 0.05), 5.0, 5.0))
    m = s.materials
    m[9] = textured(masked(m[9]), 2, len(s.textures) - 1)
    m[6] = masked(m[6], (0.55, 0.6, 0.7))
    m.append(difftrans((0.5, 0.8, 0.9)))
    s.shapes[4].material = len(m) - 1
    s.textures.append(grid_texture((0.8, 0.3, 0.25), (0.1, 0.1, 0.4), 0.06, 2.5, 3.0, 0.2, 0.1))      # gridtexture.cpp on the red wall
    m[1] = textured(m[1], 0, len(s.textures) - 1)
    return s


def cylinders_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """The textured scene plus analytic cylinders (src/shapes/cylinder.cpp): an upright checkerboard column, a tilted
    rough-metal pipe the camera looks into (open ends: the inside is visible), a glass rod and a thin Ward-anisotropic bar."""
    s = textured_test_scene(width, height, spp, mode, max_depth)
    m = s.materials
    m.append(textured(lambertian((0.5, 0.5, 0.5)), 0, 2))
    col = len(m) - 1
    rx90 = _xform(rot_x=-np.pi / 2 + 0.06, rot_y=0.04)  # object z -> (almost) world y; see the note on axis-parallel cylinders in DESIGN.md
    s.shapes.append(cylinder_shape((_xform(translate=(1.0, 0.0, 1.2)) @ rx90 @ _xform((0.35, 0.35, 2.4))).astype(F32), material=col))
    s.shapes.append(cylinder_shape(_xform((0.5, 0.5, 1.6), 0.5, 0.35, (3.2, 1.6, 0.3)), material=6))
    s.shapes.append(cylinder_shape(_xform((0.2, 0.2, 2.0), 1.2, -0.2, (3.6, 2.9, 2.6)), material=7))
    s.shapes.append(cylinder_shape(_xform((0.12, 0.12, 2.5), -0.8, 0.9, (1.2, 3.9, 3.2)), material=11))
    return s


def sphere_lights_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """The sphere scene plus two spheres that carry area luminaires (area.cpp:71-82 -> Sphere::sampleSolidAngle,
    sphere.cpp:191-229): a small ball with a lambertian BSDF — shading points on its own surface take the
    "inside the sphere" branch — and a bare emitting ball without a BSDF (not an occluder)."""
    s = spheres_test_scene(width, height, spp, mode, max_depth)
    s.emitters.append(Emitter((9.0, 7.5, 6.0), 1.0))
    s.shapes.append(sphere_shape((2.2, 1.9, 1.6), 0.3, material=0, emitter=len(s.emitters) - 1))
    s.emitters.append(Emitter((3.0, 5.0, 8.0), 2.5))
    s.shapes.append(sphere_shape((4.0, 3.8, 3.0), 0.45, material=-1, emitter=len(s.emitters) - 1))
    return s


def lights_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6, background=True) -> SceneDesc:
    """The all-BSDF scene opened up (no ceiling, no side walls: a floor and a back wall remain) and lit by every
    supported luminaire type: the two area lights, a point light, a spot light, a directional light and — with
    `background` — a constant environment that camera rays and escaping bounce rays see."""
    s = material_test_scene(width, height, spp, mode, max_depth)
    floor = quad((8.0, 0, -2.0), (-2.5, 0, -2.0), (-2.5, 0, 5.592), (8.0, 0, 5.592), toward=(2.78, 2.73, 2.8))
    back = quad((8.0, 0, 5.592), (-2.5, 0, 5.592), (-2.5, 5.488, 5.592), (8.0, 5.488, 5.592), toward=(2.78, 2.73, 2.8))
    p, t, n = merge([floor, back])
    s.shapes[0] = Shape(p, t, n, material=s.shapes[0].material)
    del s.shapes[1:3]                                   # side walls; no shape below index 3 carried an emitter
    add_light(s, point_light((0.8, 4.2, 1.0), (9.0, 7.0, 5.0)))
    add_light(s, spot_light((5.2, 4.8, -0.5), (3.6, 1.2, 3.4), (0, 1, 0), (60.0, 70.0, 90.0), 25.0, 12.0, sampling_weight=1.5))
    add_light(s, directional_light((-0.35, -1.0, 0.45), (0.9, 0.8, 0.7)))
    if background:
        add_light(s, constant_light((0.25, 0.3, 0.4), sampling_weight=0.75))
    return s


# --------------------------------------------------------------------------------------------
# C1 — Cornell box
# --------------------------------------------------------------------------------------------
def cornell_box(width=512, height=512, spp=64, max_depth=8, scale=0.01) -> SceneDesc:
    """Cornell box from the published Cornell measurements (authored here; the reference ships no
    Cornell scene). All-lambertian, quad area light that also carries a BSDF (so it occludes),
    `path` maxDepth 8, box filter, perspective fov 39.3 deg."""
    C = np.array([278.0, 274.4, 279.6])
    white, red, green = lambertian((.73, .73, .73)), lambertian((.63, .065, .05)), lambertian((.14, .45, .091))
    light_bsdf = lambertian((.78, .78, .78))
    walls = merge([
        quad((552.8, 0, 0), (0, 0, 0), (0, 0, 559.2), (549.6, 0, 559.2), toward=C),              # floor
        quad((556, 548.8, 0), (556, 548.8, 559.2), (0, 548.8, 559.2), (0, 548.8, 0), toward=C),  # ceiling
        quad((549.6, 0, 559.2), (0, 0, 559.2), (0, 548.8, 559.2), (556, 548.8, 559.2), toward=C),  # back
    ])
    right = quad((0, 0, 559.2), (0, 0, 0), (0, 548.8, 0), (0, 548.8, 559.2), toward=C)
    left = quad((552.8, 0, 0), (549.6, 0, 559.2), (556, 548.8, 559.2), (556, 548.8, 0), toward=C)
    light = quad((343, 548.3, 227), (343, 548.3, 332), (213, 548.3, 332), (213, 548.3, 227), toward=C)
    short_top = [(130, 165, 65), (82, 165, 225), (240, 165, 272), (290, 165, 114)]
    tall_top = [(423, 330, 247), (265, 330, 296), (314, 330, 456), (472, 330, 406)]
    short = merge(box_faces(None, short_top, np.array([185.5, 82.5, 169.0])))
    tall = merge(box_faces(None, tall_top, np.array([368.5, 165.0, 351.25])))

    def shape(m, mat, emit=-1):
        return Shape((m[0] * F32(scale)).astype(F32), m[1], m[2], material=mat, emitter=emit)

    s = SceneDesc()
    s.materials = [white, red, green, light_bsdf]
    s.emitters = [Emitter((17.0, 12.0, 4.0), 1.0)]
    s.shapes = [shape(walls, 0), shape(right, 2), shape(left, 1), shape(short, 0), shape(tall, 0),
                shape(light, 3, 0)]
    s.camera_to_world = look_at(np.array([278, 273, -800]) * scale, np.array([278, 273, 0]) * scale, (0, 1, 0))
    s.fov = 39.3077
    s.near_clip, s.far_clip, s.focus_depth = 1e-2, 1e4, 1e4
    s.width, s.height, s.spp = width, height, spp
    s.filter_type = FILTER_BOX
    s.mode, s.max_depth, s.rr_depth = MODE_PATH, max_depth, 10
    return s


# --------------------------------------------------------------------------------------------
# C2 — ray-intersection microbench
# --------------------------------------------------------------------------------------------
def terrain_mesh(cells=707, pad_cells=151, seed=1):
    """cells^2 displaced grid (2 tris/cell) + a pad strip so the total is exactly 1 000 000
    triangles for cells=707, pad_cells=151 (999 698 + 302)."""
    parts = [displaced_grid(cells, cells, seed=seed)]
    if pad_cells > 0:
        parts.append(displaced_grid(pad_cells, 1, seed=seed + 7, amplitude=0.05,
                                    x0=(-0.5, 0.5), z0=(-0.02, 0.02), y0=0.6))
    return merge(parts)


def microbench_scene(cells=707, pad_cells=151, seed=1) -> SceneDesc:
    """C2: one lambertian 1M-triangle mesh in [-1,1]^3 (BSDF present, so it occludes)."""
    pos, tri, nrm = terrain_mesh(cells, pad_cells, seed)
    s = SceneDesc()
    s.materials = [lambertian((0.5, 0.5, 0.5))]
    s.shapes = [Shape(pos, tri, nrm, material=0)]
    s.camera_to_world = look_at((0.0, 1.6, -2.4), (0.0, 0.0, 0.0), (0, 1, 0))
    s.fov = 45.0
    s.width = s.height = 64
    s.spp = 1
    return s


def scene_bounds(scene: SceneDesc):
    lo = np.min([s.positions.min(0) for s in scene.shapes], 0)
    hi = np.max([s.positions.max(0) for s in scene.shapes], 0)
    return lo.astype(np.float64), hi.astype(np.float64)


def coherent_rays(scene: SceneDesc, res: int) -> np.ndarray:
    """res x res pinhole camera rays through pixel centres: rows [o.xyz, mint, d.xyz, maxt]
    with mint = near/d_z, maxt = far/d_z like perspective.cpp:104-108."""
    m = scene.camera_to_world.astype(np.float64)
    tan = np.tan(np.deg2rad(scene.fov) / 2.0)
    px = (np.arange(res) + 0.5) / res * 2.0 - 1.0
    X, Y = np.meshgrid(px * tan, -px * tan, indexing="xy")
    d = np.stack([X, Y, np.ones_like(X)], -1).reshape(-1, 3)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    inv_z = 1.0 / d[:, 2]
    dw = d @ m[:3, :3].T
    rays = np.empty((res * res, 8), F32)
    rays[:, 0:3] = m[:3, 3]
    rays[:, 3] = scene.near_clip * inv_z
    rays[:, 4:7] = dw
    rays[:, 7] = scene.far_clip * inv_z
    return rays


def _square_to_sphere(s):
    z = 1.0 - 2.0 * s[:, 1]
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    phi = 2.0 * np.pi * s[:, 0]
    return np.stack([r * np.cos(phi), r * np.sin(phi), z], -1)


def incoherent_rays(scene: SceneDesc, n: int, seed: int = 2) -> np.ndarray:
    """kdbench protocol (src/utils/kdbench.cpp:221-231): two uniform points on the bounding
    sphere, direction normalised, mint = Epsilon (1e-4, adaptive), maxt = inf."""
    lo, hi = scene_bounds(scene)
    c, rad = (lo + hi) / 2.0, np.linalg.norm(hi - (lo + hi) / 2.0)
    rng = np.random.Generator(np.random.PCG64(seed))
    p1 = c + _square_to_sphere(rng.random((n, 2))) * rad
    p2 = c + _square_to_sphere(rng.random((n, 2))) * rad
    d = p2 - p1
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.empty((n, 8), F32)
    rays[:, 0:3] = p1; rays[:, 3] = 1e-4; rays[:, 4:7] = d; rays[:, 7] = np.inf
    return rays


def shadow_segments(rays: np.ndarray, scene: SceneDesc, seed: int = 3) -> np.ndarray:
    """Shadow-ray form of a ray set, as Scene::isOccluded builds it (scene.h:241-246):
    origin p1, un-normalised direction p2 - p1, t in [ShadowEpsilon, 1 - ShadowEpsilon]."""
    lo, hi = scene_bounds(scene)
    length = 2.0 * np.linalg.norm(hi - lo)
    rng = np.random.Generator(np.random.PCG64(seed))
    seg = rays.copy()
    seg[:, 4:7] = rays[:, 4:7] * (length * (0.25 + 0.75 * rng.random((len(rays), 1)))).astype(F32)
    seg[:, 3] = 1e-3
    seg[:, 7] = F32(1.0) - F32(1e-3)
    return seg


# --------------------------------------------------------------------------------------------
# C3 — `direct` on a 1M-triangle scene, several BSDFs and area lights
# --------------------------------------------------------------------------------------------
def _split_mesh(pos, tri, nrm, parts):
    """Split a mesh into `parts` shapes by triangle range (vertices re-indexed per shape)."""
    out = []
    bounds = np.linspace(0, len(tri), parts + 1).astype(np.int64)
    for a, b in zip(bounds[:-1], bounds[1:]):
        t = tri[a:b]
        used, inv = np.unique(t.ravel(), return_inverse=True)
        out.append((pos[used], inv.reshape(-1, 3).astype(np.uint32), nrm[used]))
    return out


def _quad_lights(n, y, half, ring, intensity):
    shapes = []
    for i in range(n):
        ang = 2 * np.pi * i / n
        cx, cz = ring * np.cos(ang), ring * np.sin(ang)
        q = quad((cx - half, y, cz - half), (cx + half, y, cz - half), (cx + half, y, cz + half),
                 (cx - half, y, cz + half), toward=(cx, y - 1.0, cz))
        shapes.append(q)
    return shapes


def direct_scene(cells=707, width=1920, height=1080, spp=256, n_lights=8, seed=1) -> SceneDesc:
    """C3: the 1M terrain split into 3 shapes (lambertian / microfacet / roughmetal) plus a small
    mirror patch, n_lights quad area lights above, `direct` with 1+1 samples."""
    pos, tri, nrm = terrain_mesh(cells, 0, seed)
    s = SceneDesc()
    s.materials = [lambertian((0.6, 0.55, 0.5)), microfacet((0.4, 0.2, 0.1), (1, 1, 1), 1.0, 1.0, 0.15),
                   roughmetal((1, 1, 1), 0.2), mirror((0.8, 0.8, 0.8)), lambertian((0.5, 0.5, 0.5))]
    for i, m in enumerate(_split_mesh(pos, tri, nrm, 3)):
        s.shapes.append(Shape(m[0], m[1], m[2], material=i))
    mp = quad((-0.3, 0.45, -0.3), (0.3, 0.45, -0.3), (0.3, 0.5, 0.3), (-0.3, 0.5, 0.3), toward=(0, 2, 0))
    s.shapes.append(Shape(mp[0], mp[1], mp[2], material=3))
    for i, q in enumerate(_quad_lights(n_lights, 1.5, 0.15, 0.8, 20.0)):
        s.emitters.append(Emitter((20.0, 18.0, 15.0), 1.0))
        s.shapes.append(Shape(q[0], q[1], q[2], material=4, emitter=i))
    s.camera_to_world = look_at((0.0, 1.2, -2.2), (0.0, 0.0, 0.0), (0, 1, 0))
    s.fov = 45.0
    s.width, s.height, s.spp = width, height, spp
    s.mode, s.luminaire_samples, s.bsdf_samples = MODE_DIRECT, 1, 1
    s.filter_type = FILTER_BOX
    return s


# --------------------------------------------------------------------------------------------
# C4 / C5 — `path` on a procedural 10M-triangle scene (roughmetal + dielectric)
# --------------------------------------------------------------------------------------------
def path_scene(terrain_cells=2000, n_spheres=6250, width=1920, height=1080, spp=1024, max_depth=16,
               seed=1) -> SceneDesc:
    """C4/C5: terrain_cells^2 terrain (2 tris/cell, roughmetal) + n_spheres copied subdivision-2
    icospheres (320 tris each; half roughmetal, half dielectric) + 4 quad area lights.
    Defaults give 8.0M + 2.0M = 10.0M triangles."""
    s = SceneDesc()
    s.materials = [roughmetal((0.9, 0.9, 0.9), 0.25), roughmetal((1.0, 0.8, 0.4), 0.1),
                   dielectric(1.5046, 1.0), lambertian((0.5, 0.5, 0.5))]
    pos, tri, nrm = displaced_grid(terrain_cells, terrain_cells, seed=seed)
    s.shapes.append(Shape(pos, tri, nrm, material=0))
    if n_spheres > 0:
        v, f = icosphere(2)
        rng = np.random.Generator(np.random.PCG64(seed + 11))
        centers = np.stack([rng.uniform(-0.95, 0.95, n_spheres), rng.uniform(0.25, 0.7, n_spheres),
                            rng.uniform(-0.95, 0.95, n_spheres)], -1)
        radii = rng.uniform(0.004, 0.012, n_spheres)
        for half, mat in ((slice(0, n_spheres // 2), 1), (slice(n_spheres // 2, n_spheres), 2)):
            c, r = centers[half], radii[half]
            P = (v[None] * r[:, None, None] + c[:, None, :]).reshape(-1, 3).astype(F32)
            N = np.tile(v.astype(F32), (len(c), 1))
            T = (f[None] + (np.arange(len(c)) * len(v))[:, None, None]).reshape(-1, 3).astype(np.uint32)
            s.shapes.append(Shape(P, T, N, material=mat))
    for i, q in enumerate(_quad_lights(4, 1.6, 0.25, 0.7, 30.0)):
        s.emitters.append(Emitter((30.0, 28.0, 25.0), 1.0))
        s.shapes.append(Shape(q[0], q[1], q[2], material=3, emitter=i))
    s.camera_to_world = look_at((0.0, 1.1, -2.1), (0.0, 0.1, 0.0), (0, 1, 0))
    s.fov = 45.0
    s.width, s.height, s.spp = width, height, spp
    s.mode, s.max_depth, s.rr_depth = MODE_PATH, max_depth, 10
    s.filter_type = FILTER_BOX
    return s


# --------------------------------------------------------------------------------------------
# small scenes for tests
# --------------------------------------------------------------------------------------------
def material_test_scene(width=96, height=96, spp=16, mode=MODE_PATH, max_depth=6) -> SceneDesc:
    """Small box scene exercising every supported BSDF, a BSDF-less emitter mesh and two lights."""
    s = cornell_box(width, height, spp, max_depth)
    s.materials += [microfacet((0.5, 0.3, 0.2), (1, 1, 1), 0.5, 0.5, 0.1), mirror((0.8, 0.8, 0.8)),
                    roughmetal((1, 1, 1), 0.15), dielectric()]
    s.shapes[0].material = 4      # walls: microfacet
    s.shapes[3].material = 6      # short block: roughmetal
    s.shapes[4].material = 5      # tall block: mirror
    v, f = icosphere(2)
    c = np.array([1.9, 3.2, 1.6])
    s.shapes.append(Shape((v * 0.7 + c).astype(F32), f.astype(np.uint32), v.astype(F32), material=7))
    q = quad((0.4, 2.0, 5.5), (1.4, 2.0, 5.5), (1.4, 3.0, 5.5), (0.4, 3.0, 5.5), toward=(1, 2.5, 0))
    s.emitters.append(Emitter((4.0, 6.0, 9.0), 2.0))
    s.shapes.append(Shape(q[0], q[1], q[2], material=-1, emitter=1))   # emitter without a BSDF
    # a free-standing two-sided panel (seen from both sides) and a one-sided one next to it
    s.materials += [twosided(lambertian((0.7, 0.6, 0.2))), phong((0.4, 0.3, 0.2), (0.3, 0.3, 0.3), 1.0, 1.0, 25.0)]
    pa = quad((3.6, 0.4, 1.2), (4.6, 0.4, 2.2), (4.6, 2.2, 2.2), (3.6, 2.2, 1.2), toward=(10, 1, 10))
    pb = quad((0.6, 3.2, 3.8), (1.6, 3.2, 3.2), (1.6, 4.6, 3.2), (0.6, 4.6, 3.8), toward=(10, 4, 10))
    s.shapes.append(Shape(pa[0], pa[1], pa[2], material=8))
    s.shapes.append(Shape(pb[0], pb[1], pb[2], material=9))
    # a rough glass ball (GGX) in front
    s.materials.append(roughglass((1, 1, 1), (0.9, 1.0, 0.9), 0.2, 1.5046, 1.0, "ggx"))
    c2 = np.array([4.1, 1.3, 4.3])
    s.shapes.append(Shape((v * 0.55 + c2).astype(F32), f.astype(np.uint32), v.astype(F32), material=10))
    s.mode = mode
    return s
