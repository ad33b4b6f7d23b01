"""Scene description shared by the host-side API, the synthetic scene generators and the
oracle harness.

A `SceneDesc` is the flattened form of what Mitsuba's `Scene` holds after XML parsing for
the path-tracing hot path: triangle meshes in shape order
(`Scene::addShape`, src/librender/scene.cpp:656-707), one BSDF parameter block per material,
area luminaires bound to shapes, one perspective camera, film geometry and the integrator
properties of `path` / `direct` (src/librender/integrator.cpp:270-295,
src/integrators/direct/direct.cpp:34-36).

`write_scene()` serialises it in the little-endian interchange layout read by
oracle/harness/scene_io.h, so that the compiled reference renders exactly the same input.
"""
from __future__ import annotations

import math
import struct
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

MAT_LAMBERTIAN, MAT_MICROFACET, MAT_MIRROR, MAT_ROUGHMETAL, MAT_DIELECTRIC, MAT_PHONG, MAT_ROUGHGLASS, MAT_WARD, MAT_DIFFTRANS = range(9)
BSDF_TWOSIDED = 0x100     # OR-ed into Material.type: wrapped in the `twosided` adapter (src/bsdfs/twosided.cpp)
MODE_PATH, MODE_DIRECT = 0, 1
FILTER_BOX, FILTER_GAUSSIAN, FILTER_MITCHELL, FILTER_CATMULLROM, FILTER_WSINC = 0, 1, 2, 3, 4     # src/rfilters/*.cpp


@dataclass
class Material:
    """BSDF parameter block; `p` uses the slot layout of `mtsgpu_material` (include/mtsgpu.h)."""
    type: int
    p: np.ndarray  # float32[16]
    tex: tuple = (0, 0, 0)   # per spectrum slot (p[0..2], p[3..5], p[13..15] = mask opacity): 0 constant, k > 0: SceneDesc.textures[k-1]


@dataclass
class Texture:
    """2D texture; `p` uses the layout of `mtsgpu_texture` (include/mtsgpu.h). A bitmap texture also carries its
    pixels (uint8 [height, width, 3], top row first) and the gamma of LDRTexture (-1: sRGB); float32 pixels make it an
    EXRTexture (linear, no gamma)."""
    type: int
    p: np.ndarray  # float32[12]
    pixels: Optional[np.ndarray] = None
    gamma: float = -1.0


TEX_CHECKERBOARD = 0
TEX_BITMAP = 1
MIP_EWA, MIP_TRILINEAR, MIP_NONE = 0, 1, 2                 # MIPMap::EFilterType (mipmap.h:39-46)
WRAP_REPEAT, WRAP_BLACK, WRAP_WHITE, WRAP_CLAMP = 0, 1, 2, 3   # MIPMap::EWrapMode (mipmap.h:32-37)


def bitmap_texture(pixels, gamma=-1.0, filter_type=MIP_EWA, wrap_mode=WRAP_REPEAT, max_anisotropy=8.0,
                   uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0):
    """<texture type="ldrtexture"> (src/textures/ldrtexture.cpp:37-84): gamma-corrected 8-bit RGB bitmap behind a MIPMap."""
    pixels = np.ascontiguousarray(pixels, np.uint8)
    assert pixels.ndim == 3 and pixels.shape[2] == 3
    p = np.zeros(12, np.float32)
    p[:3] = [filter_type, wrap_mode, max_anisotropy]
    p[6:10] = [uoffset, voffset, uscale, vscale]
    return Texture(TEX_BITMAP, p, pixels, float(gamma))


def checkerboard(bright=(0.4, 0.4, 0.4), dark=(0.2, 0.2, 0.2), uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0):
    """src/textures/checkerboard.cpp:30-33 + the uv transform of Texture2D (src/librender/texture.cpp:43-56)"""
    p = np.zeros(12, np.float32)
    p[:10] = list(bright) + list(dark) + [uoffset, voffset, uscale, vscale]
    return Texture(TEX_CHECKERBOARD, p)


def exr_texture(pixels, uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0):
    """<texture type="exrtexture"> (src/textures/exrtexture.cpp:34-44): linear HDR bitmap behind an EWA MIPMap with
    `repeat` wrapping (MIPMap::fromBitmap's defaults). The pixels (float [height, width, 3]) are rounded to half
    precision, which is what the file holds."""
    with np.errstate(over="ignore"):
        px = np.ascontiguousarray(pixels, np.float32).astype(np.float16).astype(np.float32)
    assert px.ndim == 3 and px.shape[2] == 3
    p = np.zeros(12, np.float32)
    p[:3] = [MIP_EWA, WRAP_REPEAT, 8.0]
    p[6:10] = [uoffset, voffset, uscale, vscale]
    return Texture(TEX_BITMAP, p, px, 1.0)


TEX_GRID = 2


def grid_texture(bright=(0.4, 0.4, 0.4), dark=(0.2, 0.2, 0.2), line_width=0.01, uscale=1.0, vscale=1.0, uoffset=0.0, voffset=0.0):
    """src/textures/gridtexture.cpp:31-35: dark lines of `line_width` on the integer uv grid"""
    p = np.zeros(12, np.float32)
    p[:11] = list(bright) + list(dark) + [uoffset, voffset, uscale, vscale, line_width]
    return Texture(TEX_GRID, p)


def textured(material: "Material", slot: int, texture_index: int) -> "Material":
    """`material` with spectrum slot 0 (p[0..2]), 1 (p[3..5]) or 2 (p[13..15]: the opacity of a mask) read from
    SceneDesc.textures[texture_index]."""
    tex = (list(material.tex) + [0, 0, 0])[:3]; tex[slot] = texture_index + 1
    return Material(material.type, material.p.copy(), tuple(tex))


def _mat(type_, values):
    p = np.zeros(16, np.float32)
    p[:len(values)] = values
    return Material(type_, p)


def lambertian(reflectance=(0.5, 0.5, 0.5)):
    """src/bsdfs/lambertian.cpp:66-74"""
    return _mat(MAT_LAMBERTIAN, list(reflectance))


def microfacet(diffuse=(0.0, 0.0, 0.0), specular=(1.0, 1.0, 1.0), kd=1.0, ks=1.0, alphaB=0.1,
               intIOR=1.5, extIOR=1.0):
    """src/bsdfs/microfacet.cpp:37-53"""
    return _mat(MAT_MICROFACET, list(diffuse) + list(specular) + [kd, ks, alphaB, intIOR, extIOR])


def mirror(reflectance=(0.8, 0.8, 0.8)):
    """src/bsdfs/mirror.cpp:28-36"""
    return _mat(MAT_MIRROR, list(reflectance))


def roughmetal(specular=(1.0, 1.0, 1.0), alphaB=0.1, ior=(0.37, 0.37, 0.37), k=(2.82, 2.82, 2.82)):
    """src/bsdfs/roughmetal.cpp:35-42"""
    return _mat(MAT_ROUGHMETAL, list(specular) + [alphaB] + list(ior) + list(k))


def dielectric(intIOR=1.5046, extIOR=1.0, reflectance=(1.0, 1.0, 1.0), transmittance=(1.0, 1.0, 1.0)):
    """src/bsdfs/dielectric.cpp:49-60"""
    return _mat(MAT_DIELECTRIC, list(reflectance) + list(transmittance) + [intIOR, extIOR])


def phong(diffuse=(0.5, 0.5, 0.5), specular=(0.2, 0.2, 0.2), kd=1.0, ks=1.0, exponent=10.0, specular_sampling_weight=None):
    """src/bsdfs/phong.cpp:32-99. Stores the values after Phong::configure: energy conservation is the caller's
    business (the harness passes verifyEnergyConservation=false) and the sampling weights are explicit."""
    f32 = np.float32
    if specular_sampling_weight is None:
        avg = lambda c: f32(f32(f32(f32(c[0]) + f32(c[1])) + f32(c[2])) * f32(1.0 / 3.0))
        ad, as_ = f32(avg(diffuse) * f32(kd)), f32(avg(specular) * f32(ks))
        specular_sampling_weight = float(f32(as_ / f32(ad + as_)))
    ssw = f32(specular_sampling_weight)
    return _mat(MAT_PHONG, list(diffuse) + list(specular) + [kd, ks, exponent, float(ssw), float(f32(1.0) - ssw)])


WARD_MODELS = {"ward": 0, "ward-duer": 1, "balanced": 2}


def difftrans(transmittance=(0.5, 0.5, 0.5)):
    """src/bsdfs/difftrans.cpp:64-73: ideally diffuse transmitter"""
    return _mat(MAT_DIFFTRANS, list(transmittance))


def ward(diffuse=(0.5, 0.5, 0.5), specular=(0.2, 0.2, 0.2), kd=1.0, ks=1.0, alpha_x=0.1, alpha_y=0.1, model="balanced",
         specular_sampling_weight=None):
    """src/bsdfs/ward.cpp:54-139, values after Ward::configure (energy conservation is the caller's business).
    alpha_x != alpha_y makes it anisotropic: its meshes then need texture coordinates and vertex normals."""
    f32 = np.float32
    if specular_sampling_weight is None:
        avg = lambda c: f32(f32(f32(f32(c[0]) + f32(c[1])) + f32(c[2])) * f32(1.0 / 3.0))
        ad, as_ = f32(avg(diffuse) * f32(kd)), f32(avg(specular) * f32(ks))
        specular_sampling_weight = float(f32(as_ / f32(ad + as_)))
    ssw = f32(specular_sampling_weight)
    return _mat(MAT_WARD, list(diffuse) + list(specular) + [kd, ks, alpha_x, alpha_y, float(ssw), float(f32(1.0) - ssw),
                                                             float(WARD_MODELS[model])])


ROUGHGLASS_DISTRIBUTIONS = {"beckmann": 0, "phong": 1, "ggx": 2}


def roughglass(specular_reflectance=(1, 1, 1), specular_transmittance=(1, 1, 1), alpha=0.1, int_ior=1.5046, ext_ior=1.0,
               distribution="beckmann"):
    """src/bsdfs/roughglass.cpp:93-146. p[6] is the value the constructor stores in the alpha texture: the
    roughness itself, or for the phong distribution the exponent 2/alpha^2 - 2 (p[10] then keeps the roughness)."""
    f32 = np.float32
    d = ROUGHGLASS_DISTRIBUTIONS[distribution]
    stored = f32(alpha)
    if d == 1:
        stored = f32(f32(f32(2) / f32(f32(alpha) * f32(alpha))) - f32(2))
    return _mat(MAT_ROUGHGLASS, list(specular_reflectance) + list(specular_transmittance)
                + [float(stored), int_ior, ext_ior, float(d), float(f32(alpha))])


BSDF_MASKED = 0x200


def masked(material: Material, opacity=(0.5, 0.5, 0.5)) -> Material:
    """<bsdf type="mask"> around `material` (src/bsdfs/mask.cpp:28-160): opacity in p[13..15], texture slot 2."""
    p = material.p.copy(); p[13:16] = opacity
    return Material(material.type | BSDF_MASKED, p, tuple((list(material.tex) + [0, 0, 0])[:3]))


def twosided(material: Material) -> Material:
    """src/bsdfs/twosided.cpp:36-143 around a BRDF (not valid for dielectric)."""
    assert (material.type & 0xFF) != MAT_DIELECTRIC
    return Material(material.type | BSDF_TWOSIDED, material.p.copy(), material.tex)


EMITTER_AREA, EMITTER_POINT, EMITTER_SPOT, EMITTER_DIRECTIONAL, EMITTER_CONSTANT, EMITTER_ENVMAP, EMITTER_COLLIMATED = range(7)


@dataclass
class Emitter:
    """A luminaire. The default is an area luminaire (src/luminaires/area.cpp:37-43), bound to the shape that
    names it; the other types are made by point_light / spot_light / directional_light / constant_light.
    `p` holds the C-ABI parameters (include/mtsgpu.h), `aux` what only the oracle harness needs to build the
    same luminaire from its properties (toWorld matrix, angles in degrees, the direction before normalisation)."""
    intensity: tuple = (1.0, 1.0, 1.0)
    sampling_weight: float = 1.0
    type: int = EMITTER_AREA
    p: tuple = ()
    aux: tuple = ()
    pixels: Optional[np.ndarray] = None     # environment map: float32 [height, width, 3], values representable in half

    def abi_params(self):
        v = np.zeros(11, np.float32); v[:len(self.p)] = self.p
        return v

    def aux_params(self):
        v = np.zeros(16, np.float32); v[:len(self.aux)] = self.aux
        return v


def invert4_f32(m):
    """Inverse of a 4x4 matrix the way the reference gets a Transform's inverse: Gauss-Jordan elimination with
    full pivoting, every operation in fp32 (Matrix::invert, include/mitsuba/core/matrix.inl:140-195). Restated so
    that values derived from an inverse (a spot light's axis, transformed normals) match the reference's bits."""
    f32 = np.float32
    a = np.array(m, f32).reshape(4, 4).copy()
    pivoted = [0] * 4
    rows, cols = [], []
    for _ in range(4):
        big, r, c = f32(0), -1, -1
        for j in range(4):
            if pivoted[j] == 1:
                continue
            for k in range(4):
                if pivoted[k] == 0 and abs(a[j, k]) >= big:
                    big, r, c = abs(a[j, k]), j, k
        pivoted[c] += 1
        if r != c:
            a[[r, c]] = a[[c, r]]
        rows.append(r); cols.append(c)
        if a[c, c] == 0:
            raise ValueError("singular matrix")
        pivinv = f32(f32(1.0) / a[c, c])
        a[c, c] = f32(1.0)
        a[c] = (a[c] * pivinv).astype(f32)
        for j in range(4):
            if j != c:
                save = a[j, c]
                a[j, c] = f32(0)
                a[j] = (a[j] - (a[c] * save).astype(f32)).astype(f32)
    for j in range(3, -1, -1):
        if rows[j] != cols[j]:
            a[:, [rows[j], cols[j]]] = a[:, [cols[j], rows[j]]]
    return a


def point_light(position, intensity=(1.0, 1.0, 1.0), sampling_weight=1.0):
    """src/luminaires/point.cpp:29-33 with toWorld = translate(position)"""
    m = np.eye(4, dtype=np.float32); m[:3, 3] = position
    return Emitter(tuple(intensity), sampling_weight, EMITTER_POINT, tuple(np.float32(position)), tuple(m.reshape(16)))


def spot_light(origin, target, up=(0, 1, 0), intensity=(1.0, 1.0, 1.0), cutoff_angle=20.0, beam_width=None, sampling_weight=1.0):
    """src/luminaires/spot.cpp:34-61 with toWorld = lookAt(origin, target, up); angles in degrees."""
    f32 = np.float32
    cutoff = f32(cutoff_angle)
    beam = f32(f32(cutoff * f32(3.0)) / f32(4.0)) if beam_width is None else f32(beam_width)
    rad = f32(np.pi) / f32(180.0)                                  # degToRad (core/util.h:302)
    m = look_at(origin, target, up)
    axis = invert4_f32(m)[2, :3]                                   # third row of worldToLuminaire
    p = list(m[:3, 3]) + list(axis) + [f32(beam * rad), f32(cutoff * rad), beam, cutoff]
    return Emitter(tuple(intensity), sampling_weight, EMITTER_SPOT, tuple(p), tuple(m.reshape(16)))


def directional_light(direction, intensity=(1.0, 1.0, 1.0), sampling_weight=1.0):
    """src/luminaires/directional.cpp:31-41 with the `direction` property (normalised there as here)."""
    f32 = np.float32
    d = np.asarray(direction, f32)
    n = d * (f32(1.0) / f32(np.sqrt(f32(f32(f32(d[0] * d[0]) + f32(d[1] * d[1])) + f32(d[2] * d[2])))))
    return Emitter(tuple(intensity), sampling_weight, EMITTER_DIRECTIONAL, tuple(n.astype(f32)) + (0.0,), tuple(d))


def constant_light(intensity=(1.0, 1.0, 1.0), sampling_weight=1.0):
    """src/luminaires/constant.cpp:29-32: constant background; its bounding sphere is derived from the scene."""
    return Emitter(tuple(intensity), sampling_weight, EMITTER_CONSTANT)


def collimated_light(origin, target, up=(0, 1, 0), radius=0.01, intensity=(1.0, 1.0, 1.0), sampling_weight=1.0):
    """src/luminaires/collimated.cpp:29-37 with toWorld = lookAt(origin, target, up): a beam of parallel light of the
    given radius along the view direction."""
    m = look_at(origin, target, up)
    return Emitter(tuple(intensity), sampling_weight, EMITTER_COLLIMATED, (float(radius),), tuple(m.reshape(16)))


def envmap_light(pixels, intensity_scale=1.0, to_world=None, sampling_weight=1.0):
    """src/luminaires/envmap.cpp:40-51: latitude-longitude environment map read from an OpenEXR file. The pixels are
    rounded to half precision here, which is what the file holds (Bitmap::loadEXR, bitmap.cpp:247-273)."""
    m = np.eye(4, dtype=np.float32) if to_world is None else np.asarray(to_world, np.float32).reshape(4, 4)
    with np.errstate(over="ignore"):
        px = np.ascontiguousarray(pixels, np.float32).astype(np.float16).astype(np.float32)
    assert px.ndim == 3 and px.shape[2] == 3
    return Emitter((1.0, 1.0, 1.0), sampling_weight, EMITTER_ENVMAP, (0.0, 0.0, 0.0, 0.0, float(intensity_scale)), tuple(m.reshape(16)), px)


def add_light(scene: "SceneDesc", emitter: Emitter) -> int:
    """Insert a non-area luminaire where Scene::getLuminaires() has it: after the other scene-level luminaires
    and before every area luminaire (those are appended by Scene::initialize, scene.cpp:296-299)."""
    assert emitter.type != EMITTER_AREA
    at = 0
    while at < len(scene.emitters) and scene.emitters[at].type != EMITTER_AREA:
        at += 1
    scene.emitters.insert(at, emitter)
    for s in scene.shapes:
        if s.emitter >= at:
            s.emitter += 1
    return at


@dataclass
class Shape:
    """One triangle mesh. `normals=None` means face normals (`faceNormals=true`)."""
    positions: np.ndarray            # float32 [nVert, 3]
    triangles: np.ndarray            # uint32  [nTri, 3]
    normals: Optional[np.ndarray]    # float32 [nVert, 3] or None
    material: int = -1               # index into SceneDesc.materials, -1: no BSDF (not an occluder)
    emitter: int = -1                # index into SceneDesc.emitters, -1: none
    texcoords: Optional[np.ndarray] = None   # float32 [nVert, 2] or None (uv = 0 everywhere, skdtree.h:414)
    sphere: Optional[tuple] = None           # (center xyz, radius, inverted): analytic sphere (src/shapes/sphere.cpp); no arrays
    cylinder: Optional[np.ndarray] = None    # toWorld 4x4 of a <shape type="cylinder"> (src/shapes/cylinder.cpp): unit tube x^2+y^2=1, 0<=z<=1
    instance: Optional[tuple] = None         # (group index, toWorld 4x4): instance of SceneDesc.groups[index] (src/shapes/instance.cpp)


def sphere_shape(center, radius, material=-1, inverted=False, emitter=-1) -> "Shape":
    """<shape type="sphere"> with `center` and `radius` (sphere.cpp:44-47): one analytic primitive. `positions` holds the
    two corners of its box so that scene-bound helpers keep working."""
    c = np.asarray(center, np.float32)
    box = np.stack([c - np.float32(abs(radius)), c + np.float32(abs(radius))]).astype(np.float32)
    return Shape(box, np.zeros((0, 3), np.uint32), None, material=material, emitter=emitter, sphere=(tuple(float(v) for v in c), float(radius), bool(inverted)))


def cylinder_params(to_world):
    """What Cylinder's constructor derives from `toWorld` (cylinder.cpp:52-61), in fp32: radius, length, objectToWorld =
    toWorld * scale(1/r, 1/r, 1/l) and its inverse (the product of the inverses, toWorld's from Gauss-Jordan)."""
    f32 = np.float32
    m = np.asarray(to_world, f32).reshape(4, 4)
    ln = lambda v: f32(np.sqrt(f32(f32(f32(v[0] * v[0]) + f32(v[1] * v[1])) + f32(v[2] * v[2]))))
    radius, length = ln(m[:3, 0]), ln(m[:3, 2])
    sv = [f32(f32(1) / radius), f32(f32(1) / radius), f32(f32(1) / length)]
    sc = np.diag(sv + [f32(1)]).astype(f32)
    sc_inv = np.diag([f32(f32(1) / v) for v in sv] + [f32(1)]).astype(f32)
    return float(radius), float(length), matmul4_f32(m, sc), matmul4_f32(sc_inv, invert4_f32(m))


def cylinder_shape(to_world, material=-1) -> "Shape":
    """<shape type="cylinder"> with a `toWorld` transform; `positions` holds a loose box for the scene-bound helpers."""
    m = np.asarray(to_world, np.float32).reshape(4, 4)
    corners = np.array([[x, y, z, 1] for x in (-1, 1) for y in (-1, 1) for z in (0, 1)], np.float64)
    w = (m.astype(np.float64) @ corners.T).T[:, :3]
    return Shape(np.stack([w.min(0), w.max(0)]).astype(np.float32), np.zeros((0, 3), np.uint32), None, material=material, cylinder=m)


def instance_shape(group, to_world, bounds) -> "Shape":
    """<shape type="instance"> of shape group `group` under `to_world`. `bounds` (two points) stands in for positions so
    that scene-bound helpers keep working."""
    m = np.asarray(to_world, np.float32).reshape(4, 4)
    return Shape(np.asarray(bounds, np.float32).reshape(2, 3), np.zeros((0, 3), np.uint32), None, instance=(int(group), m))


@dataclass
class SceneDesc:
    camera_to_world: np.ndarray = field(default_factory=lambda: np.eye(4, dtype=np.float32))
    fov: float = 45.0
    near_clip: float = 1e-2
    far_clip: float = 1e4
    aperture_radius: float = 0.0
    focus_depth: float = 1e4
    map_smaller_side: bool = True
    orthographic: bool = False       # src/cameras/orthographic.cpp instead of perspective.cpp (fov, aperture unused)
    width: int = 512
    height: int = 512
    crop: Optional[tuple] = None     # (x, y, w, h)
    filter_type: int = FILTER_BOX
    spp: int = 4
    mode: int = MODE_PATH
    max_depth: int = -1
    rr_depth: int = 10
    strict_normals: bool = False
    luminaire_samples: int = 1
    bsdf_samples: int = 1
    sampler_type: int = 0            # 0 independent, 1 ldsampler, 2 stratified, 3 halton (src/samplers/*.cpp)
    sampler_depth: int = 3           # ldsampler `depth`
    shutter_open: float = 0.0        # Camera properties shutterOpen / shutterClose (camera.cpp:28-35): a non-empty interval
    shutter_close: float = 0.0       #   makes renderBlock draw a time sample (integrator.cpp:159-160)
    importance_sample_luminaires: bool = True     # Scene property (scene.cpp:51)
    materials: List[Material] = field(default_factory=list)
    emitters: List[Emitter] = field(default_factory=list)
    textures: List[Texture] = field(default_factory=list)
    groups: List[List[Shape]] = field(default_factory=list)     # shape groups (src/shapes/shapegroup.cpp): meshes in object space
    shapes: List[Shape] = field(default_factory=list)

    @property
    def crop_rect(self):
        return self.crop if self.crop is not None else (0, 0, self.width, self.height)

    @property
    def triangle_count(self):
        return int(sum(len(s.triangles) for s in self.shapes))


def look_at(origin, target, up):
    """cameraToWorld as built by Transform::lookAt (src/libcore/transform.cpp:174-190), fp32."""
    f32 = np.float32
    p = np.asarray(origin, f32); t = np.asarray(target, f32); u = np.asarray(up, f32)
    d = t - p
    d = d / f32(np.sqrt(f32(np.dot(d, d))))
    r = np.cross(d, u).astype(f32)
    r = r / f32(np.sqrt(f32(np.dot(r, r))))
    nu = np.cross(r, d).astype(f32)
    m = np.eye(4, dtype=f32)
    m[:3, 0] = r; m[:3, 1] = nu; m[:3, 2] = d; m[:3, 3] = p
    return m


def matmul4_f32(a, b):
    """4x4 product as Matrix::operator* computes it (core/matrix.h:665-679): fp32, `sum = 0; sum += a[i][k] * b[k][j]`
    for k = 0..3, one rounding per operation."""
    f32 = np.float32
    a, b = np.asarray(a, f32), np.asarray(b, f32)
    out = np.zeros((4, 4), f32)
    for k in range(4):
        out = (out + (a[:, k:k + 1] * b[k:k + 1, :]).astype(f32)).astype(f32)
    return out


def raster_to_camera(scene: SceneDesc) -> np.ndarray:
    """m_rasterToCamera of PerspectiveCameraImpl::configure (src/cameras/perspective.cpp:43-71) restated in fp32:
    cameraToScreen.inverse() * screenToRaster.inverse(), where a Transform product keeps the product of the inverses
    in the opposite order (transform.cpp:28-31), scale / translate carry analytic inverses (transform.cpp:33-63) and
    perspective()'s matrix is inverted numerically by Transform(Matrix4x4) (transform.h:45-49, matrix.inl:140-195)."""
    f32 = np.float32
    w, h = f32(scene.width), f32(scene.height)
    aspect = f32(w / h)                                   # ProjectiveCamera::configure, camera.cpp:142

    def scale_inv(x, y, z):
        return np.diag([f32(f32(1.0) / f32(x)), f32(f32(1.0) / f32(y)), f32(f32(1.0) / f32(z)), f32(1)]).astype(f32)

    def translate_inv(x, y, z):
        m = np.eye(4, dtype=f32); m[:3, 3] = [-f32(x), -f32(y), -f32(z)]; return m

    if (aspect >= 1.0) ^ (not scene.map_smaller_side):
        s1, s2, t = (w, h, f32(1)), (f32(f32(1) / f32(f32(2) * aspect)), f32(-0.5), f32(1)), (aspect, f32(-1), f32(0))
    else:
        s1, s2, t = (w, h, f32(1)), (f32(0.5), f32(f32(-0.5) * aspect), f32(1)), (f32(1), f32(-f32(f32(1) / aspect)), f32(0))
    s2r_inv = matmul4_f32(translate_inv(*t), matmul4_f32(scale_inv(*s2), scale_inv(*s1)))
    near, far = f32(scene.near_clip), f32(scene.far_clip)
    if scene.orthographic:
        # Transform::orthographic = scale(1, 1, 1 / (far - near)) * translate(0, 0, -near) (transform.cpp:155-158); the
        # camera's scale applies to x and y only ("no non-unity scale factors in the Z direction", orthographic.cpp:41-42)
        c2s_inv = matmul4_f32(translate_inv(f32(0), f32(0), f32(-near)), scale_inv(f32(1), f32(1), f32(f32(1.0) / f32(far - near))))
        return matmul4_f32(c2s_inv, s2r_inv)
    recip = f32(f32(1.0) / f32(far - near))               # Transform::perspective, transform.cpp:100-124
    trafo = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, f32(far * recip), f32(f32(f32(-near) * far) * recip)], [0, 0, 1, 0]], f32)
    half = f32(f32(f32(scene.fov) / f32(2.0)) * f32(f32(np.pi) / f32(180.0)))     # degToRad (core/util.h:302)
    cot = f32(f32(1.0) / f32(math.tan(float(half))))     # tanf: correctly rounded in glibc, as is double tan rounded once
    c2s_inv = matmul4_f32(invert4_f32(trafo), scale_inv(cot, cot, f32(1)))
    return matmul4_f32(c2s_inv, s2r_inv)


def write_scene(path: str, scene: SceneDesc) -> None:
    """Write the interchange file for the oracle harness (layout: oracle/harness/scene_io.h)."""
    cx, cy, cw, ch = scene.crop_rect
    with open(path, "wb") as f:
        f.write(b"MTSB200\0")
        f.write(struct.pack("<I", 11))
        f.write(np.ascontiguousarray(scene.camera_to_world, np.float32).tobytes())
        f.write(struct.pack("<5fI", scene.fov, scene.near_clip, scene.far_clip, scene.aperture_radius,
                            scene.focus_depth, int(scene.map_smaller_side) | (2 if scene.orthographic else 0)))
        f.write(struct.pack("<7I", scene.width, scene.height, cx, cy, cw, ch, scene.filter_type))
        f.write(struct.pack("<2I2i3I", scene.spp, scene.mode, scene.max_depth, scene.rr_depth,
                            int(scene.strict_normals), scene.luminaire_samples, scene.bsdf_samples))
        f.write(struct.pack("<2I", scene.sampler_type, scene.sampler_depth))
        f.write(struct.pack("<2fI", scene.shutter_open, scene.shutter_close, int(scene.importance_sample_luminaires)))
        f.write(struct.pack("<I", len(scene.textures)))
        for k, t in enumerate(scene.textures):
            f.write(struct.pack("<I", t.type))
            f.write(np.ascontiguousarray(t.p, np.float32).tobytes())
            if t.type == TEX_BITMAP and t.pixels.dtype == np.float32:     # exrtexture: travels as the OpenEXR file it is read from
                from . import api
                exr = (path + ".tex%d.exr" % k).encode()
                api.write_exr(exr.decode(), np.concatenate([t.pixels, np.ones(t.pixels.shape[:2] + (1,), np.float32)], axis=2))
                f.write(struct.pack("<I", 1)); f.write(struct.pack("<I", len(exr))); f.write(exr)
            elif t.type == TEX_BITMAP:
                f.write(struct.pack("<I", 0))
                f.write(struct.pack("<f2I", t.gamma, t.pixels.shape[1], t.pixels.shape[0]))
                f.write(t.pixels.tobytes())
        f.write(struct.pack("<I", len(scene.materials)))
        for m in scene.materials:
            f.write(struct.pack("<I", m.type))
            f.write(np.ascontiguousarray(m.p, np.float32).tobytes())
            f.write(struct.pack("<3I", *((list(m.tex) + [0, 0, 0])[:3])))
        f.write(struct.pack("<I", len(scene.emitters)))
        for k, e in enumerate(scene.emitters):
            f.write(struct.pack("<4fI", *e.intensity, e.sampling_weight, e.type))
            f.write(e.abi_params().tobytes()); f.write(e.aux_params().tobytes())
            if e.type == EMITTER_ENVMAP:            # the map travels as the OpenEXR file the reference's plugin reads
                from . import api
                exr = (path + ".env%d.exr" % k).encode()
                rgba = np.concatenate([e.pixels, np.ones(e.pixels.shape[:2] + (1,), np.float32)], axis=2)
                api.write_exr(exr.decode(), rgba)
                f.write(struct.pack("<I", len(exr))); f.write(exr)
        def write_mesh(s):
            f.write(struct.pack("<2I2i2I", len(s.positions), len(s.triangles), s.material, s.emitter,
                                0 if s.normals is None else 1, 0 if s.texcoords is None else 1))
            f.write(np.ascontiguousarray(s.positions, np.float32).tobytes())
            if s.normals is not None:
                f.write(np.ascontiguousarray(s.normals, np.float32).tobytes())
            if s.texcoords is not None:
                f.write(np.ascontiguousarray(s.texcoords, np.float32).tobytes())
            f.write(np.ascontiguousarray(s.triangles, np.uint32).tobytes())
        f.write(struct.pack("<I", len(scene.groups)))
        for g in scene.groups:
            f.write(struct.pack("<I", len(g)))
            for s in g:
                write_mesh(s)
        f.write(struct.pack("<I", len(scene.shapes)))
        for s in scene.shapes:
            if s.cylinder is not None:
                f.write(struct.pack("<Ii", 3, s.material))
                f.write(np.ascontiguousarray(s.cylinder, np.float32).tobytes())
                continue
            if s.instance is not None:
                f.write(struct.pack("<2I", 2, s.instance[0]))
                f.write(np.ascontiguousarray(s.instance[1], np.float32).tobytes())
                continue
            f.write(struct.pack("<I", 0 if s.sphere is None else 1))
            if s.sphere is not None:
                (cx, cy, cz), radius, inverted = s.sphere
                f.write(struct.pack("<2i4fI", s.material, s.emitter, cx, cy, cz, radius, int(inverted)))
                continue
            write_mesh(s)
