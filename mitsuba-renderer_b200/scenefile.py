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

import struct
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

MAT_LAMBERTIAN, MAT_MICROFACET, MAT_MIRROR, MAT_ROUGHMETAL, MAT_DIELECTRIC, MAT_PHONG, MAT_ROUGHGLASS = range(7)
BSDF_TWOSIDED = 0x100     # OR-ed into Material.type: wrapped in the `twosided` adapter (src/bsdfs/twosided.cpp)
MODE_PATH, MODE_DIRECT = 0, 1
FILTER_BOX, FILTER_GAUSSIAN = 0, 1


@dataclass
class Material:
    """BSDF parameter block; `p` uses the slot layout of `mtsgpu_material` (include/mtsgpu.h)."""
    type: int
    p: np.ndarray  # float32[16]


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


def twosided(material: Material) -> Material:
    """src/bsdfs/twosided.cpp:36-143 around a BRDF (not valid for dielectric)."""
    assert (material.type & 0xFF) != MAT_DIELECTRIC
    return Material(material.type | BSDF_TWOSIDED, material.p.copy())


@dataclass
class Emitter:
    """Area luminaire (src/luminaires/area.cpp:37-43); bound to the shape that names it."""
    intensity: tuple = (1.0, 1.0, 1.0)
    sampling_weight: float = 1.0


@dataclass
class Shape:
    """One triangle mesh. `normals=None` means face normals (`faceNormals=true`)."""
    positions: np.ndarray            # float32 [nVert, 3]
    triangles: np.ndarray            # uint32  [nTri, 3]
    normals: Optional[np.ndarray]    # float32 [nVert, 3] or None
    material: int = -1               # index into SceneDesc.materials, -1: no BSDF (not an occluder)
    emitter: int = -1                # index into SceneDesc.emitters, -1: none


@dataclass
class SceneDesc:
    camera_to_world: np.ndarray = field(default_factory=lambda: np.eye(4, dtype=np.float32))
    fov: float = 45.0
    near_clip: float = 1e-2
    far_clip: float = 1e4
    aperture_radius: float = 0.0
    focus_depth: float = 1e4
    map_smaller_side: bool = True
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
    materials: List[Material] = field(default_factory=list)
    emitters: List[Emitter] = field(default_factory=list)
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


def raster_to_camera(scene: SceneDesc) -> np.ndarray:
    """m_rasterToCamera of PerspectiveCameraImpl::configure (src/cameras/perspective.cpp:43-71):
    inverse(perspective(fov, near, far)) * inverse(screenToRaster), computed in float64 and
    rounded to fp32 (the reference inverts in fp32; agreement is ~1 ulp, tested at 1e-5)."""
    w, h = float(scene.width), float(scene.height)
    aspect = np.float32(w) / np.float32(h)
    aspect = float(aspect)

    def scale(x, y, z):
        return np.diag([x, y, z, 1.0])

    def translate(x, y, z):
        m = np.eye(4); m[:3, 3] = [x, y, z]; return m

    if (aspect >= 1.0) ^ (not scene.map_smaller_side):
        s2r = scale(w, h, 1.0) @ scale(1.0 / (2.0 * aspect), -0.5, 1.0) @ translate(aspect, -1.0, 0.0)
    else:
        s2r = scale(w, h, 1.0) @ scale(0.5, -0.5 * aspect, 1.0) @ translate(1.0, -1.0 / aspect, 0.0)
    near, far = float(np.float32(scene.near_clip)), float(np.float32(scene.far_clip))
    recip = 1.0 / (far - near)
    persp = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, far * recip, -near * far * recip], [0, 0, 1, 0]], np.float64)
    cot = 1.0 / np.tan(np.deg2rad(float(np.float32(scene.fov)) / 2.0))
    c2s = scale(cot, cot, 1.0) @ persp
    return (np.linalg.inv(c2s) @ np.linalg.inv(s2r)).astype(np.float32)


def write_scene(path: str, scene: SceneDesc) -> None:
    """Write the interchange file for the oracle harness (layout: oracle/harness/scene_io.h)."""
    cx, cy, cw, ch = scene.crop_rect
    with open(path, "wb") as f:
        f.write(b"MTSB200\0")
        f.write(struct.pack("<I", 1))
        f.write(np.ascontiguousarray(scene.camera_to_world, np.float32).tobytes())
        f.write(struct.pack("<5fI", scene.fov, scene.near_clip, scene.far_clip, scene.aperture_radius,
                            scene.focus_depth, int(scene.map_smaller_side)))
        f.write(struct.pack("<7I", scene.width, scene.height, cx, cy, cw, ch, scene.filter_type))
        f.write(struct.pack("<2I2i3I", scene.spp, scene.mode, scene.max_depth, scene.rr_depth,
                            int(scene.strict_normals), scene.luminaire_samples, scene.bsdf_samples))
        f.write(struct.pack("<I", len(scene.materials)))
        for m in scene.materials:
            f.write(struct.pack("<I", m.type))
            f.write(np.ascontiguousarray(m.p, np.float32).tobytes())
        f.write(struct.pack("<I", len(scene.emitters)))
        for e in scene.emitters:
            f.write(struct.pack("<4f", *e.intensity, e.sampling_weight))
        f.write(struct.pack("<I", len(scene.shapes)))
        for s in scene.shapes:
            f.write(struct.pack("<2I2iI", len(s.positions), len(s.triangles), s.material, s.emitter,
                                0 if s.normals is None else 1))
            f.write(np.ascontiguousarray(s.positions, np.float32).tobytes())
            if s.normals is not None:
                f.write(np.ascontiguousarray(s.normals, np.float32).tobytes())
            f.write(np.ascontiguousarray(s.triangles, np.uint32).tobytes())
