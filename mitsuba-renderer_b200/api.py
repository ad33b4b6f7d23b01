"""ctypes binding of the C ABI in include/mtsgpu.h (libmtscuda.so).

This is the Python host side of the drop-in boundary: the same calls, in the same order, that
the C++ `gpupath` plugin makes from `Integrator::render`. There is no CPU
fallback: if the shared library is missing or no CUDA device is usable, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from .scenefile import SceneDesc, raster_to_camera, MODE_PATH, MODE_DIRECT, FILTER_BOX, FILTER_GAUSSIAN

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmtscuda.so")

TRACE_CLOSEST, TRACE_ANY = 0, 1

# every symbol include/mtsgpu.h declares (checked by the CPU test-suite against the header)
EXPORTED_SYMBOLS = [
    "mtsgpu_create", "mtsgpu_destroy", "mtsgpu_last_error", "mtsgpu_version", "mtsgpu_upload_materials",
    "mtsgpu_upload_emitters", "mtsgpu_upload_mesh", "mtsgpu_set_camera", "mtsgpu_set_film", "mtsgpu_build_accel",
    "mtsgpu_render", "mtsgpu_render_device", "mtsgpu_trace", "mtsgpu_trace_device", "mtsgpu_bsdf_eval",
    "mtsgpu_emitter_sample", "mtsgpu_camera_rays", "mtsgpu_film_splat", "mtsgpu_intersect_full",
    "mtsgpu_get_triaccel", "mtsgpu_accel_info_get", "mtsgpu_set_counters", "mtsgpu_get_counters", "mtsgpu_set_progress",
    "mtsgpu_film_splat_static", "mtsgpu_render_multi", "mtsgpu_serialized_count", "mtsgpu_serialized_load",
    "mtsgpu_serialized_free", "mtsgpu_serialized_save", "mtsgpu_upload_serialized", "mtsgpu_film_develop",
    "mtsgpu_film_develop_device", "mtsgpu_film_save", "mtsgpu_write_exr", "mtsgpu_write_mfilm", "mtsgpu_sampler_values",
    "mtsgpu_upload_textures", "mtsgpu_intersect_uv", "mtsgpu_upload_sphere", "mtsgpu_upload_bitmap", "mtsgpu_texture_lookup", "mtsgpu_upload_envmap", "mtsgpu_upload_group_mesh", "mtsgpu_upload_instance", "mtsgpu_set_camera_orthographic", "mtsgpu_upload_cylinder", "mtsgpu_upload_emitter_transform",
]
SAMPLER_INDEPENDENT, SAMPLER_LD, SAMPLER_STRATIFIED, SAMPLER_HALTON, SAMPLER_HAMMERSLEY = 0, 1, 2, 3, 4
FILM_EXR, FILM_M = 0, 1
BITMAP_U8, BITMAP_F32 = 0, 1
STATIC_SCRATCH_BYTES = 4 << 20


class MtsGpuError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"mtsgpu error {code}: {message}")
        self.code = code


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("p", C.c_float * 16), ("texture", C.c_uint32 * 3)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_uint32), ("p", C.c_float * 12)]


class Emitter(C.Structure):
    _fields_ = [("intensity", C.c_float * 3), ("samplingWeight", C.c_float), ("type", C.c_uint32), ("p", C.c_float * 11)]


class RenderParams(C.Structure):
    _fields_ = [("mode", C.c_int32), ("maxDepth", C.c_int32), ("rrDepth", C.c_int32), ("strictNormals", C.c_int32),
                ("luminaireSamples", C.c_int32), ("bsdfSamples", C.c_int32), ("sppTotal", C.c_uint32),
                ("sppBegin", C.c_uint32), ("sppEnd", C.c_uint32), ("seed", C.c_uint32),
                ("importanceSampleLuminaires", C.c_uint32), ("samplerType", C.c_uint32), ("samplerDepth", C.c_uint32),
                ("reserved", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("cameraRays", C.c_uint64), ("normalRays", C.c_uint64), ("shadowRays", C.c_uint64),
                ("samples", C.c_uint64), ("kernelLaunches", C.c_uint64), ("traceClosestMs", C.c_double),
                ("traceAnyMs", C.c_double), ("shadeMs", C.c_double), ("totalMs", C.c_double),
                ("closestNodeVisits", C.c_uint64), ("closestTriTests", C.c_uint64), ("shadowNodeVisits", C.c_uint64),
                ("shadowTriTests", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


PROGRESS_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_double, C.POINTER(C.c_float))


class AccelInfo(C.Structure):
    _fields_ = [("triangles", C.c_uint64), ("nodes", C.c_uint64), ("nodeBytes", C.c_uint64),
                ("triBytes", C.c_uint64), ("buildMs", C.c_double), ("depth", C.c_uint32), ("builder", C.c_uint32),
                ("buildLevels", C.c_uint32), ("reserved", C.c_uint32)]


class TraceCounters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodeVisits", C.c_uint64), ("triTests", C.c_uint64), ("maxStackEntries", C.c_uint64)]


HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("shape", np.uint32), ("prim", np.uint32)])
TRIACCEL_DTYPE = np.dtype([("k", np.uint32), ("n_u", np.float32), ("n_v", np.float32), ("n_d", np.float32),
                           ("a_u", np.float32), ("a_v", np.float32), ("b_nu", np.float32), ("b_nv", np.float32),
                           ("c_nu", np.float32), ("c_nv", np.float32), ("shapeIndex", np.uint32), ("primIndex", np.uint32)])

_lib = None


def load_library(path: Optional[str] = None):
    """dlopen libmtscuda.so and declare the prototypes. Raises if the library is not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise MtsGpuError(-1, f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
    if "MTSGPU_HOST_THREADS" not in os.environ and int(os.environ.get("LOCAL_WORLD_SIZE", "1")) > 1:
        # one process per GPU (torchrun): the ranks of a node share its cores for the host-side mesh copies
        os.environ["MTSGPU_HOST_THREADS"] = str(max(1, (os.cpu_count() or 1) // int(os.environ["LOCAL_WORLD_SIZE"])))
    lib = C.CDLL(path)
    vp, fp, u32p = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    lib.mtsgpu_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.mtsgpu_destroy.argtypes = [vp]; lib.mtsgpu_destroy.restype = None
    lib.mtsgpu_last_error.argtypes = [vp]; lib.mtsgpu_last_error.restype = C.c_char_p
    lib.mtsgpu_version.restype = C.c_char_p
    lib.mtsgpu_upload_materials.argtypes = [vp, C.POINTER(Material), C.c_int]
    lib.mtsgpu_upload_emitters.argtypes = [vp, C.POINTER(Emitter), C.c_int]
    lib.mtsgpu_upload_textures.argtypes = [vp, C.POINTER(Texture), C.c_int]
    lib.mtsgpu_intersect_uv.argtypes = [vp, vp, C.c_size_t, vp]
    lib.mtsgpu_upload_group_mesh.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, C.c_uint32, vp, C.c_uint32, C.c_int]
    lib.mtsgpu_upload_cylinder.argtypes = [vp, C.c_int, vp, vp, C.c_float, C.c_float, C.c_int]
    lib.mtsgpu_upload_instance.argtypes = [vp, C.c_int, C.c_int, vp, vp]
    lib.mtsgpu_upload_emitter_transform.argtypes = [vp, C.c_int, vp, vp]
    lib.mtsgpu_upload_envmap.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp]
    lib.mtsgpu_upload_bitmap.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_float]
    lib.mtsgpu_upload_sphere.argtypes = [vp, C.c_int, vp, vp, C.c_float, vp, C.c_int, C.c_int, C.c_int]
    lib.mtsgpu_upload_mesh.argtypes = [vp, C.c_int, vp, vp, vp, C.c_uint32, vp, C.c_uint32, C.c_int, C.c_int]
    lib.mtsgpu_set_camera.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]
    lib.mtsgpu_set_camera_orthographic.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_float, C.c_float]
    lib.mtsgpu_set_film.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_float, C.c_float]
    lib.mtsgpu_build_accel.argtypes = [vp]
    lib.mtsgpu_render.argtypes = [vp, C.POINTER(RenderParams), vp, vp, C.POINTER(Stats)]
    lib.mtsgpu_render_device.argtypes = [vp, C.POINTER(RenderParams), vp, vp, C.POINTER(Stats)]
    lib.mtsgpu_render_multi.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(RenderParams), vp, vp, C.POINTER(Stats)]
    lib.mtsgpu_trace.argtypes = [vp, vp, C.c_size_t, C.c_int, vp]
    lib.mtsgpu_trace_device.argtypes = [vp, vp, C.c_size_t, C.c_int, vp, C.POINTER(C.c_float)]
    for name in ("mtsgpu_bsdf_eval", "mtsgpu_emitter_sample", "mtsgpu_camera_rays", "mtsgpu_film_splat",
                 "mtsgpu_film_splat_static"):
        getattr(lib, name).argtypes = [vp, vp, C.c_size_t, vp]
    lib.mtsgpu_texture_lookup.argtypes = [vp, vp, C.c_size_t, C.c_int, vp]
    lib.mtsgpu_intersect_full.argtypes = [vp, vp, C.c_size_t, vp, vp]
    lib.mtsgpu_get_triaccel.argtypes = [vp, vp, C.c_size_t, vp]
    lib.mtsgpu_accel_info_get.argtypes = [vp, C.POINTER(AccelInfo)]
    lib.mtsgpu_set_counters.argtypes = [vp, C.c_int]
    lib.mtsgpu_set_progress.argtypes = [vp, PROGRESS_FN, vp, C.c_double, C.c_int]
    lib.mtsgpu_get_counters.argtypes = [vp, C.POINTER(TraceCounters)]
    lib.mtsgpu_sampler_values.argtypes = [vp, C.POINTER(RenderParams), C.c_uint32, C.c_uint32, vp]
    lib.mtsgpu_film_develop.argtypes = [vp, vp, C.c_int, vp]
    lib.mtsgpu_film_develop_device.argtypes = [vp, vp, C.c_int, vp]
    lib.mtsgpu_film_save.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_char_p]
    lib.mtsgpu_write_exr.argtypes = [C.c_char_p, vp, C.c_int, C.c_int]
    lib.mtsgpu_write_mfilm.argtypes = [C.c_char_p, vp, C.c_int, C.c_int, C.c_int]
    for name in EXPORTED_SYMBOLS:
        fn = getattr(lib, name)
        if name not in ("mtsgpu_destroy", "mtsgpu_last_error", "mtsgpu_version", "mtsgpu_serialized_free"):
            fn.restype = C.c_int
    if path == LIB_PATH:
        _lib = lib
    return lib


def gaussian_table():
    """TabulatedFilter of the gaussian rfilter (src/rfilters/gaussian.cpp:30-43: halfSize 2, sigma 0.5;
    tabulated by src/librender/rfilter.cpp:40-69), fp32 arithmetic in the reference's order."""
    f32 = np.float32
    size = f32(2.0)
    sigma = f32(0.5)
    alpha = f32(-1.0) / (f32(2.0) * sigma * sigma)
    const = f32(np.exp(alpha * size * size, dtype=np.float32))

    def ev(x, y):
        return max(f32(0), f32(np.exp(alpha * x * x, dtype=np.float32)) - const) * \
               max(f32(0), f32(np.exp(alpha * y * y, dtype=np.float32)) - const)
    vals = np.zeros((16, 16), np.float32)
    total = f32(0)
    for y in range(16):
        ypos = (f32(y) + f32(0.5)) / f32(15) * size
        for x in range(16):
            if x != 15 and y != 15:
                vals[y, x] = ev((f32(x) + f32(0.5)) / f32(15) * size, ypos)
            total = f32(total + vals[y, x])
    total = f32(total * (f32(4) * size * size / f32(15 * 15)))
    return (vals / total).astype(np.float32), 2.0


def box_table():
    """TabulatedFilter of the box rfilter (src/rfilters/box.cpp:26-47: size 0.5, value 1)."""
    vals = np.zeros((16, 16), np.float32)
    vals[:15, :15] = 1.0
    total = np.float32(225.0) * np.float32(4 * 0.5 * 0.5 / 225.0)
    return (vals / total).astype(np.float32), 0.5


def tabulate(ev, size):
    """TabulatedFilter::TabulatedFilter (src/librender/rfilter.cpp:40-69) in fp32: 15 x 15 samples of one quadrant plus a
    zero border, normalised by 4 * size^2 / 225 times their sum."""
    f32 = np.float32
    size = f32(size)
    vals = np.zeros((16, 16), np.float32)
    total = f32(0)
    for y in range(16):
        ypos = f32(f32(f32(y) + f32(0.5)) / f32(15)) * size
        for x in range(16):
            if x != 15 and y != 15:
                vals[y, x] = ev(f32(f32(f32(f32(x) + f32(0.5)) / f32(15)) * size), f32(ypos))
            total = f32(total + vals[y, x])
    total = f32(total * f32(f32(f32(f32(4) * size) * size) / f32(15 * 15)))
    return (vals / total).astype(np.float32), float(size)


def mitchell_table(B=1.0 / 3.0, C=1.0 / 3.0, half_size=2.0):
    """mitchell rfilter (src/rfilters/mitchell.cpp:32-73; catmullrom.cpp is the same with B = 0, C = 0.5)"""
    f32 = np.float32
    B, C, size = f32(B), f32(C), f32(half_size)

    def mn(x):
        x = f32(abs(x)); x2 = f32(x * x); x3 = f32(x2 * x)
        if x < 1:
            a = f32(f32(f32(12) - f32(f32(9) * B)) - f32(f32(6) * C)); b = f32(f32(f32(-18) + f32(f32(12) * B)) + f32(f32(6) * C))
            c = f32(f32(6) - f32(f32(2) * B))
            return f32(f32(f32(1.0) / f32(6.0)) * f32(f32(f32(a * x3) + f32(b * x2)) + c))
        if x < 2:
            a = f32(-B - f32(f32(6) * C)); b = f32(f32(f32(6) * B) + f32(f32(30) * C))
            c = f32(f32(f32(-12) * B) - f32(f32(48) * C)); d = f32(f32(f32(8) * B) + f32(f32(24) * C))
            return f32(f32(f32(1.0) / f32(6.0)) * f32(f32(f32(f32(a * x3) + f32(b * x2)) + f32(c * x)) + d))
        return f32(0)
    return tabulate(lambda x, y: f32(mn(f32(f32(f32(2.0) * x) / size)) * mn(f32(f32(f32(2.0) * y) / size))), size)


def wsinc_table(half_size=3.0, cycles=3.0):
    """windowed sinc rfilter (src/rfilters/wsinc.cpp:30-54 over lanczosSinc, src/libcore/util.cpp:664-674)"""
    import math
    f32 = np.float32
    size, tau = f32(half_size), f32(cycles)
    pi = f32(3.14159265358979323846)

    def lanczos(t):
        t = f32(abs(t))
        if t < f32(1e-4):
            return f32(1)
        if t > 1:
            return f32(0)
        t = f32(t * pi)
        tt = f32(t * tau)
        return f32(f32(f32(math.sin(float(tt))) / tt) * f32(f32(math.sin(float(t))) / t))       # sinf: correctly rounded in glibc
    return tabulate(lambda x, y: f32(lanczos(f32(x / size)) * lanczos(f32(y / size))), size)


def filter_table(filter_type):
    from .scenefile import FILTER_MITCHELL, FILTER_CATMULLROM, FILTER_WSINC
    if filter_type == FILTER_MITCHELL:
        return mitchell_table()
    if filter_type == FILTER_CATMULLROM:
        return mitchell_table(0.0, 0.5)
    if filter_type == FILTER_WSINC:
        return wsinc_table()
    return box_table() if filter_type == FILTER_BOX else gaussian_table()


def render_multi(contexts, params: RenderParams):
    """mtsgpu_render_multi: contexts on different GPUs holding the same scene render disjoint sample ranges;
    returns (film summed over NVLink peer memory, stats)."""
    lib = contexts[0].lib
    arr = (C.c_void_p * len(contexts))(*[c.h for c in contexts])
    out = np.empty(contexts[0].film_shape, np.float32)
    st = Stats()
    rc = lib.mtsgpu_render_multi(arr, len(contexts), C.byref(params), None, out.ctypes.data_as(C.c_void_p), C.byref(st))
    if rc != 0:
        raise MtsGpuError(rc, lib.mtsgpu_last_error(contexts[0].h).decode())
    return out, st.as_dict()


class Context:
    """One GPU context (`mtsgpu_ctx`). Mirrors the call sequence of the plugin's render()."""

    def __init__(self, device: int = 0, lib_path: Optional[str] = None):
        self.lib = load_library(lib_path)
        h = C.c_void_p()
        rc = self.lib.mtsgpu_create(device, C.byref(h))
        if rc != 0:
            raise MtsGpuError(rc, self.lib.mtsgpu_last_error(None).decode())
        self.h = h
        self.device = device
        self.film_shape = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.mtsgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise MtsGpuError(rc, self.lib.mtsgpu_last_error(self.h).decode())

    @staticmethod
    def _ptr(a):
        return None if a is None else a.ctypes.data_as(C.c_void_p)

    # ---- scene upload -----------------------------------------------------------------------
    def upload_scene(self, scene: SceneDesc, build: bool = True):
        texs = (Texture * max(1, len(scene.textures)))()
        for i, t in enumerate(scene.textures):
            texs[i].type = t.type
            texs[i].p[:] = [float(v) for v in t.p]
        self._check(self.lib.mtsgpu_upload_textures(self.h, texs, len(scene.textures)))
        for i, t in enumerate(scene.textures):
            if t.pixels is not None and t.pixels.dtype == np.float32:      # exrtexture: linear floats
                px = np.ascontiguousarray(t.pixels, np.float32)
                self._check(self.lib.mtsgpu_upload_bitmap(self.h, i, px.shape[1], px.shape[0], px.shape[2], self._ptr(px),
                                                          BITMAP_F32, 1.0))
            elif t.pixels is not None:
                px = np.ascontiguousarray(t.pixels, np.uint8)
                self._check(self.lib.mtsgpu_upload_bitmap(self.h, i, px.shape[1], px.shape[0], px.shape[2], self._ptr(px),
                                                          BITMAP_U8, t.gamma))
        mats = (Material * max(1, len(scene.materials)))()
        for i, m in enumerate(scene.materials):
            mats[i].type = m.type
            mats[i].p[:] = [float(v) for v in m.p]
            mats[i].texture[:] = ([int(v) for v in m.tex] + [0, 0, 0])[:3]
        self._check(self.lib.mtsgpu_upload_materials(self.h, mats, len(scene.materials)))
        ems = (Emitter * max(1, len(scene.emitters)))()
        for i, e in enumerate(scene.emitters):
            ems[i].intensity[:] = [float(v) for v in e.intensity]
            ems[i].samplingWeight = e.sampling_weight
            ems[i].type = e.type
            ems[i].p[:] = [float(v) for v in e.abi_params()]
        self._check(self.lib.mtsgpu_upload_emitters(self.h, ems, len(scene.emitters)))
        for i, e in enumerate(scene.emitters):
            if e.type == 6:                        # collimated beam: toWorld and its fp32 Gauss-Jordan inverse
                from .scenefile import invert4_f32
                l2w = np.ascontiguousarray(np.asarray(e.aux, np.float32).reshape(4, 4))
                w2l = np.ascontiguousarray(invert4_f32(l2w))
                self._check(self.lib.mtsgpu_upload_emitter_transform(self.h, i, self._ptr(l2w), self._ptr(w2l)))
            if e.pixels is not None:               # environment map: toWorld and its fp32 Gauss-Jordan inverse (Transform(Matrix4x4))
                from .scenefile import invert4_f32
                l2w = np.ascontiguousarray(np.asarray(e.aux, np.float32).reshape(4, 4))
                w2l = np.ascontiguousarray(invert4_f32(l2w))
                px = np.ascontiguousarray(e.pixels, np.float32)
                self._check(self.lib.mtsgpu_upload_envmap(self.h, i, px.shape[1], px.shape[0], px.shape[2], self._ptr(px),
                                                          self._ptr(l2w), self._ptr(w2l)))
        for g, group in enumerate(scene.groups):   # shape groups: meshes in object space, their own acceleration structure
            for k, s in enumerate(group):
                pos = np.ascontiguousarray(s.positions, np.float32)
                nrm = None if s.normals is None else np.ascontiguousarray(s.normals, np.float32)
                tri = np.ascontiguousarray(s.triangles, np.uint32)
                uv = None if s.texcoords is None else np.ascontiguousarray(s.texcoords, np.float32)
                self._check(self.lib.mtsgpu_upload_group_mesh(self.h, g, k, self._ptr(pos), self._ptr(nrm), self._ptr(uv), len(pos),
                                                              self._ptr(tri), len(tri), s.material))
        for i, s in enumerate(scene.shapes):
            if s.cylinder is not None:
                from .scenefile import cylinder_params
                radius, length, o2w, w2o = cylinder_params(s.cylinder)
                o2w, w2o = np.ascontiguousarray(o2w, np.float32), np.ascontiguousarray(w2o, np.float32)
                self._check(self.lib.mtsgpu_upload_cylinder(self.h, i, self._ptr(o2w), self._ptr(w2o), radius, length, s.material))
                continue
            if s.instance is not None:              # toWorld and its fp32 Gauss-Jordan inverse (Transform(Matrix4x4))
                from .scenefile import invert4_f32
                o2w = np.ascontiguousarray(s.instance[1], np.float32)
                w2o = np.ascontiguousarray(invert4_f32(o2w))
                self._check(self.lib.mtsgpu_upload_instance(self.h, i, s.instance[0], self._ptr(o2w), self._ptr(w2o)))
                continue
            if s.sphere is not None:                # center + radius: objectToWorld = translate(center) (sphere.cpp:44-47)
                center, radius, inverted = s.sphere
                o2w = np.eye(4, dtype=np.float32); o2w[:3, 3] = center
                w2o = np.eye(4, dtype=np.float32); w2o[:3, 3] = -np.asarray(center, np.float32)
                c = np.asarray(center, np.float32)
                self._check(self.lib.mtsgpu_upload_sphere(self.h, i, self._ptr(o2w), self._ptr(w2o), radius, self._ptr(c),
                                                          int(inverted), s.material, s.emitter))
                continue
            pos = np.ascontiguousarray(s.positions, np.float32)
            nrm = None if s.normals is None else np.ascontiguousarray(s.normals, np.float32)
            tri = np.ascontiguousarray(s.triangles, np.uint32)
            uv = None if s.texcoords is None else np.ascontiguousarray(s.texcoords, np.float32)
            self._check(self.lib.mtsgpu_upload_mesh(self.h, i, self._ptr(pos), self._ptr(nrm), self._ptr(uv), len(pos),
                                                    self._ptr(tri), len(tri), s.material, s.emitter))
        self.set_camera(scene)
        self.set_film(scene)
        if build:
            self.build_accel()

    def set_camera(self, scene: SceneDesc):
        r2c = np.ascontiguousarray(raster_to_camera(scene), np.float32)
        c2w = np.ascontiguousarray(scene.camera_to_world, np.float32)
        if scene.orthographic:
            self._check(self.lib.mtsgpu_set_camera_orthographic(self.h, self._ptr(r2c), self._ptr(c2w), scene.near_clip,
                                                                scene.far_clip, scene.shutter_open, max(scene.shutter_close - scene.shutter_open, 0.0)))
            return
        self._check(self.lib.mtsgpu_set_camera(self.h, self._ptr(r2c), self._ptr(c2w), scene.near_clip, scene.far_clip,
                                               scene.aperture_radius, scene.focus_depth, scene.shutter_open,
                                               max(scene.shutter_close - scene.shutter_open, 0.0)))

    def set_film(self, scene: SceneDesc, table=None, size=None):
        cx, cy, cw, ch = scene.crop_rect
        if table is None:
            table, size = filter_table(scene.filter_type)
        table = np.ascontiguousarray(table, np.float32)
        self._check(self.lib.mtsgpu_set_film(self.h, scene.width, scene.height, cx, cy, cw, ch, self._ptr(table), size, size))
        self.film_shape = (ch, cw, 5)

    def build_accel(self):
        self._check(self.lib.mtsgpu_build_accel(self.h))

    def accel_info(self):
        info = AccelInfo()
        self._check(self.lib.mtsgpu_accel_info_get(self.h, C.byref(info)))
        return {n: getattr(info, n) for n, _ in info._fields_}

    # ---- hot path ---------------------------------------------------------------------------
    @staticmethod
    def params_from_scene(scene: SceneDesc, spp_begin=0, spp_end=None, seed=0) -> RenderParams:
        p = RenderParams()
        p.mode = scene.mode; p.maxDepth = scene.max_depth; p.rrDepth = scene.rr_depth
        p.strictNormals = int(scene.strict_normals)
        p.luminaireSamples = scene.luminaire_samples; p.bsdfSamples = scene.bsdf_samples
        p.sppTotal = scene.spp; p.sppBegin = spp_begin; p.sppEnd = scene.spp if spp_end is None else spp_end
        p.seed = seed; p.importanceSampleLuminaires = int(getattr(scene, "importance_sample_luminaires", True))
        p.samplerType = getattr(scene, "sampler_type", 0); p.samplerDepth = getattr(scene, "sampler_depth", 3)
        return p

    def render(self, params: RenderParams):
        """mtsgpu_render: returns (film[cropH, cropW, 5] host array, stats dict)."""
        out = np.empty(self.film_shape, np.float32)
        st = Stats()
        self._check(self.lib.mtsgpu_render(self.h, C.byref(params), None, self._ptr(out), C.byref(st)))
        return out, st.as_dict()

    def render_device(self, params: RenderParams, d_film_ptr: int):
        st = Stats()
        self._check(self.lib.mtsgpu_render_device(self.h, C.byref(params), None, C.c_void_p(d_film_ptr), C.byref(st)))
        return st.as_dict()

    def trace(self, rays: np.ndarray, mode: int = TRACE_CLOSEST):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        out = np.empty(n, HIT_DTYPE if mode == TRACE_CLOSEST else np.uint32)
        self._check(self.lib.mtsgpu_trace(self.h, self._ptr(rays), n, mode, self._ptr(out)))
        return out

    def trace_device(self, d_rays_ptr: int, n: int, mode: int, d_out_ptr: int) -> float:
        ms = C.c_float(0)
        self._check(self.lib.mtsgpu_trace_device(self.h, C.c_void_p(d_rays_ptr), n, mode, C.c_void_p(d_out_ptr), C.byref(ms)))
        return ms.value

    def set_progress(self, fn=None, min_interval=0.0, want_film=True):
        """mtsgpu_set_progress: fn(fraction, film or None) is called between wavefront batches on the rendering thread; film is a
        (cropH, cropW, 5) copy of the accumulators at that point."""
        if fn is None:
            self._progress_cb = None
            self._check(self.lib.mtsgpu_set_progress(self.h, C.cast(None, PROGRESS_FN), None, 0.0, 0))
            return
        shape = self.film_shape

        def trampoline(_user, fraction, film_ptr):
            film = None
            if film_ptr:
                film = np.ctypeslib.as_array(film_ptr, shape=(shape[0] * shape[1] * 5,)).reshape(shape).copy()
            fn(float(fraction), film)
        self._progress_cb = PROGRESS_FN(trampoline)          # keep the ctypes thunk alive
        self._check(self.lib.mtsgpu_set_progress(self.h, self._progress_cb, None, float(min_interval), int(want_film)))

    def set_counters(self, enabled: bool):
        self._check(self.lib.mtsgpu_set_counters(self.h, int(enabled)))

    def counters(self):
        c = TraceCounters()
        self._check(self.lib.mtsgpu_get_counters(self.h, C.byref(c)))
        return {n: getattr(c, n) for n, _ in c._fields_}

    # ---- test hooks -------------------------------------------------------------------------
    def _hook(self, fn, q, cols_in, cols_out):
        q = np.ascontiguousarray(q, np.float32).reshape(-1, cols_in)
        out = np.zeros((len(q), cols_out), np.float32)
        self._check(fn(self.h, self._ptr(q), len(q), self._ptr(out)))
        return out

    def bsdf_eval(self, queries):
        return self._hook(self.lib.mtsgpu_bsdf_eval, queries, 9, 16)

    def emitter_sample(self, queries):
        return self._hook(self.lib.mtsgpu_emitter_sample, queries, 5, 16)

    def camera_rays(self, samples):
        return self._hook(self.lib.mtsgpu_camera_rays, samples, 4, 8)

    def film_splat(self, samples):
        q = np.ascontiguousarray(samples, np.float32).reshape(-1, 6)
        out = np.zeros(self.film_shape, np.float32)
        self._check(self.lib.mtsgpu_film_splat(self.h, self._ptr(q), len(q), self._ptr(out)))
        return out

    def film_splat_static(self, samples):
        """mtsgpu_film_splat_static: the splat for callers that may not allocate GPU memory (fixed built-in
        scratch block; see include/mtsgpu.h for the size limit)."""
        q = np.ascontiguousarray(samples, np.float32).reshape(-1, 6)
        out = np.zeros(self.film_shape, np.float32)
        self._check(self.lib.mtsgpu_film_splat_static(self.h, self._ptr(q), len(q), self._ptr(out)))
        return out

    def intersect_full(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.empty(len(rays), HIT_DTYPE)
        out = np.zeros((len(rays), 12), np.float32)
        self._check(self.lib.mtsgpu_intersect_full(self.h, self._ptr(rays), len(rays), self._ptr(hits), self._ptr(out)))
        return hits, out

    def sampler_values(self, params: RenderParams, pixel: int, draw: int):
        """2D draw number `draw` of every sample of one pixel: [sppTotal, 2]."""
        out = np.empty((params.sppTotal, 2), np.float32)
        self._check(self.lib.mtsgpu_sampler_values(self.h, C.byref(params), pixel, draw, self._ptr(out)))
        return out

    def film_develop(self, film, has_alpha=True):
        """EXRFilm::develop's normalisation on the GPU: accumulators [H,W,5] -> RGBA [H,W,4]."""
        film = np.ascontiguousarray(film, np.float32)
        out = np.empty(film.shape[:2] + (4,), np.float32)
        self._check(self.lib.mtsgpu_film_develop(self.h, self._ptr(film), int(has_alpha), self._ptr(out)))
        return out

    def film_save(self, film, path, fmt=FILM_EXR, option=1, device_ptr=None):
        """Film::develop(destFile): `film` is a host array, or pass `device_ptr` for a film left on the GPU."""
        if device_ptr is not None:
            self._check(self.lib.mtsgpu_film_save(self.h, C.c_void_p(device_ptr), 1, fmt, int(option), path.encode()))
        else:
            film = np.ascontiguousarray(film, np.float32)
            self._check(self.lib.mtsgpu_film_save(self.h, self._ptr(film), 0, fmt, int(option), path.encode()))

    def texture_lookup(self, samples, sample_count=1):
        """Camera samples {x, y, lens x, lens y} -> rows {uv, dudx, dudy, dvdx, dvdy, slot 0 rgb, slot 1 rgb} at the hit
        of the camera's ray differential (mtsgpu_texture_lookup)."""
        samples = np.ascontiguousarray(samples, np.float32).reshape(-1, 4)
        out = np.zeros((len(samples), 12), np.float32)
        self._check(self.lib.mtsgpu_texture_lookup(self.h, self._ptr(samples), len(samples), sample_count, self._ptr(out)))
        return out

    def intersect_uv(self, rays):
        """Closest hit -> rows {uv, texture slot 0 rgb, slot 1 rgb} of the hit's material at that uv."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out = np.zeros((len(rays), 8), np.float32)
        self._check(self.lib.mtsgpu_intersect_uv(self.h, self._ptr(rays), len(rays), self._ptr(out)))
        return out

    def triaccel(self, n):
        rec = np.empty(n, TRIACCEL_DTYPE)
        box = np.zeros(6, np.float32)
        self._check(self.lib.mtsgpu_get_triaccel(self.h, self._ptr(rec), n, self._ptr(box)))
        return rec, box


def write_exr(path, rgba):
    """Bitmap::saveEXR layout (half A,B,G,R, ZIP) for an [H,W,4] float array; host only."""
    lib = load_library()
    a = np.ascontiguousarray(rgba, np.float32)
    rc = lib.mtsgpu_write_exr(path.encode(), a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0])
    if rc != 0:
        raise MtsGpuError(rc, lib.mtsgpu_last_error(None).decode())


def write_mfilm(path, values):
    """MFilm::develop text layout for [H,W] (luminance) or [H,W,3] (spectra) values; host only."""
    lib = load_library()
    a = np.ascontiguousarray(values, np.float32)
    ch = 1 if a.ndim == 2 else a.shape[2]
    rc = lib.mtsgpu_write_mfilm(path.encode(), a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0], ch)
    if rc != 0:
        raise MtsGpuError(rc, lib.mtsgpu_last_error(None).decode())


def read_exr(path):
    """Minimal reader for the files write_exr produces (scanline, HALF channels, ZIP / ZIPS / uncompressed): returns
    {channel name: [H,W] float16}. Used by the tests to check the writer; not a general OpenEXR reader."""
    import struct
    import zlib
    d = open(path, "rb").read()
    magic, version = struct.unpack_from("<ii", d, 0)
    if magic != 20000630 or (version & 0xFF) != 2 or (version & ~0xFF):
        raise ValueError("not a plain scanline OpenEXR file")
    pos, attrs = 8, {}
    while d[pos] != 0:
        e = d.index(b"\0", pos); name = d[pos:e].decode(); pos = e + 1
        e = d.index(b"\0", pos); typ = d[pos:e].decode(); pos = e + 1
        size, = struct.unpack_from("<i", d, pos); pos += 4
        attrs[name] = (typ, d[pos:pos + size]); pos += size
    pos += 1
    chans, c = [], attrs["channels"][1]
    p = 0
    while c[p] != 0:
        e = c.index(b"\0", p); nm = c[p:e].decode(); p = e + 1
        ptype, _, xs, ys = struct.unpack_from("<i4sii", c, p); p += 16
        if ptype != 1 or xs != 1 or ys != 1:
            raise ValueError("only HALF channels without subsampling")
        chans.append(nm)
    comp = attrs["compression"][1][0]
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    lines = {0: 1, 2: 1, 3: 16}[comp]
    nblk = (h + lines - 1) // lines
    offs = struct.unpack_from(f"<{nblk}Q", d, pos)
    out = {nm: np.zeros((h, w), np.float16) for nm in chans}
    for off in offs:
        y, size = struct.unpack_from("<ii", d, off)
        blob = d[off + 8:off + 8 + size]
        rows = min(lines, y1 + 1 - y)
        raw_size = rows * len(chans) * w * 2
        if comp != 0 and size < raw_size:
            t = np.frombuffer(zlib.decompress(blob), np.uint8).astype(np.int64)
            t = np.cumsum(np.concatenate([t[:1], t[1:] - 128])) & 0xFF          # undo the delta predictor
            half = (raw_size + 1) // 2
            raw = np.empty(raw_size, np.uint8); raw[0::2] = t[:half]; raw[1::2] = t[half:]
        else:
            raw = np.frombuffer(blob, np.uint8)
        v = raw.view(np.float16).reshape(rows, len(chans), w)
        for k, nm in enumerate(chans):
            out[nm][y - y0:y - y0 + rows] = v[:, k]
    return out
