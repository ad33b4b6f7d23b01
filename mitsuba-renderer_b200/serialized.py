"""`.serialized` triangle-mesh files through the library's native reader / writer (csrc/serialized.cpp).

Mirrors the reference's `serialized` shape (src/shapes/serialized.cpp:30-91: filename, shapeIndex, toWorld,
faceNormals, flipNormals) — the mesh comes back exactly as that shape would hand it to the kd-tree. Needs no GPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import api
from .scenefile import Shape


class MeshData(C.Structure):
    _fields_ = [("nVert", C.c_uint32), ("nTri", C.c_uint32), ("flags", C.c_uint32),
                ("positions", C.POINTER(C.c_float)), ("normals", C.POINTER(C.c_float)),
                ("texcoords", C.POINTER(C.c_float)), ("colors", C.POINTER(C.c_float)),
                ("triangles", C.POINTER(C.c_uint32))]


def _lib():
    lib = api.load_library()
    if not getattr(lib, "_serialized_ready", False):
        lib.mtsgpu_serialized_count.argtypes = [C.c_char_p, C.POINTER(C.c_int)]
        lib.mtsgpu_serialized_load.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(MeshData)]
        lib.mtsgpu_serialized_free.argtypes = [C.POINTER(MeshData)]; lib.mtsgpu_serialized_free.restype = None
        lib.mtsgpu_serialized_save.argtypes = [C.c_char_p, C.POINTER(MeshData), C.c_int]
        lib.mtsgpu_upload_serialized.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.c_void_p, C.c_void_p,
                                                 C.c_int, C.c_int, C.c_int, C.c_int]
        lib._serialized_ready = True
    return lib


def _check(lib, rc):
    if rc != 0:
        raise api.MtsGpuError(rc, lib.mtsgpu_last_error(None).decode())


def _mat(m):
    return None if m is None else np.ascontiguousarray(m, np.float32).reshape(4, 4)


def count(path: str) -> int:
    lib = _lib()
    n = C.c_int(0)
    _check(lib, lib.mtsgpu_serialized_count(path.encode(), C.byref(n)))
    return n.value


def load(path: str, shape_index: int = 0, to_world=None, to_world_inv=None, face_normals=False, flip_normals=False):
    """-> dict(positions [V,3], normals [V,3] | None, texcoords [V,2] | None, colors | None, triangles [T,3], flags)"""
    lib = _lib()
    m = MeshData()
    tw, ti = _mat(to_world), _mat(to_world_inv)
    _check(lib, lib.mtsgpu_serialized_load(path.encode(), shape_index, None if tw is None else tw.ctypes.data,
                                           None if ti is None else ti.ctypes.data, int(face_normals), int(flip_normals), C.byref(m)))
    try:
        def arr(ptr, n, k, dt):
            return None if not ptr else np.ctypeslib.as_array(ptr, shape=(n * k,)).astype(dt).reshape(n, k).copy()
        return {"positions": arr(m.positions, m.nVert, 3, np.float32), "normals": arr(m.normals, m.nVert, 3, np.float32),
                "texcoords": arr(m.texcoords, m.nVert, 2, np.float32), "colors": arr(m.colors, m.nVert, 3, np.float32),
                "triangles": arr(m.triangles, m.nTri, 3, np.uint32), "flags": int(m.flags)}
    finally:
        lib.mtsgpu_serialized_free(C.byref(m))


def load_shape(path: str, shape_index: int = 0, material: int = 0, emitter: int = -1, **kw) -> Shape:
    d = load(path, shape_index, **kw)
    return Shape(d["positions"], d["triangles"], d["normals"], material=material, emitter=emitter, texcoords=d["texcoords"])


def save(path: str, meshes: Sequence, face_normals: Optional[Sequence[bool]] = None):
    """meshes: `Shape`s or dicts with positions / triangles / normals (/ texcoords / colors)."""
    lib = _lib()
    arr = (MeshData * len(meshes))()
    keep = []
    for i, s in enumerate(meshes):
        get = (lambda k: s.get(k)) if isinstance(s, dict) else (lambda k: getattr(s, k, None))
        pos = np.ascontiguousarray(get("positions"), np.float32); tri = np.ascontiguousarray(get("triangles"), np.uint32)
        arr[i].nVert, arr[i].nTri = len(pos), len(tri)
        arr[i].flags = 0x10 if (face_normals and face_normals[i]) else 0
        arr[i].positions = pos.ctypes.data_as(C.POINTER(C.c_float)); arr[i].triangles = tri.ctypes.data_as(C.POINTER(C.c_uint32))
        keep += [pos, tri]
        for key in ("normals", "texcoords", "colors"):
            v = get(key)
            if v is not None:
                v = np.ascontiguousarray(v, np.float32); keep.append(v)
                setattr(arr[i], key, v.ctypes.data_as(C.POINTER(C.c_float)))
    _check(lib, lib.mtsgpu_serialized_save(path.encode(), arr, len(meshes)))


def upload(ctx: "api.Context", shape_slot: int, path: str, shape_index: int = 0, to_world=None, to_world_inv=None,
           face_normals=False, flip_normals=False, material: int = 0, emitter: int = -1):
    """File -> GPU scene (mtsgpu_upload_serialized): shape `shape_slot` of the context's scene."""
    lib = _lib()
    tw, ti = _mat(to_world), _mat(to_world_inv)
    rc = lib.mtsgpu_upload_serialized(ctx.h, shape_slot, path.encode(), shape_index, None if tw is None else tw.ctypes.data,
                                      None if ti is None else ti.ctypes.data, int(face_normals), int(flip_normals), material, emitter)
    ctx._check(rc)
