"""TEST INFRASTRUCTURE: drives oracle/_ref/<variant>/mts_oracle (the reference's own code, compiled
by oracle/Makefile) on scenes written by mitsuba_renderer_b200.scenefile. Only tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs may import this."""
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("shape", np.uint32), ("prim", np.uint32)])


def binary(variant="parity"):
    return os.path.join(ROOT, "oracle", "_ref", variant, "mts_oracle")


def available(variant="parity"):
    return os.path.exists(binary(variant))


def _run(args, variant="parity", timeout=900, env=None):
    e = dict(os.environ)
    if env:
        e.update(env)
    p = subprocess.run([binary(variant)] + [str(a) for a in args], capture_output=True, text=True, timeout=timeout, env=e)
    if p.returncode != 0:
        raise RuntimeError(f"mts_oracle {' '.join(map(str, args))} failed ({p.returncode}):\n{p.stdout}\n{p.stderr}")
    for line in reversed(p.stdout.strip().splitlines()):
        i = line.find('{"cmd"')
        if i >= 0:
            return json.loads(line[i:])
    raise RuntimeError("mts_oracle printed no result line:\n" + p.stdout + p.stderr)


def load_serialized(path, shape_index=0, face_normals=False, flip_normals=False, to_world=None, variant="parity"):
    """One mesh of a `.serialized` file as the reference's `serialized` shape plugin produces it."""
    with tempfile.TemporaryDirectory(prefix="mtsoracle_") as d:
        out = os.path.join(d, "mesh.bin")
        args = ["mesh", path, shape_index, out, int(face_normals), int(flip_normals)]
        if to_world is not None:
            args += [repr(float(v)) for v in np.asarray(to_world, np.float32).reshape(16)]
        _run(args, variant)
        raw = np.fromfile(out, np.uint8)
    nv, nt, hn, ht = raw[:16].view(np.uint32)
    o = 16
    def take(count, dt):
        nonlocal o
        a = raw[o:o + 4 * count].view(dt).copy(); o += 4 * count
        return a
    pos = take(3 * nv, np.float32).reshape(nv, 3)
    nrm = take(3 * nv, np.float32).reshape(nv, 3) if hn else None
    uv = take(2 * nv, np.float32).reshape(nv, 2) if ht else None
    tri = take(3 * nt, np.uint32).reshape(nt, 3)
    return {"positions": pos, "normals": nrm, "texcoords": uv, "triangles": tri}


def develop_mfilm(accum, spectra=False, variant="parity"):
    """The text file the reference's MFilm::develop writes for the accumulators [H,W,5]."""
    a = np.ascontiguousarray(accum, np.float32)
    with tempfile.TemporaryDirectory(prefix="mtsoracle_") as d:
        inp, out = os.path.join(d, "acc.bin"), os.path.join(d, "film.m")
        a.tofile(inp)
        _run(["develop", inp, a.shape[1], a.shape[0], out, int(spectra)], variant)
        return open(out, "rb").read()


def save_serialized(scene, path, variant="parity"):
    """Every mesh of `scene` written by the reference's TriMesh::serialize + offset table."""
    from mitsuba_renderer_b200 import scenefile
    with tempfile.TemporaryDirectory(prefix="mtsoracle_") as d:
        sp = os.path.join(d, "scene.bin")
        scenefile.write_scene(sp, scene)
        return _run(["meshsave", sp, path], variant)


class Session:
    """A scene written once to a temp dir; each call runs one harness command on it."""

    def __init__(self, scene, variant="parity"):
        from mitsuba_renderer_b200 import scenefile
        self.variant = variant
        self.dir = tempfile.TemporaryDirectory(prefix="mtsoracle_")
        self.scene_path = os.path.join(self.dir.name, "scene.bin")
        scenefile.write_scene(self.scene_path, scene)
        self.scene = scene

    def _io(self, cmd, arr, extra=()):
        inp = os.path.join(self.dir.name, "in.bin")
        out = os.path.join(self.dir.name, "out.bin")
        np.ascontiguousarray(arr, np.float32).tofile(inp)
        info = _run([cmd, self.scene_path, inp, out] + list(extra), self.variant)
        return out, info

    def trace(self, rays, mode="closest", threads=1, repeat=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out, info = self._io("trace", rays, [mode, threads, repeat])
        n = len(rays)
        raw = np.fromfile(out, np.uint32)
        if mode == "any":
            return raw.astype(np.uint32), info
        hits = np.empty(n, HIT_DTYPE)
        tuv = raw[:3 * n].view(np.float32).reshape(n, 3)
        ids = raw[3 * n:5 * n].reshape(n, 2)
        hits["t"], hits["u"], hits["v"] = tuv[:, 0], tuv[:, 1], tuv[:, 2]
        hits["shape"], hits["prim"] = ids[:, 0], ids[:, 1]
        if mode == "closest_full":
            full = raw[5 * n:].view(np.float32).reshape(n, 12)
            return hits, full, info
        return hits, info

    def kd_stats(self, rays, threads=1):
        """Per-ray work of the reference's kd-tree on these closest-hit queries, through its own instrumented traversal
        (rayIntersectHavranCollectStatistics, sahkdtree3.h:310-410): (nodes traversed, primitives intersected) per ray.
        Only the counts are used: that function reuses its output variable `t` for split distances (sahkdtree3.h:370)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out, info = self._io("trace", rays, ["stats", threads, 1])
        n = len(rays)
        raw = np.fromfile(out, np.uint32)
        return raw[3 * n:5 * n].reshape(n, 2).copy(), info

    def triaccel(self):
        out = os.path.join(self.dir.name, "ta.bin")
        info = _run(["triaccel", self.scene_path, out], self.variant)
        raw = np.fromfile(out, np.uint8)
        box = raw[:24].view(np.float32).copy()
        from mitsuba_renderer_b200.api import TRIACCEL_DTYPE
        return raw[24:].view(TRIACCEL_DTYPE).copy(), box, info

    def bsdf(self, queries):
        out, _ = self._io("bsdf", queries)
        return np.fromfile(out, np.float32).reshape(-1, 16)

    def uv(self, rays):
        """rows {its.uv, BSDF::getDiffuseReflectance(its)} of the closest hit; zeros on a miss."""
        out, _ = self._io("uv", np.ascontiguousarray(rays, np.float32).reshape(-1, 8))
        return np.fromfile(out, np.float32).reshape(-1, 5)

    def texlookup(self, samples):
        """rows {its.uv, dudx, dudy, dvdx, dvdy, BSDF::getDiffuseReflectance(its), 0 0 0} at the hit of the camera's
        ray differential through raster position / lens sample {x, y, lx, ly}; zeros on a miss."""
        out, _ = self._io("texlookup", np.ascontiguousarray(samples, np.float32).reshape(-1, 4))
        return np.fromfile(out, np.float32).reshape(-1, 12)

    def emitter(self, queries):
        out, _ = self._io("emitter", queries)
        return np.fromfile(out, np.float32).reshape(-1, 16)

    def camera(self, samples):
        out, _ = self._io("camera", samples)
        return np.fromfile(out, np.float32).reshape(-1, 8)

    def splat(self, samples):
        table = os.path.join(self.dir.name, "table.bin")
        out, info = self._io("splat", samples, [table])
        cx, cy, cw, ch = self.scene.crop_rect
        return np.fromfile(out, np.float32).reshape(ch, cw, 5), np.fromfile(table, np.float32).reshape(16, 16), info

    def render(self, threads=None, integrator=None, env=None, repeat=1):
        """repeat > 1: timing runs (info["render_s_all"]); the returned film then holds the sum of all passes."""
        out = os.path.join(self.dir.name, "film.bin")
        threads = threads or os.cpu_count() or 1
        args = ["render", self.scene_path, out, threads] + ([integrator] if integrator else (["-"] if repeat > 1 else []))
        if repeat > 1:
            args.append(repeat)
        e = dict(os.environ)
        if env:
            e.update(env)
        p = subprocess.run([binary(self.variant)] + [str(a) for a in args], capture_output=True, text=True, timeout=7200, env=e)
        if p.returncode != 0:
            raise RuntimeError(f"mts_oracle render failed:\n{p.stdout}\n{p.stderr}")
        info = None
        for line in p.stdout.splitlines():
            i = line.find('{"cmd"')
            if i >= 0:
                info = json.loads(line[i:])
        cx, cy, cw, ch = self.scene.crop_rect
        return np.fromfile(out, np.float32).reshape(ch, cw, 5), info


def parse_count(s):
    """Statistics prints '1.234 M' style numbers."""
    parts = str(s).split()
    if not parts:
        return 0.0
    v = float(parts[0])
    mult = {"K": 1e3, "M": 1e6, "G": 1e9, "T": 1e12}
    return v * mult.get(parts[1], 1.0) if len(parts) > 1 else v
