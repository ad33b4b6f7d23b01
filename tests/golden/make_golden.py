#!/usr/bin/env python
"""Regenerates tests/golden/*.npz from the compiled reference (oracle/_ref/parity, built from
/root/reference by oracle/Makefile). Run in the build container:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import golden_cases as gc, oracle_util  # noqa: E402
from mitsuba_renderer_b200 import scenes  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def save(name, **arrays):
    np.savez_compressed(os.path.join(OUT, name), **arrays)
    print(name, {k: getattr(v, "shape", v) for k, v in arrays.items()})


def main():
    assert oracle_util.available(), "build oracle/_ref first (make -C oracle VARIANT=parity)"
    # ---- ray queries -------------------------------------------------------------------------------
    for name, scene in (("trace_materials", gc.scene_materials()), ("trace_grid", gc.scene_grid())):
        rays = gc.trace_rays(scene)
        sess = oracle_util.Session(scene)
        hits, full, _ = sess.trace(rays, "closest_full")
        seg = scenes.shadow_segments(rays, scene)
        save(name, rays_digest=gc.digest(rays), hits=hits.view(np.uint8), full=full,
             any_segments=sess.trace(seg, "any")[0].astype(np.uint8), any_rays=sess.trace(rays, "any")[0].astype(np.uint8))
    # spheres, sphere luminaires and instances of shape groups: non-triangle primitives of the scene's structure
    for name, scene in (("trace_spheres", gc.scene_spheres()), ("trace_instances", gc.scene_instances())):
        rays = gc.trace_rays(scene)
        sess = oracle_util.Session(scene)
        hits, full, _ = sess.trace(rays, "closest_full")
        nontri = hits["prim"] == 0xFFFFFFFF
        hits["u"][nontri] = 0; hits["v"][nontri] = 0         # left unset by the reference for non-triangles
        seg = scenes.shadow_segments(rays, scene)
        save(name, rays_digest=gc.digest(rays), hits=hits.view(np.uint8), full=full,
             any_segments=sess.trace(seg, "any")[0].astype(np.uint8), any_rays=sess.trace(rays, "any")[0].astype(np.uint8))
    scene = gc.scene_smoke()
    rays = scenes.incoherent_rays(scene, 20000, 7)
    sess = oracle_util.Session(scene)
    save("smoke_cornell", rays_digest=gc.digest(rays), hits=sess.trace(rays, "closest")[0].view(np.uint8),
         occluded=sess.trace(scenes.shadow_segments(rays, scene), "any")[0])
    # ---- `.serialized` meshes: the reference's loader on its own fixture and on files written by our writer ----
    import tempfile
    from mitsuba_renderer_b200 import serialized
    fx = {}
    for i in range(3):
        m = oracle_util.load_serialized(gc.SERIALIZED_FIXTURE, i)
        fx[f"fixture{i}"] = np.frombuffer(gc.mesh_digest(m).encode(), np.uint8)
        fx[f"fixture{i}_counts"] = np.array([len(m["positions"]), len(m["triangles"])])
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "two.serialized")
        serialized.save(path, gc.serialized_meshes())
        for idx in (0, 1):
            for fn, fl in ((0, 0), (0, 1), (1, 0), (1, 1)):
                m = oracle_util.load_serialized(path, idx, fn, fl)
                fx[f"mesh{idx}_{fn}{fl}_normals"] = np.zeros((0, 3), np.float32) if m["normals"] is None else m["normals"]
                fx[f"mesh{idx}_{fn}{fl}_triangles"] = m["triangles"]
        m = oracle_util.load_serialized(path, 1, to_world=gc.SERIALIZED_TO_WORLD)
        fx["mesh1_xf_positions"], fx["mesh1_xf_normals"], fx["mesh1_xf_triangles"] = m["positions"], m["normals"], m["triangles"]
    save("serialized", **fx)
    # ---- MFilm::develop text for fixed accumulators ------------------------------------------------------------------
    acc = gc.film_accumulators()
    save("mfilm", acc_digest=gc.digest(acc), luminance=np.frombuffer(oracle_util.develop_mfilm(acc, False), np.uint8),
         spectra=np.frombuffer(oracle_util.develop_mfilm(acc, True), np.uint8))
    # ---- texture coordinates + textured reflectances at the hit ---------------------------------------------------------
    ts = gc.scene_textured()
    rays = gc.trace_rays(ts)
    tsess = oracle_util.Session(ts)
    thits, tfull, _ = tsess.trace(rays, "closest_full")       # `wi` in the tangent frames of the anisotropic Ward meshes
    save("uv_textured", rays_digest=gc.digest(rays), out=tsess.uv(rays), hits=thits.view(np.uint8), full=tfull)
    # ---- environment map: importance sampling and pdf -------------------------------------------------------------------
    es = gc.scene_envmap()
    q = gc.emitter_queries(es, 6000)
    save("emitter_envmap", queries_digest=gc.digest(q), out=oracle_util.Session(es).emitter(q))
    # ---- bitmap textures filtered with the uv partials of the camera's ray differentials -------------------------------------
    bs_ = gc.scene_bitmaps()
    q = gc.lens_camera_samples(bs_, 6000, 31)
    save("texlookup_bitmaps", samples_digest=gc.digest(q), out=oracle_util.Session(bs_).texlookup(q))
    # ---- deterministic render: halton sampler (the same sample values on every pixel, no random state) ----------------------
    hs = gc.scene_halton()
    save("render_halton", film=oracle_util.Session(hs).render()[0])
    # ---- TriAccel records + enlarged scene box --------------------------------------------------------
    rec, box, _ = oracle_util.Session(gc.scene_materials()).triaccel()
    save("triaccel_materials", records=rec.view(np.uint8), box=box)
    # ---- BSDF f / pdf / sample -------------------------------------------------------------------------
    bs = gc.bsdf_scene()
    q = gc.bsdf_queries(len(bs.materials), 4000, 21)
    save("bsdf", queries_digest=gc.digest(q), out=oracle_util.Session(bs).bsdf(q))
    # ---- luminaire sampling ---------------------------------------------------------------------------
    ms = gc.scene_materials()
    q = gc.emitter_queries(ms, 4000)
    save("emitter", queries_digest=gc.digest(q), out=oracle_util.Session(ms).emitter(q))
    ls = gc.scene_lights()                             # point, spot, directional, constant + the two area lights
    q = gc.emitter_queries(ls, 6000)
    save("emitter_lights", queries_digest=gc.digest(q), out=oracle_util.Session(ls).emitter(q))
    # ---- camera -------------------------------------------------------------------------------------
    for ap in (0.0, 0.05):
        cs = gc.camera_scene(ap)
        q = gc.camera_samples(cs, 4000)
        save(f"camera_{'pinhole' if ap == 0 else 'thinlens'}", samples_digest=gc.digest(q), rays=oracle_util.Session(cs).camera(q))
    # ---- film splat + tabulated filters ----------------------------------------------------------------
    for name, (ft, crop) in gc.SPLAT_CASES.items():
        ss = gc.splat_scene(ft, crop)
        q = gc.splat_samples(ss, 8000)
        film, table, info = oracle_util.Session(ss).splat(q)
        save(f"splat_{name}", samples_digest=gc.digest(q), film=film, table=table, filter_size=np.float32(info["filter_size"][0]))
    # ---- image: the reference's own path tracer at high spp ---------------------------------------------
    rs = gc.render_scene()
    import copy
    hi = copy.copy(rs); hi.spp = 4096
    ref, _ = oracle_util.Session(hi).render()
    cpu, _ = oracle_util.Session(rs).render()
    save("render_cornell48", ref_4096=ref, cpu_64=cpu)


if __name__ == "__main__":
    main()
