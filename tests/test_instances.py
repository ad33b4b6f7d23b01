"""Instancing (src/shapes/shapegroup.cpp, src/shapes/instance.cpp): shape groups with an acceleration
structure of their own, shown through transformed instances that are single primitives of the scene's structure —
ray queries, intersection records and deterministic renders against the compiled reference."""
import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile
from mitsuba_renderer_b200.api import Context, TRACE_CLOSEST, TRACE_ANY
from tests import golden_cases as gc
from tests.test_trace_parity import check_closest
from tests.test_shading_parity import close

pytestmark = pytest.mark.gpu
INSTANCES = [11, 12, 13, 14, 15]        # shape indices of the instances in scenes.instances_test_scene


def test_ray_queries_and_intersection_records(oracle, gpu_ctx_factory):
    scene = scenes.instances_test_scene(32, 32, 1)
    rays = gc.trace_rays(scene, n=8000)
    sess = oracle.Session(scene)
    ref, ref_full, _ = sess.trace(rays, "closest_full")
    ctx = gpu_ctx_factory(scene)
    hits = ctx.trace(rays, TRACE_CLOSEST)
    on_inst = np.isin(ref["shape"], INSTANCES)
    ref["u"][on_inst] = 0; ref["v"][on_inst] = 0           # the outer intersection cache leaves u, v unset for non-triangles
    got = hits.copy(); got["u"][np.isin(got["shape"], INSTANCES)] = 0; got["v"][np.isin(got["shape"], INSTANCES)] = 0
    check_closest(got, ref, "instances")
    for k in INSTANCES:
        assert (ref["shape"] == k).sum() > 100, k
    assert (hits["prim"][on_inst] == 0xFFFFFFFF).all()      # KNoTriangleFlag
    # t comes from the TriAccel test on the object-space ray, with the reference's own float sequence: bit-identical
    assert np.array_equal(hits["t"].view(np.uint32), ref["t"].view(np.uint32))
    seg = scenes.shadow_segments(rays, scene)
    assert np.array_equal(ctx.trace(seg, TRACE_ANY), sess.trace(seg, "any")[0])
    assert np.array_equal(ctx.trace(rays, TRACE_ANY), sess.trace(rays, "any")[0])
    # Instance::fillIntersectionRecord: p on the world ray, frames through the instance transform
    _, full = ctx.intersect_full(rays)
    same = (hits["shape"] == ref["shape"]) & (ref["shape"] != 0xFFFFFFFF)
    ok = close(full[same], ref_full[same], 2e-5, 2e-6)
    assert ok.all(), f"{np.count_nonzero(~ok.all(axis=1))} records differ, e.g. {full[same][~ok.all(axis=1)][0]} vs {ref_full[same][~ok.all(axis=1)][0]}"
    uv_ref = sess.uv(rays)
    uv = ctx.intersect_uv(rays)
    assert close(uv[same, :2], uv_ref[same, :2], 2e-5, 2e-6).all()


@pytest.mark.parametrize("mode", [scenefile.MODE_PATH, scenefile.MODE_DIRECT])
def test_halton_render_with_instances_matches_reference(oracle, gpu_ctx_factory, mode):
    scene = scenes.instances_test_scene(64, 64, 16, mode, max_depth=7)
    scene.rr_depth = 3
    scene.sampler_type = api.SAMPLER_HALTON
    cpu = oracle.Session(scene).render()[0]
    gpu, _ = gpu_ctx_factory(scene).render(Context.params_from_scene(scene, seed=1))
    assert np.array_equal(gpu[..., 3:], cpu[..., 3:])
    w = np.maximum(cpu[..., 4:5], 1e-20)
    c, g = cpu[..., :3] / w, gpu[..., :3] / w
    err = np.abs(g - c).max(axis=-1) / (np.abs(c).max(axis=-1) + 1e-3)
    print(f"mode {mode}: within 2e-3: {(err < 2e-3).mean():.4f}, median {np.median(err):.2e}, means {c.mean():.6f} {g.mean():.6f}")
    assert (err < 2e-3).mean() > 0.995 and np.median(err) < 1e-6


def test_instance_refusals():
    scene = scenes.instances_test_scene(16, 16, 1)
    scene.shapes[11].instance = (5, scene.shapes[11].instance[1])
    ctx = Context(0)
    with pytest.raises(api.MtsGpuError, match="shape group 5 was not uploaded"):
        ctx.upload_scene(scene)
    ctx.close()
