"""GPU parity: closest-hit / any-hit through the C ABI (mtsgpu_trace) vs the compiled reference's
kd-tree (ShapeKDTree::rayIntersect, skdtree.cpp:108-132,180-199) on the same seeded rays.

Bar (BASELINE.json north_star): shape/primitive ids bit-exact except exact-t ties, t within 1e-5
relative. We additionally record that t/u/v are bit-identical, which the design guarantees."""
import numpy as np
import pytest

from mitsuba_renderer_b200 import scenes, scenefile
from mitsuba_renderer_b200.api import TRACE_CLOSEST, TRACE_ANY

pytestmark = pytest.mark.gpu
T_REL_TOL = 1e-5


def check_closest(gpu, ref, label=""):
    n = len(ref)
    hit_r = ref["shape"] != 0xFFFFFFFF
    hit_g = gpu["shape"] != 0xFFFFFFFF
    assert np.array_equal(hit_r, hit_g), f"{label}: hit/miss differs for {np.count_nonzero(hit_r != hit_g)} of {n} rays"
    same_id = (gpu["shape"] == ref["shape"]) & (gpu["prim"] == ref["prim"])
    # documented tie class: two triangles yield a bit-identical t; the kd-tree keeps the last one
    # tested in its leaf order, the BVH another one
    tie = ~same_id & (gpu["t"].view(np.uint32) == ref["t"].view(np.uint32))
    bad = ~same_id & ~tie
    assert not bad.any(), f"{label}: {np.count_nonzero(bad)} id mismatches that are not exact-t ties"
    h = hit_r
    rel = np.abs(gpu["t"][h] - ref["t"][h]) / np.maximum(np.abs(ref["t"][h]), 1e-30)
    assert rel.size == 0 or rel.max() <= T_REL_TOL, f"{label}: max relative t error {rel.max()}"
    exact = same_id & h
    assert np.array_equal(gpu["t"][exact].view(np.uint32), ref["t"][exact].view(np.uint32)), f"{label}: t not bit-identical"
    assert np.array_equal(gpu["u"][exact].view(np.uint32), ref["u"][exact].view(np.uint32)), f"{label}: u not bit-identical"
    assert np.array_equal(gpu["v"][exact].view(np.uint32), ref["v"][exact].view(np.uint32)), f"{label}: v not bit-identical"
    return int(np.count_nonzero(tie))


def ray_sets(scene, n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    inc = scenes.incoherent_rays(scene, n, seed)
    coh = scenes.coherent_rays(scene, int(np.sqrt(n)))
    lo, hi = scenes.scene_bounds(scene)
    # rays starting inside the scene box, random directions, explicit mint/maxt
    o = lo + (hi - lo) * rng.random((n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    inside = np.empty((n, 8), np.float32)
    inside[:, :3] = o; inside[:, 3] = 1e-4; inside[:, 4:7] = d; inside[:, 7] = np.inf
    # axis-parallel rays (zero direction components) and rays that miss the box
    axis = inside[: n // 4].copy()
    ax = rng.integers(0, 3, len(axis))
    axis[:, 4:7] = 0
    axis[np.arange(len(axis)), 4 + ax] = rng.choice([-1.0, 1.0], len(axis))
    miss = inc[: n // 4].copy()
    miss[:, :3] += (hi - lo).max() * 10
    finite = inc[: n // 4].copy()
    finite[:, 7] = rng.random(len(finite)).astype(np.float32) * np.float32(np.linalg.norm(hi - lo))
    finite[:, 3] = 1e-3
    return {"incoherent": inc, "coherent": coh, "inside": inside, "axis": axis, "miss": miss, "finite": finite}


SCENES = {
    "cornell": lambda: scenes.cornell_box(32, 32, 1),
    "materials": lambda: scenes.material_test_scene(32, 32, 1),
    "grid64": lambda: scenes.microbench_scene(cells=64, pad_cells=9),
    "grid300": lambda: scenes.microbench_scene(cells=300, pad_cells=31),
}


@pytest.mark.parametrize("name", list(SCENES))
def test_closest_and_any_hit_match_reference(name, oracle, gpu_ctx_factory):
    scene = SCENES[name]()
    ctx = gpu_ctx_factory(scene)
    sess = oracle.Session(scene)
    ties = 0
    for label, rays in ray_sets(scene, 40000, 11).items():
        ref, _ = sess.trace(rays, "closest")
        gpu = ctx.trace(rays, TRACE_CLOSEST)
        ties += check_closest(gpu, ref, f"{name}/{label}")
        seg = scenes.shadow_segments(rays, scene)
        ref_any, _ = sess.trace(seg, "any")
        gpu_any = ctx.trace(seg, TRACE_ANY)
        assert np.array_equal(ref_any, gpu_any), f"{name}/{label}: any-hit differs for {np.count_nonzero(ref_any != gpu_any)} rays"
        # same rays as un-normalised any-hit queries with the original interval
        ref_any2, _ = sess.trace(rays, "any")
        assert np.array_equal(ref_any2, ctx.trace(rays, TRACE_ANY)), f"{name}/{label}: any-hit (original interval) differs"
    print(f"{name}: exact-t ties {ties}")


def test_triaccel_records_bit_exact(oracle, gpu_ctx_factory):
    """TriAccel::load (triaccel.h:63-96) and the enlarged scene box (gkdtree.h:1172-1176)."""
    scene = scenes.material_test_scene(16, 16, 1)
    # add a degenerate triangle (k = 3, never hit)
    scene.shapes[0].triangles = np.vstack([scene.shapes[0].triangles, np.array([[0, 0, 1]], np.uint32)])
    ctx = gpu_ctx_factory(scene)
    ref, box_ref, _ = oracle.Session(scene).triaccel()
    rec, box = ctx.triaccel(len(ref))
    assert np.array_equal(box.view(np.uint32), box_ref.view(np.uint32))
    assert np.array_equal(rec["k"] & 3, ref["k"] & 3)
    ok = (ref["k"] & 3) != 3
    assert (~ok).sum() >= 1
    for f in ("n_u", "n_v", "n_d", "a_u", "a_v", "b_nu", "b_nv", "c_nu", "c_nv"):
        assert np.array_equal(rec[f][ok].view(np.uint32), ref[f][ok].view(np.uint32)), f
    assert np.array_equal(rec["shapeIndex"], ref["shapeIndex"]) and np.array_equal(rec["primIndex"], ref["primIndex"])


def test_edge_cases(oracle, gpu_ctx_factory):
    """Single triangle, empty ray batch, BSDF-less shapes do not occlude (shape.h:361)."""
    s = scenefile.SceneDesc()
    s.materials = [scenefile.lambertian()]
    tri = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    s.shapes = [scenefile.Shape(tri, np.array([[0, 1, 2]], np.uint32), None, material=-1)]
    ctx = gpu_ctx_factory(s)
    assert len(ctx.trace(np.zeros((0, 8), np.float32))) == 0
    rays = np.array([[0.2, 0.2, 1, 1e-4, 0, 0, -1, np.inf], [2, 2, 1, 1e-4, 0, 0, -1, np.inf],
                     [0.2, 0.2, 1, 1e-4, 0, 0, -1, 0.5]], np.float32)
    sess = oracle.Session(s)
    ref, _ = sess.trace(rays, "closest")
    gpu = ctx.trace(rays)
    check_closest(gpu, ref, "single")
    assert gpu["shape"][0] == 0 and gpu["shape"][1] == 0xFFFFFFFF
    assert np.array_equal(ctx.trace(rays, TRACE_ANY), sess.trace(rays, "any")[0])
    assert ctx.trace(rays, TRACE_ANY)[0] == 0      # no BSDF => not an occluder


def test_closest_any_consistency_full_size(gpu_ctx_factory):
    """Size-independent properties at BASELINE config C2 scale (1M triangles, 4M+4M rays here; the
    bench runs 16M+16M): determinism, closest/any consistency, permutation invariance."""
    scene = scenes.microbench_scene()
    assert scene.triangle_count == 1_000_000
    ctx = gpu_ctx_factory(scene)
    for rays in (scenes.incoherent_rays(scene, 1 << 22, 2), scenes.coherent_rays(scene, 2048)):
        h1 = ctx.trace(rays)
        h2 = ctx.trace(rays)
        assert np.array_equal(h1.view(np.uint8), h2.view(np.uint8))
        hit = h1["shape"] != 0xFFFFFFFF
        assert 0.05 < hit.mean() < 1.0
        # a ray cut just beyond its closest hit must be occluded; cut just before it must not
        seg = rays.copy(); seg[:, 7] = np.where(hit, h1["t"] * np.float32(1 + 1e-5), rays[:, 7])
        occ = ctx.trace(seg, TRACE_ANY)
        assert np.array_equal(occ.astype(bool), hit)
        seg[:, 7] = np.where(hit, h1["t"] * np.float32(1 - 1e-5), rays[:, 7])
        assert not ctx.trace(seg, TRACE_ANY)[hit].any()
        perm = np.random.Generator(np.random.PCG64(5)).permutation(len(rays))
        assert np.array_equal(ctx.trace(rays[perm]).view(np.uint8), h1[perm].view(np.uint8))


def test_full_size_matches_reference(oracle, gpu_ctx_factory):
    """C2 scene at full size (1M triangles) against the reference kd-tree on 1M-ray subsets."""
    import os
    scene = scenes.microbench_scene()
    ctx = gpu_ctx_factory(scene)
    sess = oracle.Session(scene)
    threads = os.cpu_count() or 1
    for label, rays in (("incoherent", scenes.incoherent_rays(scene, 1 << 20, 2)), ("coherent", scenes.coherent_rays(scene, 1024))):
        ref, info = sess.trace(rays, "closest", threads)
        ties = check_closest(ctx.trace(rays), ref, label)
        seg = scenes.shadow_segments(rays, scene)
        assert np.array_equal(sess.trace(seg, "any", threads)[0], ctx.trace(seg, TRACE_ANY))
        print(label, "ties", ties, info)


@pytest.mark.gpu
def test_ploc_builder_gives_the_same_hits(monkeypatch):
    """MTSGPU_BUILDER=ploc (agglomerative build, csrc/bvh_build.cu) must not change any answer: same golden vectors."""
    import os
    from mitsuba_renderer_b200 import api
    from tests import golden_cases as gc
    monkeypatch.setenv("MTSGPU_BUILDER", "ploc")
    for name, make in (("trace_materials", gc.scene_materials), ("trace_grid", gc.scene_grid)):
        scene = make()
        g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
        rays = gc.trace_rays(scene)
        ctx = api.Context(0)
        ctx.upload_scene(scene)
        check_closest(ctx.trace(rays, TRACE_CLOSEST), g["hits"].view(api.HIT_DTYPE).reshape(-1), name + " (ploc)")
        assert np.array_equal(ctx.trace(scenes.shadow_segments(rays, scene), TRACE_ANY), g["any_segments"])
        ctx.close()


def test_traversal_stack_overflow_is_reported(monkeypatch):
    """A ray that needs more stack entries than there are must make the call fail (MTSGPU_ERR_LIMIT), not lose a subtree
    silently. The real limit is 96 entries against at most 3 per level of the tree (accel_info: depth); the debug switch
    shrinks it so that an ordinary scene overflows."""
    from mitsuba_renderer_b200 import api
    scene = scenes.microbench_scene(cells=120, pad_cells=9)
    rays = scenes.incoherent_rays(scene, 20000, 3)
    monkeypatch.setenv("MTSGPU_DEBUG_STACK_ENTRIES", "2")
    ctx = api.Context(0)
    ctx.upload_scene(scene)
    with pytest.raises(api.MtsGpuError, match="stack overflow"):
        ctx.trace(rays)
    with pytest.raises(api.MtsGpuError, match="stack overflow"):
        ctx.trace(scenes.shadow_segments(rays, scene), TRACE_ANY)
    ctx.close()
    monkeypatch.delenv("MTSGPU_DEBUG_STACK_ENTRIES")
    ctx = api.Context(0)
    ctx.upload_scene(scene)
    info = ctx.accel_info()
    assert 3 * info["depth"] <= 96
    assert (ctx.trace(rays)["shape"] != 0xFFFFFFFF).any()
    ctx.close()


def test_rays_from_far_outside_the_scene(oracle, gpu_ctx_factory):
    """Origins hundreds of scene extents away (a distant camera): the box test's rounding grows with |o / d|, and the
    traversal widens its slabs accordingly (trace.cuh: nodeTest). Every hit the kd-tree reports must be reported identically;
    where the two differ at all the GPU's hit is the nearer one (the kd-tree's own split-plane arithmetic loses a few
    hits at that distance)."""
    scene = scenes.microbench_scene(cells=200, pad_cells=11)
    ctx = gpu_ctx_factory(scene)
    sess = oracle.Session(scene)
    lo, hi = scenes.scene_bounds(scene)
    base = scenes.incoherent_rays(scene, 200000, 9)
    for factor in (50.0, 400.0):
        rays = base.copy()
        rays[:, 0:3] -= rays[:, 4:7] * np.float32(factor * np.linalg.norm(hi - lo))
        rays[:, 3] = 0.0
        ref, _ = sess.trace(rays, "closest")
        gpu = ctx.trace(rays)
        rhit, ghit = ref["shape"] != 0xFFFFFFFF, gpu["shape"] != 0xFFFFFFFF
        same = (gpu["shape"] == ref["shape"]) & (gpu["prim"] == ref["prim"])
        assert np.array_equal(gpu["t"][same & rhit].view(np.uint32), ref["t"][same & rhit].view(np.uint32))
        differ = ~same
        assert differ.mean() < 2e-3, differ.mean()
        assert not (rhit & ~ghit).any(), "a hit of the reference is missing"
        assert (gpu["t"][differ & ghit] <= ref["t"][differ & ghit]).all()
        print(f"origins {factor} extents away: {rhit.sum()} reference hits, {int(differ.sum())} differ (GPU nearer or reference missed)")
