"""`.serialized` mesh files: the library's native reader / writer against the reference's
loader (golden vectors from tests/golden/make_golden.py, the reference's own fixture file when it is
present, and the compiled reference itself when oracle/_ref is built). Host code only — no GPU needed,
except for the last test, which traces rays through a scene uploaded straight from a file."""
import os

import numpy as np
import pytest

from tests import golden_cases as gc
from tests import oracle_util
from mitsuba_renderer_b200 import api, scenes, serialized

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "serialized.npz")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def two_mesh_file(tmp_path_factory):
    path = str(tmp_path_factory.mktemp("ser") / "two.serialized")
    serialized.save(path, gc.serialized_meshes())
    return path


def test_writer_reader_round_trip(two_mesh_file):
    assert serialized.count(two_mesh_file) == 2
    src = gc.serialized_meshes()
    for i, m in enumerate(src):
        got = serialized.load(two_mesh_file, i, face_normals=True)     # face normals: stored normals are dropped, nothing computed
        assert np.array_equal(bits(got["positions"]), bits(m["positions"]))
        assert np.array_equal(got["triangles"], m["triangles"])
        assert got["normals"] is None
        assert got["flags"] & 0x1000                                   # single precision
    got = serialized.load(two_mesh_file, 1)
    assert np.array_equal(bits(got["normals"]), bits(src[1]["normals"]))
    assert np.array_equal(bits(got["texcoords"]), bits(src[1]["texcoords"]))


def test_normals_windings_and_transform_match_reference_golden(two_mesh_file):
    g = np.load(GOLD)
    for idx in (0, 1):
        for fn, fl in ((0, 0), (0, 1), (1, 0), (1, 1)):
            got = serialized.load(two_mesh_file, idx, face_normals=bool(fn), flip_normals=bool(fl))
            ref_n, ref_t = g[f"mesh{idx}_{fn}{fl}_normals"], g[f"mesh{idx}_{fn}{fl}_triangles"]
            assert np.array_equal(got["triangles"], ref_t), (idx, fn, fl)
            if ref_n.shape[0] == 0:
                assert got["normals"] is None
            else:                                                      # computeNormals: bit-exact, same accumulation order
                assert np.array_equal(bits(got["normals"]), bits(ref_n)), (idx, fn, fl)
    got = serialized.load(two_mesh_file, 1, to_world=gc.SERIALIZED_TO_WORLD)
    assert np.array_equal(bits(got["positions"]), bits(g["mesh1_xf_positions"]))
    assert np.array_equal(got["triangles"], g["mesh1_xf_triangles"])              # det < 0 swaps the winding
    assert np.allclose(got["normals"], g["mesh1_xf_normals"], rtol=0, atol=1e-6)   # inverse computed here in fp64, there in fp32


@pytest.mark.skipif(not os.path.exists(gc.SERIALIZED_FIXTURE), reason="the reference's fixture file is not on this machine")
def test_reference_fixture_file_matches_golden_digests():
    g = np.load(GOLD)
    assert serialized.count(gc.SERIALIZED_FIXTURE) == 3
    for i in range(3):
        m = serialized.load(gc.SERIALIZED_FIXTURE, i)
        assert [len(m["positions"]), len(m["triangles"])] == list(g[f"fixture{i}_counts"])
        assert gc.mesh_digest(m) == bytes(g[f"fixture{i}"]).decode()


@pytest.mark.skipif(not oracle_util.available(), reason="oracle/_ref not built")
def test_files_written_by_the_reference_are_read_identically(tmp_path):
    scene = scenes.material_test_scene(16, 16, 1)
    path = str(tmp_path / "ref.serialized")
    oracle_util.save_serialized(scene, path)                           # TriMesh::serialize + offset table
    assert serialized.count(path) == len(scene.shapes)
    for i, sh in enumerate(scene.shapes):
        a = serialized.load(path, i, face_normals=sh.normals is None)
        b = oracle_util.load_serialized(path, i, face_normals=sh.normals is None)
        assert np.array_equal(bits(a["positions"]), bits(sh.positions)) and np.array_equal(a["triangles"], sh.triangles)
        assert np.array_equal(bits(a["positions"]), bits(b["positions"])) and np.array_equal(a["triangles"], b["triangles"])
        if sh.normals is None:
            assert a["normals"] is None and b["normals"] is None
        else:
            assert np.array_equal(bits(a["normals"]), bits(b["normals"]))


def test_error_behaviour(tmp_path, two_mesh_file):
    with pytest.raises(api.MtsGpuError, match="out of range"):         # trimesh.cpp:165-169
        serialized.load(two_mesh_file, 5)
    bad = tmp_path / "bad.serialized"
    bad.write_bytes(b"\x00\x01\x02\x03" + b"\x00" * 64)
    with pytest.raises(api.MtsGpuError, match="invalid file format"):  # trimesh.cpp:185-186
        serialized.load(str(bad), 0)
    old = tmp_path / "old.serialized"
    old.write_bytes(b"\x04\x1c\x03\x00" + b"\x00" * 64)
    with pytest.raises(api.MtsGpuError, match="old version"):          # trimesh.cpp:180-183
        serialized.load(str(old), 0)
    data = open(two_mesh_file, "rb").read()
    cut = tmp_path / "cut.serialized"
    cut.write_bytes(data[:60])
    with pytest.raises(api.MtsGpuError):
        serialized.load(str(cut), 0)
    with pytest.raises(api.MtsGpuError, match="cannot open"):
        serialized.load(str(tmp_path / "missing.serialized"), 0)
    with pytest.raises(api.MtsGpuError):
        serialized.save(str(tmp_path / "empty.serialized"), [{"positions": np.zeros((0, 3), np.float32), "triangles": np.zeros((0, 3), np.uint32)}])


@pytest.mark.gpu
def test_scene_uploaded_from_file_traces_like_the_inline_scene(tmp_path):
    from tests.test_trace_parity import check_closest
    scene = gc.scene_materials()
    path = str(tmp_path / "scene.serialized")
    serialized.save(path, scene.shapes)
    rays = gc.trace_rays(scene)
    inline = api.Context(0)
    inline.upload_scene(scene)
    want = inline.trace(rays, api.TRACE_CLOSEST)
    want_full = inline.intersect_full(rays)[1]
    inline.close()
    ctx = api.Context(0)
    ctx.upload_scene(scene, build=False)
    for i, sh in enumerate(scene.shapes):                              # replace every inline mesh by its file version
        serialized.upload(ctx, i, path, i, face_normals=sh.normals is None, material=sh.material, emitter=sh.emitter)
    ctx.build_accel()
    got = ctx.trace(rays, api.TRACE_CLOSEST)
    got_full = ctx.intersect_full(rays)[1]
    ctx.close()
    assert np.array_equal(got.view(np.uint8), want.view(np.uint8))
    assert np.array_equal(bits(got_full), bits(want_full))
    g = np.load(os.path.join(os.path.dirname(GOLD), "trace_materials.npz"))
    check_closest(got, g["hits"].view(api.HIT_DTYPE).reshape(-1), "serialized upload")   # and like the reference's kd-tree
