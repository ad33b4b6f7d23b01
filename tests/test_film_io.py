"""Film development and image writers: MFilm::develop text against the reference's own output
(golden, and the compiled reference when present), the OpenEXR writer against an independent reader and numpy's
float16 rounding, and the GPU development kernels against the restated arithmetic."""
import os

import numpy as np
import pytest

from tests import golden_cases as gc, oracle_util
from mitsuba_renderer_b200 import api, scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mfilm.npz")
F = np.float32


def develop_m_numpy(acc, spectra):
    """mfilm.cpp:186-210 in fp32, operation by operation."""
    w = acc[..., 4]
    with np.errstate(divide="ignore"):
        inv = np.where(w > 0, (F(1) / w).astype(F), F(1)).astype(F)
    if spectra:
        return (inv[..., None] * acc[..., :3]).astype(F)
    lum = ((acc[..., 0] * F(0.212671)).astype(F) + (acc[..., 1] * F(0.715160)).astype(F)).astype(F)
    lum = (lum + (acc[..., 2] * F(0.072169)).astype(F)).astype(F)
    return (inv * lum).astype(F)


def develop_exr_numpy(acc, has_alpha):
    """exrfilm.cpp:163-176"""
    w = acc[..., 4]
    with np.errstate(divide="ignore"):
        inv = np.where(w != 0, (F(1) / w).astype(F), F(1)).astype(F)
    out = np.empty(acc.shape[:2] + (4,), F)
    out[..., :3] = np.maximum(F(0), (acc[..., :3] * inv[..., None]).astype(F))
    out[..., 3] = (acc[..., 3] * inv).astype(F) if has_alpha else F(1)
    return out


def test_mfilm_text_matches_reference_golden(tmp_path):
    acc = gc.film_accumulators()
    g = np.load(GOLD)
    assert gc.digest(acc) == str(g["acc_digest"])
    for spectra, key in ((False, "luminance"), (True, "spectra")):
        path = str(tmp_path / f"{key}.m")
        api.write_mfilm(path, develop_m_numpy(acc, spectra))
        assert open(path, "rb").read() == bytes(g[key]), key        # byte for byte what MFilm::develop wrote


@pytest.mark.skipif(not oracle_util.available(), reason="oracle/_ref not built")
def test_mfilm_text_matches_compiled_reference(tmp_path):
    acc = gc.film_accumulators(17, 5, seed=77)
    for spectra in (False, True):
        path = str(tmp_path / "x.m")
        api.write_mfilm(path, develop_m_numpy(acc, spectra))
        assert open(path, "rb").read() == oracle_util.develop_mfilm(acc, spectra)


def test_exr_writer_round_trips_through_an_independent_reader(tmp_path):
    acc = gc.film_accumulators(37, 29, seed=5)       # 37 rows: two full ZIP blocks of 16 lines and a short one
    rgba = develop_exr_numpy(acc, True)
    rgba[5, 6] = (np.inf, 65520.0, 65519.9, 5.9e-8)  # overflow edge, and the smallest subnormal's neighbourhood
    rgba[5, 7] = (2.98e-8, 2.9802322e-8, 6.1e-5, 6.2e-5)
    path = str(tmp_path / "a.exr")
    api.write_exr(path, rgba)
    ch = api.read_exr(path)
    assert sorted(ch) == ["A", "B", "G", "R"]
    with np.errstate(over="ignore"):
        want = rgba.astype(np.float16)               # IEEE round-to-nearest-even, like Imf's half
    for k, name in enumerate("RGBA"):
        assert np.array_equal(ch[name].view(np.uint16), want[..., k].view(np.uint16)), name
    # a noise image does not compress: blocks are then stored raw, which the reader must see too
    noise = np.random.Generator(np.random.PCG64(1)).random((20, 16, 4)).astype(F)
    api.write_exr(path, noise)
    ch = api.read_exr(path)
    assert np.array_equal(ch["G"].view(np.uint16), noise[..., 1].astype(np.float16).view(np.uint16))
    with pytest.raises(api.MtsGpuError):
        api.write_exr(str(tmp_path / "no_such_dir" / "a.exr"), noise)


@pytest.mark.gpu
def test_gpu_development_and_film_save(tmp_path):
    scene = scenes.cornell_box(29, 37, 4)
    ctx = api.Context(0)
    ctx.upload_scene(scene)
    film, _ = ctx.render(api.Context.params_from_scene(scene, seed=3))
    film[0, 0] = 0                                   # an untouched pixel
    for has_alpha in (True, False):
        assert np.array_equal(ctx.film_develop(film, has_alpha).view(np.uint32), develop_exr_numpy(film, has_alpha).view(np.uint32))
    exr, m, ms = (str(tmp_path / n) for n in ("f.exr", "f.m", "fs.m"))
    ctx.film_save(film, exr, api.FILM_EXR, 1)
    ctx.film_save(film, m, api.FILM_M, 0)
    ctx.film_save(film, ms, api.FILM_M, 1)
    ch = api.read_exr(exr)
    want = develop_exr_numpy(film, True).astype(np.float16)
    for k, name in enumerate("RGBA"):
        assert np.array_equal(ch[name].view(np.uint16), want[..., k].view(np.uint16))
    for path, spectra in ((m, False), (ms, True)):
        ref = str(tmp_path / "ref.m")
        api.write_mfilm(ref, develop_m_numpy(film, spectra))
        assert open(path, "rb").read() == open(ref, "rb").read()
    # a film left on the device by mtsgpu_render_device
    import torch
    dfilm = torch.zeros(film.size, dtype=torch.float32, device="cuda")
    ctx.render_device(api.Context.params_from_scene(scene, seed=3), dfilm.data_ptr())
    dev = str(tmp_path / "dev.exr")
    ctx.film_save(None, dev, api.FILM_EXR, 1, device_ptr=dfilm.data_ptr())
    host = dfilm.cpu().numpy().reshape(film.shape)
    want = develop_exr_numpy(host, True).astype(np.float16)
    assert np.array_equal(api.read_exr(dev)["R"].view(np.uint16), want[..., 0].view(np.uint16))
    ctx.close()
