"""Two GPUs in ONE process (the `mitsuba` CLI case): mtsgpu_render_multi splits the sample range over the
contexts and sums the films on GPU 0 with a kernel that reads GPU 1's accumulators over NVLink peer memory.
Needs 2 GPUs (`gpurun --gpus 2`); skipped otherwise."""
import os

import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile

pytestmark = pytest.mark.gpu


def two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")


def test_render_multi_equals_single_gpu():
    two_gpus()
    s = scenes.material_test_scene(64, 64, 32, scenefile.MODE_PATH, max_depth=8)
    s.filter_type = scenefile.FILTER_GAUSSIAN
    ctxs = [api.Context(0), api.Context(1)]
    try:
        for c in ctxs:
            c.upload_scene(s)
        p = api.Context.params_from_scene(s, seed=5)
        single, st1 = ctxs[0].render(p)
        multi, st2 = api.render_multi(ctxs, p)
        # same sample set (RNG keyed by global sample index), different fp32 summation order
        assert np.allclose(multi, single, rtol=2e-5, atol=2e-5)
        assert st2["samples"] == st1["samples"] and st2["normalRays"] == st1["normalRays"] and st2["shadowRays"] == st1["shadowRays"]
        odd = api.Context.params_from_scene(s, 3, 10, seed=5)       # an uneven range
        a, _ = ctxs[0].render(odd)
        b, _ = api.render_multi(ctxs, odd)
        assert np.allclose(a, b, rtol=2e-5, atol=2e-5)
    finally:
        for c in ctxs:
            c.close()


def test_render_multi_reports_progress_on_the_calling_thread():
    """The callback contract of mtsgpu_set_progress holds for the multi-GPU call too: Mitsuba's ProgressReporter and RenderQueue
    (renderproc.cpp:123-130) must be driven from a registered thread, never from the per-GPU worker threads."""
    import threading
    two_gpus()
    s = scenes.material_test_scene(128, 128, 64, scenefile.MODE_PATH, max_depth=8)
    ctxs = [api.Context(0), api.Context(1)]
    try:
        for c in ctxs:
            c.upload_scene(s)
        calls = []
        ctxs[0].set_progress(lambda fraction, film: calls.append((fraction, film, threading.get_ident())), 0.0, True)
        p = api.Context.params_from_scene(s, seed=5)
        multi, _ = api.render_multi(ctxs, p)
        ctxs[0].set_progress(None)
        assert calls and calls[-1][0] == 1.0 and all(f is None for _, f, _ in calls)
        assert {t for _, _, t in calls} == {threading.get_ident()}
        assert [f for f, _, _ in calls] == sorted(f for f, _, _ in calls)
        single, _ = ctxs[0].render(p)
        assert np.allclose(multi, single, rtol=2e-5, atol=2e-5)
    finally:
        for c in ctxs:
            c.close()


def test_plugin_with_two_gpus(oracle):
    two_gpus()
    plugin_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mitsuba-renderer_b200")
    if not os.path.exists(os.path.join(plugin_dir, "plugins", "gpupath.so")):
        pytest.skip("plugin not built")
    s = scenes.cornell_box(64, 64, 128)
    sess = oracle.Session(s)
    one, _ = sess.render(2, "gpupath", env={"MTS_PLUGIN_PATH": plugin_dir})
    two, _ = sess.render(2, "gpupath", env={"MTS_PLUGIN_PATH": plugin_dir, "MTS_ORACLE_GPUS": "2"})
    assert one[..., 4].sum() > 0 and np.allclose(one, two, rtol=2e-5, atol=2e-5)
