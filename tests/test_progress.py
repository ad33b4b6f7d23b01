"""Progress and preview while a render runs (mtsgpu_set_progress; in the plugin: ProgressReporter + RenderQueue::signalRefresh,
what BlockedRenderProcess::processResult does after every block, src/librender/renderproc.cpp:123-130)."""
import os

import numpy as np
import pytest

from mitsuba_renderer_b200 import api, scenes, scenefile
from mitsuba_renderer_b200.api import Context

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN_DIR = os.path.join(ROOT, "mitsuba-renderer_b200")


def test_progress_callback_sees_the_film_grow(monkeypatch):
    """Small wavefront batches (one tile each) so that a 64 x 64 render has 16 of them: the callback must report rising
    fractions, films whose weight grows and never exceeds the final one, end with (1, None), and leave the result untouched."""
    monkeypatch.setenv("MTSGPU_BATCH_LOG2", "16")
    s = scenes.cornell_box(64, 64, 256, max_depth=5)
    ctx = api.Context(0)
    try:
        ctx.upload_scene(s)
        p = Context.params_from_scene(s, seed=3)
        plain, _ = ctx.render(p)
        calls = []
        ctx.set_progress(lambda f, film: calls.append((f, None if film is None else film.copy())), 0.0, True)
        watched, _ = ctx.render(p)
        ctx.set_progress(None)
        assert np.array_equal(watched, plain)
        assert len(calls) >= 4 and calls[-1][0] == 1.0 and calls[-1][1] is None
        fr = [c[0] for c in calls]
        assert all(a < b for a, b in zip(fr, fr[1:])) and fr[0] > 0
        weights = [float(c[1][..., 4].sum()) for c in calls[:-1]]
        assert all(a <= b for a, b in zip(weights, weights[1:])) and 0 < weights[0] and weights[-1] < plain[..., 4].sum()
        for f, film in calls[:-1]:
            done = film[..., 4] > 0                       # tiles rendered so far carry their final values
            assert np.array_equal(film[done], plain[done])
        # progress only: no film is copied
        seen = []
        ctx.set_progress(lambda f, film: seen.append(film), 0.0, False)
        ctx.render(p)
        ctx.set_progress(None)
        assert len(seen) >= 4 and all(x is None for x in seen)
        # and without a callback nothing is called
        n = len(seen)
        ctx.render(p)
        assert len(seen) == n
    finally:
        ctx.close()


def test_plugin_signals_refresh_with_a_growing_film(oracle):
    """Through the real RenderQueue of the compiled reference: a listener (what the GUI and the `-r` flush thread are) receives
    refresh events while the plugin renders, sees a partly filled film, and the final film is what a render without previews gives."""
    if not os.path.exists(os.path.join(PLUGIN_DIR, "plugins", "gpupath.so")):
        pytest.skip("plugins/gpupath.so is not built (needs the reference headers)")
    s = scenes.cornell_box(64, 64, 256, max_depth=5)
    s.sampler_type = 3
    sess = oracle.Session(s)
    env = {"MTS_PLUGIN_PATH": PLUGIN_DIR, "MTSGPU_BATCH_LOG2": "16"}
    quiet, qi = sess.render(2, "gpupath", env=dict(env, MTS_ORACLE_REFRESH="1000"))
    shown, si = sess.render(2, "gpupath", env=dict(env, MTS_ORACLE_REFRESH="0"))
    assert qi["refresh_events"] == 0 and si["refresh_events"] >= 3
    assert 0 < si["film_weight_at_last_refresh"] < float(shown[..., 4].sum())
    assert np.array_equal(quiet, shown)
