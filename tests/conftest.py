import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def oracle():
    """The compiled reference (oracle/_ref/parity/mts_oracle). Test infrastructure only."""
    from tests import oracle_util
    if not oracle_util.available():
        pytest.skip("oracle/_ref is not built (needs /root/reference); golden fixtures still run")
    return oracle_util


@pytest.fixture(scope="session")
def gpu_ctx_factory():
    """Creates contexts through the C ABI; fails loudly (no fallback) if the library is missing."""
    from mitsuba_renderer_b200 import api
    made = []

    def make(scene=None):
        ctx = api.Context(0)
        if scene is not None:
            ctx.upload_scene(scene)
        made.append(ctx)
        return ctx
    yield make
    for c in made:
        c.close()
