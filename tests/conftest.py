import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def c1_frames():
    """Two consecutive frames of the C1-shaped synthetic sequence (1241x376)."""
    from svo_b200 import synth
    canvas = synth.make_canvas(1241, 376, 11)
    return synth.render_frame(canvas, 1241, 376, 0), synth.render_frame(canvas, 1241, 376, 1)


@pytest.fixture(scope="session")
def small_frames():
    """Two consecutive frames of a small synthetic sequence (320x240), cheap enough for the scalar oracle."""
    from svo_b200 import synth
    canvas = synth.make_canvas(320, 240, 5)
    return synth.render_frame(canvas, 320, 240, 0), synth.render_frame(canvas, 320, 240, 1)
