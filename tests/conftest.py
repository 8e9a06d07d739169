import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no GPU in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import tmoracle
    return tmoracle


@pytest.fixture(scope="session")
def small_case(oracle):
    """seeded synthetic case: 1 palette x 16 colours, 512 tiles (N = 2048), 1500 queries"""
    o = oracle
    pal, s = o.gen_palette(o.BASE_SEED)
    tiles, s = o.gen_tileset(s, 512)
    rgb, s = o.gen_queries(s, 1500, tiles, pal)
    return dict(pal=pal, tiles=tiles, rgb=rgb)
