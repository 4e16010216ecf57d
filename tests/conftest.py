import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (HERE, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def tinyr():
    """The product package with its CUDA library built; fails loudly if it cannot be built/loaded."""
    import tiny_renderer_b200 as tr
    if not os.path.exists(tr.LIB_PATH):
        tr.build()
    tr.lib()
    return tr


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


_MODELS = {}


def get_model(name):
    """Reference asset (staged copy) when present, else the procedural stand-in."""
    import assets
    if name not in _MODELS:
        _MODELS[name] = assets.get_model(name)
    return _MODELS[name]
