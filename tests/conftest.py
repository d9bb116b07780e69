import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, 'kiosk-autoscaler_b200')
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason='no CUDA device')
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope='session', autouse=True)
def _built():
    """Build the oracle and (if stale) the library before any test runs."""
    from oracle import build as obuild
    obuild.build()
    import importlib.util
    spec = importlib.util.spec_from_file_location('kas_build', os.path.join(PKG, 'build.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if os.path.exists(mod.NVCC):
        mod.build()          # no-op when the library is newer than its sources; a failing nvcc fails the session
    elif not os.path.exists(mod.OUT):
        raise RuntimeError('libkas.so is missing and there is no nvcc to build it')
    elif mod._stale():
        raise RuntimeError('libkas.so is older than its sources and there is no nvcc to rebuild it')
    yield
