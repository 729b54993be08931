# -*- coding: utf-8 -*-
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


def pytest_collection_modifyitems(config, items):
    # keep test_zz_* files last regardless of how paths were given on the command line
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)).startswith('test_zz_'), ))


@pytest.fixture(scope='session')
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.skip('CUDA is not available on this machine')
    return torch.device('cuda', 0)


@pytest.fixture(autouse=True)
def _no_pipeline_timeouts(request):
    """After every GPU test: no in-kernel wait may have given up (the kernels never hang, they record it)."""
    yield
    if request.node.get_closest_marker('gpu') is None:
        return
    import torch
    if not torch.cuda.is_available():
        return
    from distributed_dot_product.parallel import check_status
    check_status()
