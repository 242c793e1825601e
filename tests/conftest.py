import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')
    config.addinivalue_line('markers', 'multigpu: needs at least 2 CUDA devices')


def pytest_collection_modifyitems(config, items):
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    skip_gpu = pytest.mark.skip(reason='no CUDA device')
    skip_multi = pytest.mark.skip(reason='needs >= 2 CUDA devices')
    for item in items:
        if 'gpu' in item.keywords and n == 0:
            item.add_marker(skip_gpu)
        if 'multigpu' in item.keywords and n < 2:
            item.add_marker(skip_multi)


@pytest.fixture(autouse=True)
def manual_seed_zero():
    torch.manual_seed(0)


@pytest.fixture(scope='session')
def cuda_sleep():
    """``cuda_sleep(seconds)`` enqueues a busy-wait kernel of roughly that length."""
    torch.empty(1, device='cuda')
    start = torch.cuda.Event(enable_timing=True)
    end = torch.cuda.Event(enable_timing=True)
    start.record()
    torch.cuda._sleep(1000000)
    end.record()
    end.synchronize()
    cycles_per_ms = 1000000 / start.elapsed_time(end)

    def sleep(seconds):
        torch.cuda._sleep(int(seconds * cycles_per_ms * 1000))

    return sleep
