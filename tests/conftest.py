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
def _same_seed_for_every_test():
    torch.manual_seed(0)


class _DeviceSleeper:
    """``sleeper(seconds)`` enqueues a spin kernel of about that length on the current stream.

    ``torch.cuda._sleep`` counts clock cycles, so the rate is calibrated once per session: the
    median of three timed probes (the first one also absorbs context creation)."""

    PROBE_CYCLES = 2_000_000

    def __init__(self):
        torch.cuda.init()
        rates = []
        for _ in range(3):
            before = torch.cuda.Event(enable_timing=True)
            after = torch.cuda.Event(enable_timing=True)
            before.record()
            torch.cuda._sleep(self.PROBE_CYCLES)
            after.record()
            after.synchronize()
            rates.append(self.PROBE_CYCLES / before.elapsed_time(after))    # cycles per ms
        self.cycles_per_second = 1e3 * sorted(rates)[1]

    def __call__(self, seconds):
        torch.cuda._sleep(int(seconds * self.cycles_per_second))


@pytest.fixture(scope='session')
def cuda_sleep():
    return _DeviceSleeper()
