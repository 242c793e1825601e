import time

import pytest
import torch
from torch import nn

from torchgpipe_b200.balance import balance_by_size, balance_by_time, blockpartition
from torchgpipe_b200.balance.profile import layerwise_sandbox


def test_blockpartition():
    assert blockpartition.solve([1, 2, 3, 4, 5, 6], partitions=2) == [[1, 2, 3, 4], [5, 6]]
    assert blockpartition.solve([1, 1, 1, 1], partitions=2) == [[1, 1], [1, 1]]
    assert blockpartition.solve([5], partitions=1) == [[5]]
    out = blockpartition.solve([9, 1, 1, 1, 1, 1, 1, 9], partitions=4)
    assert sum(out, []) == [9, 1, 1, 1, 1, 1, 1, 9] and len(out) == 4
    assert all(len(b) > 0 for b in out)


def test_blockpartition_errors():
    with pytest.raises(ValueError, match='positive integer'):
        blockpartition.solve([42], partitions=0)
    with pytest.raises(ValueError, match='shorter than intended partitions'):
        blockpartition.solve([42], partitions=2)


def test_blockpartition_short_sequence():
    assert blockpartition.solve([1, 2], partitions=2) == [[1], [2]]


def test_balance_by_time_on_cpu():
    class Delay(nn.Module):
        def __init__(self, seconds):
            super().__init__()
            self.seconds = seconds

        def forward(self, x):
            time.sleep(self.seconds)
            return x

    model = nn.Sequential(*[Delay(i / 100) for i in [1, 2, 3, 4, 5, 6]])
    sample = torch.rand(1)
    balance = balance_by_time(2, model, sample, device='cpu', timeout=0.5)
    assert balance == [4, 2]


def test_balance_by_time_tuple_io():
    class Twin(nn.Module):
        def forward(self, x):
            return x, x.detach()

    class Add(nn.Module):
        def forward(self, ab):
            a, b = ab
            return a + b

    model = nn.Sequential(Twin(), Add())
    balance = balance_by_time(1, model, torch.rand(1, requires_grad=True), device='cpu',
                              timeout=0.05)
    assert balance == [2]


def test_sandbox_leaves_module_untouched():
    model = nn.Sequential(nn.BatchNorm1d(1))
    before = {k: v.clone() for k, v in model.state_dict().items()}
    model.eval()
    balance_by_time(1, model, torch.rand(8, 1), device='cpu', timeout=0.05)
    assert not model.training
    for k, v in model.state_dict().items():
        assert torch.equal(v, before[k])
    assert all(p.grad is None for p in model.parameters())
    clones = list(layerwise_sandbox(model, torch.device('cpu')))
    assert clones[0] is not model[0] and clones[0].training


def test_already_has_grad():
    model = nn.Sequential(nn.Conv2d(3, 3, 1))
    model(torch.rand(1, 3, 4, 4)).mean().backward()
    with pytest.raises(ValueError, match='some parameter already has gradient'):
        balance_by_time(1, model, torch.rand(1, 3, 4, 4), device='cpu', timeout=0.05)


def test_balance_by_size_needs_cuda():
    with pytest.raises(ValueError, match='only CUDA'):
        balance_by_size(1, nn.Sequential(nn.Linear(1, 1)), torch.rand(1, 1), device='cpu')


def test_deprecated_module_name():
    with pytest.raises(ImportError, match='balance'):
        import torchgpipe_balancing  # noqa: F401


@pytest.mark.gpu
def test_balance_by_time_cuda_events():
    model = nn.Sequential(*[nn.Linear(64, 64) for _ in range(6)])
    balance = balance_by_time(2, model, torch.rand(32, 64), device='cuda', timeout=0.2)
    assert sum(balance) == 6 and len(balance) == 2


@pytest.mark.gpu
def test_time_profiler_measures_device_time_on_graph_replays():
    """Layers much faster than their Python launch cost: eager event timing sees the host,
    graph-replay timing sees the device. A 16x wider layer must look clearly more expensive in
    graph mode; skip connections (stash / pop run twice per layer) must not break the capture."""
    from torchgpipe_b200 import ops
    from torchgpipe_b200.balance.profile import LayerLab, _fresh_leaves, _graphed_times
    from torchgpipe_b200.models import unet
    device = torch.device('cuda')
    model = ops.convert_to_sm100(unet(depth=2, num_convs=2, base_channels=32))
    lab = LayerLab(model, device)
    seed = _fresh_leaves(torch.rand(4, 3, 64, 64), device)
    times = _graphed_times(lab, seed, replays=3)
    assert times is not None and len(times) == len(model)       # no silent fallback
    assert all(len(t) == 3 and min(t) >= 0 for t in times)

    small, big = nn.Conv2d(16, 16, 3, padding=1), nn.Conv2d(256, 256, 3, padding=1)
    pair = nn.Sequential(small, nn.Conv2d(16, 256, 1), big)
    us = balance_profile_times(pair, torch.rand(8, 16, 64, 64), device)
    assert us[2] > 4 * us[0]
    del lab, times, seed
    import gc
    gc.collect()
    torch.cuda.empty_cache()


def balance_profile_times(model, sample, device):
    from torchgpipe_b200.balance.profile import profile_times
    return profile_times(model, sample, 0.1, device)


@pytest.mark.gpu
def test_balance_by_size_latent():
    class Expand(nn.Module):
        def __init__(self, times):
            super().__init__()
            self.times = times

        def forward(self, x):
            for _ in range(self.times):
                x = x + torch.rand_like(x, requires_grad=True)
            return x

    sample = torch.rand(10, 100, 100)
    model = nn.Sequential(*[Expand(i) for i in [1, 2, 3, 4, 5, 6]])
    assert balance_by_size(2, model, sample, device='cuda') == [4, 2]
    model = nn.Sequential(*[Expand(i) for i in [6, 5, 4, 3, 2, 1]])
    assert balance_by_size(2, model, sample, device='cuda') == [2, 4]


@pytest.mark.gpu
def test_balance_by_size_param_and_scale():
    class Tradeoff(nn.Module):
        def __init__(self, param_size, latent_size):
            super().__init__()
            self.fc = nn.Linear(param_size, param_size)
            self.latent_size = latent_size

        def forward(self, x):
            for _ in range(self.latent_size):
                x = x + torch.rand_like(x, requires_grad=True)
            return x

    model = nn.Sequential(*[nn.Linear(i + 1, i + 2) for i in range(6)])
    assert balance_by_size(2, model, torch.rand(7, 1), param_scale=100, device='cuda') == [4, 2]
    model = nn.Sequential(Tradeoff(param_size=1, latent_size=6),
                          Tradeoff(param_size=2, latent_size=5),
                          Tradeoff(param_size=3, latent_size=4),
                          Tradeoff(param_size=4, latent_size=3),
                          Tradeoff(param_size=5, latent_size=2),
                          Tradeoff(param_size=6, latent_size=1))
    sample = torch.rand(1, requires_grad=True)
    assert balance_by_size(2, model, sample, param_scale=0, device='cuda') == [2, 4]
    assert balance_by_size(2, model, sample, param_scale=100, device='cuda') == [4, 2]


def test_optimal_solver_minimises_the_heaviest_block():
    import itertools
    import random
    from torchgpipe_b200.balance import balance_cost, blockpartition
    random.seed(5)
    for _ in range(60):
        n = random.randint(1, 9)
        seq = [random.choice([1, 2, 3, 5, 8, 0.5]) for _ in range(n)]
        k = random.randint(1, n)
        blocks = blockpartition.solve_optimal(seq, k)
        assert len(blocks) == k and all(blocks) and sum(blocks, []) == seq
        # brute force over all cut positions
        best = min(max(sum(seq[a:z]) for a, z in zip((0,) + cuts, cuts + (n,)))
                   for cuts in itertools.combinations(range(1, n), k - 1))
        assert max(map(sum, blocks)) == pytest.approx(best)
    # the input on which the reference's heuristic never terminates
    hard = [8, 2, 8, 3, 2, 13, 2, 5, 2, 2, 3, 13]
    assert max(map(sum, blockpartition.solve_optimal(hard, 10))) == 13
    assert balance_cost([1, 1, 1, 10], 2, solver='optimal') == [3, 1]
    with pytest.raises(ValueError):
        balance_cost([1, 2], 1, solver='greedy')
