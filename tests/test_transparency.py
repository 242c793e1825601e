import pytest
import torch
from torch import nn

from torchgpipe_b200 import GPipe


def _grads(model):
    return torch.cat([p.grad.flatten() for p in model.parameters()])


@pytest.mark.parametrize('checkpoint', ['always', 'except_last', 'never'])
@pytest.mark.parametrize('balance', [[2, 2], [1, 3], [4], [1, 1, 1, 1]])
def test_gradients_match_unpipelined_model(balance, checkpoint):
    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(3, 5), nn.Tanh(), nn.Linear(5, 5), nn.Linear(5, 1))
    x = torch.rand(8, 3)

    model(x).sum().backward()
    expected = _grads(model).clone()
    model.zero_grad()

    pipe = GPipe(model, balance, devices=['cpu'] * len(balance), chunks=4, checkpoint=checkpoint)
    out = pipe(x)
    out.sum().backward()
    assert torch.allclose(_grads(pipe), expected, atol=1e-6)


def test_baseline_config_toy_pipeline_on_cpu():
    # BASELINE.json config 1: 4x(Linear+ReLU), balance [2,2], chunks 4, CPU devices.
    torch.manual_seed(0)
    layers = []
    for _ in range(4):
        layers += [nn.Linear(16, 16), nn.ReLU()]
    model = nn.Sequential(*layers)
    x = torch.rand(32, 16)
    ref = model(x)
    pipe = GPipe(model, [4, 4], devices=['cpu', 'cpu'], chunks=4)
    out = pipe(x)
    assert torch.allclose(out, ref, atol=1e-6)
    out.mean().backward()
