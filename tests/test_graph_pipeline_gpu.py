"""``GPipe(..., cuda_graphs=True)``: the single-process API on the CUDA-graph stage engine."""
import copy

import pytest
import torch
from torch import nn
import torch.nn.functional as F


def rel_err(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def test_graphs_refuse_deferred_batch_norm_and_cpu_devices():
    from torchgpipe_b200 import GPipe
    model = nn.Sequential(nn.Conv2d(3, 4, 1), nn.BatchNorm2d(4))
    with pytest.raises(ValueError, match='deferred_batch_norm'):
        GPipe(model, [2], devices=['cpu'], chunks=2, deferred_batch_norm=True, cuda_graphs=True)
    with pytest.raises(ValueError, match='CUDA device'):
        GPipe(model, [2], devices=['cpu'], chunks=2, cuda_graphs=True)


def _no_dropout(model):
    for m in model.modules():
        if isinstance(getattr(m, 'p', None), float):
            m.p = 0.0
    return model


def _unet():
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models import unet
    torch.manual_seed(3)
    return _no_dropout(ops.convert_to_sm100(unet(depth=2, num_convs=2, base_channels=32)))


def _train_steps(pipe, x, steps=2):
    opt = torch.optim.SGD(pipe.parameters(), lr=0.05)
    losses, grads = [], None
    for _ in range(steps):
        out = pipe(x)
        loss = F.binary_cross_entropy_with_logits(out, torch.ones_like(out))
        loss.backward()
        losses.append(loss.detach())
        grads = [p.grad.clone() for p in pipe.parameters()]
        opt.step()
        opt.zero_grad()
    return torch.stack(losses), grads, out.detach()


@pytest.mark.gpu
@pytest.mark.parametrize('stages', [1, 2, 3])
def test_graph_mode_matches_eager_gpipe(stages):
    """Same API, same numbers: loss trajectory and gradients of GPipe(cuda_graphs=True) against
    the eager engine - with skip connections crossing the stage boundaries, two micro-batch
    shapes (5 samples in 2 chunks) and several stages as threads on one device."""
    from torchgpipe_b200 import GPipe
    model = _unet()
    n = len(model)
    balance = {1: [n], 2: [n // 2, n - n // 2], 3: [n // 3, n // 3, n - 2 * (n // 3)]}[stages]
    devices = ['cuda:0'] * stages
    x = torch.rand(5, 3, 32, 32, device='cuda:0')
    eager = GPipe(copy.deepcopy(model), balance, devices=devices, chunks=2,
                  checkpoint='except_last')
    graphed = GPipe(copy.deepcopy(model), balance, devices=devices, chunks=2,
                    checkpoint='except_last', cuda_graphs=True)
    le, ge, oe = _train_steps(eager, x)
    lg, gg, og = _train_steps(graphed, x)
    assert graphed._graph_engine.native_launches_replayed() > 0
    assert rel_err(lg, le) < 1e-2
    assert rel_err(og, oe) < 5e-2
    for a, b in zip(gg, ge):
        assert F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item() > 0.97


@pytest.mark.gpu
def test_graph_mode_eval_after_train_uses_eval_kernels():
    """A training step and an evaluation step of the same shape must not share a captured graph
    (Dropout off, BatchNorm on running statistics, running statistics untouched)."""
    from torchgpipe_b200 import GPipe, ops
    torch.manual_seed(4)
    model = ops.convert_to_sm100(nn.Sequential(
        nn.Conv2d(8, 16, 3, padding=1, bias=False), nn.BatchNorm2d(16), nn.ReLU(),
        nn.Dropout2d(0.5), nn.Conv2d(16, 8, 1, bias=False)))
    x = torch.randn(4, 8, 12, 12, device='cuda:0')
    eager = GPipe(copy.deepcopy(model), [5], devices=['cuda:0'], chunks=2)
    graphed = GPipe(copy.deepcopy(model), [5], devices=['cuda:0'], chunks=2, cuda_graphs=True)
    for pipe in (eager, graphed):
        pipe.train()
        pipe(x).square().mean().backward()          # updates running statistics once
        pipe.eval()
    bn_g, bn_e = graphed.partitions[0][1], eager.partitions[0][1]
    before = bn_g.running_mean.clone()
    with torch.no_grad():
        out_g, out_e = graphed(x), eager(x)
        out_g2 = graphed(x)
    assert torch.equal(bn_g.running_mean, before)                   # eval left them alone
    assert torch.allclose(bn_g.running_mean, bn_e.running_mean, atol=1e-4)
    assert rel_err(out_g, out_e) < 5e-3                             # no dropout, running stats
    assert torch.equal(out_g, out_g2)                               # deterministic


class _Boom(nn.Module):
    def forward(self, x):
        raise RuntimeError('boom in stage 1')


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_graph_mode_failure_in_one_stage_ends_the_step_everywhere():
    from torchgpipe_b200 import GPipe
    model = nn.Sequential(nn.Conv2d(4, 4, 1), _Boom(), nn.Conv2d(4, 4, 1))
    pipe = GPipe(model, [1, 1, 1], devices=['cuda:0'] * 3, chunks=3, cuda_graphs=True)
    with pytest.raises(RuntimeError, match='boom in stage 1'):
        pipe(torch.randn(6, 4, 5, 5, device='cuda:0'))
    # the engine is still usable for a model that works
    good = GPipe(nn.Sequential(nn.Conv2d(4, 4, 1), nn.ReLU()), [1, 1], devices=['cuda:0'] * 2,
                 chunks=2, cuda_graphs=True)
    out = good(torch.randn(4, 4, 5, 5, device='cuda:0'))
    out.mean().backward()
    assert all(p.grad is not None for p in good.parameters())


@pytest.mark.gpu
def test_graph_mode_refuses_input_gradients():
    from torchgpipe_b200 import GPipe
    pipe = GPipe(nn.Sequential(nn.Conv2d(4, 4, 1)), [1], devices=['cuda:0'], chunks=1,
                 cuda_graphs=True)
    x = torch.randn(2, 4, 5, 5, device='cuda:0', requires_grad=True)
    with pytest.raises(NotImplementedError, match='pipeline input'):
        pipe(x)


@pytest.mark.gpu
@pytest.mark.multigpu
def test_graph_mode_on_two_devices():
    from torchgpipe_b200 import GPipe
    model = _unet()
    n = len(model)
    x = torch.rand(6, 3, 32, 32, device='cuda:0')
    eager = GPipe(copy.deepcopy(model), [n // 2, n - n // 2], devices=['cuda:0', 'cuda:1'],
                  chunks=3)
    graphed = GPipe(copy.deepcopy(model), [n // 2, n - n // 2], devices=['cuda:0', 'cuda:1'],
                    chunks=3, cuda_graphs=True)
    le, ge, _ = _train_steps(eager, x)
    lg, gg, og = _train_steps(graphed, x)
    assert og.device == torch.device('cuda', 1)
    assert rel_err(lg, le) < 1e-2
    for a, b in zip(gg, ge):
        assert F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item() > 0.97
