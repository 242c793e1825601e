"""Pipeline x data-parallel hybrid on CPU (gloo, 4 processes = 2 stages x 2 replicas): gradients
after the per-stage all-reduce equal the mean over replicas of the un-pipelined model's."""
import copy
import os

import pytest
import torch
from torch import nn
import torch.multiprocessing as mp


def _model():
    torch.manual_seed(0)
    return nn.Sequential(nn.Linear(4, 8), nn.Tanh(), nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 2))


def _data(replica):
    return torch.rand(6, 4, generator=torch.Generator().manual_seed(10 + replica))


def _worker(rank, world, port, ret):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from torchgpipe_b200.parallel import HybridLayout, HybridPipeline
        layout = HybridLayout(stages=2, replicas=2)
        assert (layout.replica, layout.stage) == divmod(rank, 2)
        hybrid = HybridPipeline(_model(), [2, 3], 3, layout, torch.device('cpu'),
                                checkpoint='except_last', use_symmetric=False)
        x = _data(layout.replica) if layout.stage == 0 else None
        outs = hybrid.forward(x)
        if layout.stage == 1:
            hybrid.backward([o.pow(2).mean() / 3 for o in outs])
        else:
            hybrid.backward(None)
        ret[rank] = [p.grad.clone() for p in hybrid.model().parameters()]
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_hybrid_gradients_are_replica_means():
    world = 4
    port = 29800 + os.getpid() % 1000
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)

    per_replica = []
    for replica in range(2):
        model = _model()
        loss = sum(model(c).pow(2).mean() / 3 for c in _data(replica).chunk(3))
        loss.backward()
        per_replica.append([p.grad.clone() for p in model.parameters()])
    want = [(a + b) / 2 for a, b in zip(*per_replica)]
    for replica in range(2):
        got = ret[replica * 2] + ret[replica * 2 + 1]      # stage 0 params then stage 1 params
        for g, w in zip(got, want):
            assert torch.allclose(g, w, atol=1e-6)
