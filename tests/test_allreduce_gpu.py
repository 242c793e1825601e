"""Fused NVLink all-reduce vs NCCL (needs >= 2 GPUs; one process per GPU)."""
import os

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _worker(rank, world, port, ret):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world)
    try:
        from torchgpipe_b200.parallel import GradientBucket, SymmetricBuffer, fused_allreduce_
        n = (1 << 22) + 8 * world            # not a round size on purpose
        buf = SymmetricBuffer(n, device=torch.device('cuda', rank))
        results = {}
        variants = [False] + ([True] if buf.has_multicast else [])
        for use_mc in variants:
            for it in range(3):              # several launches: barrier tickets must keep working
                g = torch.Generator(device='cuda').manual_seed(1000 * it + rank)
                x = torch.randn(buf.numel, device='cuda', generator=g)
                ref = x.clone()
                dist.all_reduce(ref)
                ref.mul_(0.5)
                buf.tensor.copy_(x)
                cast = torch.empty(buf.numel, dtype=torch.bfloat16, device='cuda')
                fused_allreduce_(buf, 0.5, cast_out=cast, multicast=use_mc)
                torch.cuda.synchronize()
                assert int(buf.error_flag.item()) == 0, 'barrier timed out'
                err = (buf.tensor - ref).abs().max().item()
                cerr = (cast.float() - ref).abs().max().item()
                results[(use_mc, it)] = (err, cerr)
        # gradient bucket: grads are views into the symmetric buffer
        lin = torch.nn.Linear(64, 32).cuda()
        bucket = GradientBucket(lin.parameters())
        lin(torch.full((4, 64), float(rank + 1), device='cuda')).sum().backward()
        own = torch.cat([p.grad.flatten().clone() for p in lin.parameters()])
        ref = own.clone()
        dist.all_reduce(ref)
        ref /= world
        bucket.allreduce()
        torch.cuda.synchronize()
        got = torch.cat([p.grad.flatten() for p in lin.parameters()])
        results['bucket'] = ((got - ref).abs().max().item(), 0.0)
        results['multicast'] = buf.has_multicast
        ret[rank] = results
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_fused_allreduce_matches_nccl():
    world = 2
    port = 29700 + os.getpid() % 1000
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    for rank in range(world):
        res = ret[rank]
        for key, value in res.items():
            if key == 'multicast':
                continue
            err, cerr = value
            assert err < 1e-5, (rank, key, err)
            assert cerr < 5e-2, (rank, key, cerr)


def _hybrid_worker(rank, world, port, ret):
    """1 pipeline stage x 2 replicas on 2 GPUs: the fused kernel on the symmetric bucket IS the
    data-parallel all-reduce (use_symmetric=True), through the CUDA-graph stage engine."""
    import torch.distributed as dist
    from torch import nn
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world)
    try:
        from torchgpipe_b200 import ops
        from torchgpipe_b200.parallel import HybridLayout, HybridPipeline
        device = torch.device('cuda', rank)
        torch.manual_seed(0)
        plain = nn.Sequential(nn.Conv2d(8, 16, 3, padding=1, bias=False), nn.BatchNorm2d(16),
                              nn.ReLU(), nn.Conv2d(16, 8, 1, bias=False))
        import copy
        model = ops.convert_to_sm100(copy.deepcopy(plain))
        layout = HybridLayout(stages=1, replicas=world)
        hybrid = HybridPipeline(model, [4], 2, layout, device, checkpoint='except_last',
                                cuda_graphs=True, use_symmetric=True)
        assert hybrid.bucket.symmetric is not None
        x = torch.randn(4, 8, 10, 10, generator=torch.Generator().manual_seed(7 + rank)).to(device)
        for step in range(2):                   # second step: graphs replay, bucket reused
            hybrid.zero_grad()
            outs = hybrid.forward(x)
            hybrid.backward([o.square().mean() / 2 for o in outs])
        torch.cuda.synchronize()
        got = [p.grad.detach().clone().cpu() for p in hybrid.model().parameters()]
        # oracle: the plain model on this replica's data, gradients averaged with NCCL
        plain = plain.to(device)
        loss = sum(plain(c).square().mean() / 2 for c in x.chunk(2))
        loss.backward()
        want = []
        for p in plain.parameters():
            g = p.grad.detach().clone()
            dist.all_reduce(g)
            want.append((g / world).cpu())
        ret[rank] = [(float((a - b).norm() / b.norm().clamp_min(1e-12))) for a, b in zip(got, want)]
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_hybrid_pipeline_uses_the_fused_allreduce_on_gpus():
    world = 2
    port = 29900 + os.getpid() % 1000
    ret = mp.Manager().dict()
    mp.spawn(_hybrid_worker, args=(world, port, ret), nprocs=world, join=True)
    for rank in range(world):
        assert all(err < 2e-2 for err in ret[rank]), (rank, ret[rank])
