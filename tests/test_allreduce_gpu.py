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
