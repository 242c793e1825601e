"""Pipeline x data-parallel hybrid: S pipeline stages x R replicas with the fused NVLink gradient
all-reduce (``BASELINE.json`` config 5: U-Net (24,160) as 4 stages x 2 replicas on 8 GPUs).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29530 benchmarks/ppdp.py --stages 4 --replicas 2 [--model 24,160]
        [--allreduce fused|nccl] [--steps 3 --warmup 2]

Each replica is a ``DistributedGPipe`` (NCCL p2p between its stages); stage s of every replica
shares a ``GradientBucket`` in symmetric memory and all-reduces it with ``tgb_allreduce_scale_f32``
right after that stage's backward pass (later stages finish first, so their all-reduce overlaps the
backward drain of earlier stages). ``--allreduce nccl`` is the baseline: same pipeline, NCCL
all_reduce + scale. Prints one JSON line: samples/s (max over ranks), per-stage all-reduce time,
bytes and achieved fraction of the link roofline.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.optim import SGD

from torchgpipe_b200.utils import link_bandwidth_gbs  # noqa: E402

LINK_GBS = link_bandwidth_gbs()      # measured by benchmarks/p2p_bandwidth.py


def even_balance(n_layers, stages):
    base = n_layers // stages
    return [base] * (stages - 1) + [n_layers - base * (stages - 1)]


def main():
    p = argparse.ArgumentParser()
    p.add_argument('--stages', type=int, default=4)
    p.add_argument('--replicas', type=int, default=2)
    p.add_argument('--model', default='24,160', help='U-Net (B,C)')
    p.add_argument('--balance', default=None, help='comma list; default: reference [472,54,36,515] '
                   'for (24,160) with 4 stages, else even')
    p.add_argument('--batch', type=int, default=32, help='mini-batch per replica')
    p.add_argument('--chunks', type=int, default=8)
    p.add_argument('--hw', type=int, default=192)
    p.add_argument('--steps', type=int, default=3)
    p.add_argument('--warmup', type=int, default=2)
    p.add_argument('--allreduce', choices=['fused', 'nccl'], default='fused')
    p.add_argument('--no-graphs', action='store_true')
    args = p.parse_args()

    local = int(os.environ.get('LOCAL_RANK', 0))
    device = torch.device('cuda', local)
    torch.cuda.set_device(device)
    dist.init_process_group('nccl')
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models import unet
    from torchgpipe_b200.parallel import HybridLayout, HybridPipeline

    layout = HybridLayout(args.stages, args.replicas)
    B, C = (int(v) for v in args.model.split(','))
    # Built on the meta device: every rank materialises only its own stage (a 7.8 B-parameter
    # model is 31 GB in fp32 - eight full host copies would not fit).
    with torch.device('meta'):
        model = unet(depth=5, num_convs=B, base_channels=C)
    model = ops.convert_to_sm100(model)
    if args.balance:
        balance = [int(v) for v in args.balance.split(',')]
    elif (B, C, args.stages) == (24, 160, 4):
        balance = [472, 54, 36, 515]            # reference unet-memory pipeline-4
    else:
        balance = even_balance(len(model), args.stages)
    hybrid = HybridPipeline(model, balance, args.chunks, layout, device,
                            checkpoint='except_last', cuda_graphs=not args.no_graphs,
                            use_symmetric=args.allreduce == 'fused')
    # identical initial weights in every replica: seed by stage, fan-in scaled normal
    gen = torch.Generator(device=device).manual_seed(1234 + layout.stage)
    with torch.no_grad():
        for prm in hybrid.model().parameters():
            fan_in = prm[0].numel() if prm.dim() > 1 else prm.numel()
            prm.normal_(0.0, (2.0 / max(fan_in, 1)) ** 0.5, generator=gen)
    opt = SGD(hybrid.model().parameters(), lr=0.01)
    first, last = layout.stage == 0, layout.stage == args.stages - 1
    torch.manual_seed(100 + layout.replica)     # different data per replica
    x = torch.rand(args.batch, 3, args.hw, args.hw, device=device) if first else None
    total = float(args.batch * args.hw * args.hw)
    ar_events = []

    def step(record=False):
        outs = hybrid.forward(x)
        if last:
            losses = [F.binary_cross_entropy_with_logits(o, torch.ones_like(o), reduction='sum')
                      / total for o in outs]
            hybrid.pipe.backward(losses)
        else:
            hybrid.pipe.backward(None)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        hybrid.bucket.allreduce()
        b.record()
        if record:
            ar_events.append((a, b))
        opt.step()
        hybrid.zero_grad()

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.steps):
        step(record=True)
    t1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = t0.elapsed_time(t1) / args.steps
    ar_ms = sum(a.elapsed_time(b) for a, b in ar_events) / max(len(ar_events), 1)
    nbytes = hybrid.bucket.flat.numel() * 4
    # library baseline on a bucket of the same size: NCCL all_reduce + scale, same group
    nccl_ms = None
    try:
        scratch = torch.zeros_like(hybrid.bucket.flat)
        for _ in range(2):
            dist.all_reduce(scratch, group=layout.dp_group)
            scratch.mul_(1.0 / args.replicas)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            dist.all_reduce(scratch, group=layout.dp_group)
            scratch.mul_(1.0 / args.replicas)
        b.record()
        torch.cuda.synchronize()
        nccl_ms = a.elapsed_time(b) / 3
        del scratch
    except RuntimeError:
        pass
    stats = torch.tensor([ms, ar_ms], device=device, dtype=torch.float64)
    dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    gathered = [None] * dist.get_world_size()
    dist.all_gather_object(gathered, {'stage': layout.stage, 'replica': layout.replica,
                                      'allreduce_ms': ar_ms, 'grad_bytes': nbytes,
                                      'nccl_ms': nccl_ms,
                                      'peak_gib': torch.cuda.max_memory_allocated() / 2 ** 30})
    if dist.get_rank() == 0:
        R = args.replicas
        per_stage = []
        for s in range(args.stages):
            rows = [g for g in gathered if g['stage'] == s]
            t = max(r['allreduce_ms'] for r in rows)
            wire = 2 * rows[0]['grad_bytes'] * (R - 1) / R
            ideal = wire / (LINK_GBS * 1e9) * 1e3
            per_stage.append({'stage': s, 'grad_mib': rows[0]['grad_bytes'] / 2 ** 20,
                              'allreduce_ms': round(t, 3),
                              'nccl_plus_scale_ms': (round(max(r['nccl_ms'] for r in rows), 3)
                                                     if rows[0]['nccl_ms'] else None),
                              'peak_gib': round(max(r['peak_gib'] for r in rows), 1),
                              'roofline_ms': round(ideal, 3),
                              'fraction_of_roofline': round(ideal / t, 3) if t > 0 else None})
        print(json.dumps({'benchmark': 'ppdp', 'model': f'unet({B},{C})', 'stages': args.stages,
                          'replicas': R, 'balance': balance, 'allreduce': args.allreduce,
                          'global_batch': args.batch * R, 'chunks': args.chunks,
                          'samples_per_s': args.batch * R / stats[0].item() * 1e3,
                          'ms_per_step': stats[0].item(), 'per_stage_allreduce': per_stage}),
              flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
