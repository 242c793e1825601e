"""Kernel-time breakdown of one training step of the flagship benchmark (torch.profiler / CUPTI).

    python tools/profile_step.py [--no-kernels] [--gpus 1] > gpurun_out/step_profile.txt

Prints the top CUDA kernels by total device time for ONE step (after warm-up). Used to decide
what to optimise next; ncu captures of individual kernels live under profiles/.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.nn.functional as F
from torch.optim import SGD
from torch.profiler import ProfilerActivity, profile


def main():
    p = argparse.ArgumentParser()
    p.add_argument('--no-kernels', action='store_true')
    p.add_argument('--model', choices=['unet', 'resnet101', 'amoebanetd'], default='unet')
    p.add_argument('--batch', type=int, default=None)
    p.add_argument('--chunks', type=int, default=2)
    p.add_argument('--rows', type=int, default=45)
    p.add_argument('--graphs', action='store_true',
                   help='the graph-replayed stage engine (what bench.py runs) instead of eager GPipe')
    args = p.parse_args()

    from torchgpipe_b200 import GPipe, ops
    from torchgpipe_b200 import models
    torch.manual_seed(0)
    if args.model == 'unet':
        model = models.unet(depth=5, num_convs=5, base_channels=64)
        batch, hw = args.batch or 80, 192
    elif args.model == 'resnet101':
        model = models.resnet101(num_classes=1000)
        batch, hw = args.batch or 220, 224
    else:
        model = models.amoebanetd(num_classes=1000, num_layers=18, num_filters=256)
        batch, hw = args.batch or 40, 224
    if not args.no_kernels:
        model = ops.convert_to_sm100(model)
    x = torch.rand(batch, 3, hw, hw, device='cuda')
    if args.model == 'unet':
        t = torch.ones(batch, 1, hw, hw, device='cuda')
        loss_fn = F.binary_cross_entropy_with_logits
    else:
        t = torch.randint(1000, (batch,), device='cuda')
        loss_fn = F.cross_entropy

    if args.graphs:
        from torchgpipe_b200.distributed import DistributedGPipe, context
        context.worker('solo', args.chunks).__enter__()
        pipe = DistributedGPipe(model, 0, {0: 'solo'}, [len(model)], args.chunks,
                                device=torch.device('cuda:0'), checkpoint='except_last',
                                cuda_graphs=True)
        opt = SGD(pipe.model().parameters(), lr=0.1)
        scale = 1.0 / args.chunks

        def step():
            outs = pipe.forward(x)
            pipe.backward([loss_fn(o, tt) * scale for o, tt in zip(outs, t.chunk(args.chunks))])
            opt.step()
            opt.zero_grad()
    else:
        pipe = GPipe(model, [len(model)], devices=['cuda:0'], chunks=args.chunks,
                     checkpoint='except_last')
        opt = SGD(pipe.parameters(), lr=0.1)

        def step():
            loss = loss_fn(pipe(x), t)
            loss.backward()
            opt.step()
            opt.zero_grad()

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step()
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=args.rows,
                                    max_name_column_width=90))


if __name__ == '__main__':
    main()
