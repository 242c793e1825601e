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
    p.add_argument('--batch', type=int, default=80)
    p.add_argument('--chunks', type=int, default=2)
    p.add_argument('--rows', type=int, default=45)
    args = p.parse_args()

    from torchgpipe_b200 import GPipe, ops
    from torchgpipe_b200.models import unet
    torch.manual_seed(0)
    model = unet(depth=5, num_convs=5, base_channels=64)
    if not args.no_kernels:
        model = ops.convert_to_sm100(model)
    pipe = GPipe(model, [241], devices=['cuda:0'], chunks=args.chunks, checkpoint='except_last')
    opt = SGD(pipe.parameters(), lr=0.1)
    x = torch.rand(args.batch, 3, 192, 192, device='cuda')
    t = torch.ones(args.batch, 1, 192, 192, device='cuda')

    def step():
        loss = F.binary_cross_entropy_with_logits(pipe(x), t)
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
