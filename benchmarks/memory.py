"""Memory benchmarks: the largest U-Net (B,C) / AmoebaNet-D (L,D) per partition count, as in the
reference's ``benchmarks/{unet,amoebanetd}-memory/main.py``.

    python benchmarks/memory.py unet pipeline-4 [--impl ours|reference] [--smoke]

Runs two training steps and reports parameter count, parameter bytes x param_scale and the peak
reserved memory per device minus parameters ("peak activation memory" in the reference tables).
On 180 GB B200s the P40-era model sizes fit with a wide margin; ``--scale`` multiplies the base
channels to find the B200 limit.
"""
import argparse

import torch
import torch.nn.functional as F
from torch.optim import RMSprop, SGD

from common import AMOEBANETD_MEMORY, UNET_MEMORY, emit, load_impl

TABLES = {'unet': UNET_MEMORY, 'amoebanetd': AMOEBANETD_MEMORY}


def main():
    p = argparse.ArgumentParser()
    p.add_argument('model', choices=sorted(TABLES))
    p.add_argument('experiment')
    p.add_argument('--impl', choices=['ours', 'reference'], default='ours')
    p.add_argument('--scale', type=float, default=1.0)
    p.add_argument('--devices', type=str, default=None)
    p.add_argument('--no-kernels', action='store_true')
    p.add_argument('--smoke', action='store_true')
    p.add_argument('--out', default=None)
    args = p.parse_args()
    cfg = dict(TABLES[args.model][args.experiment])
    GPipe, builders, convert = load_impl(args.impl)

    a, b = cfg['model']
    if args.smoke:
        a, b = (2, 16) if args.model == 'unet' else (3, 64)
    if args.model == 'unet':
        model = builders['unet'](depth=5, num_convs=a, base_channels=int(b * args.scale))
        hw, param_scale, make_opt = 192 if not args.smoke else 64, 5.0, \
            lambda ps: RMSprop(ps, lr=0.1, momentum=0.9, centered=True)     # reference: RMSprop
    else:
        model = builders['amoebanetd'](num_classes=1000, num_layers=a,
                                       num_filters=int(b * args.scale))
        hw, param_scale, make_opt = 224, 3.0, lambda ps: SGD(ps, lr=0.1, momentum=0.9)
    if convert is not None and not args.no_kernels:
        model = convert(model)
    n_params = sum(p.numel() for p in model.parameters())
    batch = cfg['batch'] if not args.smoke else 4
    chunks = cfg['chunks'] if not args.smoke else (2 if cfg['chunks'] else None)

    n = len(cfg['balance']) if cfg['balance'] else 1
    devices = ([torch.device(d if d.startswith('c') else f'cuda:{d}')
                for d in args.devices.split(',')] if args.devices
               else [torch.device('cuda', i) for i in range(n)])
    if cfg['balance'] is None:
        pipe = model.to(devices[0])
        in_dev = out_dev = devices[0]
    else:
        balance = cfg['balance']
        if sum(balance) != len(model):          # smoke / scaled models: even split
            balance = [len(model) // n] * (n - 1) + [len(model) - len(model) // n * (n - 1)]
        pipe = GPipe(model, balance, devices=devices, chunks=chunks, checkpoint='always')
        in_dev, out_dev = pipe.devices[0], pipe.devices[-1]
    opt = make_opt(pipe.parameters())
    cuda = [d for d in dict.fromkeys(devices[:n]) if d.type == 'cuda']
    for d in cuda:
        torch.cuda.reset_peak_memory_stats(d)
    x = torch.rand(batch, 3, hw, hw, device=in_dev)
    for _ in range(2):
        out = pipe(x)
        if args.model == 'unet':
            loss = F.binary_cross_entropy_with_logits(out, torch.ones_like(out))
        else:
            loss = F.cross_entropy(out, torch.randint(1000, (batch,), device=out_dev))
        loss.backward()
        opt.step()
        opt.zero_grad()
    peak = [torch.cuda.max_memory_reserved(d) for d in cuda]
    param_bytes = n_params * 4 * param_scale
    emit({'benchmark': f'{args.model}-memory', 'experiment': args.experiment, 'impl': args.impl,
          'model': [a, int(b * args.scale)], 'parameters': n_params,
          'param_bytes_scaled_gib': param_bytes / 2 ** 30, 'param_scale': param_scale,
          'peak_reserved_gib_per_device': [v / 2 ** 30 for v in peak],
          'peak_activation_gib': max(0.0, (sum(peak) - param_bytes) / 2 ** 30),
          'batch': batch, 'chunks': chunks}, args.out)


if __name__ == '__main__':
    main()
