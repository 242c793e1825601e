"""Speed benchmarks: U-Net (5,64), ResNet-101, AmoebaNet-D (18,256) - the reference's
``benchmarks/{unet,resnet101,amoebanetd}-speed/main.py`` experiments in one driver.

    python benchmarks/speed.py unet pipeline-4 [--steps 5 --warmup 3] [--impl ours|reference]
    python benchmarks/speed.py amoebanetd n8m32 --balance auto     # balance_by_time

Single-process mode (one process drives all stages, like the reference CLIs); for the
one-process-per-GPU NCCL mode see ``bench.py``. Differences from the reference drivers, on
purpose: timing uses CUDA events on *every* pipeline device and takes the slowest (the reference
reads the host clock after synchronising the input device only); a step-based loop replaces the
10 000-sample "epoch"; results are JSON lines with bubble and copy time from the per-cell timeline.
"""
import argparse

import torch
import torch.nn.functional as F
from torch.optim import SGD

from common import (AMOEBANETD_SPEED, RESNET101_SPEED, UNET_SPEED, device_timed, emit, load_impl)

TABLES = {'unet': UNET_SPEED, 'resnet101': RESNET101_SPEED, 'amoebanetd': AMOEBANETD_SPEED}


def build(model_name, builders, smoke):
    if model_name == 'unet':
        model = builders['unet'](depth=5, num_convs=5, base_channels=16 if smoke else 64)
        return model, (3, 64 if smoke else 192), 'bce'
    if model_name == 'resnet101':
        return builders['resnet101'](num_classes=1000), (3, 64 if smoke else 224), 'ce'
    model = builders['amoebanetd'](num_classes=1000, num_layers=18,
                                   num_filters=64 if smoke else 256)
    return model, (3, 224), 'ce'


def main():
    p = argparse.ArgumentParser()
    p.add_argument('model', choices=sorted(TABLES))
    p.add_argument('experiment')
    p.add_argument('--impl', choices=['ours', 'reference'], default='ours')
    p.add_argument('--steps', type=int, default=5)
    p.add_argument('--warmup', type=int, default=3)
    p.add_argument('--devices', type=str, default=None, help='e.g. "0,1,2,3" or "cpu,cpu"')
    p.add_argument('--balance', default='reference', help='"reference" or "auto" (balance_by_time)')
    p.add_argument('--no-kernels', action='store_true')
    p.add_argument('--graphs', action='store_true',
                   help='ours: GPipe(cuda_graphs=True), every stage pass replayed from CUDA graphs')
    p.add_argument('--smoke', action='store_true', help='tiny model / batch (plumbing check)')
    p.add_argument('--out', default=None, help='append the JSON line to this file')
    args = p.parse_args()

    table = TABLES[args.model]
    if args.experiment not in table:
        raise SystemExit(f'experiment must be one of {sorted(table)}')
    cfg = dict(table[args.experiment])
    GPipe, builders, convert = load_impl(args.impl)
    torch.manual_seed(0)
    model, (channels, hw), loss_kind = build(args.model, builders, args.smoke)
    if convert is not None and not args.no_kernels:
        model = convert(model)
    chunks = cfg['chunks']
    if args.smoke and chunks:
        chunks = min(chunks, 2)
    batch = cfg['batch'] if not args.smoke else 4

    if args.devices:
        devices = [torch.device(d if d.startswith('c') else f'cuda:{d}')
                   for d in args.devices.split(',')]
    else:
        n = len(cfg['balance']) if cfg['balance'] else 1
        devices = [torch.device('cuda', i) for i in range(n)]

    timeline = None
    if cfg['balance'] is None:
        model = model.to(devices[0])
        pipe, in_dev, out_dev = model, devices[0], devices[0]
    else:
        balance = cfg['balance']
        if args.smoke:
            n = len(balance)
            balance = [len(model) // n] * (n - 1) + [len(model) - len(model) // n * (n - 1)]
        if args.balance == 'auto' and args.impl == 'ours':
            from torchgpipe_b200.balance import balance_by_time
            sample = torch.rand(max(batch // chunks, 1), channels, hw, hw)
            balance = balance_by_time(len(balance), model, sample, device=devices[0])
        extra = {'cuda_graphs': True} if (args.graphs and args.impl == 'ours') else {}
        pipe = GPipe(model, balance, devices=devices, chunks=chunks,
                     checkpoint=cfg.get('checkpoint', 'except_last'), **extra)
        in_dev, out_dev = pipe.devices[0], pipe.devices[-1]
        cfg['balance'] = list(balance)
        if args.impl == 'ours' and not args.graphs:
            from torchgpipe_b200.runtime.timeline import Timeline
            timeline = Timeline()

    opt = SGD(pipe.parameters(), lr=0.1)
    x = torch.rand(batch, channels, hw, hw, device=in_dev)
    if loss_kind == 'bce':
        target = torch.ones(batch, 1, hw, hw, device=out_dev)
    else:
        target = torch.randint(1000, (batch,), device=out_dev)

    def step():
        out = pipe(x)
        loss = F.binary_cross_entropy_with_logits(out, target) if loss_kind == 'bce' \
            else F.cross_entropy(out, target)
        loss.backward()
        opt.step()
        opt.zero_grad()

    used = list(dict.fromkeys(devices[:len(cfg['balance'])] if cfg['balance'] else devices[:1]))
    ms = device_timed(used, step, args.steps, args.warmup)
    record = {'benchmark': f'{args.model}-speed', 'experiment': args.experiment, 'impl': args.impl,
              'samples_per_s': batch / ms * 1e3, 'ms_per_step': ms, 'batch': batch,
              'chunks': chunks, 'balance': cfg['balance'], 'devices': [str(d) for d in used],
              'checkpoint': cfg.get('checkpoint', 'except_last'), 'steps': args.steps,
              'dtype': 'fp32 storage / tf32 conv', 'data': 'synthetic',
              'kernels': 'sm100' if (convert is not None and not args.no_kernels) else 'torch',
              'engine': 'GPipe(cuda_graphs=True)' if (args.graphs and args.impl == 'ours')
              else 'GPipe'}
    if timeline is not None:
        pipe.timeline = timeline
        step()
        for d in used:
            if d.type == 'cuda':
                torch.cuda.synchronize(d)
        rep = timeline.report(len(cfg['balance']))
        record['forward_bubble_frac'] = rep['bubble_frac']
        record['forward_copy_ms'] = rep['copy_ms']
        record['forward_exposed_comm_ms'] = rep['exposed_comm_ms']
    emit(record, args.out)


if __name__ == '__main__':
    main()
