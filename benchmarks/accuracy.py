"""ResNet-101 accuracy benchmark (reference ``benchmarks/resnet101-accuracy/main.py``): train with
GPipe at large batch sizes and check that accuracy is not hurt, against naive and DataParallel runs.

    python benchmarks/accuracy.py gpipe-1k --data /path/to/imagenet [--epochs 90]
    python benchmarks/accuracy.py gpipe-256 --synthetic --epochs 1 --steps-per-epoch 20

Experiments (same batch sizes, chunks and balances as the reference, ``main.py:23-82``):
``naive-128``, ``dataparallel-256``, ``dataparallel-1k``, ``gpipe-256`` (2 GPUs, m=4, [140,230]),
``gpipe-1k`` / ``gpipe-4k`` (8 GPUs, m=64, [26,22,33,44,44,66,66,69]).
Optimiser and schedule follow Goyal et al. as in the reference (``main.py:225-245``): SGD
momentum 0.9 nesterov, weight decay 1e-4, base LR 0.1 scaled linearly with batch/256 after a
4-epoch gradual warm-up, x0.1 at epochs 30 / 60 / 80. ImageNet needs ``torchvision`` and a local
copy; with ``--synthetic`` (no dataset on the box) random data exercises the same loop and the
output is throughput only. One JSON line per epoch.
"""
import argparse
import time

import torch
import torch.nn.functional as F
from torch import nn
from torch.optim import SGD
from torch.optim.lr_scheduler import LambdaLR

from common import emit, load_impl

EXPERIMENTS = {
    'naive-128': dict(batch=128, mode='naive', gpus=1),
    'dataparallel-256': dict(batch=256, mode='dp', gpus=2),
    'dataparallel-1k': dict(batch=1024, mode='dp', gpus=4),
    'gpipe-256': dict(batch=256, mode='gpipe', gpus=2, chunks=4, balance=[140, 230]),
    'gpipe-1k': dict(batch=1024, mode='gpipe', gpus=8, chunks=64,
                     balance=[26, 22, 33, 44, 44, 66, 66, 69]),
    'gpipe-4k': dict(batch=4096, mode='gpipe', gpus=8, chunks=64,
                     balance=[26, 22, 33, 44, 44, 66, 66, 69]),
}


def lr_lambda(steps_per_epoch: int, batch: int):
    scale = max(1.0, batch / 256)

    def schedule(step: int) -> float:
        epoch = step / steps_per_epoch
        warm = min(4.0, epoch) / 4.0
        mult = warm * (scale - 1.0) + 1.0
        decay = 1.0 if epoch < 30 else 0.1 if epoch < 60 else 0.01 if epoch < 80 else 0.001
        return decay * mult
    return schedule


class Synthetic:
    """``steps`` identical random batches (the speed benchmarks' data model)."""

    def __init__(self, batch, steps, hw=224, classes=1000):
        g = torch.Generator().manual_seed(0)
        self.x = torch.rand(batch, 3, hw, hw, generator=g)
        self.y = torch.randint(classes, (batch,), generator=g)
        self.steps = steps

    def __len__(self):
        return self.steps

    def __iter__(self):
        for _ in range(self.steps):
            yield self.x, self.y


def imagenet_loaders(root, batch, workers):
    import os
    import torchvision
    from torchvision import transforms as T
    norm = T.Normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])
    train = torchvision.datasets.ImageFolder(os.path.join(root, 'train'), T.Compose(
        [T.RandomResizedCrop(224), T.RandomHorizontalFlip(), T.ToTensor(), norm]))
    valid = torchvision.datasets.ImageFolder(os.path.join(root, 'val'), T.Compose(
        [T.Resize(256), T.CenterCrop(224), T.ToTensor(), norm]))
    mk = torch.utils.data.DataLoader
    return (mk(train, batch, shuffle=True, num_workers=workers, pin_memory=True, drop_last=True),
            mk(valid, batch, shuffle=False, num_workers=workers, pin_memory=True))


def evaluate(model, loader, in_dev, out_dev):
    model.eval()
    correct = total = 0
    loss_sum = 0.0
    with torch.no_grad():
        for x, y in loader:
            out = model(x.to(in_dev, non_blocking=True))
            y = y.to(out_dev, non_blocking=True)
            loss_sum += F.cross_entropy(out, y, reduction='sum').item()
            correct += (out.argmax(1) == y).sum().item()
            total += y.numel()
    model.train()
    return loss_sum / max(total, 1), correct / max(total, 1)


def main():
    p = argparse.ArgumentParser()
    p.add_argument('experiment', choices=sorted(EXPERIMENTS))
    p.add_argument('--impl', choices=['ours', 'reference'], default='ours')
    p.add_argument('--data', default=None, help='ImageNet root with train/ and val/')
    p.add_argument('--synthetic', action='store_true')
    p.add_argument('--epochs', type=int, default=90)
    p.add_argument('--steps-per-epoch', type=int, default=50, help='synthetic mode only')
    p.add_argument('--workers', type=int, default=8)
    p.add_argument('--devices', default=None)
    p.add_argument('--no-kernels', action='store_true')
    p.add_argument('--deferred-batch-norm', action='store_true')
    p.add_argument('--batch', type=int, default=None, help='override (smoke runs)')
    p.add_argument('--out', default=None)
    args = p.parse_args()
    cfg = dict(EXPERIMENTS[args.experiment])
    if args.batch:
        cfg['batch'] = args.batch
    if not args.synthetic and not args.data:
        raise SystemExit('give --data IMAGENET_ROOT or --synthetic')

    GPipe, builders, convert = load_impl(args.impl)
    # in-place ReLU must stay off under GPipe checkpointing (reference main.py:211)
    model = builders['resnet101'](num_classes=1000, inplace=cfg['mode'] != 'gpipe')
    if convert is not None and not args.no_kernels and cfg['mode'] == 'gpipe':
        model = convert(model)
    devices = ([torch.device(d if d.startswith('c') else f'cuda:{d}')
                for d in args.devices.split(',')] if args.devices
               else [torch.device('cuda', i) for i in range(cfg['gpus'])])
    if cfg['mode'] == 'gpipe':
        balance = cfg['balance']
        if len(devices) != len(balance):            # smoke runs on fewer devices
            n = len(devices)
            balance = [len(model) // n] * (n - 1) + [len(model) - len(model) // n * (n - 1)]
        model = GPipe(model, balance, devices=devices, chunks=cfg['chunks'],
                      deferred_batch_norm=args.deferred_batch_norm)
        in_dev, out_dev = model.devices[0], model.devices[-1]
    elif cfg['mode'] == 'dp':
        model = nn.DataParallel(model.to(devices[0]), device_ids=[d.index for d in devices],
                                output_device=devices[-1].index)
        in_dev, out_dev = devices[0], devices[-1]
    else:
        model = model.to(devices[0])
        in_dev = out_dev = devices[0]

    if args.synthetic:
        train, valid = Synthetic(cfg['batch'], args.steps_per_epoch), None
    else:
        train, valid = imagenet_loaders(args.data, cfg['batch'], args.workers)
    opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    sched = LambdaLR(opt, lr_lambda(len(train), cfg['batch']))

    for epoch in range(args.epochs):
        tick = time.time()
        seen, loss_sum = 0, 0.0
        for x, y in train:
            out = model(x.to(in_dev, non_blocking=True))
            loss = F.cross_entropy(out, y.to(out_dev, non_blocking=True))
            loss.backward()
            opt.step()
            sched.step()
            opt.zero_grad()
            seen += x.size(0)
            loss_sum += float(loss.detach()) * x.size(0)
        for d in devices:
            if d.type == 'cuda':
                torch.cuda.synchronize(d)
        elapsed = time.time() - tick
        rec = {'benchmark': 'resnet101-accuracy', 'experiment': args.experiment, 'impl': args.impl,
               'epoch': epoch + 1, 'lr': sched.get_last_lr()[0], 'train_loss': loss_sum / seen,
               'samples_per_s': seen / elapsed, 'data': 'synthetic' if args.synthetic else 'imagenet'}
        if valid is not None:
            rec['valid_loss'], acc = evaluate(model, valid, in_dev, out_dev)
            rec['top1_error'] = 1.0 - acc
        emit(rec, args.out)


if __name__ == '__main__':
    main()
