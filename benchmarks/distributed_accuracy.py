"""Multi-process accuracy demo (reference ``benchmarks/distributed/accuracy/main.py``): one rank per
pipeline stage trains VGG-16 / coarse ResNet on CIFAR-10 resized to 224x224, logs loss, validation
accuracy and samples/s per epoch.

    torchrun --nproc-per-node 4 --master-addr 127.0.0.1 benchmarks/distributed_accuracy.py \
        --model vgg16 --balance 10,10,10,9 --batch-size 128 --chunks 4 --epochs 10

Differences from the reference driver, on purpose: ranks come from ``torchrun`` (not ``--rank /
--world`` plus RPC), activations / gradients / targets travel over ``torch.distributed`` p2p (NCCL
on GPUs, gloo with ``--cpu``), ``--cuda-graphs`` replays each stage from CUDA graphs, and the
validation split is CIFAR-10's test set (the reference validates on the training split,
``main.py:62-63``). Without a local CIFAR-10 copy (``--data``) or with ``--synthetic`` a fixed
random data set of the same shape is used - the box has no network - and the accuracy column
is then only a plumbing check. Optimiser / LR schedule: ``main.py:217-237``.
"""
import argparse
import os
import sys
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.optim import SGD
from torch.optim.lr_scheduler import LambdaLR

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from torchgpipe_b200 import models, ops  # noqa: E402
from torchgpipe_b200.balance import balance_by_time  # noqa: E402
from torchgpipe_b200.distributed import (DistributedGPipe, DistributedGPipeDataLoader,  # noqa: E402
                                         ProcessGroupTransport)

MODELS = {'vgg16': models.vgg16, 'resnet101': models.resnet101_coarse,
          'resnet50': models.resnet50_coarse, 'resnet101-fine': models.resnet101}


class SyntheticSet(torch.utils.data.Dataset):
    """Class-dependent blobs + noise: learnable, so a falling loss still means something."""

    def __init__(self, n, hw, classes, seed):
        g = torch.Generator().manual_seed(seed)
        self.proto = torch.randn(classes, 3, 8, 8, generator=torch.Generator().manual_seed(7))
        self.y = torch.randint(classes, (n,), generator=g)
        self.noise_seed = seed
        self.hw = hw

    def __len__(self):
        return self.y.numel()

    def __getitem__(self, i):
        y = int(self.y[i])
        g = torch.Generator().manual_seed(self.noise_seed * 1000003 + i)
        x = F.interpolate(self.proto[y][None], size=(self.hw, self.hw), mode='nearest')[0]
        return x + 0.5 * torch.randn(3, self.hw, self.hw, generator=g), y


def datasets(args, classes):
    if args.data and not args.synthetic:
        import torchvision
        from torchvision import transforms as T
        tf = T.Compose([T.Resize((args.image_size, args.image_size)), T.ToTensor(),
                        T.Normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])])
        return (torchvision.datasets.CIFAR10(args.data, train=True, transform=tf, download=False),
                torchvision.datasets.CIFAR10(args.data, train=False, transform=tf, download=False),
                'cifar10')
    return (SyntheticSet(args.synthetic_size, args.image_size, classes, 1),
            SyntheticSet(max(args.batch_size, args.synthetic_size // 5), args.image_size, classes, 2),
            'synthetic')


def lr_schedule(steps_per_epoch, batch):
    scale = max(1.0, batch / 256)

    def f(step):
        epoch = step / steps_per_epoch
        mult = min(4.0, epoch) / 4.0 * (scale - 1.0) + 1.0
        return mult * (1.0 if epoch < 30 else 0.1 if epoch < 60 else 0.01 if epoch < 80 else 0.001)
    return f


def main():
    p = argparse.ArgumentParser()
    p.add_argument('--model', '-m', choices=sorted(MODELS), default='vgg16')
    p.add_argument('--balance', '-b', default=None, help='comma list; default: balance_by_time')
    p.add_argument('--batch-size', '-s', type=int, default=128)
    p.add_argument('--chunks', '-c', type=int, default=4)
    p.add_argument('--epochs', '-e', type=int, default=10)
    p.add_argument('--skip-epochs', '-k', type=int, default=1)
    p.add_argument('--classes', type=int, default=10)
    p.add_argument('--image-size', type=int, default=224)
    p.add_argument('--lr', type=float, default=0.1)
    p.add_argument('--data', default=None, help='directory holding cifar-10-batches-py')
    p.add_argument('--synthetic', action='store_true')
    p.add_argument('--synthetic-size', type=int, default=2560)
    p.add_argument('--checkpoint', default='never', choices=['always', 'except_last', 'never'])
    p.add_argument('--cuda-graphs', action='store_true')
    p.add_argument('--no-kernels', action='store_true')
    p.add_argument('--cpu', action='store_true', help='gloo + CPU tensors (plumbing check)')
    args = p.parse_args()
    if args.skip_epochs > args.epochs:
        p.error('--skip-epochs must not exceed --epochs')
    if args.batch_size % args.chunks:
        p.error('--batch-size must be a multiple of --chunks')

    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29533')
    if args.cpu:
        device = torch.device('cpu')
        dist.init_process_group('gloo', rank=rank, world_size=world)
    else:
        device = torch.device('cuda', int(os.environ.get('LOCAL_RANK', 0)))
        torch.cuda.set_device(device)
        dist.init_process_group('nccl', rank=rank, world_size=world)   # lazy: per-pair p2p comms
    first, last = rank == 0, rank == world - 1

    torch.manual_seed(0)
    model = MODELS[args.model](num_classes=args.classes, inplace=False)
    if args.balance:
        balance = [int(x) for x in args.balance.split(',')]
    else:
        sample = torch.rand(args.batch_size // args.chunks, 3, args.image_size, args.image_size)
        # every rank must derive the same balance: rank 0 profiles, the others take its answer
        box = [balance_by_time(world, model, sample, timeout=0.2 if args.cpu else 1.0,
                               device=device)] if first else [None]
        dist.broadcast_object_list(box, src=0)
        balance = box[0]
    if not args.no_kernels and not args.cpu:
        model = ops.convert_to_sm100(model)
    workers = {r: f'worker{r}' for r in range(world)}
    transport = ProcessGroupTransport(device)
    pipe = DistributedGPipe(model, rank, workers, balance, args.chunks, device=device,
                            checkpoint=args.checkpoint, transport=transport,
                            cuda_graphs=args.cuda_graphs)

    train_set, valid_set, data_name = datasets(args, args.classes)
    mk = torch.utils.data.DataLoader
    # only rank 0 reads data; drop_last keeps every mini-batch divisible into equal micro-batches
    train_src = mk(train_set, args.batch_size, shuffle=True, drop_last=True) if first else None
    valid_src = mk(valid_set, args.batch_size, shuffle=False, drop_last=True) if first else None
    n_train, n_valid = len(train_set) // args.batch_size, len(valid_set) // args.batch_size
    loader = dict(transport=transport, last_stage_rank=world - 1)
    train = DistributedGPipeDataLoader(train_src, rank, args.chunks, n_train, last,
                                       workers[world - 1], **loader)
    valid = DistributedGPipeDataLoader(valid_src, rank, args.chunks, n_valid, last,
                                       workers[world - 1], **loader)

    opt = SGD(pipe.model().parameters(), lr=args.lr, momentum=0.9, weight_decay=1e-4,
              nesterov=True)
    sched = LambdaLR(opt, lr_schedule(max(n_train, 1), args.batch_size))

    def say(msg):
        if last:
            print(f'{time.time() - t0:8.1f}s | {msg}', flush=True)

    def sync():
        if device.type == 'cuda':
            torch.cuda.synchronize(device)
        dist.barrier()

    def evaluate():
        pipe.model().eval()
        loss_sum = torch.zeros((), device=device)
        correct = torch.zeros((), device=device)
        seen = 0
        with torch.no_grad():
            for x, y in valid:
                outs = pipe.forward(x.to(device, non_blocking=True) if first else None)
                if last:
                    out = torch.cat(list(outs))
                    y = y.to(device)
                    loss_sum += F.cross_entropy(out, y, reduction='sum')
                    correct += (out.argmax(1) == y).sum()
                    seen += y.numel()
        pipe.model().train()
        return (loss_sum.item() / max(seen, 1), correct.item() / max(seen, 1)) if last else (0., 0.)

    def run_epoch(epoch):
        sync()
        tick = time.time()
        loss_sum = torch.zeros((), device=device)
        seen = 0
        micro = args.batch_size // args.chunks
        for x, y in train:
            outs = pipe.forward(x.to(device, non_blocking=True) if first else None)
            if last:
                y = y.to(device, non_blocking=True)
                # mean over the mini-batch = sum of per-micro-batch means / chunks
                losses = [F.cross_entropy(o, t) / args.chunks
                          for o, t in zip(outs, y.chunk(args.chunks))]
                loss_sum += sum(l.detach() for l in losses) * micro * args.chunks
                pipe.backward(losses)
            else:
                pipe.backward(None)
            opt.step()
            opt.zero_grad()
            sched.step()
            seen += args.batch_size
        sync()
        elapsed = time.time() - tick
        valid_loss, accuracy = evaluate()
        say('%d/%d epoch | lr:%.5f | train loss:%.3f %.3f samples/sec | valid loss:%.3f '
            'accuracy:%.3f' % (epoch + 1, args.epochs, sched.get_last_lr()[0],
                               loss_sum.item() / max(seen, 1), seen / elapsed, valid_loss,
                               accuracy))
        return seen / elapsed, elapsed, accuracy

    t0 = time.time()
    say('%s, %d stages, balance %s, batch %d, chunks %d, data %s, torch %s' % (
        args.model, world, balance, args.batch_size, args.chunks, data_name, torch.__version__))
    throughputs, times, accuracy = [], [], 0.0
    for epoch in range(args.epochs):
        thr, sec, accuracy = run_epoch(epoch)
        if epoch >= args.skip_epochs:
            throughputs.append(thr)
            times.append(sec)
    n = max(len(throughputs), 1)
    say('%s | valid accuracy: %.4f | %.3f samples/sec, %.3f sec/epoch (average)' % (
        args.model, accuracy, sum(throughputs) / n, sum(times) / n))
    transport.flush()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
