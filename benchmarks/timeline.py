"""U-Net timeline / ablation benchmark (reference ``benchmarks/unet-timeline/main.py``): how much
each of GPipe's three mechanisms contributes - backward dependency edges, side copy streams, and
portals for skip tensors.

    python benchmarks/timeline.py dep-str-ptl   # everything on (the product)
    python benchmarks/timeline.py dep-str-x     # skips travel through every stage as tuples
    python benchmarks/timeline.py dep-x-x       # ... and copies run on the compute streams
    python benchmarks/timeline.py baseline      # ... and no backward ordering edges
    python benchmarks/timeline.py product       # all mechanisms + sm_100a operators + CUDA graphs

The four ablation rows run the same PyTorch operators (the ablation is about the pipeline's
mechanisms, as in the reference); ``product`` is the path a user of this repository gets.

4 stages, batch 128, 8 micro-batches, balance [34,76,70,61] (reference main.py:146-166). Reports
samples/s, mean GPU utilisation sampled with nvidia-smi (reference gpu_utils.py) and peak memory.
"""
import argparse
import subprocess
import threading
import time

import torch
import torch.nn.functional as F
from torch import nn
from torch.optim import SGD

from common import device_timed, emit, load_impl

EXPERIMENTS = {'baseline': (False, False, False), 'dep-x-x': (True, False, False),
               'dep-str-x': (True, True, False), 'dep-str-ptl': (True, True, True),
               # everything on, plus what this repository adds: the sm_100a operators and
               # CUDA-graph replay of every stage pass (GPipe(cuda_graphs=True))
               'product': (True, True, True)}


class UtilSampler:
    def __init__(self, ids):
        self.ids, self.samples, self.stop_flag = ids, [], False

    def _run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(['nvidia-smi', '--query-gpu=utilization.gpu',
                                      '--format=csv,noheader,nounits', '-i',
                                      ','.join(map(str, self.ids))], capture_output=True,
                                     text=True, timeout=5).stdout
                vals = [float(v) for v in out.split()]
                if vals:
                    self.samples.append(sum(vals) / len(vals))
            except Exception:
                pass
            time.sleep(0.05)

    def __enter__(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *exc):
        self.stop_flag = True
        self.thread.join(timeout=6)

    @property
    def mean(self):
        return sum(self.samples) / len(self.samples) if self.samples else None


def tuplify_skips(model: nn.Sequential) -> nn.Sequential:
    """Rewrites ``@skippable`` stash/pop layers so skip tensors ride along the main lane as extra
    tuple members through every layer in between (what a pipeline without portals must do)."""
    from torchgpipe_b200.skip.skippable import Skippable

    class Carry(nn.Module):
        """Applies ``inner`` to the first tuple member and carries the rest."""

        def __init__(self, inner, push=False, pop=False):
            super().__init__()
            self.inner, self.push, self.pop = inner, push, pop

        def forward(self, xs):
            xs = xs if isinstance(xs, tuple) else (xs,)
            x, rest = xs[0], xs[1:]
            if self.push:                       # Stash: x is both output and new skip
                return (x,) + rest + (x,)
            if self.pop:                        # PopCat: consume the most recent skip
                skip, rest = rest[-1], rest[:-1]
                dh, dw = skip.size(2) - x.size(2), skip.size(3) - x.size(3)
                if dh or dw:
                    x = F.pad(x, [0, dw, 0, dh])
                x = torch.cat((x, skip), dim=1)
                return (x,) + rest if rest else x
            y = self.inner(x)
            return (y,) + rest if rest else y

    layers = []
    for name, layer in model.named_children():
        if isinstance(layer, Skippable):
            layers.append((name, Carry(None, push=bool(layer.stashable_names),
                                       pop=bool(layer.poppable_names))))
        else:
            layers.append((name, Carry(layer)))
    from collections import OrderedDict
    return nn.Sequential(OrderedDict(layers))


def main():
    p = argparse.ArgumentParser()
    p.add_argument('experiment', choices=sorted(EXPERIMENTS))
    p.add_argument('--steps', type=int, default=5)
    p.add_argument('--warmup', type=int, default=3)
    p.add_argument('--devices', type=str, default=None)
    p.add_argument('--smoke', action='store_true')
    p.add_argument('--out', default=None)
    args = p.parse_args()
    use_dep, use_streams, use_portals = EXPERIMENTS[args.experiment]

    GPipe, builders, convert = load_impl('ours')
    import torchgpipe_b200.runtime.pipeline as rp
    torch.manual_seed(0)
    model = builders['unet'](depth=5, num_convs=5, base_channels=16 if args.smoke else 64)
    balance = [34, 76, 70, 61]
    product = args.experiment == 'product'
    if not use_portals:
        model = tuplify_skips(model)
    elif product and not args.smoke:
        model = convert(model)
    batch, chunks, hw = (8, 4, 64) if args.smoke else (128, 8, 192)
    devices = ([torch.device(d if d.startswith('c') else f'cuda:{d}')
                for d in args.devices.split(',')] if args.devices
               else [torch.device('cuda', i) for i in range(4)])
    pipe = GPipe(model, balance, devices=devices, chunks=chunks,
                 **({'cuda_graphs': True} if product and devices[0].type == 'cuda' else {}))
    if not use_dep:
        rp.depend = lambda fork_from, join_to: None         # no backward ordering edges
    if not use_streams:
        from torchgpipe_b200.runtime.streams import current_stream
        pipe._copy_streams = [[current_stream(d)] * chunks for d in pipe.devices]
    opt = SGD(pipe.parameters(), lr=0.1)
    x = torch.rand(batch, 3, hw, hw, device=pipe.devices[0])
    target = torch.ones(batch, 1, hw, hw, device=pipe.devices[-1])

    def step():
        loss = F.binary_cross_entropy_with_logits(pipe(x), target)
        loss.backward()
        opt.step()
        opt.zero_grad()

    cuda = [d for d in dict.fromkeys(pipe.devices) if d.type == 'cuda']
    for d in cuda:
        torch.cuda.reset_peak_memory_stats(d)
    with UtilSampler([d.index for d in cuda]) as util:
        ms = device_timed(list(dict.fromkeys(pipe.devices)), step, args.steps, args.warmup)
    emit({'benchmark': 'unet-timeline', 'experiment': args.experiment,
          'samples_per_s': batch / ms * 1e3, 'ms_per_step': ms,
          'gpu_utilization_pct': util.mean,
          'peak_reserved_gib': sum(torch.cuda.max_memory_reserved(d) for d in cuda) / 2 ** 30,
          'dependency': use_dep, 'streams': use_streams, 'portals': use_portals}, args.out)


if __name__ == '__main__':
    main()
