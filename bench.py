#!/usr/bin/env python
"""Headline benchmark: U-Net (5,64) training samples/s as an N-stage GPipe pipeline.

Metric and configs are the reference's ``benchmarks/unet-speed`` experiments
(``/root/reference/benchmarks/unet-speed/main.py:24-68``): ``pipeline-1`` (batch 80,
2 micro-batches, balance [241]), ``pipeline-2`` (512 / 32 / [104,137]), ``pipeline-4``
(512 / 16 / [30,66,84,61]), ``pipeline-8`` (640 / 40 / [16,27,31,44,22,57,27,17]);
input 3x192x192, target ones 1x192x192, BCE-with-logits, SGD(lr=0.1),
checkpoint='except_last', fp32 storage with TF32 tensor-core convolutions (what cuDNN
does for the reference under PyTorch defaults).

One step = forward + loss + backward + optimizer.step + zero_grad over one mini-batch.

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 4 --steps 5 --warmup 3
    python bench.py --impl reference ...   # the unmodified reference from baseline/_ref

``--impl ours`` (default): this repo. One rank per GPU; for N>1 the stages talk NCCL
point-to-point (``torchgpipe_b200.distributed``). ``--impl reference``: the reference
library installed under ``baseline/_ref`` driven through its own public API
(``torchgpipe.GPipe`` in one process spanning N GPUs - its only packaged multi-GPU mode;
under torchrun rank 0 drives it and the other ranks only join the barriers).

Rank 0 prints ONE JSON line. ``value`` is the U-Net experiment of that N; the AmoebaNet-D
(18,256) half of the metric (reference ``benchmarks/amoebanetd-speed/main.py:50-96``: n2m32,
n4m32, n8m32 - it defines no 1-GPU run) is measured right after it in a child process per rank
(own rendezvous port, so a failure there cannot take the headline down) and reported under
``secondary``. Both arms do the same.

Before anything is timed, the arm under test runs one eager step of the converted model and one
of the plain PyTorch model (same weights, dropout off) through the same pipeline and asserts that
the losses and the per-stage gradient norms agree (``numerics``).
"""
import argparse
import copy
import json
import os
import socket
import statistics
import subprocess
import sys
import threading
import time

EXPERIMENTS = {
    'unet': {
        1: dict(batch=80, chunks=2, balance=[241]),
        2: dict(batch=512, chunks=32, balance=[104, 137]),
        4: dict(batch=512, chunks=16, balance=[30, 66, 84, 61]),
        8: dict(batch=640, chunks=40, balance=[16, 27, 31, 44, 22, 57, 27, 17]),
    },
    # reference benchmarks/amoebanetd-speed/main.py:50-96 (n{2,4,8}m32); it defines no 1-GPU run
    'amoebanetd': {
        2: dict(batch=1280, chunks=32, balance=[9, 15]),
        4: dict(batch=1152, chunks=32, balance=[3, 6, 7, 8]),
        8: dict(batch=1280, chunks=32, balance=[2, 2, 2, 3, 3, 4, 4, 4]),
    },
    # reference benchmarks/resnet101-speed/main.py:30-67
    'resnet101': {
        1: dict(batch=220, chunks=2, balance=[370]),
        2: dict(batch=25000, chunks=1667, balance=[135, 235]),
        4: dict(batch=5632, chunks=256, balance=[44, 92, 124, 110]),
        8: dict(batch=5400, chunks=150, balance=[26, 22, 33, 44, 44, 66, 66, 69]),
    },
}
# Reference's published P40 numbers for the same experiments.
PUBLISHED = {'unet': {1: 24.456, 2: 35.502, 4: 67.042, 8: 88.497},
             'amoebanetd': {2: 47.386, 4: 72.412, 8: 132.413},
             'resnet101': {1: 81.796, 2: 135.539, 4: 265.958, 8: 411.662}}
INPUT_HW = 192
MODEL_INFO = {'unet': dict(hw=192, name='U-Net (B=5,C=64) 241 layers 232.7M params',
                           metric='U-Net(5,64)', loss='bce_with_logits'),
              'amoebanetd': dict(hw=224, name='AmoebaNet-D (L=18,D=256) 24 layers 122.5M params',
                                 metric='AmoebaNet-D(18,256)', loss='cross_entropy'),
              'resnet101': dict(hw=224, name='ResNet-101 370 layers 44.5M params',
                                metric='ResNet-101', loss='cross_entropy')}


def build_model(builders, name):
    if name == 'unet':
        return builders['unet'](depth=5, num_convs=5, base_channels=64, input_channels=3,
                                output_channels=1)
    if name == 'resnet101':
        return builders['resnet101'](num_classes=1000)
    return builders['amoebanetd'](num_classes=1000, num_layers=18, num_filters=256)


def make_target(torch, name, batch, hw, device):
    if name == 'unet':
        return torch.ones(batch, 1, hw, hw, device=device)
    g = torch.Generator().manual_seed(1)
    return torch.randint(1000, (batch,), generator=g).to(device)


def loss_mean(F, name, output, target):
    if name == 'unet':
        return F.binary_cross_entropy_with_logits(output, target)
    return F.cross_entropy(output, target)


def loss_sum(F, name, output, target):
    """Sum-reduced loss of one micro-batch (divide by the mini-batch normaliser outside)."""
    if name == 'unet':
        return F.binary_cross_entropy_with_logits(output, target, reduction='sum')
    return F.cross_entropy(output, target, reduction='sum')


# --------------------------------------------------------------------------- helpers
class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index=0):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', f'--query-gpu={self.QUERY}', '--format=csv,noheader,nounits',
                 '-lms', '200', '-i', str(self.gpu_index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (OSError, ValueError):
            self.proc = None
            return

        def pump():
            assert self.proc is not None and self.proc.stdout is not None
            for line in self.proc.stdout:
                self.lines.append(line.strip())
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, sm_max, power = [], [], []
        reasons = set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.lines:
            parts = [p.strip() for p in line.split(',')]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                sm_max.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, flag in zip(names, parts[5:9]):
                if flag.lower().startswith('active'):
                    reasons.add(name)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['no samples']}
        # under load = upper half of the samples by power draw
        order = sorted(range(len(sm)), key=lambda k: power[k])
        loaded = [sm[k] for k in order[len(order) // 2:]] or sm
        return {'sm_mhz': statistics.median(loaded), 'sm_max_mhz': max(sm_max),
                'power_w_max': max(power), 'samples': len(sm), 'reasons': sorted(reasons)}


def dist_env():
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', str(rank)))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    return rank, local_rank, world


class L2Flusher:
    """Overwrites a buffer larger than the 126 MB L2 so no step starts with warm lines."""

    def __init__(self, torch, device, mib=256):
        self.buf = torch.empty(mib * 1024 * 1024 // 4, dtype=torch.float32, device=device)

    def __call__(self):
        self.buf.zero_()


def timed_loop(torch, dist, step_fn, steps, warmup, devices, flush, world, reduce_device):
    """W untimed steps, then exactly K steps between barrier + synchronize on both sides.

    Device time: one CUDA-event pair per participating device of this process (this arm's rank
    owns one; the reference arm's driver rank owns all N). The end event of a device is recorded
    once every stream of that device has drained, so work on side / communication streams is
    inside the interval; the result is the longest interval, then the MAX over ranks. Both arms
    are timed by this one function."""
    def sync_all():
        for d in devices:
            torch.cuda.synchronize(d)

    for _ in range(warmup):
        step_fn()
        flush()
    sync_all()
    if world > 1:
        dist.barrier()
    sync_all()
    starts = [torch.cuda.Event(enable_timing=True) for _ in devices]
    ends = [torch.cuda.Event(enable_timing=True) for _ in devices]
    wall0 = time.perf_counter()
    for d, e in zip(devices, starts):
        e.record(torch.cuda.current_stream(d))
    for _ in range(steps):
        step_fn()
        flush()
    for d, e in zip(devices, ends):
        torch.cuda.synchronize(d)
        e.record(torch.cuda.current_stream(d))
    sync_all()
    wall = (time.perf_counter() - wall0) * 1e3
    if world > 1:
        dist.barrier()
    ms = max([s.elapsed_time(e) for s, e in zip(starts, ends)] or [0.0])
    if world > 1:
        t = torch.tensor([ms, wall], device=reduce_device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, wall = t.tolist()
    return ms, wall


def resolve_balance(args, cfg, torch, dist, model, device, world, rank):
    """The published balance (default, the headline), a comma list, or ``auto``: rank 0 times the
    layers of the model it is about to run (``balance_by_time`` on its GPU, one micro-batch) and
    every rank uses that split. A non-default balance is recorded in the JSON line
    (``config.balance_source``); it is a second row next to the headline, never the headline."""
    spec = args.balance
    if spec in (None, '', 'reference'):
        return list(cfg['balance']), 'reference'
    n = len(cfg['balance'])
    if spec != 'auto':
        balance = [int(v) for v in spec.split(',')]
        assert len(balance) == n and sum(balance) == sum(cfg['balance']), \
            f'--balance needs {n} entries summing to {sum(cfg["balance"])}'
        return balance, 'given'
    box = [None]
    if rank == 0:
        from torchgpipe_b200.balance import balance_by_time
        hw = MODEL_INFO[args.model]['hw']
        sample = torch.rand(max(cfg['batch'] // cfg['chunks'], 1), 3, hw, hw)
        box[0] = [int(v) for v in balance_by_time(n, model, sample, timeout=2.0, device=device,
                                                 solver='optimal')]
        torch.cuda.empty_cache()
    if world > 1:
        dist.broadcast_object_list(box, src=0)
    return box[0], 'balance_by_time on the sm_100a operators (this run)'


def metric_name(model):
    """Constant across N (the experiment name pipeline-N lives in config)."""
    return '%s GPipe training throughput' % MODEL_INFO[model]['metric']


def free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def run_secondary(args, torch, dist, rank, local_rank, world, device):
    """AmoebaNet-D(18,256) n{N}m32 of the same arm in a child process per rank.

    Called after the headline measurement is finished and its memory released. The children form
    their own process group on a fresh port; rank 0's child prints the JSON line this returns.
    Any failure (exception, crash, timeout) becomes {'error': ...} - the headline is safe."""
    n = args.gpus
    if args.no_secondary or args.child or args.model != 'unet':
        return None
    if n not in EXPERIMENTS['amoebanetd']:
        return {'skipped': 'the reference defines no %d-GPU AmoebaNet-D speed experiment '
                           '(benchmarks/amoebanetd-speed/main.py: n2, n4, n8 only)' % n}
    port = [free_port()]
    if world > 1:
        dist.broadcast_object_list(port, src=0)
    env = dict(os.environ)
    env.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port[0]), RANK=str(rank),
               LOCAL_RANK=str(local_rank), WORLD_SIZE=str(world))
    for key in ('TORCHELASTIC_RUN_ID', 'TORCHELASTIC_USE_AGENT_STORE', 'GROUP_RANK',
                'ROLE_RANK', 'TORCHELASTIC_MAX_RESTARTS', 'TORCHELASTIC_RESTART_COUNT'):
        env.pop(key, None)
    cmd = [sys.executable, os.path.abspath(__file__), '--child', '--model', 'amoebanetd',
           '--impl', args.impl, '--gpus', str(n), '--steps', str(min(args.steps, 5)),
           '--warmup', str(max(3, min(args.warmup, 3)))]
    if args.no_graphs:
        cmd.append('--no-graphs')
    if args.no_kernels:
        cmd.append('--no-kernels')
    try:
        proc = subprocess.run(cmd, env=env, capture_output=True, text=True,
                              timeout=args.secondary_timeout)
        out, err, rc = proc.stdout, proc.stderr, proc.returncode
    except subprocess.TimeoutExpired as exc:
        out = exc.stdout.decode() if isinstance(exc.stdout, bytes) else (exc.stdout or '')
        err, rc = 'timed out after %d s' % args.secondary_timeout, -1
    if rank != 0:
        return None
    for line in reversed(out.splitlines()):
        line = line.strip()
        if line.startswith('{'):
            try:
                return json.loads(line)
            except ValueError:
                continue
    return {'error': 'child exited %d: %s' % (rc, (err or '')[-600:])}


# --------------------------------------------------------------------------- our arm
def set_dropout(model, p=None, saved=None):
    """p given: turn every (fused or plain) channel dropout off, returning what to restore;
    saved given: restore."""
    if saved is not None:
        for m, value in saved:
            m.p = value
        return None
    saved = []
    for m in model.modules():
        if hasattr(m, 'p') and isinstance(getattr(m, 'p'), float) and m.p != p:
            saved.append((m, m.p))
            m.p = p
    return saved


def check_numerics(torch, dist, F, args, make_pipe, model, plain, host_input, target, cfg,
                   rank, world, device):
    """One eager training step of the converted model and one of the plain PyTorch model (same
    weights, dropout off, same pipeline layout) - every rank compares what it owns.

    Returns {'loss_rel_diff', 'grad_norm_rel_diff', ...}; raises if they disagree beyond what
    TF32 convolutions explain, so a run whose kernels are wrong on some rank never reports a
    throughput."""
    first, last = rank == 0, rank == world - 1
    chunks, batch = cfg['chunks'], cfg['batch']
    hw = MODEL_INFO[args.model]['hw']
    total = float(batch * hw * hw) if args.model == 'unet' else float(batch)
    # a slice of the mini-batch is enough: two micro-batches' worth per chunk count would change
    # the micro-batch shapes, so keep the full shape but only one step
    results = []
    for tag, net in (('plain', plain), ('sm100', model)):
        saved = set_dropout(net, 0.0)
        pipe = make_pipe(net, graphs=False, name='check-' + tag)
        stage = pipe.model()
        stage.train()
        dev_in = host_input.to(device) if first else None
        outs = pipe.forward(dev_in)
        if last:
            losses = [loss_sum(F, args.model, o, t) / total
                      for o, t in zip(outs, target.chunk(chunks))]
            pipe.backward(losses)
            loss = float(torch.stack([x.detach() for x in losses]).sum().item())
        else:
            pipe.backward(None)
            loss = 0.0
        sq = torch.zeros((), device=device, dtype=torch.float64)
        for prm in stage.parameters():
            if prm.grad is not None:
                sq += prm.grad.double().pow(2).sum()
                prm.grad = None
        results.append((loss, float(sq.sqrt().item())))
        set_dropout(net, saved=saved)
        del pipe, outs
    (loss_a, g_a), (loss_b, g_b) = results
    mine = [abs(loss_a - loss_b) / max(abs(loss_a), 1e-12) if last else 0.0,
            abs(g_a - g_b) / max(g_a, 1e-12)]
    if world > 1:
        t = torch.tensor(mine, device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        mine = t.tolist()
        lt = torch.tensor([loss_a, loss_b], device=device, dtype=torch.float64)
        dist.all_reduce(lt)                 # only the last rank contributes non-zeros
        loss_a, loss_b = lt.tolist()
    report = {'loss_plain_torch': loss_a, 'loss_sm100': loss_b, 'loss_rel_diff': mine[0],
              'grad_norm_rel_diff_max_over_ranks': mine[1],
              'tolerance': {'loss': 1e-2, 'grad_norm': 1e-1},
              'how': 'one eager step each, same weights and input, dropout off, before timing'}
    if not (mine[0] < 1e-2 and mine[1] < 1e-1):
        raise AssertionError('numerics check failed: %s' % json.dumps(report))
    return report


def check_numerics_single(torch, F, args, GPipe, model, plain, host_input, target, cfg, devices):
    """check_numerics for the single-process engine (eager GPipe for both models)."""
    results = []
    x = host_input.to(devices[0])
    for net in (plain, model):
        saved = set_dropout(net, 0.0)
        pipe = GPipe(net, cfg['balance'], devices=devices, chunks=cfg['chunks'],
                     checkpoint='except_last')
        pipe.train()
        loss = loss_mean(F, args.model, pipe(x), target)
        loss.backward()
        sq = sum(float(p.grad.double().pow(2).sum().item()) for p in pipe.parameters()
                 if p.grad is not None)
        for p in pipe.parameters():
            p.grad = None
        results.append((float(loss.item()), sq ** 0.5))
        set_dropout(net, saved=saved)
        del pipe, loss
    (loss_a, g_a), (loss_b, g_b) = results
    report = {'loss_plain_torch': loss_a, 'loss_sm100': loss_b,
              'loss_rel_diff': abs(loss_a - loss_b) / max(abs(loss_a), 1e-12),
              'grad_norm_rel_diff_max_over_ranks': abs(g_a - g_b) / max(g_a, 1e-12),
              'tolerance': {'loss': 1e-2, 'grad_norm': 1e-1},
              'how': 'one eager step each, same weights and input, dropout off, before timing'}
    if not (report['loss_rel_diff'] < 1e-2 and report['grad_norm_rel_diff_max_over_ranks'] < 1e-1):
        raise AssertionError('numerics check failed: %s' % json.dumps(report))
    return report


def run_ours(args):
    import gc
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from torch.optim import SGD

    rank, local_rank, world = dist_env()
    n = args.gpus
    cfg = EXPERIMENTS[args.model][n]
    device = torch.device('cuda', local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        # Lazy initialisation on purpose: point-to-point traffic then gets one NCCL communicator
        # (and stream) per rank pair instead of being serialised through a single one.
        dist.init_process_group('nccl')
    assert world in (1, n), f'launch with --nproc-per-node {n} (got WORLD_SIZE={world})'

    from torchgpipe_b200 import GPipe
    from torchgpipe_b200 import ops
    from torchgpipe_b200.models import amoebanetd, resnet101, unet
    info = MODEL_INFO[args.model]
    hw = info['hw']

    torch.manual_seed(0)
    model = build_model(dict(unet=unet, amoebanetd=amoebanetd, resnet101=resnet101), args.model)
    plain = None
    if not args.no_kernels:
        if not args.no_check:
            plain = copy.deepcopy(model)
        model = ops.convert_to_sm100(model)

    batch, chunks = cfg['batch'], cfg['chunks']
    balance, balance_source = resolve_balance(args, cfg, torch, dist, model, device, world, rank)
    cfg = dict(cfg, balance=balance)
    host_input = torch.rand(batch, 3, hw, hw).pin_memory()
    loss_host = torch.zeros(1).pin_memory()

    first, last = rank == 0, rank == world - 1
    # N > 1: one rank per stage (DistributedGPipe over NCCL p2p). One process (N = 1, or all N
    # GPUs driven by a single process when launched without torchrun): the public GPipe class,
    # with every stage pass replayed from CUDA graphs unless --no-graphs.
    stage_mode = world > 1
    numerics = None
    target = None
    if not stage_mode:
        pipe_devices = [device] if n == 1 else [torch.device('cuda', k) for k in range(n)]
        target = make_target(torch, args.model, batch, hw, pipe_devices[-1])
        if plain is not None:
            numerics = check_numerics_single(torch, F, args, GPipe, model, plain, host_input,
                                             target, cfg, pipe_devices)
            del plain
            gc.collect()
            torch.cuda.empty_cache()
        pipe = GPipe(model, balance, devices=pipe_devices, chunks=chunks,
                     checkpoint='except_last', cuda_graphs=not args.no_graphs)
        params = list(pipe.parameters())
        dev_input = host_input.to(device)

        def step(inp=None):
            output = pipe(dev_input if inp is None else inp)
            loss = loss_mean(F, args.model, output, target)
            loss.backward()
            opt.step()
            opt.zero_grad()
            return loss
    else:
        from torchgpipe_b200.distributed import DistributedGPipe, ProcessGroupTransport, context
        entered = []

        def make_pipe(net, graphs, name='stage'):
            workers = {r: f'{name}{r}' for r in range(world)}
            if world > 1:
                tp = ProcessGroupTransport(device)
            else:
                tp = None                               # in-process channels; nothing is sent
                ctx = context.worker(workers[0], chunks)
                ctx.__enter__()
                entered.append(ctx)
            return DistributedGPipe(net, rank, workers, balance, chunks, device=device,
                                    checkpoint='except_last', transport=tp, cuda_graphs=graphs)

        target = make_target(torch, args.model, batch, hw, device) if last else None
        if plain is not None:
            numerics = check_numerics(torch, dist, F, args, make_pipe, model, plain, host_input,
                                      target, cfg, rank, world, device)
            del plain
            gc.collect()
            torch.cuda.empty_cache()
        pipe = make_pipe(model, graphs=not args.no_graphs)
        params = list(pipe.model().parameters())
        dev_input = host_input.to(device) if first else None
        # normaliser of the mini-batch mean loss: elements for BCE, samples for cross-entropy
        total = float(batch * hw * hw) if args.model == 'unet' else float(batch)

        def step(inp=None):
            outs = pipe.forward((dev_input if inp is None else inp) if first else None)
            if last:
                losses = [loss_sum(F, args.model, o, t) / total
                          for o, t in zip(outs, target.chunk(chunks))]
                pipe.backward(losses)
                loss = torch.stack([x.detach() for x in losses]).sum()
            else:
                pipe.backward(None)
                loss = None
            opt.step()
            opt.zero_grad()
            return loss

    opt = SGD(params, lr=0.1)
    flush = L2Flusher(torch, device)
    my_devices = [device] if stage_mode else list(dict.fromkeys(pipe.devices))

    sampler = ClockSampler(local_rank) if rank == 0 else None
    ops.reset_launch_count()

    # device-timed, inputs resident on the device
    for _ in range(args.warmup):
        step()
        flush()

    def total_launches():
        replayed = pipe.native_launches_replayed() if hasattr(pipe, 'native_launches_replayed') \
            else 0
        return ops.launch_count() + replayed

    if sampler:
        sampler.start()
    launches_before = total_launches()
    ms, _ = timed_loop(torch, dist, lambda: step(), args.steps, 0, my_devices, flush, world, device)
    launches = total_launches() - launches_before
    clocks = sampler.stop() if sampler else None

    # end to end: H2D of the step's input from pinned memory + D2H of the loss, every step
    stage_in = torch.empty_like(dev_input) if first else None

    def e2e_step():
        if first:
            stage_in.copy_(host_input, non_blocking=True)
        loss = step(stage_in)
        if loss is not None:
            loss_host.copy_(loss.detach().reshape(1), non_blocking=False)

    _, e2e_wall = timed_loop(torch, dist, e2e_step, args.steps, 1, my_devices, flush, world,
                             device)

    pipeline_report = None
    if world > 1:
        t = torch.tensor([float(launches)], device=device, dtype=torch.float64)
        dist.all_reduce(t)
        launches = int(t.item())
        # one extra (untimed) step with per-pass CUDA events: bubble and exposed communication
        pipe.timeline = True
        step()
        mine = pipe.timeline_report()
        pipe.timeline = False
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        pipeline_report = {'ideal_bubble_frac': (n - 1) / (chunks + n - 1),
                           'per_rank': [{k: round(v, 3) for k, v in (g or {}).items()}
                                        for g in gathered]}

    samples = batch * args.steps
    value = samples / (ms / 1e3)
    e2e_value = samples / (e2e_wall / 1e3)
    result = {
        'metric': metric_name(args.model),
        'impl': 'ours',
        'value': value, 'unit': 'samples/s', 'n_gpus': n, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': ms / args.steps, 'higher_is_better': True,
        'scaling': 'weak',
        # the only published numbers are the reference authors' Tesla P40
        # runs; the same-box reference arm (--impl reference) is the comparison that matters
        'vs_baseline': value / PUBLISHED[args.model][n],
        'baseline_hardware': 'Tesla P40 (published); same-box B200 ratio: run --impl reference',
        'dtype': 'fp32 storage / tf32 tensor-core conv', 'data': 'synthetic',
        'config': {'model': MODEL_INFO[args.model]['name'],
                   'experiment': 'pipeline-%d' % n,
                   'global_batch': batch, 'chunks': chunks, 'balance': balance,
                   'balance_source': balance_source,
                   'input': [3, MODEL_INFO[args.model]['hw'], MODEL_INFO[args.model]['hw']],
                   'checkpoint': 'except_last',
                   'optimizer': 'SGD(lr=0.1)', 'loss': MODEL_INFO[args.model]['loss'],
                   'parallelism': 'pp%d' % n,
                   'scaling_note': 'batch and micro-batch count per N are the reference\'s own '
                                   'experiment definitions (80/512/512/640 for U-Net)',
                   'transport': 'single-process' if world == 1 else 'nccl-p2p',
                   'engine': 'DistributedGPipe (one rank per stage)' if stage_mode
                   else 'GPipe (single process, public API)',
                   'l2': 'flushed between steps (256 MiB write)',
                   'kernels': 'off' if args.no_kernels else 'sm100',
                   'cuda_graphs': not args.no_graphs},
        'clocks': clocks,
        'e2e': {'value': e2e_value, 'unit': 'samples/s',
                'h2d_bytes_per_step': host_input.numel() * 4, 'd2h_bytes_per_step': 4,
                'ms_per_step': e2e_wall / args.steps},
        'gpu_launches': launches,
        'numerics': numerics,
        'pipeline': pipeline_report,
    }
    # release everything before the secondary model runs in its own processes
    del pipe, opt, params, step, e2e_step, model, flush, dev_input, stage_in, target
    gc.collect()
    torch.cuda.empty_cache()
    secondary = run_secondary(args, torch, dist, rank, local_rank, world, device)
    if secondary is not None:
        result['secondary'] = secondary
    if rank == 0:
        print(json.dumps(result), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------- reference arm
def ensure_reference(here):
    """baseline/_ref is staged by __graft_entry__.build() on the build box and travels with the
    working tree; if it is missing, try the same staging script here (works wherever
    /root/reference and /opt/wheelhouse exist). Returns (ref_root or None, reason)."""
    ref_root = os.path.join(here, 'baseline', '_ref')
    marker = os.path.join(ref_root, 'torchgpipe', '__init__.py')
    if os.path.exists(marker):
        return ref_root, ''
    if not os.path.isdir('/root/reference'):
        return None, ('baseline/_ref is missing and /root/reference is not on this machine '
                      '(python -c "import __graft_entry__ as g; g.build()" stages it)')
    rank, _, world = dist_env()
    if rank == 0:
        proc = subprocess.run(['bash', os.path.join(here, 'baseline', 'stage_reference.sh')],
                              capture_output=True, text=True)
        why = (proc.stdout + proc.stderr)[-300:].replace('\n', ' ')
    else:
        why = ''
        for _ in range(600):                # rank 0 is staging
            if os.path.exists(marker):
                break
            time.sleep(0.5)
    if os.path.exists(marker):
        return ref_root, ''
    return None, 'baseline/stage_reference.sh failed: ' + why


def run_reference(args):
    rank, local_rank, world = dist_env()
    here = os.path.dirname(os.path.abspath(__file__))
    ref_root, why = ensure_reference(here)
    if ref_root is None:
        if rank == 0:
            print(json.dumps({'impl': 'reference', 'unavailable': why}), flush=True)
        return
    # Reference first on the path; this repo's package is never imported on this arm.
    sys.path.insert(0, os.path.join(ref_root, 'benchmarks_models'))
    sys.path.insert(0, ref_root)

    import gc
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from torch.optim import SGD
    import torchgpipe
    from torchgpipe import GPipe
    assert os.path.realpath(torchgpipe.__file__).startswith(os.path.realpath(ref_root))
    from unet import unet  # reference model builders (benchmarks/models/*)
    from amoebanet import amoebanetd
    from resnet import resnet101
    hw = MODEL_INFO[args.model]['hw']

    n = args.gpus
    cfg = EXPERIMENTS[args.model][n]
    device = torch.device('cuda', local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group('nccl', device_id=device)

    batch, chunks, balance = cfg['batch'], cfg['chunks'], cfg['balance']
    balance_source = 'reference'
    if args.balance not in (None, '', 'reference'):
        assert args.balance != 'auto', 'give the reference arm the list the other arm printed'
        balance = [int(v) for v in args.balance.split(',')]
        assert len(balance) == n and sum(balance) == sum(cfg['balance'])
        balance_source = 'given'
    driver = rank == 0
    flush = L2Flusher(torch, device)
    host_input = torch.rand(batch, 3, hw, hw).pin_memory()
    loss_host = torch.zeros(1).pin_memory()

    devices = []
    if driver:
        torch.manual_seed(0)
        model = build_model(dict(unet=unet, amoebanetd=amoebanetd, resnet101=resnet101),
                            args.model)
        devices = [torch.device('cuda', k) for k in range(n)]
        pipe = GPipe(model, balance, devices=devices, chunks=chunks)   # checkpoint default
        opt = SGD(pipe.parameters(), lr=0.1)
        dev_input = host_input.to(devices[0])
        target = make_target(torch, args.model, batch, hw, devices[-1])
        stage_in = torch.empty_like(dev_input)

        def step(inp=None):
            output = pipe(dev_input if inp is None else inp)
            loss = loss_mean(F, args.model, output, target)
            loss.backward()
            opt.step()
            opt.zero_grad()
            return loss
    else:
        def step(inp=None):
            return None

    sampler = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(args.warmup):
        step()
        flush()
    if sampler:
        sampler.start()
    ms, _ = timed_loop(torch, dist, lambda: step(), args.steps, 0, devices, flush, world, device)
    clocks = sampler.stop() if sampler else None

    def e2e_step():
        if driver:
            stage_in.copy_(host_input, non_blocking=True)
            loss = step(stage_in)
            loss_host.copy_(loss.detach().reshape(1).to('cpu'))

    _, e2e_wall = timed_loop(torch, dist, e2e_step, args.steps, 1, devices, flush, world, device)

    samples = batch * args.steps
    value = samples / (ms / 1e3)
    result = {
        'metric': metric_name(args.model),
        'impl': 'reference',
        'value': value, 'unit': 'samples/s', 'n_gpus': n, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': ms / args.steps, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': value / PUBLISHED[args.model][n],
        'baseline_hardware': 'Tesla P40 (published)',
        'dtype': 'fp32 storage / tf32 cuDNN conv (PyTorch defaults)', 'data': 'synthetic',
        'config': {'model': MODEL_INFO[args.model]['name'],
                   'experiment': 'pipeline-%d' % n,
                   'global_batch': batch, 'chunks': chunks, 'balance': balance,
                   'balance_source': balance_source,
                   'input': [3, MODEL_INFO[args.model]['hw'], MODEL_INFO[args.model]['hw']],
                   'checkpoint': 'except_last',
                   'optimizer': 'SGD(lr=0.1)', 'loss': MODEL_INFO[args.model]['loss'],
                   'parallelism': 'pp%d' % n, 'transport': 'single-process (reference GPipe)',
                   'l2': 'flushed between steps (256 MiB write)',
                   'torchgpipe': torchgpipe.__version__},
        'clocks': clocks,
        'e2e': {'value': samples / (e2e_wall / 1e3), 'unit': 'samples/s',
                'h2d_bytes_per_step': host_input.numel() * 4, 'd2h_bytes_per_step': 4,
                'ms_per_step': e2e_wall / args.steps},
        'gpu_launches': 0,
    }
    if driver:
        del pipe, opt, model, dev_input, stage_in, target, step, e2e_step
    del flush
    gc.collect()
    torch.cuda.empty_cache()
    secondary = run_secondary(args, torch, dist, rank, local_rank, world, device)
    if secondary is not None:
        result['secondary'] = secondary
    if rank == 0:
        print(json.dumps(result), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    p = argparse.ArgumentParser()
    p.add_argument('--gpus', type=int, default=1)
    p.add_argument('--steps', type=int, default=5)
    p.add_argument('--warmup', type=int, default=3)
    p.add_argument('--impl', choices=['ours', 'reference'], default='ours')
    p.add_argument('--model', choices=['unet', 'amoebanetd', 'resnet101'], default='unet',
                   help='unet (default, headline), amoebanetd (N in 2,4,8) or resnet101')
    p.add_argument('--no-graphs', action='store_true',
                   help='ours arm, N>1: issue stage kernels eagerly instead of replaying CUDA graphs')
    p.add_argument('--no-kernels', action='store_true',
                   help='ours arm: keep torch/cuDNN ops (diagnostic; not the product path)')
    p.add_argument('--no-check', action='store_true',
                   help='ours arm: skip the plain-PyTorch numerics comparison before timing')
    p.add_argument('--no-secondary', action='store_true',
                   help='do not measure AmoebaNet-D(18,256) n{N}m32 after the U-Net headline')
    p.add_argument('--secondary-timeout', type=int, default=400,
                   help='seconds; keeps headline + secondary inside the per-N limit of the driver')
    p.add_argument('--balance', default='reference',
                   help="'reference' (the published balance, default), 'auto' (ours arm: "
                        "balance_by_time on the operators that run) or a comma list")
    p.add_argument('--child', action='store_true', help=argparse.SUPPRESS)
    args = p.parse_args()
    if args.gpus not in EXPERIMENTS[args.model]:
        raise SystemExit('--gpus must be one of %s' % sorted(EXPERIMENTS[args.model]))
    if args.impl == 'reference':
        try:
            run_reference(args)
        except ImportError as exc:
            print(json.dumps({'impl': 'reference', 'unavailable': f'import failed: {exc}'}))
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
