"""Fused NVLink/NVSwitch gradient all-reduce vs NCCL on the same bucket.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node R --master-addr 127.0.0.1 \
        --master-port 29520 benchmarks/allreduce_bench.py [--mib 64 256 1024]

For every size prints one JSON line (rank 0): time (CUDA events, max over ranks), achieved
bytes/s per direction per GPU and the fraction of the roofline
``max(flops / peak, wire_bytes / measured link bandwidth)`` - the add+scale is 2 flops per element, so the
link term always dominates. Wire bytes per direction per GPU: P2P 2N(R-1)/R (pulls in, pushes
out, and the same again for the peers' traffic), NCCL ring 2N(R-1)/R, multicast (NVLS)
N(R+1)/R: every GPU's whole bucket goes up to the switch once for the in-switch reduction (N
outbound), the reduced slice comes back (N/R inbound) and the multicast store delivers all R
slices to every GPU (N inbound).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from torchgpipe_b200.utils import link_bandwidth_gbs  # noqa: E402

LINK_GBS = link_bandwidth_gbs()      # measured by benchmarks/p2p_bandwidth.py


def timed(fn, iters, device, world):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    p = argparse.ArgumentParser()
    p.add_argument('--mib', type=int, nargs='+', default=[64, 256, 1024])
    p.add_argument('--iters', type=int, default=10)
    args = p.parse_args()
    rank = int(os.environ.get('LOCAL_RANK', 0))
    device = torch.device('cuda', rank)
    torch.cuda.set_device(device)
    dist.init_process_group('nccl')
    world = dist.get_world_size()
    from torchgpipe_b200.parallel import SymmetricBuffer, fused_allreduce_

    for mib in args.mib:
        n = mib * 1024 * 1024 // 4
        buf = SymmetricBuffer(n, device=device)
        plain = torch.randn(buf.numel, device=device)
        buf.tensor.copy_(plain)
        nbytes = buf.numel * 4
        rows = []
        ms = timed(lambda: (dist.all_reduce(plain), plain.mul_(1.0 / world)), args.iters, device,
                   world)
        rows.append(('nccl+scale', ms, 2 * nbytes * (world - 1) / world))
        ms = timed(lambda: fused_allreduce_(buf, 1.0 / world, multicast=False), args.iters,
                   device, world)
        rows.append(('fused_p2p', ms, 2 * nbytes * (world - 1) / world))
        os.environ['TGB_ALLREDUCE_PULL'] = '1'      # the earlier pull-only two-phase variant
        ms = timed(lambda: fused_allreduce_(buf, 1.0 / world, multicast=False), args.iters,
                   device, world)
        os.environ['TGB_ALLREDUCE_PULL'] = '0'
        rows.append(('fused_p2p_pull_only', ms, 2 * nbytes * (world - 1) / world))
        if buf.has_multicast:
            ms = timed(lambda: fused_allreduce_(buf, 1.0 / world, multicast=True), args.iters,
                       device, world)
            rows.append(('fused_nvls', ms, nbytes * (world + 1) / world))
        assert int(buf.error_flag.item()) == 0
        if dist.get_rank() == 0:
            for name, ms, wire in rows:
                ideal_ms = wire / (LINK_GBS * 1e9) * 1e3
                print(json.dumps({'op': name, 'mib': mib, 'replicas': world, 'ms': round(ms, 4),
                                  'wire_gbs_per_dir': round(wire / ms / 1e6, 1),
                                  'algbw_gbs': round(nbytes / ms / 1e6, 1),
                                  'roofline_ms': round(ideal_ms, 4), 'link_gbs': LINK_GBS,
                                  'fraction_of_roofline': round(ideal_ms / ms, 3)}), flush=True)
        del buf
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
