"""NVLink / NVSwitch peer-copy bandwidth between the GPUs of this box.

    python benchmarks/p2p_bandwidth.py [--mib 1024] [--out profiles/p2p_bandwidth.json]

One process, every visible GPU. Measures with CUDA events (best of 5 after 2 warm-ups):

* ``unidirectional_gbs``  - GPU 0 -> GPU 1, one copy of ``--mib`` MiB (copy engine);
* ``bidirectional_gbs``   - 0 -> 1 and 1 -> 0 at the same time, per direction;
* ``fan_in_gbs``          - every other GPU -> GPU 0 at the same time (inbound total at GPU 0);
* ``all_pairs``           - the unidirectional figure for every ordered pair (uniformity through
  the switch).

``unidirectional_gbs`` is the denominator of every all-reduce roofline and of the pipeline's
``comm_ms_bound`` (``benchmarks/allreduce_bench.py``, ``DistributedGPipe.timeline_report``).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch


def copy_ms(pairs, nbytes, iters=5):
    """pairs: list of (src_dev, dst_dev); all copies of one round run concurrently."""
    bufs = []
    for s, d in pairs:
        src = torch.empty(nbytes, dtype=torch.uint8, device=f'cuda:{s}')
        dst = torch.empty(nbytes, dtype=torch.uint8, device=f'cuda:{d}')
        stream = torch.cuda.Stream(device=f'cuda:{s}')
        bufs.append((src, dst, stream))
    best = float('inf')
    for it in range(iters + 2):
        for dev in range(torch.cuda.device_count()):
            torch.cuda.synchronize(dev)
        events = []
        for src, dst, stream in bufs:
            with torch.cuda.device(src.device), torch.cuda.stream(stream):
                a = torch.cuda.Event(enable_timing=True)
                b = torch.cuda.Event(enable_timing=True)
                a.record(stream)
                dst.copy_(src, non_blocking=True)
                b.record(stream)
                events.append((a, b))
        for dev in range(torch.cuda.device_count()):
            torch.cuda.synchronize(dev)
        ms = max(a.elapsed_time(b) for a, b in events)
        if it >= 2:
            best = min(best, ms)
    return best


def main():
    p = argparse.ArgumentParser()
    p.add_argument('--mib', type=int, default=1024)
    p.add_argument('--out', default=None)
    args = p.parse_args()
    n = torch.cuda.device_count()
    if n < 2:
        raise SystemExit('needs at least 2 GPUs')
    nbytes = args.mib << 20
    gbs = lambda ms, copies=1: copies * nbytes / ms / 1e6      # noqa: E731
    uni = gbs(copy_ms([(0, 1)], nbytes))
    bidi = gbs(copy_ms([(0, 1), (1, 0)], nbytes))
    fan_pairs = [(s, 0) for s in range(1, n)]
    fan = gbs(copy_ms(fan_pairs, nbytes), len(fan_pairs))
    pairs = {}
    if n <= 4:
        for s in range(n):
            for d in range(n):
                if s != d:
                    pairs[f'{s}->{d}'] = round(gbs(copy_ms([(s, d)], nbytes, iters=3)), 1)
    can = all(torch.cuda.can_device_access_peer(0, d) for d in range(1, n))
    record = {'gpus': n, 'gpu': torch.cuda.get_device_name(0), 'mib': args.mib,
              'peer_access': can, 'unidirectional_gbs': round(uni, 1),
              'bidirectional_gbs_per_direction': round(bidi, 1),
              'fan_in_gbs_total_at_gpu0': round(fan, 1), 'all_pairs': pairs,
              'how': 'tensor.copy_ between devices on a side stream, CUDA events, best of 5'}
    print(json.dumps(record), flush=True)
    if args.out:
        with open(args.out, 'w') as f:
            json.dump(record, f, indent=1)


if __name__ == '__main__':
    main()
