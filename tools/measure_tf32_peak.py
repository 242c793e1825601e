"""Measures the TF32 tensor-core roof the convolution kernels are judged against: cuBLAS
``torch.matmul`` of two fp32 8192^2 matrices with TF32 allowed (2*N^3 FLOP), best of 10 (burst)
and back to back for 3 s (sustained), CUDA events. Writes profiles/measured_tf32_peak.json.

    python tools/measure_tf32_peak.py [--out profiles/measured_tf32_peak.json]
"""
import argparse
import json
import time

import torch


def main():
    p = argparse.ArgumentParser()
    p.add_argument('--out', default=None)
    p.add_argument('--n', type=int, default=8192)
    args = p.parse_args()
    torch.backends.cuda.matmul.allow_tf32 = True
    n = args.n
    a = torch.randn(n, n, device='cuda')
    b = torch.randn(n, n, device='cuda')
    flop = 2.0 * n ** 3
    for _ in range(3):
        a @ b
    torch.cuda.synchronize()
    best = float('inf')
    for _ in range(10):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        a @ b
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters, t0 = 0, time.time()
    s.record()
    while time.time() - t0 < 3.0:
        for _ in range(20):
            a @ b
        iters += 20
        torch.cuda.synchronize()
    e.record()
    torch.cuda.synchronize()
    sustained = s.elapsed_time(e) / iters
    rec = {'gpu': torch.cuda.get_device_name(0), 'n': n,
           'tf32_tflops_burst': round(flop / best / 1e9, 1),
           'tf32_tflops_sustained': round(flop / sustained / 1e9, 1),
           'how': 'torch.matmul fp32 with allow_tf32 (cuBLAS), 2*N^3 FLOP, CUDA events; burst = '
                  'best of 10, sustained = back to back for 3 s'}
    print(json.dumps(rec), flush=True)
    if args.out:
        with open(args.out, 'w') as f:
            json.dump(rec, f, indent=1)


if __name__ == '__main__':
    main()
