"""Times the native convolution paths (forward, data gradient, weight gradient) against
PyTorch/cuDNN (TF32 on, the reference's arithmetic) for a list of layer shapes.

    python tools/conv_microbench.py [resnet|amoeba|unet16|stems] > gpurun_out/conv_microbench.txt

CUDA events, 3 warm-up + 10 timed launches, a 256 MiB L2 flush between launches.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.nn.functional as F

from torchgpipe_b200.ops import functional as TF

CL = torch.channels_last
SHAPES = {
    # name: N, C, H, W, K, (R, S), stride, pad
    'resnet': [
        ('l2.conv2 3x3/2', 110, 128, 56, 56, 128, (3, 3), (2, 2), (1, 1)),
        ('l3.conv2 3x3/2', 110, 256, 28, 28, 256, (3, 3), (2, 2), (1, 1)),
        ('l4.conv2 3x3/2', 110, 512, 14, 14, 512, (3, 3), (2, 2), (1, 1)),
        ('l2.down 1x1/2', 110, 256, 56, 56, 512, (1, 1), (2, 2), (0, 0)),
        ('l3.down 1x1/2', 110, 512, 28, 28, 1024, (1, 1), (2, 2), (0, 0)),
        ('l4.down 1x1/2', 110, 1024, 14, 14, 2048, (1, 1), (2, 2), (0, 0)),
        ('l3.conv2 3x3/1', 110, 256, 14, 14, 256, (3, 3), (1, 1), (1, 1)),
    ],
    'amoeba': [
        ('1x7 s(1,2)', 40, 64, 56, 56, 64, (1, 7), (1, 2), (0, 3)),
        ('7x1 s(2,1)', 40, 64, 56, 28, 64, (7, 1), (2, 1), (3, 0)),
        ('3x3/2', 40, 128, 28, 28, 128, (3, 3), (2, 2), (1, 1)),
        ('1x1/2', 40, 512, 28, 28, 512, (1, 1), (2, 2), (0, 0)),
        ('1x7 s1', 40, 128, 28, 28, 128, (1, 7), (1, 1), (0, 3)),
        ('1x1 s1', 40, 1024, 14, 14, 256, (1, 1), (1, 1), (0, 0)),
    ],
}


def timed(fn, flush, iters=10):
    for _ in range(3):
        fn()
    times = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    times.sort()
    return times[len(times) // 2]


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else 'resnet'
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cudnn.benchmark = True
    flush = torch.empty(64 << 20, device='cuda')
    print(f'{"layer":<18}{"GFLOP":>8} | {"fwd+stats":>9}{"fwd":>8}{"cudnn":>8} | {"dgrad ours":>11}'
          f'{"cudnn":>8} | {"wgrad ours":>11}{"cudnn":>8}   (ms)')
    for name, N, C, H, W, K, (R, S), stride, pad in SHAPES[which]:
        x = torch.randn(N, C, H, W, device='cuda').contiguous(memory_format=CL)
        w = (torch.randn(K, C, R, S, device='cuda') / (C * R * S) ** 0.5).contiguous(
            memory_format=CL)
        y = F.conv2d(x, w, stride=stride, padding=pad)
        dy = torch.randn_like(y)
        cache = {}
        gflop = 2.0 * y.numel() * C * R * S / 1e9
        f_o = timed(lambda: TF.conv2d_forward(x, w, pad, 2, stride), flush)
        f_n = timed(lambda: TF.conv2d_forward(x, w, pad, 0, stride), flush)
        f_c = timed(lambda: F.conv2d(x, w, stride=stride, padding=pad), flush)
        d_o = timed(lambda: TF.conv2d_dgrad(dy, w, x.shape, pad, stride, cache), flush)
        d_c = timed(lambda: torch.nn.grad.conv2d_input(x.shape, w, dy, stride=stride,
                                                       padding=pad), flush)
        w_o = timed(lambda: TF._conv2d_wgrad(dy, x, w.shape, pad, stride), flush)
        w_c = timed(lambda: torch.nn.grad.conv2d_weight(x, w.shape, dy, stride=stride,
                                                        padding=pad), flush)
        print(f'{name:<18}{gflop:8.1f} | {f_o:9.3f}{f_n:8.3f}{f_c:8.3f} | {d_o:11.3f}{d_c:8.3f} | '
              f'{w_o:11.3f}{w_c:8.3f}')


if __name__ == '__main__':
    main()
