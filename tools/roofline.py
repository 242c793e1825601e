"""Roofline table for the per-layer timings of ``build/selftest unet`` (profiles/*.txt).

    python tools/roofline.py profiles/unet_layers_selftest.txt > profiles/roofline_unet_layers.md

Denominators are MEASURED on the box: HBM copy bandwidth from ``MEASURED_PEAKS.json`` (driver-
written) and the TF32 roof from ``profiles/measured_tf32_peak.json`` (``tools/
measure_tf32_peak.py``: cuBLAS fp32 matmul with TF32 allowed, burst figure); when that file is
missing, half of the cuBLAS bf16 burst throughput. A 3x3 stride-1 convolution layer does
2*N*H*W*C*K*9 FLOP and must move at least its input, its output and its weights once (fp32).
"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main() -> None:
    table = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'profiles',
                                                               'unet_layers_selftest.txt')
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except OSError:
        peaks = {'hbm_gbs': 6557.8, 'bf16_tflops': 1651.1}
    tf32_peak = peaks['bf16_tflops'] / 2.0          # TFLOP/s
    tf32_note = f'half of the cuBLAS bf16 burst {peaks["bf16_tflops"]:.0f}'
    try:
        measured = json.load(open(os.path.join(ROOT, 'profiles', 'measured_tf32_peak.json')))
        tf32_peak = measured['tf32_tflops_burst']
        tf32_note = (f'measured: cuBLAS TF32 matmul 8192^3, burst; sustained '
                     f'{measured["tf32_tflops_sustained"]:.0f}')
    except (OSError, KeyError, ValueError):
        pass
    hbm = peaks['hbm_gbs']                          # GB/s
    rows = {}
    for line in open(table):
        m = re.match(r'\[(wgrad )?unet_(\d+)_(\d+)_(\d+)( stats=(\d))?\] ([\d.]+) ms', line)
        if not m:
            continue
        kind = 'wgrad' if m.group(1) else ('fprop+stats' if m.group(6) == '1' else 'fprop')
        key = tuple(int(m.group(i)) for i in (2, 3, 4))
        rows.setdefault(key, {})[kind] = float(m.group(7))
    print(f'Roofline against measured peaks: TF32 {tf32_peak:.0f} TFLOP/s ({tf32_note}), '
          f'HBM {hbm:.0f} GB/s; micro-batch {batch}.\n')
    print('| C -> K @ HxW | GFLOP | min MB | roof ms (compute / memory) | fprop ms (% of roof) | '
          'fprop+stats | wgrad |')
    print('|---|---|---|---|---|---|---|')
    for (c, k, hw), t in rows.items():
        flop = 2.0 * batch * hw * hw * c * k * 9
        bytes_ = 4.0 * (batch * hw * hw * (c + k) + 9 * c * k)
        t_c = flop / (tf32_peak * 1e12) * 1e3
        t_m = bytes_ / (hbm * 1e9) * 1e3
        roof = max(t_c, t_m)

        def cell(kind: str) -> str:
            if kind not in t:
                return '-'
            return f'{t[kind]:.3f} ({100 * roof / t[kind]:.0f} %)'
        print(f'| {c} -> {k} @ {hw}x{hw} | {flop / 1e9:.1f} | {bytes_ / 1e6:.0f} | '
              f'{t_c:.3f} / {t_m:.3f} | {cell("fprop")} | {cell("fprop+stats")} | {cell("wgrad")} |')


if __name__ == '__main__':
    main()
