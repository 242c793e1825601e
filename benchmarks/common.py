"""Shared pieces of the benchmark drivers: experiment tables (same names, batch sizes, chunks,
balances and checkpoint modes as the reference's ``benchmarks/*/main.py``), device-side timing and
JSON-lines output."""
import json
import os
import sys
import time
from typing import Any, Callable, Dict, List, Optional

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

# name -> dict(batch, chunks, balance, checkpoint) ; 'baseline' runs without GPipe on one device.
# reference: benchmarks/unet-speed/main.py:24-68
UNET_SPEED = {
    'baseline': dict(batch=40, chunks=None, balance=None),
    'pipeline-1': dict(batch=80, chunks=2, balance=[241]),
    'pipeline-2': dict(batch=512, chunks=32, balance=[104, 137]),
    'pipeline-4': dict(batch=512, chunks=16, balance=[30, 66, 84, 61]),
    'pipeline-8': dict(batch=640, chunks=40, balance=[16, 27, 31, 44, 22, 57, 27, 17]),
}
# reference: benchmarks/resnet101-speed/main.py:23-67
RESNET101_SPEED = {
    'baseline': dict(batch=118, chunks=None, balance=None),
    'pipeline-1': dict(batch=220, chunks=2, balance=[370]),
    'pipeline-2': dict(batch=25000, chunks=1667, balance=[135, 235]),
    'pipeline-4': dict(batch=5632, chunks=256, balance=[44, 92, 124, 110]),
    'pipeline-8': dict(batch=5400, chunks=150, balance=[26, 22, 33, 44, 44, 66, 66, 69]),
}
# reference: benchmarks/amoebanetd-speed/main.py:36-96 (m=1 uses checkpoint='always')
AMOEBANETD_SPEED = {
    'n2m1': dict(batch=96, chunks=1, balance=[7, 17], checkpoint='always'),
    'n2m4': dict(batch=256, chunks=4, balance=[9, 15]),
    'n2m32': dict(batch=1280, chunks=32, balance=[9, 15]),
    'n4m1': dict(batch=160, chunks=1, balance=[3, 4, 5, 12], checkpoint='always'),
    'n4m4': dict(batch=360, chunks=4, balance=[3, 6, 7, 8]),
    'n4m32': dict(batch=1152, chunks=32, balance=[3, 6, 7, 8]),
    'n8m1': dict(batch=196, chunks=1, balance=[2, 2, 2, 2, 2, 3, 4, 7], checkpoint='always'),
    'n8m4': dict(batch=480, chunks=4, balance=[2, 2, 2, 3, 3, 4, 4, 4]),
    'n8m32': dict(batch=1280, chunks=32, balance=[2, 2, 2, 3, 3, 4, 4, 4]),
}
# reference: benchmarks/unet-memory/main.py:22-78  (B, C) grows with the number of partitions
UNET_MEMORY = {
    'baseline': dict(model=(6, 72), batch=32, chunks=None, balance=None),
    'pipeline-1': dict(model=(11, 128), batch=32, chunks=32, balance=[472]),
    'pipeline-2': dict(model=(24, 128), batch=64, chunks=32, balance=[500, 577]),
    'pipeline-4': dict(model=(24, 160), batch=128, chunks=32, balance=[472, 54, 36, 515]),
    'pipeline-8': dict(model=(48, 160), batch=32, chunks=32,
                       balance=[800, 140, 85, 63, 71, 62, 99, 837]),
}
# reference: benchmarks/amoebanetd-memory/main.py (L, D) per experiment
AMOEBANETD_MEMORY = {
    'baseline': dict(model=(18, 208), batch=128, chunks=None, balance=None),
    'pipeline-1': dict(model=(18, 416), batch=128, chunks=128, balance=[24]),
    'pipeline-2': dict(model=(18, 544), batch=128, chunks=128, balance=[10, 14]),
    'pipeline-4': dict(model=(36, 544), batch=128, chunks=128, balance=[10, 10, 10, 12]),
    'pipeline-8': dict(model=(72, 512), batch=128, chunks=128, balance=[10] * 7 + [8]),
}


def emit(record: Dict[str, Any], path: Optional[str] = None) -> None:
    """One JSON line to stdout (and appended to ``path``)."""
    line = json.dumps(record)
    print(line, flush=True)
    if path:
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        with open(path, 'a') as f:
            f.write(line + '\n')


def device_timed(devices: List[torch.device], fn: Callable[[], None], steps: int,
                 warmup: int) -> float:
    """Milliseconds per step: CUDA events on every device, max over devices (CPU: wall clock)."""
    for _ in range(warmup):
        fn()
    cuda = [d for d in devices if d.type == 'cuda']
    for d in cuda:
        torch.cuda.synchronize(d)
    if not cuda:
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        return (time.perf_counter() - t0) * 1e3 / steps
    starts = [torch.cuda.Event(enable_timing=True) for _ in cuda]
    ends = [torch.cuda.Event(enable_timing=True) for _ in cuda]
    for d, e in zip(cuda, starts):
        e.record(torch.cuda.current_stream(d))
    for _ in range(steps):
        fn()
    for d, e in zip(cuda, ends):
        e.record(torch.cuda.current_stream(d))
    for d in cuda:
        torch.cuda.synchronize(d)
    return max(s.elapsed_time(e) for s, e in zip(starts, ends)) / steps


def load_impl(impl: str):
    """Returns (GPipe class, model builders dict, convert function or None)."""
    if impl == 'reference':
        root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                            'baseline', '_ref')
        sys.path.insert(0, os.path.join(root, 'benchmarks_models'))
        sys.path.insert(0, root)
        from torchgpipe import GPipe
        from amoebanet import amoebanetd
        from resnet import resnet101
        from unet import unet
        return GPipe, dict(unet=unet, resnet101=resnet101, amoebanetd=amoebanetd), None
    from torchgpipe_b200 import GPipe, ops
    from torchgpipe_b200.models import amoebanetd, resnet101, unet
    return GPipe, dict(unet=unet, resnet101=resnet101, amoebanetd=amoebanetd), \
        ops.convert_to_sm100
