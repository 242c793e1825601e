"""GPU sanity check: converted AmoebaNet-D (fused conv+BN, NHWC pooling) against the plain model,
both judged against a float64 run of the plain model (TF32 noise through a deep BatchNorm net is
large at batch 8, so "plain fp32/TF32 vs fp64" is the yardstick for "fused vs fp64")."""
import copy
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchgpipe_b200 import ops  # noqa: E402
from torchgpipe_b200.models import amoebanetd  # noqa: E402


def grads(model, x):
    model.zero_grad()
    out = model(x)
    out.square().mean().backward()
    return out.detach(), [p.grad.detach().double().flatten() for p in model.parameters()]


def worst_cos(a, b):
    return min(F.cosine_similarity(p, q, dim=0).item() for p, q in zip(a, b))


torch.manual_seed(0)
plain = amoebanetd(num_classes=10, num_layers=3, num_filters=64).cuda()
fused = ops.convert_to_sm100(copy.deepcopy(plain))
exact = copy.deepcopy(plain).double()
x = torch.randn(8, 3, 224, 224, device='cuda')
o64, g64 = grads(exact, x.double())
of, gf = grads(fused, x)
op, gp = grads(plain, x)
print('output rel err  fused vs fp64 %.4f   plain(tf32) vs fp64 %.4f' % (
    ((of - o64).norm() / o64.norm()).item(), ((op - o64).norm() / o64.norm()).item()))
print('worst grad cosine  fused vs fp64 %.4f   plain(tf32) vs fp64 %.4f   fused vs plain %.4f' % (
    worst_cos(gf, g64), worst_cos(gp, g64), worst_cos(gf, gp)))
print('native launches', ops.launch_count())
assert worst_cos(gf, g64) > min(0.9, worst_cos(gp, g64) - 0.05)
