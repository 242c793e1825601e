import json
import os
import subprocess
import sys

import pytest
import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_compat_install_aliases_reference_import_paths():
    code = (
        'import sys; sys.path.insert(0, %r)\n'
        'import torchgpipe_b200.compat as c; c.install()\n'
        'import torchgpipe\n'
        'from torchgpipe import GPipe, is_checkpointing, is_recomputing\n'
        'from torchgpipe.balance import balance_by_time, balance_by_size, blockpartition\n'
        'from torchgpipe.skip import skippable, stash, pop, Namespace, verify_skippables\n'
        'from torchgpipe.batchnorm import DeferredBatchNorm\n'
        'from torchgpipe.distributed import context\n'
        'from torchgpipe.distributed.gpipe import DistributedGPipe, DistributedGPipeDataLoader\n'
        'from torchgpipe.pipeline import clock_cycles\n'
        'from torchgpipe.stream import CPUStream\n'
        'import torchgpipe_b200\n'
        'assert GPipe is torchgpipe_b200.GPipe\n'
        'print("ok")\n' % ROOT)
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip().endswith('ok')


def test_version_and_public_names():
    import torchgpipe_b200
    assert torchgpipe_b200.__version__
    assert set(torchgpipe_b200.__all__) == {'GPipe', 'is_checkpointing', 'is_recomputing'}


@pytest.mark.timeout(240)
def test_speed_driver_smoke_on_cpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'benchmarks', 'speed.py'), 'unet',
                          'pipeline-2', '--smoke', '--devices', 'cpu,cpu', '--steps', '1',
                          '--warmup', '0'], capture_output=True, text=True, timeout=200,
                         cwd=os.path.join(ROOT, 'benchmarks'))
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    assert rec['benchmark'] == 'unet-speed' and rec['samples_per_s'] > 0
    assert sum(rec['balance']) == 241 and 'forward_bubble_frac' in rec


def test_bench_reference_arm_reports_unavailable_without_install(tmp_path):
    # run a copy of bench.py from a directory that has no baseline/_ref next to it
    src = open(os.path.join(ROOT, 'bench.py')).read()
    (tmp_path / 'bench.py').write_text(src)
    out = subprocess.run([sys.executable, str(tmp_path / 'bench.py'), '--impl', 'reference'],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    assert rec['impl'] == 'reference' and 'unavailable' in rec


def test_demo_models_shapes_and_layer_counts():
    # reference benchmarks/distributed/accuracy/{vgg,resnet}: 39-child VGG-16, 11-child ResNet
    from torchgpipe_b200.models import resnet101_coarse, resnet50_coarse, vgg16
    v, r101, r50 = vgg16(10), resnet101_coarse(10), resnet50_coarse(10)
    assert len(v) == 39 and len(r101) == 11 and len(r50) == 11
    assert sum(isinstance(m, nn.Conv2d) for m in v) == 13
    assert sum(p.numel() for p in vgg16(1000).parameters()) == 138357544      # torchvision's vgg16
    assert sum(p.numel() for p in resnet101_coarse(1000).parameters()) == 44549160
    with torch.no_grad():
        assert r50.eval()(torch.rand(2, 3, 64, 64)).shape == (2, 10)


@pytest.mark.timeout(300)
def test_distributed_accuracy_driver_two_ranks_on_gloo():
    port = 29600 + os.getpid() % 300
    out = subprocess.run(
        [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2',
         '--master-addr', '127.0.0.1', '--master-port', str(port),
         os.path.join(ROOT, 'benchmarks', 'distributed_accuracy.py'), '--cpu', '--model',
         'resnet50', '--balance', '5,6', '--batch-size', '4', '--chunks', '2', '--epochs', '1',
         '--skip-epochs', '0', '--image-size', '32', '--synthetic', '--synthetic-size', '8',
         '--lr', '0.001'], capture_output=True, text=True, timeout=280, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert 'valid accuracy' in out.stdout and 'samples/sec' in out.stdout


@pytest.mark.timeout(240)
def test_accuracy_driver_smoke_on_cpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'benchmarks', 'accuracy.py'),
                          'gpipe-256', '--synthetic', '--epochs', '1', '--steps-per-epoch', '1',
                          '--devices', 'cpu,cpu', '--batch', '4'], capture_output=True, text=True,
                         timeout=200, cwd=os.path.join(ROOT, 'benchmarks'))
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    assert rec['experiment'] == 'gpipe-256' and rec['epoch'] == 1


@pytest.mark.timeout(300)
@pytest.mark.parametrize('experiment', ['baseline', 'dep-str-ptl'])
def test_timeline_ablation_driver_smoke_on_cpu(experiment):
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'benchmarks', 'timeline.py'),
                          experiment, '--smoke', '--devices', 'cpu,cpu,cpu,cpu', '--steps', '1',
                          '--warmup', '0'], capture_output=True, text=True, timeout=280,
                         cwd=os.path.join(ROOT, 'benchmarks'))
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    assert rec['benchmark'] == 'unet-timeline' and rec['samples_per_s'] > 0
    assert rec['portals'] == (experiment == 'dep-str-ptl')


@pytest.mark.timeout(300)
def test_memory_driver_smoke_on_cpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'benchmarks', 'memory.py'), 'unet',
                          'pipeline-2', '--smoke', '--devices', 'cpu,cpu'], capture_output=True,
                         text=True, timeout=280, cwd=os.path.join(ROOT, 'benchmarks'))
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    assert rec['benchmark'] == 'unet-memory' and rec['parameters'] > 0


def test_bench_experiment_tables_match_the_reference_configs():
    sys.path.insert(0, ROOT)
    import bench
    # reference benchmarks/{unet,amoebanetd,resnet101}-speed/main.py
    assert bench.EXPERIMENTS['unet'][8] == dict(batch=640, chunks=40,
                                                balance=[16, 27, 31, 44, 22, 57, 27, 17])
    assert bench.EXPERIMENTS['amoebanetd'][8] == dict(batch=1280, chunks=32,
                                                      balance=[2, 2, 2, 3, 3, 4, 4, 4])
    assert bench.EXPERIMENTS['resnet101'][4] == dict(batch=5632, chunks=256,
                                                     balance=[44, 92, 124, 110])
    for model, sizes in (('unet', 241), ('amoebanetd', 24), ('resnet101', 370)):
        for cfg in bench.EXPERIMENTS[model].values():
            assert sum(cfg['balance']) == sizes
            assert cfg['batch'] % cfg['chunks'] == 0 or model == 'resnet101'


def test_bench_helpers_metric_balance_and_dropout_switch():
    """Host-side pieces of bench.py that the driver's records depend on."""
    import argparse
    import torch
    from torch import nn
    sys.path.insert(0, ROOT)
    import bench
    # the metric string must not change with N (the driver checks consistency across the curve)
    assert bench.metric_name('unet') == 'U-Net(5,64) GPipe training throughput'
    assert 'pipeline' not in bench.metric_name('unet')

    cfg = bench.EXPERIMENTS['unet'][4]
    args = argparse.Namespace(balance='reference', model='unet')
    assert bench.resolve_balance(args, cfg, torch, None, None, None, 1, 0) == \
        ([30, 66, 84, 61], 'reference')
    args.balance = '28,61,99,53'
    assert bench.resolve_balance(args, cfg, torch, None, None, None, 1, 0) == \
        ([28, 61, 99, 53], 'given')
    args.balance = '1,2,3'
    with pytest.raises(AssertionError):
        bench.resolve_balance(args, cfg, torch, None, None, None, 1, 0)

    # the numerics check turns every channel dropout off and restores it afterwards
    from torchgpipe_b200.ops.modules import ConvInstNormLeaky
    model = nn.Sequential(nn.Dropout2d(0.1),
                          ConvInstNormLeaky(nn.Conv2d(4, 4, 3, padding=1, bias=False), 0.1, 1e-5,
                                            0.01))
    saved = bench.set_dropout(model, 0.0)
    assert model[0].p == 0.0 and model[1].p == 0.0 and len(saved) == 2
    bench.set_dropout(model, saved=saved)
    assert model[0].p == 0.1 and model[1].p == 0.1

    # AmoebaNet-D has no 1-GPU experiment in the reference: the secondary block says so
    args = argparse.Namespace(gpus=1, no_secondary=False, child=False, model='unet')
    note = bench.run_secondary(args, torch, None, 0, 0, 1, None)
    assert 'skipped' in note and 'n2, n4, n8' in note['skipped']
    args.child = True
    assert bench.run_secondary(args, torch, None, 0, 0, 1, None) is None
