"""Cross-implementation parity on CPU: the UNMODIFIED reference (installed under
``baseline/_ref`` by ``baseline/stage_reference.sh``) and this package, on the reference's own
model builders - same ``state_dict`` keys, checkpoints interchangeable in both directions, same
outputs and gradients through the pipeline with skips, checkpointing and BatchNorm.

The reference runs in a subprocess (its import name ``torchgpipe`` is what
``torchgpipe_b200.compat`` aliases). Skipped when the reference is not staged.
"""
import os
import subprocess
import sys

import pytest
import torch

from torchgpipe_b200 import GPipe
from torchgpipe_b200.models import unet
from torchgpipe_b200.models.resnet import build_resnet

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, 'baseline', '_ref')

pytestmark = pytest.mark.skipif(
    not os.path.isdir(os.path.join(REF, 'torchgpipe')) or
    not os.path.isdir(os.path.join(REF, 'benchmarks_models')),
    reason='reference not staged (run baseline/stage_reference.sh)')

# Runs inside the reference's environment: build the model, optionally load a checkpoint, do one
# forward/backward through torchgpipe.GPipe on CPU devices, save everything.
_DRIVER = r'''
import sys, torch
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + '/benchmarks_models')
model_name, balance, chunks, checkpoint, out_path, load_path = sys.argv[2:8]
balance = [int(b) for b in balance.split(',')]
import torchgpipe
from torchgpipe import GPipe
assert torchgpipe.__file__.startswith(sys.argv[1])
torch.manual_seed(7)
if model_name == 'resnet':
    from resnet import build_resnet
    model = build_resnet([1, 1, 1, 1], num_classes=10)
    x = torch.rand(4, 3, 32, 32)
else:
    from unet import unet
    model = unet(depth=2, num_convs=2, base_channels=8, input_channels=3, output_channels=1)
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout2d):
            m.p = 0.0
    x = torch.rand(4, 3, 32, 32)
pipe = GPipe(model, balance, devices=['cpu'] * len(balance), chunks=int(chunks),
             checkpoint=checkpoint)
if load_path != '-':
    pipe.load_state_dict(torch.load(load_path))
initial = {k: v.clone() for k, v in pipe.state_dict().items()}
out = pipe(x)
out.square().mean().backward()
torch.save({'x': x, 'initial': initial, 'out': out.detach(),
            'grads': {k: p.grad.clone() for k, p in pipe.named_parameters()},
            'final': pipe.state_dict()}, out_path)
'''


def _run_reference(tmp_path, model_name, balance, chunks, checkpoint, load_path='-'):
    script = tmp_path / 'ref_driver.py'
    script.write_text(_DRIVER)
    out_path = tmp_path / f'ref_{model_name}.pt'
    env = {k: v for k, v in os.environ.items() if k != 'PYTHONPATH'}
    done = subprocess.run([sys.executable, str(script), REF, model_name,
                           ','.join(map(str, balance)), str(chunks), checkpoint, str(out_path),
                           str(load_path)], capture_output=True, text=True, timeout=280, env=env,
                          cwd=str(tmp_path))
    assert done.returncode == 0, done.stderr[-3000:]
    return torch.load(out_path)


def _ours(model_name):
    torch.manual_seed(7)
    if model_name == 'resnet':
        return build_resnet([1, 1, 1, 1], num_classes=10)
    model = unet(depth=2, num_convs=2, base_channels=8, input_channels=3, output_channels=1)
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout2d):
            m.p = 0.0
    return model


CASES = [('resnet', [20, 31], 2, 'always'), ('resnet', [51], 1, 'never'),
         ('unet', [17, 15, 17], 2, 'except_last')]


@pytest.mark.timeout(300)
@pytest.mark.parametrize('model_name,balance,chunks,checkpoint', CASES)
def test_reference_checkpoint_loads_here_and_training_step_matches(tmp_path, model_name, balance,
                                                                   chunks, checkpoint):
    model = _ours(model_name)
    assert sum(balance) == len(model)
    ref = _run_reference(tmp_path, model_name, balance, chunks, checkpoint)
    pipe = GPipe(model, balance, devices=['cpu'] * len(balance), chunks=chunks,
                 checkpoint=checkpoint)
    # identical key sets and order (same builders, same seed -> also identical initial values)
    assert list(pipe.state_dict().keys()) == list(ref['initial'].keys())
    pipe.load_state_dict(ref['initial'])                    # strict
    out = pipe(ref['x'])
    assert torch.allclose(out, ref['out'], rtol=1e-4, atol=1e-5)
    out.square().mean().backward()
    for name, p in pipe.named_parameters():
        g = ref['grads'][name]
        assert torch.allclose(p.grad, g, rtol=1e-3, atol=1e-5 + 1e-3 * g.abs().max().item()), name
    # buffers after the step (BatchNorm running statistics; the reference updates them a second
    # time when it recomputes, this implementation does not - compare without recomputation)
    if checkpoint == 'never':
        for name, v in pipe.state_dict().items():
            assert torch.allclose(v.float(), ref['final'][name].float(), rtol=1e-4, atol=1e-5), name


@pytest.mark.timeout(300)
def test_checkpoint_written_here_loads_into_the_reference(tmp_path):
    model = _ours('resnet')
    balance = [20, 31]
    pipe = GPipe(model, balance, devices=['cpu', 'cpu'], chunks=2, checkpoint='always')
    with torch.no_grad():
        for p in pipe.parameters():
            p.add_(0.01 * torch.randn_like(p))              # not the seed-7 initialisation
    path = tmp_path / 'ours.pt'
    torch.save(pipe.state_dict(), path)
    ref = _run_reference(tmp_path, 'resnet', balance, 2, 'always', load_path=path)
    for name, v in pipe.state_dict().items():
        assert torch.equal(v, ref['initial'][name]), name
    out = pipe(ref['x'])
    assert torch.allclose(out, ref['out'], rtol=1e-4, atol=1e-5)


@pytest.mark.timeout(300)
def test_sequential_state_dict_moves_a_reference_checkpoint_to_another_balance(tmp_path):
    ref = _run_reference(tmp_path, 'resnet', [20, 31], 2, 'never')
    first = GPipe(_ours('resnet'), [20, 31], devices=['cpu', 'cpu'], chunks=2)
    first.load_state_dict(ref['final'])
    other = GPipe(_ours('resnet'), [10, 10, 31], devices=['cpu'] * 3, chunks=4)
    other.load_sequential_state_dict(first.sequential_state_dict())
    first.eval()
    other.eval()
    x = ref['x']
    assert torch.allclose(first(x), other(x), atol=1e-6)


_PURE_DRIVER = r'''
import sys, json, random, signal
sys.path.insert(0, sys.argv[1])
from torchgpipe.balance import blockpartition
from torchgpipe.pipeline import clock_cycles
random.seed(11)

class Hang(Exception):
    pass

def on_alarm(*args):
    raise Hang()

signal.signal(signal.SIGALRM, on_alarm)
cases = []
for _ in range(300):
    n = random.randint(1, 14)
    seq = [random.choice([1, 1, 2, 3, 5, 8, 13, 0.5, 2.25]) for _ in range(n)]
    k = random.randint(1, n)
    signal.alarm(2)
    try:
        blocks = blockpartition.solve(seq, k)
    except Hang:
        blocks = None             # the reference's sliding loop never terminates on this input
    finally:
        signal.alarm(0)
    cases.append((seq, k, blocks))
clocks = {f'{m},{n}': [list(map(list, s)) for s in clock_cycles(m, n)]
          for m in range(1, 7) for n in range(1, 6)}
json.dump({'cases': cases, 'clocks': clocks}, open(sys.argv[2], 'w'))
'''


@pytest.mark.timeout(120)
def test_block_partition_and_clock_cycles_agree_with_the_reference(tmp_path):
    import json
    from torchgpipe_b200.balance import blockpartition
    from torchgpipe_b200.runtime.schedule import clock_cycles
    script = tmp_path / 'pure_driver.py'
    script.write_text(_PURE_DRIVER)
    out = tmp_path / 'pure.json'
    env = {k: v for k, v in os.environ.items() if k != 'PYTHONPATH'}
    done = subprocess.run([sys.executable, str(script), REF, str(out)], capture_output=True,
                          text=True, timeout=100, env=env, cwd=str(tmp_path))
    assert done.returncode == 0, done.stderr[-3000:]
    ref = json.load(open(out))
    hangs = 0
    for seq, k, blocks in ref['cases']:
        mine = blockpartition.solve(seq, k)
        if blocks is None:
            # own solver must still answer with k non-empty contiguous blocks
            hangs += 1
            assert len(mine) == k and all(mine) and sum(mine, []) == seq, (seq, k)
        else:
            assert mine == blocks, (seq, k)
    assert hangs < len(ref['cases']) // 10
    for key, schedule in ref['clocks'].items():
        m, n = map(int, key.split(','))
        assert [[list(cell) for cell in tick] for tick in clock_cycles(m, n)] == schedule, key


# Each snippet must raise; the reference and this package have to agree on the exception type
# and (modulo the package name) on the message. (Deliberate difference, not listed: a 2-D input
# to DeferredBatchNorm raises here - the reference defines the check but never calls it.)
_ERROR_SNIPPETS = [
    "GPipe(nn.Sequential(nn.Linear(1, 1)))",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[2], devices=['cpu'])",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[0, 1], devices=['cpu', 'cpu'])",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[-1, 2], devices=['cpu', 'cpu'])",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu'], chunks=0)",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu'], checkpoint='sometimes')",
    "GPipe(nn.Linear(1, 1), balance=[1], devices=['cpu'])",
    "GPipe(nn.Sequential(shared, shared), balance=[2], devices=['cpu'])",
    "GPipe(nn.Sequential(nn.Linear(1, 1), nn.Linear(1, 1)), balance=[1, 1], devices=['cpu'])",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu']).cpu()",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu']).to('cpu')",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu'])('not a tensor')",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu'])(torch.rand(1, 1), 2)",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu'])[3]",
    "balance_by_time(1, nn.Sequential(with_grad), torch.rand(1, 1), device='cpu')",
    "blockpartition.solve([1, 2], 3)",
    "blockpartition.solve([1, 2], 0)",
    "verify_skippables(nn.Sequential(StashX()))",
    "verify_skippables(nn.Sequential(PopX()))",
    "verify_skippables(nn.Sequential(StashX(), StashX(), PopX()))",
    "nn.Sequential(StashX(), PopX())(torch.rand(1))  if False else Bad()(torch.rand(1))",
    "PopX()(torch.rand(1))",
    "Undeclared()(torch.rand(1))",
    "DeferredBatchNorm(3)(torch.rand(2, 4, 5, 5))",
    "microbatch.check('x')",
    "microbatch.check((torch.rand(1), 2))",
    "microbatch.Batch((torch.rand(2),)).tensor",
    "microbatch.Batch(torch.rand(2)).tensors",
    "microbatch.scatter(torch.rand(0, 1), 2) and None.x",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu']).to(torch.device('cpu'))",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu']).to(tensor=torch.rand(1))",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu']).cuda()",
    "GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu']).no_such_attribute",
    "balance_by_size(1, nn.Sequential(nn.Linear(1, 1)), torch.rand(1, 1), device='cpu')",
    "GPipe(nn.Sequential(OrderedDict(a=shared, b=nn.Sequential(shared))), balance=[2], devices=['cpu'])",
]

_ERROR_DRIVER = r'''
import sys, json
if sys.argv[1] != '-':
    sys.path.insert(0, sys.argv[1])
    pkg = 'torchgpipe'
else:
    sys.path.insert(0, sys.argv[3])
    pkg = 'torchgpipe_b200'
import importlib, torch
from torch import nn
mod = importlib.import_module(pkg)
GPipe = mod.GPipe
balance = importlib.import_module(pkg + '.balance')
balance_by_time = balance.balance_by_time
blockpartition = importlib.import_module(pkg + '.balance.blockpartition')
skip = importlib.import_module(pkg + '.skip')
verify_skippables, skippable, stash, pop = skip.verify_skippables, skip.skippable, skip.stash, skip.pop

from collections import OrderedDict
microbatch = importlib.import_module(pkg + '.microbatch')
DeferredBatchNorm = importlib.import_module(pkg + '.batchnorm').DeferredBatchNorm
balance_by_size = balance.balance_by_size

shared = nn.Linear(1, 1)
with_grad = nn.Linear(1, 1)
with_grad.weight.grad = torch.zeros_like(with_grad.weight)

@skippable(stash=['x'])
class StashX(nn.Module):
    def forward(self, t):
        yield stash('x', t)
        return t

@skippable(pop=['x'])
class PopX(nn.Module):
    def forward(self, t):
        x = yield pop('x')
        return t + x

@skippable(stash=['x'])
class Bad(nn.Module):          # declares a stash it never performs
    def forward(self, t):
        return t

@skippable()
class Undeclared(nn.Module):   # stashes a name it did not declare
    def forward(self, t):
        yield stash('y', t)
        return t

out = []
for snippet in json.loads(sys.argv[2]):
    try:
        eval(snippet)
        out.append(['no error', ''])
    except Exception as exc:
        out.append([type(exc).__name__, str(exc).replace('torchgpipe_b200', 'torchgpipe')])
print('RESULT' + json.dumps(out))
'''


@pytest.mark.timeout(200)
def test_error_types_and_messages_agree_with_the_reference(tmp_path):
    import json
    script = tmp_path / 'err_driver.py'
    script.write_text(_ERROR_DRIVER)
    env = {k: v for k, v in os.environ.items() if k != 'PYTHONPATH'}

    def run(first):
        done = subprocess.run([sys.executable, str(script), first, json.dumps(_ERROR_SNIPPETS),
                               ROOT], capture_output=True, text=True, timeout=150, env=env,
                              cwd=str(tmp_path))
        assert done.returncode == 0, done.stderr[-3000:]
        line = [l for l in done.stdout.splitlines() if l.startswith('RESULT')][-1]
        return json.loads(line[len('RESULT'):])

    theirs, ours = run(REF), run('-')
    for snippet, a, b in zip(_ERROR_SNIPPETS, theirs, ours):
        assert a[0] != 'no error', snippet
        assert a[0] == b[0], (snippet, a, b)
        assert a[1] == b[1], (snippet, a, b)


_DBN_DRIVER = r'''
import sys, importlib, torch
from torch import nn
if sys.argv[1] != '-':
    sys.path.insert(0, sys.argv[1]); pkg = 'torchgpipe'
else:
    sys.path.insert(0, sys.argv[3]); pkg = 'torchgpipe_b200'
DBN = importlib.import_module(pkg + '.batchnorm').DeferredBatchNorm
torch.manual_seed(3)
results = {}
for momentum in (0.1, None):
    net = nn.Sequential(nn.Conv2d(3, 4, 3, padding=1), nn.BatchNorm2d(4, momentum=momentum), nn.ReLU())
    net = DBN.convert_deferred_batch_norm(net, chunks=3)
    x = torch.rand(9, 3, 6, 6)
    outs = []
    for step in range(2):                       # two mini-batches of three micro-batches
        for chunk in x.chunk(3):
            out = net(chunk + step)
            out.mean().backward()
            outs.append(out.detach())
    net.eval()
    results[str(momentum)] = dict(
        outs=torch.cat(outs), eval=net(x).detach(),
        state={k: v.clone() for k, v in net.state_dict().items()},
        grads=[p.grad.clone() for p in net.parameters()], kind=type(net[1]).__name__)
torch.save(results, sys.argv[2])
'''


@pytest.mark.timeout(200)
def test_deferred_batch_norm_matches_the_reference(tmp_path):
    script = tmp_path / 'dbn_driver.py'
    script.write_text(_DBN_DRIVER)
    env = {k: v for k, v in os.environ.items() if k != 'PYTHONPATH'}
    got = {}
    for tag, first in (('ref', REF), ('ours', '-')):
        out = tmp_path / f'dbn_{tag}.pt'
        done = subprocess.run([sys.executable, str(script), first, str(out), ROOT],
                              capture_output=True, text=True, timeout=150, env=env,
                              cwd=str(tmp_path))
        assert done.returncode == 0, done.stderr[-3000:]
        got[tag] = torch.load(out)
    for momentum in ('0.1', 'None'):
        a, b = got['ref'][momentum], got['ours'][momentum]
        assert a['kind'] == b['kind'] == 'DeferredBatchNorm'
        assert list(a['state'].keys()) == list(b['state'].keys())
        for k in a['state']:
            assert torch.allclose(a['state'][k].float(), b['state'][k].float(), atol=1e-6), k
        assert torch.allclose(a['outs'], b['outs'], atol=1e-6)
        assert torch.allclose(a['eval'], b['eval'], atol=1e-5)
        for g, h in zip(a['grads'], b['grads']):
            assert torch.allclose(g, h, atol=1e-6)


# The reference's multi-process mode, run as two threads in one process with its RPC send
# replaced by a direct call (what its own test-suite does with mock.patch), against this
# package's DistributedGPipe on the in-process channel transport.
_DIST_DRIVER = r'''
import sys, importlib, threading, torch
from torch import nn
ref = sys.argv[1] != '-'
sys.path.insert(0, sys.argv[1] if ref else sys.argv[3])
pkg = 'torchgpipe' if ref else 'torchgpipe_b200'
context = importlib.import_module(pkg + '.distributed.context')
gp = importlib.import_module(pkg + '.distributed.gpipe')
DistributedGPipe = gp.DistributedGPipe
if ref:
    def direct_put(name, id, values, backward=False):
        (context.put_backward if backward else context.put_forward)(name, id, values)
    DistributedGPipe._put = staticmethod(direct_put)

torch.manual_seed(5)
model = nn.Sequential(nn.Linear(4, 8), nn.Tanh(), nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 2))
x, balance, chunks = torch.rand(6, 4), [2, 3], 3
workers = {0: 'w0', 1: 'w1'}
losses_out, errors = [], []

def run(rank):
    try:
        with context.worker(workers[rank], chunks):
            barrier.wait()
            pipe = DistributedGPipe(model, rank, workers, balance, chunks,
                                    device=torch.device('cpu'))
            outs = pipe.forward(x if rank == 0 else None)
            if rank == 1:
                losses = [o.square().sum() / 6 for o in outs]
                losses_out.extend(l.detach() for l in losses)
                pipe.backward(losses)
            else:
                pipe.backward(None)
            barrier.wait()
    except Exception as exc:
        errors.append(repr(exc))
        barrier.abort()

barrier = threading.Barrier(2)
threads = [threading.Thread(target=run, args=(r,)) for r in (0, 1)]
[t.start() for t in threads]
[t.join(60) for t in threads]
assert not errors, errors
torch.save({'grads': [p.grad.clone() for p in model.parameters()],
            'losses': torch.stack(losses_out)}, sys.argv[2])
'''


@pytest.mark.timeout(200)
def test_distributed_mode_matches_the_reference_distributed_mode(tmp_path):
    if not os.path.isdir(os.path.join(REF, 'torchgpipe', 'distributed')):
        pytest.skip('reference distributed package not staged')
    script = tmp_path / 'dist_driver.py'
    script.write_text(_DIST_DRIVER)
    env = {k: v for k, v in os.environ.items() if k != 'PYTHONPATH'}
    got = {}
    for tag, first in (('ref', REF), ('ours', '-')):
        out = tmp_path / f'dist_{tag}.pt'
        done = subprocess.run([sys.executable, str(script), first, str(out), ROOT],
                              capture_output=True, text=True, timeout=150, env=env,
                              cwd=str(tmp_path))
        assert done.returncode == 0, done.stderr[-3000:]
        got[tag] = torch.load(out)
    assert torch.allclose(got['ref']['losses'], got['ours']['losses'], atol=1e-6)
    for a, b in zip(got['ref']['grads'], got['ours']['grads']):
        assert torch.allclose(a, b, atol=1e-6)


_NAMESPACE_PROBE = r'''
import copy, json, pickle, sys
if sys.argv[1] != '-':
    sys.path.insert(0, sys.argv[1])
pkg = __import__(sys.argv[2], fromlist=['skip'])
Namespace = __import__(sys.argv[2] + '.skip', fromlist=['Namespace']).Namespace
a, b = Namespace(), Namespace()
lo, hi = sorted([a, b])
twin = copy.deepcopy(lo)
thawed = pickle.loads(pickle.dumps(hi))
def attempt(fn):
    try:
        return bool(fn())
    except Exception:
        return 'raises'
facts = {
    'distinct': lo != hi, 'self_equal': lo == lo, 'copy_equal': twin == lo,
    'copy_same_hash': hash(twin) == hash(lo), 'pickle_equal': thawed == hi,
    'usable_as_dict_key': {lo: 1}.get(twin) == 1,
    'lt': lo < hi, 'le': lo <= hi, 'gt': hi > lo, 'ge': hi >= lo, 'not_lt_self': not (lo < lo),
    'none_is_instance': isinstance(None, Namespace), 'int_is_not': not isinstance(3, Namespace),
    'eq_none': attempt(lambda: lo == None), 'ne_none': attempt(lambda: lo != None),
    'lt_none': attempt(lambda: lo < None), 'gt_none': attempt(lambda: lo > None),
    'le_none': attempt(lambda: lo <= None), 'ge_none': attempt(lambda: lo >= None),
    'sort_with_none_first': attempt(lambda: sorted([(hi, 'x'), (None, 'x'), (lo, 'x')],
                                                   key=lambda p: (p[0] is not None, ))[0][0] is None),
    'eq_foreign': attempt(lambda: lo == 'foo'),
}
print(json.dumps(facts, sort_keys=True))
'''


def test_namespace_truth_table_equals_the_reference(tmp_path):
    """Equality, hashing, ordering, copies and the ``None`` conventions of ``skip.Namespace``
    (ours is built on (creation index, salt) keys, the reference on UUIDs)."""
    import json
    script = tmp_path / 'ns_probe.py'
    script.write_text(_NAMESPACE_PROBE)
    env = {k: v for k, v in os.environ.items() if k != 'PYTHONPATH'}

    def probe(path, package):
        done = subprocess.run([sys.executable, str(script), path, package], capture_output=True,
                              text=True, timeout=120, env=env, cwd=str(tmp_path))
        assert done.returncode == 0, done.stderr[-2000:]
        return json.loads(done.stdout.strip().splitlines()[-1])

    theirs = probe(REF, 'torchgpipe')
    ours = probe(ROOT, 'torchgpipe_b200')
    assert ours.keys() == theirs.keys()
    for fact, value in theirs.items():
        if value == 'raises':
            # the reference trips over its own "None is a Namespace" rule in comparisons with
            # None (AttributeError on None.id); here those comparisons have an answer
            assert ours[fact] in (True, False), fact
        else:
            assert ours[fact] == value, fact
    assert ours['distinct'] and ours['copy_equal'] and ours['none_is_instance']
    assert ours['eq_none'] is False and ours['gt_none'] is True and ours['lt_none'] is False
