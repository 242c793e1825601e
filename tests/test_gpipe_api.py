from collections import OrderedDict
import copy
import threading
import time

import pytest
import torch
from torch import nn

from torchgpipe_b200 import GPipe


def test_parameters_visible():
    model = nn.Sequential(nn.Linear(1, 1))
    pipe = GPipe(model, balance=[1], devices=['cpu'], chunks=1)
    assert list(pipe.parameters()) != []


def test_public_attrs_and_coercion():
    class MyString:
        def __init__(self, v):
            self.v = v

        def __str__(self):
            return self.v

    model = nn.Sequential(nn.Linear(1, 1))
    pipe = GPipe(model, balance=(1,), devices=('cpu',), chunks=42.000,
                 checkpoint=MyString('always'))
    assert pipe.balance == [1]
    assert pipe.devices == [torch.device('cpu')]
    assert pipe.chunks == 42 and isinstance(pipe.chunks, int)
    assert pipe.checkpoint == 'always' and isinstance(pipe.checkpoint, str)


def test_class_level_defaults():
    assert GPipe.balance == [] and GPipe.devices == []
    assert GPipe.chunks == 1 and GPipe.checkpoint == 'except_last'


def test_sequence_protocol():
    a, b = nn.Linear(1, 1), nn.Linear(1, 1)
    pipe = GPipe(nn.Sequential(a, b), balance=[1, 1], devices=['cpu', 'cpu'])
    assert len(pipe) == 2
    assert list(pipe) == [a, b]
    assert pipe[0] is a and pipe[-1] is b and pipe[1] is b and pipe[-2] is a
    with pytest.raises(IndexError):
        pipe[2]
    with pytest.raises(IndexError):
        pipe[-3]


def test_balance_errors():
    a, b = nn.Linear(1, 1), nn.Linear(1, 1)
    model = nn.Sequential(a, b)
    with pytest.raises(ValueError, match='torchgpipe.balance'):
        GPipe(model)
    with pytest.raises(ValueError, match='different length'):
        GPipe(model, balance=[1], devices=['cpu'])
    with pytest.raises(ValueError, match='different length'):
        GPipe(model, balance=[3], devices=['cpu'])
    with pytest.raises(ValueError, match='positive integer'):
        GPipe(model, balance=[0, 2], devices=['cpu', 'cpu'])
    with pytest.raises(ValueError, match='positive integer'):
        GPipe(model, balance=[-1, 3], devices=['cpu', 'cpu'])
    with pytest.raises(ValueError, match='torchgpipe.balance'):
        GPipe(model, balance=[3], devices=['cpu'])


def test_chunks_and_mode_errors():
    model = nn.Sequential(nn.Linear(1, 1))
    with pytest.raises(ValueError, match='number of chunks must be positive integer'):
        GPipe(model, balance=[1], devices=['cpu'], chunks=0)
    with pytest.raises(ValueError):
        GPipe(model, balance=[1], devices=['cpu'], chunks=-1)
    with pytest.raises(ValueError, match="checkpoint is not one of 'always', 'except_last', or 'never'"):
        GPipe(model, balance=[1], devices=['cpu'], checkpoint='INVALID')


def test_too_few_devices():
    model = nn.Sequential(nn.Linear(1, 1), nn.Linear(1, 1), nn.Linear(1, 1))
    with pytest.raises(IndexError, match='too few devices'):
        GPipe(model, balance=[1, 1, 1], devices=['cpu'])


def test_surplus_devices_are_dropped():
    model = nn.Sequential(nn.Linear(1, 1), nn.Linear(1, 1))
    pipe = GPipe(model, balance=[1, 1], devices=['cpu'] * 5)
    assert len(pipe.devices) == 2


def test_indivisible_and_small_batches_are_silent(recwarn):
    model = nn.Sequential(nn.Linear(1, 1))
    pipe = GPipe(model, balance=[1], devices=['cpu'], chunks=4)
    assert pipe(torch.rand(7, 1)).shape == (7, 1)
    assert pipe(torch.rand(2, 1)).shape == (2, 1)
    assert len(recwarn) == 0


def _count_checkpoints(output):
    seen, stack, count = set(), [output.grad_fn], 0
    while stack:
        fn = stack.pop()
        if fn is None or fn in seen:
            continue
        seen.add(fn)
        if fn.__class__.__name__ == 'CheckpointBackward':
            count += 1
        stack.extend(f for f, _ in fn.next_functions)
    return count


@pytest.mark.parametrize('mode,expected', [('always', 2), ('except_last', 1), ('never', 0)])
def test_checkpoint_modes(mode, expected):
    model = nn.Sequential(nn.Linear(1, 1))
    pipe = GPipe(model, balance=[1], devices=['cpu'], chunks=2, checkpoint=mode)
    out = pipe(torch.rand(2, 1))
    assert _count_checkpoints(out) == expected


def test_checkpoint_disabled_in_eval_and_no_grad():
    model = nn.Sequential(nn.Linear(1, 1))
    pipe = GPipe(model, balance=[1], devices=['cpu'], chunks=2, checkpoint='always')
    pipe.eval()
    assert _count_checkpoints(pipe(torch.rand(2, 1))) == 0
    pipe.train()
    with torch.no_grad():
        out = pipe(torch.rand(2, 1))
    assert out.grad_fn is None and not out.requires_grad


def test_no_grad_builds_no_graph_in_partitions():
    seen = []

    class Probe(nn.Module):
        def forward(self, x):
            y = x * 2
            seen.append(y.requires_grad)
            return y

    pipe = GPipe(nn.Sequential(nn.Linear(1, 1), Probe()), balance=[1, 1],
                 devices=['cpu', 'cpu'], chunks=2, checkpoint='never')
    with torch.no_grad():
        pipe(torch.rand(2, 1))
    assert seen == [False, False]


def test_exception_propagates_with_type():
    class ExpectedException(Exception):
        pass

    class Raise(nn.Module):
        def forward(self, *_):
            raise ExpectedException('boom')

    pipe = GPipe(nn.Sequential(Raise()), balance=[1], devices=['cpu'], chunks=1)
    with pytest.raises(ExpectedException, match='boom'):
        pipe(torch.rand(1))
    # the module stays usable: persistent workers survived
    with pytest.raises(ExpectedException):
        pipe(torch.rand(1))


def test_exception_stops_later_ticks_and_does_not_hang():
    class ExpectedException(Exception):
        pass

    class Pass(nn.Module):
        def forward(self, x):
            return x

    counter = 0
    lock = threading.Lock()

    class Counter(nn.Module):
        def forward(self, x):
            nonlocal counter
            time.sleep(0.05)
            with lock:
                counter += 1
            return x

    class Raise(nn.Module):
        def forward(self, x):
            raise ExpectedException()

    model = nn.Sequential(Pass(), Pass(), Counter(), Raise())
    pipe = GPipe(model, [1, 1, 1, 1], devices=['cpu'] * 4, chunks=3)
    with pytest.raises(ExpectedException):
        pipe(torch.rand(3))
    # tick where Raise first runs also ran Counter for micro-batch 1; no later tick ran
    assert counter == 2


def test_tuple_input_output():
    class Two(nn.Module):
        def forward(self, ab):
            a, b = ab
            return a + 1, b * 2

    pipe = GPipe(nn.Sequential(Two(), Two()), balance=[1, 1], devices=['cpu', 'cpu'], chunks=2)
    a, b = torch.zeros(4, 1), torch.ones(4, 1)
    x, y = pipe((a, b))
    assert torch.equal(x, torch.full((4, 1), 2.)) and torch.equal(y, torch.full((4, 1), 4.))


def test_singleton_tuple():
    class One(nn.Module):
        def forward(self, t):
            (a,) = t
            return (a + 1,)

    pipe = GPipe(nn.Sequential(One()), balance=[1], devices=['cpu'], chunks=2)
    out = pipe((torch.zeros(2, 1),))
    assert isinstance(out, tuple) and len(out) == 1


def test_non_tensor_inputs_and_outputs():
    class NonTensor(nn.Module):
        def forward(self, _):
            return 'hello'

    class Pass(nn.Module):
        def forward(self, x):
            return x

    pipe = GPipe(nn.Sequential(Pass()), balance=[1], devices=['cpu'])
    with pytest.raises(TypeError):
        pipe('hello')
    with pytest.raises(TypeError):
        pipe((torch.rand(1), 'hello'))
    with pytest.raises(TypeError):
        pipe(torch.rand(1), torch.rand(1))          # extra positional argument
    bad = GPipe(nn.Sequential(NonTensor(), Pass()), balance=[1, 1], devices=['cpu', 'cpu'])
    with pytest.raises(TypeError):
        bad(torch.rand(1))


def test_deferred_batch_norm_flag():
    bn = nn.BatchNorm2d(3)
    ref = copy.deepcopy(bn)
    pipe = GPipe(nn.Sequential(bn), balance=[1], devices=['cpu'], chunks=2,
                 deferred_batch_norm=True)
    x = torch.rand(4, 3, 10, 10)
    pipe(x).mean().backward()
    ref(x)
    assert torch.allclose(pipe[0].running_mean, ref.running_mean, atol=1e-4)
    assert torch.allclose(pipe[0].running_var, ref.running_var, atol=1e-4)
    keys = set(pipe.state_dict().keys())
    assert {'partitions.0.0.sum', 'partitions.0.0.sum_squares',
            'partitions.0.0.running_mean', 'partitions.0.0.num_batches_tracked'} <= keys


def test_partitions_and_state_dict_keys():
    a, b = nn.Linear(1, 1), nn.Linear(1, 1)
    model = nn.Sequential(a, nn.ReLU(), b, nn.ReLU())
    pipe = GPipe(model, [2, 2], devices=['cpu', 'cpu'])
    assert isinstance(pipe.partitions, nn.ModuleList)
    assert isinstance(pipe.partitions[0], nn.Sequential)
    assert list(pipe.state_dict().keys()) == [
        'partitions.0.0.weight', 'partitions.0.0.bias',
        'partitions.1.2.weight', 'partitions.1.2.bias']


def test_sequential_state_dict_roundtrip_across_balances():
    def make():
        return nn.Sequential(OrderedDict([('a', nn.Linear(2, 2)), ('act', nn.ReLU()),
                                          ('b', nn.Linear(2, 2))]))
    plain = make()
    p1 = GPipe(make(), [1, 2], devices=['cpu', 'cpu'])
    p1.load_sequential_state_dict(plain.state_dict())
    p2 = GPipe(make(), [2, 1], devices=['cpu', 'cpu'])
    p2.load_sequential_state_dict(p1.state_dict())         # GPipe keys with another balance
    assert list(p1.sequential_state_dict().keys()) == list(plain.state_dict().keys())
    x = torch.rand(3, 2)
    assert torch.allclose(plain(x), p1(x)) and torch.allclose(plain(x), p2(x))
    make().load_state_dict(p2.sequential_state_dict())


def test_moving_is_denied_but_dtype_cast_allowed():
    pipe = GPipe(nn.Sequential(nn.Linear(1, 1)), balance=[1], devices=['cpu'])
    for call in (lambda: pipe.cuda(), lambda: pipe.cpu(), lambda: pipe.to(torch.device('cpu')),
                 lambda: pipe.to('cpu'), lambda: pipe.to(0), lambda: pipe.to(torch.rand(1)),
                 lambda: pipe.to(device='cpu'), lambda: pipe.to(tensor=torch.rand(1)),
                 lambda: pipe.to(torch.device('cpu'), torch.float64)):
        with pytest.raises(TypeError, match='denied to move'):
            call()
    pipe.to(torch.float64)
    assert pipe[0].weight.dtype == torch.float64
    pipe.half()
    assert pipe[0].weight.dtype == torch.float16
    pipe.to(dtype=torch.float32)


def test_empty_module_is_identity():
    pipe = GPipe(nn.Sequential(), balance=[], devices=[])
    x = torch.rand(3)
    assert pipe(x) is x
    assert pipe((x, x)) == (x, x)
    with pytest.raises(TypeError):
        pipe(42)


def test_named_children_and_no_getattr_passthrough():
    model = nn.Sequential(OrderedDict([('a', nn.Linear(1, 1)), ('b', nn.Linear(1, 1))]))
    pipe = GPipe(model, [1, 1], devices=['cpu', 'cpu'])
    names = {n for n, _ in pipe.named_modules()}
    assert 'partitions.0.a' in names and 'partitions.1.b' in names
    with pytest.raises(AttributeError):
        pipe.a


def test_verify_module():
    with pytest.raises(TypeError, match='module must be nn.Sequential to be partitioned'):
        GPipe(nn.Linear(1, 1), [1], devices=['cpu'])
    conv = nn.Conv2d(1, 1, 1)
    with pytest.raises(ValueError, match='module with duplicate children is not supported'):
        GPipe(nn.Sequential(conv, conv), [1, 1], devices=['cpu', 'cpu'])

    class Shared(nn.Module):
        def __init__(self, p):
            super().__init__()
            self.p = p

        def forward(self, x):
            return x

    p = nn.Parameter(torch.rand(1))
    with pytest.raises(ValueError, match='duplicate parameters in distinct children'):
        GPipe(nn.Sequential(Shared(p), Shared(p)), [1, 1], devices=['cpu', 'cpu'])


def test_executors_and_streams_are_reused_across_calls():
    pipe = GPipe(nn.Sequential(nn.Linear(1, 1), nn.Linear(1, 1)), [1, 1],
                 devices=['cpu', 'cpu'], chunks=2)
    pipe(torch.rand(2, 1))
    pool, streams = pipe._executors, pipe._copy_streams
    n_threads = threading.active_count()
    for _ in range(3):
        pipe(torch.rand(2, 1))
    assert pipe._executors is pool and pipe._copy_streams is streams
    assert threading.active_count() == n_threads
    assert pool.num_threads == 1            # both stages on 'cpu' share one thread


def test_deepcopy_and_pickle_drop_runtime_state():
    import pickle
    pipe = GPipe(nn.Sequential(nn.Linear(1, 1)), [1], devices=['cpu'])
    pipe(torch.rand(1, 1))
    clone = copy.deepcopy(pipe)
    assert clone._executors is None
    clone(torch.rand(1, 1))
    pickle.loads(pickle.dumps(pipe))(torch.rand(1, 1))


def test_timeline_reports_bubble():
    from torchgpipe_b200.runtime.timeline import Timeline

    class Sleep(nn.Module):
        def forward(self, x):
            time.sleep(0.02)
            return x

    pipe = GPipe(nn.Sequential(Sleep(), Sleep()), [1, 1], devices=['cpu', 'cpu'], chunks=4,
                 checkpoint='never')
    pipe.timeline = Timeline()
    pipe(torch.rand(4, 1))
    rep = pipe.timeline.report(2)
    assert len(rep['busy_ms']) == 2 and rep['span_ms'] > 0
    assert 0.0 <= rep['bubble_frac'] < 1.0


@pytest.mark.gpu
def test_default_devices_are_all_cuda_devices():
    model = nn.Sequential(nn.Linear(1, 1))
    pipe = GPipe(model, [1])
    assert pipe.devices == [torch.device('cuda', 0)]
    out = pipe(torch.rand(2, 1, device='cuda'))
    assert out.device == torch.device('cuda', 0)
