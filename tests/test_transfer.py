import pytest
import torch

from torchgpipe_b200.copy import Copy, Wait
from torchgpipe_b200.stream import CPUStream, current_stream, new_stream


def test_copy_wait_cpu_cpu():
    x = torch.rand(3, requires_grad=True)
    y, = Copy.apply(CPUStream, CPUStream, x)
    y, = Wait.apply(CPUStream, CPUStream, y)
    y.sum().backward()
    assert torch.equal(x.grad, torch.ones(3))


def test_copy_moves_every_tensor_of_a_tuple():
    a = torch.rand(2, requires_grad=True)
    b = torch.rand(3, requires_grad=True)
    c, d = Copy.apply(CPUStream, CPUStream, a, b)
    assert c.grad_fn is d.grad_fn
    (c.sum() + 2 * d.sum()).backward()
    assert torch.equal(a.grad, torch.ones(2)) and torch.equal(b.grad, torch.full((3,), 2.))


def test_wait_shares_one_node_for_all_tensors():
    a = torch.rand(1, requires_grad=True)
    b = torch.rand(1, requires_grad=True)
    a, b = Wait.apply(CPUStream, CPUStream, a, b)
    assert a.grad_fn is b.grad_fn
    assert a.grad_fn.__class__.__name__ == 'WaitBackward'


def _check_copy_wait(prev_stream, next_stream, cuda_sleep=None):
    from torchgpipe_b200.stream import get_device, use_stream
    device = get_device(prev_stream)
    with use_stream(prev_stream):
        if cuda_sleep is not None and device.type == 'cuda':
            cuda_sleep(0.3)
        x = torch.ones(100, device=device, requires_grad=True)
    y, = Copy.apply(prev_stream, next_stream, x)
    y, = Wait.apply(prev_stream, next_stream, y)
    with use_stream(next_stream):
        assert torch.allclose(y.sum(), torch.tensor(100.0, device=get_device(next_stream)))
        y.norm().backward()
    with use_stream(prev_stream):
        assert torch.allclose(x.grad.sum(), torch.tensor(10.0, device=device))


@pytest.mark.gpu
def test_copy_wait_cuda_cuda(cuda_sleep):
    _check_copy_wait(current_stream(torch.device('cuda')), new_stream(torch.device('cuda')),
                     cuda_sleep)


@pytest.mark.gpu
def test_copy_wait_cpu_cuda(cuda_sleep):
    _check_copy_wait(CPUStream, current_stream(torch.device('cuda')), cuda_sleep)


@pytest.mark.gpu
def test_copy_wait_cuda_cpu(cuda_sleep):
    _check_copy_wait(current_stream(torch.device('cuda')), CPUStream, cuda_sleep)
