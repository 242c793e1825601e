import pytest
import torch

from torchgpipe_b200.microbatch import Batch, check, gather, scatter


def test_atomic_batch():
    x = torch.tensor(42)
    b = Batch(x)
    assert b.atomic
    assert b.tensor is x
    with pytest.raises(AttributeError):
        b.tensors
    assert list(b) == [x]
    assert len(b) == 1
    assert b[0] is x
    with pytest.raises(IndexError):
        b[1]


def test_tuple_batch():
    x, y = torch.tensor(1), torch.tensor(2)
    b = Batch((x, y))
    assert not b.atomic
    with pytest.raises(AttributeError):
        b.tensor
    assert b.tensors == (x, y)
    assert list(b) == [x, y]
    assert len(b) == 2
    assert b[1] is y


def test_call_wraps():
    a = Batch(torch.tensor(42))
    b = Batch((torch.tensor(1), torch.tensor(2)))
    assert a.call(lambda t: t).atomic
    assert not b.call(lambda t: t).atomic
    assert not a.call(lambda t: (t, t)).atomic


def test_setitem_index():
    a = Batch(torch.tensor(42))
    a[0] = torch.tensor(0)
    assert a.atomic and a[0].item() == 0
    with pytest.raises(IndexError):
        a[1] = torch.tensor(1)

    b = Batch((torch.tensor(1), torch.tensor(2)))
    b[0] = torch.tensor(0)
    assert not b.atomic
    assert [t.item() for t in b] == [0, 2]


def test_setitem_slice():
    a = Batch(torch.tensor(42))
    a[:] = (torch.tensor(0),)
    assert a.atomic and a.tensor.item() == 0
    with pytest.raises(IndexError):
        a[:] = (torch.tensor(0), torch.tensor(1))
    with pytest.raises(NotImplementedError):
        a[0:1] = (torch.tensor(0),)

    b = Batch((torch.tensor(1), torch.tensor(2)))
    b[:] = (torch.tensor(3), torch.tensor(4), torch.tensor(5))
    assert len(b) == 3


def test_check():
    check(torch.tensor(1))
    check((torch.tensor(1), torch.tensor(2)))
    with pytest.raises(TypeError):
        check(42)
    with pytest.raises(TypeError):
        check((torch.tensor(1), 'x'))


def test_scatter_gather_tensor():
    x = torch.arange(12.).view(6, 2)
    parts = scatter(x, 3)
    assert [p.tensor.size(0) for p in parts] == [2, 2, 2]
    assert torch.equal(gather(parts), x)


def test_scatter_uneven_and_small():
    x = torch.zeros(7, 1)
    assert [p.tensor.size(0) for p in scatter(x, 4)] == [2, 2, 2, 1]
    assert len(scatter(torch.zeros(2, 1), 8)) == 2      # fewer micro-batches than chunks


def test_scatter_gather_tuple():
    a, b = torch.zeros(4, 1), torch.ones(4, 2)
    parts = scatter((a, b), 2)
    assert len(parts) == 2 and not parts[0].atomic
    assert parts[0].tensors[0].shape == (2, 1) and parts[0].tensors[1].shape == (2, 2)
    ga, gb = gather(parts)
    assert torch.equal(ga, a) and torch.equal(gb, b)
