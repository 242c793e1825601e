import pytest
import torch

from torchgpipe_b200.stream import (CPUStream, StreamPool, current_stream, default_stream,
                                    get_device, is_cuda, new_stream, record_stream, use_device,
                                    use_stream, wait_stream)


class TestCPU:
    def test_placeholders(self):
        cpu = torch.device('cpu')
        assert new_stream(cpu) is CPUStream
        assert current_stream(cpu) is CPUStream
        assert default_stream(cpu) is CPUStream
        assert not is_cuda(CPUStream)
        assert get_device(CPUStream).type == 'cpu'

    def test_contexts_are_noops(self):
        with use_device(torch.device('cpu')):
            pass
        with use_stream(CPUStream):
            pass

    def test_wait_and_record_are_noops(self):
        wait_stream(CPUStream, CPUStream)
        record_stream(torch.ones(1), CPUStream)

    def test_stream_pool_round_robin(self):
        pool = StreamPool([torch.device('cpu')] * 2, chunks=8, width=3)
        assert len(pool) == 2
        assert pool[0][0] is CPUStream and pool[1][7] is CPUStream
        assert len(pool[0]) == 3


@pytest.mark.gpu
class TestCUDA:
    def test_new_current_default(self):
        d = torch.device('cuda')
        s = new_stream(d)
        assert isinstance(s, torch.cuda.Stream) and s != torch.cuda.default_stream()
        assert current_stream(d) == torch.cuda.current_stream()
        assert default_stream(d) == torch.cuda.default_stream()
        assert is_cuda(s) and get_device(s).type == 'cuda'

    def test_use_stream(self):
        s = new_stream(torch.device('cuda'))
        assert torch.cuda.current_stream() != s
        with use_stream(s):
            assert torch.cuda.current_stream() == s

    def test_wait_stream_cuda_cuda(self, cuda_sleep):
        source = current_stream(torch.device('cuda'))
        target = new_stream(torch.device('cuda'))
        with use_stream(target):
            cuda_sleep(0.3)
        wait_stream(source, target)
        assert not target.query() or source.query()
        torch.cuda.synchronize()

    def test_wait_stream_cpu_cuda_blocks_host(self, cuda_sleep):
        target = new_stream(torch.device('cuda'))
        with use_stream(target):
            cuda_sleep(0.3)
        wait_stream(CPUStream, target)
        assert target.query()

    def test_record_stream_defers_reuse(self, cuda_sleep):
        side = new_stream(torch.device('cuda'))
        x = torch.rand(1, device='cuda')
        ptr = x.data_ptr()
        with use_stream(side):
            cuda_sleep(0.3)
        record_stream(x, side)
        del x
        y = torch.rand(1, device='cuda')
        assert y.data_ptr() != ptr          # block still owned by the side stream
        side.synchronize()
        del y
        z = torch.rand(1, device='cuda')
        assert z.data_ptr() == ptr

    def test_record_stream_on_shifted_view(self, cuda_sleep):
        side = new_stream(torch.device('cuda'))
        x = torch.rand(2, device='cuda')
        base_ptr = x.data_ptr()
        view = x[1:]
        with use_stream(side):
            cuda_sleep(0.3)
        record_stream(view, side)           # must protect the whole block
        del x, view
        y = torch.rand(2, device='cuda')
        assert y.data_ptr() != base_ptr
        torch.cuda.synchronize()

    def test_stream_pool_has_fixed_width(self):
        pool = StreamPool([torch.device('cuda')], chunks=1000, width=4)
        streams = {id(pool[0][i]) for i in range(1000)}
        assert len(streams) == 4
