"""Why (not) NVLS: reports whether the GPUs advertise multicast objects, whether torch's
symmetric memory gets a multicast mapping, and what NCCL itself selects.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node R --master-addr 127.0.0.1 \
        --master-port 29531 tools/nvls_probe.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl')
    out = {'rank': rank, 'world': dist.get_world_size()}
    try:
        from cuda.bindings import driver as cu
        cu.cuInit(0)
        err, dev = cu.cuDeviceGet(rank)
        err, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        out['CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED'] = int(v)
        err, v = cu.cuDeviceGetAttribute(
            cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev)
        out['HANDLE_TYPE_FABRIC_SUPPORTED'] = int(v)
    except Exception as exc:        # cuda-python layout differs between releases
        out['driver_query_error'] = repr(exc)[:200]
    try:
        from torchgpipe_b200.parallel import SymmetricBuffer, fused_allreduce_
        buf = SymmetricBuffer(1 << 20, device=torch.device('cuda', rank))
        out['symmetric_memory_multicast_ptr'] = hex(buf.multicast_ptr)
        if buf.has_multicast:
            buf.tensor.fill_(float(rank + 1))
            torch.cuda.synchronize()
            dist.barrier()
            fused_allreduce_(buf, 1.0, multicast=True)
            torch.cuda.synchronize()
            want = sum(range(1, dist.get_world_size() + 1))
            out['nvls_allreduce_correct'] = bool((buf.tensor == want).all().item())
            out['nvls_error_flag'] = int(buf.error_flag.item())
    except Exception as exc:
        out['symmetric_memory_error'] = repr(exc)[:300]
    gathered = [None] * dist.get_world_size()
    dist.all_gather_object(gathered, out)
    if rank == 0:
        print(json.dumps(gathered), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
