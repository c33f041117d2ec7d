"""Multi-GPU plumbing: the scene is replicated, rank r renders a contiguous range of
global sample indices, and ONE sum-reduce of the float radiance buffers to rank 0 precedes the
resolve.  torch.distributed is used for the collective only (NCCL on GPUs, gloo in CPU tests)."""
import ctypes

import numpy as np


def sample_range(rank, world, samples):
    """Contiguous per-rank slice of [0, samples); slices are disjoint and cover the range."""
    return samples * rank // world, samples * (rank + 1) // world


class _CudaArray:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


def accum_tensor(renderer, device):
    """Zero-copy torch view of the context's accumulation buffer (for NCCL)."""
    import torch
    ptr, n = renderer.accum_device_ptr()
    return torch.as_tensor(_CudaArray(ptr, n), device=device)


def reduce_to_rank0(buf, dst=0):
    """Sum-reduce a torch tensor (CUDA: NCCL over NVLink; CPU: gloo) in place on `dst`."""
    import torch.distributed as dist
    dist.reduce(buf, dst=dst, op=dist.ReduceOp.SUM)
    return buf


def render_distributed(renderer, scene, rank, world, device=None, stream=None):
    """Render this rank's share of scene.samples and reduce; returns the RGB8 frame on rank 0."""
    import torch
    b, e = sample_range(rank, world, scene.samples)
    renderer.clear()
    if e > b:
        renderer.render(b, e)
    if world > 1:
        ext = torch.cuda.ExternalStream(renderer.stream(), device=device)
        with torch.cuda.stream(ext):
            reduce_to_rank0(accum_tensor(renderer, device))
    renderer.synchronize()
    return renderer.resolve(scene.samples) if rank == 0 else None
