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


class _CudaBytes:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


def device_bytes(ptr, n, device):
    """Zero-copy uint8 torch view of `n` device bytes at `ptr`."""
    import torch
    return torch.as_tensor(_CudaBytes(ptr, n), device=device)


def upload_replicated(renderer, scene, rank, world, device=None, src=0, frame=None):
    """Replicates the scene on every rank's GPU with ONE host-side upload: rank `src` runs b200rt_upload_scene (BVH
    build, record building, H2D) and broadcasts its device-resident scene blob over NCCL / NVLink; the other ranks
    import the copy (b200rt_import_scene_*) instead of repeating the host work on cores they would have to share.
    Only rank `src` needs `scene`; the others may pass None and `frame` = (width, height, samples, mode).
    Returns the number of bytes of the device blob (what crossed PCIe on `src`)."""
    import torch
    import torch.distributed as dist
    if world == 1:
        renderer.upload(scene)
        return renderer.export_scene()[2] if scene.mode >= 6 else 0
    nb = renderer.SCENE_IMAGE_BYTES
    hdr = torch.zeros(nb, dtype=torch.uint8, device=device)
    ptr = n = None
    if rank == src:
        renderer.upload(scene)
        h, ptr, n = renderer.export_scene()
        hdr.copy_(torch.frombuffer(bytearray(h), dtype=torch.uint8))
    ext = torch.cuda.ExternalStream(renderer.stream(), device=device)
    with torch.cuda.stream(ext):
        dist.broadcast(hdr, src=src)
        h = hdr.cpu().numpy().tobytes()
        if rank != src:
            ptr = renderer.import_scene_begin(h)
            n = int.from_bytes(h[16:24], "little")        # SceneImage::bytes
        dist.broadcast(device_bytes(ptr, n, device), src=src)
    if rank != src:
        renderer.import_scene_end(h, scene, frame)
    return n


class PeerFrame:
    """One process per GPU: rank 0 maps every other rank's accumulation buffer (CUDA IPC handles exchanged once, like
    a communicator) and combines + resolves them with the library's fused peer-memory kernel (b200rt_resolve_peers):
    the radiance buffers cross NVLink once, as loads issued by the resolving kernel -- no separate collective."""

    def __init__(self, renderer, rank, world):
        import torch.distributed as dist
        self.r, self.rank, self.world, self.ptrs = renderer, rank, world, []
        if world == 1:
            return
        handles = [None] * world
        dist.all_gather_object(handles, renderer.accum_ipc_export() if rank != 0 else None)
        if rank == 0:
            self.ptrs = [renderer.ipc_open(h) for h in handles[1:]]

    def resolve(self, total_spp, out=None):
        """Call on every rank after its render: waits for all ranks' GPUs, then rank 0 reads the peers' buffers."""
        import torch.distributed as dist
        self.r.synchronize()
        if self.world > 1:
            dist.barrier()
        img = None
        if self.rank == 0:
            img = self.r.resolve_peers(total_spp, self.ptrs, out) if self.world > 1 else self.r.resolve(total_spp, out)
        if self.world > 1:
            dist.barrier()          # the peers' buffers stay untouched until rank 0 has read them
        return img

    def close(self):
        for p in self.ptrs:
            self.r.ipc_close(p)
        self.ptrs = []


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
