"""N > 1 host logic on CPU: world_size-2 gloo ranks each take their sample range, the partial
radiance sums are reduced to rank 0 and must equal the single-range result (Philox stream is keyed
by the GLOBAL sample index, so the split does not change the samples -- only float summation order).
The per-rank 'renderer' here is the oracle port (test infrastructure) standing in for a GPU."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import oracle_util as ou
    import raytracing_course_hw_b200 as rt
    from raytracing_course_hw_b200.dist import reduce_to_rank0, sample_range

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = ou.load_scene(os.path.join(ou.SCENES, "hw8/sphere_emissive.gltf"), width=48, height=32, samples=6)
    b, e = sample_range(rank, world, sc.samples)
    part = ou.port_render(sc, 1, spp_begin=b, spp_end=e)["accum"]
    buf = torch.from_numpy(part.copy())
    reduce_to_rank0(buf)
    if rank == 0:
        np.save(out_path, buf.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_sample_ranges_partition():
    from raytracing_course_hw_b200.dist import sample_range
    for world in (1, 2, 3, 4, 8):
        for spp in (1, 5, 8, 1024):
            r = [sample_range(k, world, spp) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))


def test_two_rank_gloo_reduce_equals_single_range(tmp_path):
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util as ou
    import raytracing_course_hw_b200 as rt

    out = str(tmp_path / "sum.npy")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    sc = ou.load_scene(os.path.join(ou.SCENES, "hw8/sphere_emissive.gltf"), width=48, height=32, samples=6)
    full = ou.port_render(sc, 1)["accum"]
    assert np.allclose(got, full, rtol=1e-5, atol=1e-6)
