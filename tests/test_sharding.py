"""Host-side multi-GPU logic on CPU: image-tile sharding and the gradient sum, with a world_size-2
gloo group standing in for NCCL. Each rank runs the ORACLE on its tiles (the checker, not the product) and the
all-reduced gradient must equal the single-rank gradient."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hfb200 import workloads as wl


@pytest.mark.parametrize("world,tile,w,h", [(2, 32, 256, 256), (8, 32, 1024, 1024), (3, 16, 100, 70), (4, 64, 96, 96)])
def test_tile_pixels_partition_the_image(world, tile, w, h):
    seen = torch.zeros(w * h, dtype=torch.int32)
    counts = []
    for r in range(world):
        p = wl.tile_pixels(w, h, tile, r, world)
        seen[p] += 1
        counts.append(p.numel())
    assert bool((seen == 1).all())
    assert max(counts) - min(counts) <= tile * tile * 2


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import hf_oracle as O
    O.set_threads(2)
    h = wl.sine_heights(64).numpy()
    c = O.OracleHeightfield(h, 0.25)
    cam = wl.CONFIGS["C1"]["camera"]
    pix = wl.tile_pixels(64, 64, 16, rank, world)
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 64, 64, 4, seed=7, pixels=pix)
    r = O.Rays(o.numpy().T, d.numpy().T)
    pi = c.ray_intersect_preliminary(r)
    # adjoints keyed by GLOBAL ray id so that the shards see what the single-rank run sees
    gid = (pix.repeat_interleave(4) * 4 + torch.arange(4).repeat(pix.numel()))
    a_t = (2 * wl.hash_uniform(1, gid, 0) - 1).numpy()
    g = torch.from_numpy(c.backward(r, pi, {"t": a_t}, literal=False))
    dist.all_reduce(g, op=dist.ReduceOp.SUM)
    if rank == 0:
        torch.save(g, out)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_gradient_sums_to_the_single_rank_gradient(tmp_path):
    out = str(tmp_path / "g.pt")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    g2 = torch.load(out).numpy()
    from oracle import hf_oracle as O
    h = wl.sine_heights(64).numpy()
    c = O.OracleHeightfield(h, 0.25)
    cam = wl.CONFIGS["C1"]["camera"]
    pix = torch.arange(64 * 64)
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 64, 64, 4, seed=7, pixels=pix)
    r = O.Rays(o.numpy().T, d.numpy().T)
    pi = c.ray_intersect_preliminary(r)
    gid = (pix.repeat_interleave(4) * 4 + torch.arange(4).repeat(pix.numel()))
    g1 = c.backward(r, pi, {"t": (2 * wl.hash_uniform(1, gid, 0) - 1).numpy()}, literal=False)
    assert np.allclose(g1, g2, rtol=1e-12, atol=1e-12 * np.abs(g1).max())
