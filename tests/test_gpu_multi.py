"""The multi-GPU product path on real devices: hf_comm_init -> hf_ray_intersect -> hf_backward -> hf_allreduce_grad on disjoint
image-tile shards, through the C ABI, one process per GPU. The reference has nothing here, so the
check is self-consistency + the oracle: the all-reduced gradient grid of N ranks equals the single-rank grid (1e-4 relative L2, the
north star's gradient tolerance; fp32 atomics make both sides order-dependent) and the oracle's fp64 gradient of the full batch.
Skipped with the reason visible when the box has a single GPU (tests/test_sharding.py covers the host logic over gloo on CPU)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)          # plumbing only: ships the NCCL unique id
    import hfb200
    from hfb200 import workloads as wl
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    hs, img, spp, tile = 256, 128, 8, 16
    cam = wl.CONFIGS["C1"]["camera"]
    h = wl.sine_heights(hs, device=dev)

    _cache = {}

    def trace(shape, pix):
        """ray_intersect + backward on the rays of `pix`; inputs and records are built once per pixel list (no allocation on replay)"""
        key = (id(shape), pix.data_ptr())
        if key not in _cache:
            o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], img, img, spp, seed=3, device=dev, pixels=pix)
            ray = hfb200.Ray3f(o, d, device=dev)
            n = o.shape[1]
            adj = shape.prepare_adjoints(wl.hashed_adjoints(n, seed=1, device=dev, ids=wl.global_ray_ids(pix, spp)), n)
            _cache[key] = (ray, adj, hfb200.SurfaceInteraction3f(n, dev, shape, ("t", "p", "n", "uv")), hfb200.PreliminaryIntersection3f(n, dev, shape))
        ray, adj, si, pi = _cache[key]
        shape.ray_intersect(ray, out=(si, pi))
        shape.backward(ray, pi, adj)

    def new_id():
        ids = [hfb200.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        return ids[0]

    shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.25}, device=dev)
    ptr_before = shape.grad.data_ptr()
    shape.comm_init(rank, world, new_id())
    ptr_after = shape.grad.data_ptr()
    moved = ptr_after != ptr_before                        # the grid moves into ncclMemAlloc memory when NCCL offers it
    my_pix = wl.tile_pixels(img, img, tile, rank, world, device=dev)

    results = {}
    for attempt in ("first communicator", "second hf_comm_init on the same context", "after hf_comm_destroy"):
        if attempt.startswith("second"):
            shape.comm_init(rank, world, new_id())
        elif attempt.startswith("after"):
            from hfb200 import _load, _check
            _check(_load().hf_comm_destroy(shape._ctx))
            shape.comm_init(rank, world, new_id())
        assert shape.grad.data_ptr() == ptr_after, "the gradient grid must stay where the first hf_comm_init put it"
        shape.zero_grad()
        trace(shape, my_pix)
        torch.cuda.synchronize()
        local = shape.grad.double().clone()                # hf_grad()'s NEW pointer must be the buffer the backward kernel wrote
        assert float(local.abs().sum()) > 0, "backward did not land in the buffer hf_grad() returns after hf_comm_init"
        shape.allreduce_grad()
        torch.cuda.synchronize()
        results[attempt] = shape.grad.double().cpu()
        # every rank holds the same bits after the allreduce (replicas stay identical)
        mine = shape.grad.clone().cpu()
        gathered = [None] * world
        dist.all_gather_object(gathered, mine.numpy().tobytes())
        assert all(g == gathered[0] for g in gathered), "ranks disagree on the all-reduced gradient"
        # and the sum of the per-rank grids is what came out
        parts = [None] * world
        dist.all_gather_object(parts, local.cpu().numpy())
        total = np.sum(parts, axis=0)
        assert np.linalg.norm(results[attempt].numpy() - total) <= 1e-5 * np.linalg.norm(total)

    if rank == 0:
        one = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.25}, device=dev)
        all_pix = torch.arange(img * img, dtype=torch.int64, device=dev)
        one.zero_grad()
        trace(one, all_pix)
        torch.cuda.synchronize()
        g1 = one.grad.double().cpu().numpy()
        for k, g in results.items():
            err = np.linalg.norm(g.numpy() - g1) / np.linalg.norm(g1)
            assert err < 1e-4, f"{k}: all-reduced gradient differs from the single-rank gradient by {err:.2e}"
        # the oracle on the whole batch (fp64 accumulation)
        from oracle import hf_oracle as O
        o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], img, img, spp, seed=3, pixels=all_pix.cpu())
        adj = wl.hashed_adjoints(o.shape[1], seed=1, ids=wl.global_ray_ids(all_pix.cpu(), spp))
        ref = O.OracleHeightfield(h.cpu().numpy(), 0.25)
        r = O.Rays(o.numpy().T, d.numpy().T)
        a = {"t": adj["t"].numpy(), "px": adj["p"][0].numpy(), "py": adj["p"][1].numpy(), "pz": adj["p"][2].numpy(),
             "nx": adj["n"][0].numpy(), "ny": adj["n"][1].numpy(), "nz": adj["n"][2].numpy(),
             "uvx": adj["uv"][0].numpy(), "uvy": adj["uv"][1].numpy()}
        rg = ref.backward(r, ref.ray_intersect_preliminary(r), a, literal=False)
        err = np.linalg.norm(results["first communicator"].numpy() - rg) / np.linalg.norm(rg)
        assert err < 1e-4, f"all-reduced gradient vs oracle fp64: {err:.2e}"
        with open(os.path.join(out_dir, "ok"), "w") as f:
            f.write(f"moved={moved} err_vs_oracle={err:.3e}\n")
    dist.barrier()
    del shape
    dist.destroy_process_group()


@pytest.mark.skipif(_ngpu() < 2, reason=f"needs >= 2 GPUs on one box (found {_ngpu()}); host logic is covered by tests/test_sharding.py over gloo")
def test_allreduced_gradient_equals_single_rank_gradient(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    port = 29700 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert os.path.exists(tmp_path / "ok")
    print(open(tmp_path / "ok").read())


def _peer_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import hfb200
    from hfb200 import workloads as wl
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    hs, img, spp, tile = 301, 128, 8, 16                      # 90 601 texels: slabs are not a power of two, the last one is short and
                                                              # ends in the middle of a float4
    cam = wl.CONFIGS["C1"]["camera"]
    h = wl.fbm_heights(hs, octaves=5, seed=0xC1, device=dev)

    _cache = {}

    def trace(shape, pix):
        """ray_intersect + backward on the rays of `pix`; inputs and records are built once per pixel list (no allocation on replay)"""
        key = (id(shape), pix.data_ptr())
        if key not in _cache:
            o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], img, img, spp, seed=3, device=dev, pixels=pix)
            ray = hfb200.Ray3f(o, d, device=dev)
            n = o.shape[1]
            adj = shape.prepare_adjoints(wl.hashed_adjoints(n, seed=1, device=dev, ids=wl.global_ray_ids(pix, spp)), n)
            _cache[key] = (ray, adj, hfb200.SurfaceInteraction3f(n, dev, shape, ("t", "p", "n", "uv")), hfb200.PreliminaryIntersection3f(n, dev, shape))
        ray, adj, si, pi = _cache[key]
        shape.ray_intersect(ray, out=(si, pi))
        shape.backward(ray, pi, adj)

    def everyone_agrees(t, what):
        got = [None] * world
        dist.all_gather_object(got, t.detach().cpu().numpy().tobytes())
        assert all(g == got[0] for g in got), f"ranks disagree on {what}"

    shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.25}, device=dev)
    handles = [None] * world
    dist.all_gather_object(handles, shape.peer_export())
    shape.peer_attach(rank, world, handles)
    my_pix = wl.tile_pixels(img, img, tile, rank, world, device=dev)
    all_pix = torch.arange(img * img, dtype=torch.int64, device=dev)

    # single-rank references (a second, unattached context on this GPU)
    one = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.25}, device=dev)
    one.zero_grad(); trace(one, all_pix); torch.cuda.synchronize()
    g1 = one.grad.double().clone()

    camr = hfb200.perspective_camera(cam["origin"], cam["target"], cam["up"], cam["fov"], img, img, 16, seed=5)
    light = hfb200.directional_light((0.3, 0.2, 1.0))
    target = torch.zeros(img, img, device=dev)
    one.zero_grad(); one.render_step_l2(camr, light, target); torch.cuda.synchronize()
    gr = one.grad.double().clone()
    my_pix32 = my_pix.to(torch.int32)

    other_pix = wl.tile_pixels(img, img, tile, (rank + 1) % world, world, device=dev)

    def peer_step(pix=my_pix):
        shape.peer_zero_grad()
        trace(shape, pix)
        shape.peer_reduce_grad()

    for mode in (1, 0):      # 1: local scatter + reduce-scatter by peer stores (default); 0: scatters go to the owner with remote atomics
        shape.peer_set_mode(mode)
        # ---- (1) == the single-rank gradient, identical bits on every rank; the barrier counters only grow: repeated steps keep working
        for attempt in range(3):
            peer_step()
            torch.cuda.synchronize()
            err = float((shape.grad.double() - g1).norm() / g1.norm())
            assert err < 1e-4, f"mode {mode}: peer reduction vs single rank: {err:.2e} (attempt {attempt})"
            everyone_agrees(shape.grad, f"the reduced gradient (mode {mode})")
        # ---- (1b) the ranks swap their tiles: lines of a partial grid that were sent last time are zero now and are NOT sent (mode 1),
        # so the owner must have cleared its staging area behind the previous reduction
        peer_step(other_pix)
        torch.cuda.synchronize()
        err = float((shape.grad.double() - g1).norm() / g1.norm())
        assert err < 1e-4, f"mode {mode}: peer reduction after the ranks swapped tiles: {err:.2e}"
        everyone_agrees(shape.grad, f"the reduced gradient after a tile swap (mode {mode})")
        # ---- (2) the same step replayed from a CUDA graph (barriers are kernels)
        replay = hfb200.capture(peer_step)
        for _ in range(2):
            replay()
        torch.cuda.synchronize()
        assert float((shape.grad.double() - g1).norm() / g1.norm()) < 1e-4
        everyone_agrees(shape.grad, f"the reduced gradient after a graph replay (mode {mode})")
        del replay
        # ---- (3) the fused render step
        shape.peer_zero_grad()
        shape.render_step_l2(camr, light, target, my_pix32)
        shape.peer_reduce_grad()
        torch.cuda.synchronize()
        assert float((shape.grad.double() - gr).norm() / gr.norm()) < 1e-4, f"mode {mode}: render step"

    # ---- (4) reduce-scatter -> sharded Adam -> all-gather of the heights: every rank ends with the single-rank Adam's field
    params1, paramsN = hfb200.traverse(one), hfb200.traverse(shape)
    opt1, optN = hfb200.Adam(lr=1e-2, params=params1), hfb200.Adam(lr=1e-2, params=paramsN)
    big = g1.abs() > 1e-3 * g1.abs().max()     # Adam's first steps move a texel by ~lr whatever the gradient's size: compare where the
    for it in range(4):                        # gradient is well above its rounding noise
        shape.peer_set_mode(it % 2)
        one.zero_grad(); trace(one, all_pix); opt1.step(); params1.update()
        shape.peer_zero_grad()
        trace(shape, my_pix)
        optN.step(peer=True); paramsN.update()
        torch.cuda.synchronize()
        assert float((shape.heights - one.heights).abs()[big].max()) < 1e-5, f"heights diverge from the single-rank optimiser at step {it}"
        everyone_agrees(shape.heights, "the heights after the sharded Adam step")
        assert np.allclose(shape.bbox(), one.bbox(), atol=1e-5)
    shape.peer_detach()
    shape.zero_grad(); trace(shape, my_pix); torch.cuda.synchronize()        # detached: a plain local context again
    assert float(shape.grad.abs().sum()) > 0
    if rank == 0:
        with open(os.path.join(out_dir, "ok_peer"), "w") as f:
            f.write("ok\n")
    dist.barrier()
    del shape, one
    dist.destroy_process_group()


@pytest.mark.skipif(_ngpu() < 2, reason=f"needs >= 2 GPUs on one box (found {_ngpu()})")
def test_peer_scatter_equals_single_rank(tmp_path):
    """include/hf_b200.h "multi-GPU over peer memory": gradient scatters go to the texel's owner over NVLink, the gather is a store
    kernel, Adam is sharded and fused with the all-gather of the heights. Against one rank tracing every ray."""
    import torch.multiprocessing as mp
    world = 2
    port = 29900 + (os.getpid() % 2000)
    mp.spawn(_peer_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert os.path.exists(tmp_path / "ok_peer")


def _regroup_worker(rank, world, port, out_dir):
    """Peers attach as a group of `world`, detach, and all but the last rank form a SMALLER group on the same contexts. The barrier
    state (counter and epoch in each rank's peer-mapped counter block) counts in units of the group size, so hf_peer_export must
    hand out a clean one: with the leftovers of the larger group the smaller one walks through its barriers without waiting."""
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import hfb200
    from hfb200 import workloads as wl
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    hs, img, spp, tile = 257, 128, 8, 16
    cam = wl.CONFIGS["C1"]["camera"]
    h = wl.fbm_heights(hs, octaves=5, seed=0xC1, device=dev)

    def trace(shape, pix):
        o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], img, img, spp, seed=3, device=dev, pixels=pix)
        ray = hfb200.Ray3f(o, d, device=dev)
        n = o.shape[1]
        adj = shape.prepare_adjoints(wl.hashed_adjoints(n, seed=1, device=dev, ids=wl.global_ray_ids(pix, spp)), n)
        si, pi = shape.ray_intersect(ray, return_prelim=True, fields=("t", "p", "n", "uv"))
        shape.backward(ray, pi, adj)

    shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.25}, device=dev)
    one = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.25}, device=dev)
    one.zero_grad(); trace(one, torch.arange(img * img, dtype=torch.int64, device=dev)); torch.cuda.synchronize()
    g1 = one.grad.double().clone()

    def run_group(n_ranks, steps):
        """every rank takes part in the handle exchange; ranks >= n_ranks stay out of the group"""
        handles = [None] * world
        dist.all_gather_object(handles, shape.peer_export())
        if rank < n_ranks:
            shape.peer_attach(rank, n_ranks, handles[:n_ranks])
            pix = wl.tile_pixels(img, img, tile, rank, n_ranks, device=dev)
            for it in range(steps):
                shape.peer_zero_grad(); trace(shape, pix); shape.peer_reduce_grad()
                torch.cuda.synchronize()
                err = float((shape.grad.double() - g1).norm() / g1.norm())
                assert err < 1e-4, f"group of {n_ranks}, step {it}: reduced gradient vs single rank {err:.2e}"
        dist.barrier()
        if rank < n_ranks:
            shape.peer_detach()
        dist.barrier()

    run_group(world, 3)           # counters now stand at a multiple of `world`
    run_group(world - 1, 3)       # a smaller group on the same contexts
    run_group(world, 2)           # and a larger one again
    if rank == 0:
        with open(os.path.join(out_dir, "ok_regroup"), "w") as f:
            f.write("ok\n")
    dist.barrier()
    del shape, one
    dist.destroy_process_group()


@pytest.mark.skipif(_ngpu() < 3, reason=f"needs >= 3 GPUs on one box (found {_ngpu()})")
def test_peer_groups_of_different_sizes_on_the_same_contexts(tmp_path):
    import torch.multiprocessing as mp
    world = min(_ngpu(), 4)
    port = 30900 + (os.getpid() % 2000)
    mp.spawn(_regroup_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert os.path.exists(tmp_path / "ok_regroup")
