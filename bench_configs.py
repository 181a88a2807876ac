"""bench_configs.py -- BASELINE.json configs[3] (C4) and configs[4] (C5) as the file states them, behind `bench.py --config C4|C5`.

C4  16384^2 fBm heightfield (1 GiB of fp32 heights, seed 0xC4, max_height 0.15), grazing camera, 2048^2 image x 8 spp, path tracing
    depth 3 with gradients: one optimisation step = zero_grad -> hf_render_step_l2(max_depth = 3: forward image, L2 adjoint image,
    replayed adjoint) [-> ncclAllReduce of the 1 GiB gradient grid] -> Adam -> pyramid rebuild, at 1 / 2 / 4 / 8 GPUs (image tiles
    dealt round-robin). This is the one configuration whose texels and pyramid nodes do not fit L2.
C5  ray-count sweep: 2^20 .. 2^30 incoherent rays (origins uniform on the upper hemisphere of radius 3, targets uniform on the field)
    x heightfield sizes 512^2 .. 16384^2 (fBm seed 0xC5), intersect-only and forward + backward, in chunks of <= 2^26 rays; with N GPUs
    the chunks are dealt round-robin and the gradient grids are summed by one allreduce at the end.

Both print ONE JSON line in bench.py's contract (metric / value / unit / n_gpus / ms_per_step / roofline / cpu_baseline / e2e ...).
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))


def _peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def _traffic(kernel, units):
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[kernel]
        return float(t["dram_bytes"]) * units / float(t["rays"])
    except Exception:
        return None


def _env(rank_dev):
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the heightfield path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    return rank, world, local, dev, dist


def _timers(dev, world, dist):
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        s, e = ev(), ev()
        s.record()
        for _ in range(steps):
            fn()
        e.record()
        barrier()
        ms = torch.tensor([s.elapsed_time(e)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())
    return ev, barrier, timed


# =====================================================================================================================
# C4
# =====================================================================================================================
C4_LIGHT = dict(towards=(0.3, 0.2, 1.0), irradiance=2.0, albedo=0.8, shadow_rays=True, max_depth=3)


def c4_sizes(a):
    hs = a.heights if a.heights != 4096 else 16384          # bench.py's defaults are C3's: C4 has its own unless overridden
    img = a.image if a.image != 1024 else 2048
    spp = a.spp if a.spp != 64 else 8
    return hs, img, spp


def c4_workload(a):
    hs, img, spp = c4_sizes(a)
    full = (hs, img, spp) == (16384, 2048, 8)
    return (("C4: " if full else "reduced C4: ") + f"{hs}^2 fBm heightfield (10 octaves, seed 0xC4, max_height 0.15), grazing camera, {img}^2 image x {spp} spp, "
            "path tracing depth 3 with gradients (L2 step)")


def run_c4(a, sampler_cls):
    import hfb200
    from hfb200 import workloads as wl
    rank, world, local, dev, dist = _env(None)
    ev, barrier, timed = _timers(dev, world, dist)
    hs, img, spp = c4_sizes(a)
    cfg = wl.CONFIGS["C4"]
    cm = cfg["camera"]
    h = wl.fbm_heights(hs, octaves=10, seed=0xC4, device=dev)
    shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": cfg["max_height"]}, device=dev)
    cam = hfb200.perspective_camera(cm["origin"], cm["target"], cm["up"], cm["fov"], img, img, spp, seed=0)
    light = hfb200.directional_light(C4_LIGHT["towards"], irradiance=C4_LIGHT["irradiance"], albedo=C4_LIGHT["albedo"],
                                     shadow_rays=C4_LIGHT["shadow_rays"], max_depth=C4_LIGHT["max_depth"])
    pix = None if world == 1 else wl.tile_pixels(img, img, a.tile, rank, world, device=dev).to(torch.int32)
    npix_local = img * img if pix is None else pix.numel()
    n_total = img * img * spp
    # target: the same scene over another fBm field (seed 0x7B); every rank renders the pixels it owns
    tgt = hfb200.load_dict({"type": "heightfield", "heights": wl.fbm_heights(hs, octaves=10, seed=0x7B, device=dev), "max_height": cfg["max_height"]}, device=dev)
    target = tgt.render_forward(cam, light, pix)
    del tgt
    torch.cuda.empty_cache()
    if world > 1:
        ids = [None] * world
        dist.all_gather_object(ids, hfb200.comm_unique_id() if rank == 0 else None)
        shape.comm_init(rank, world, ids[0])
    params = hfb200.traverse(shape)
    opt = hfb200.Adam(lr=1e-4, params=params)
    image = torch.zeros_like(target)
    adj_img = torch.zeros_like(target)
    loss = torch.zeros(1, dtype=torch.float64, device=dev)

    def step():
        shape.zero_grad()
        loss.zero_()
        shape.render_step_l2(cam, light, target, pix, image=image, loss=loss)
        if world > 1:
            shape.allreduce_grad()
        opt.step()
        params.update()

    # per-kernel legs (eager, CUDA events): forward path kernel, replayed adjoint kernel, the collective, Adam + rebuild
    def leg(fn, reps):
        fn(); torch.cuda.synchronize()
        return timed(fn, reps) / reps
    reps = max(2, min(a.steps, 5))
    ms_fwd = leg(lambda: shape.render_forward(cam, light, pix, image=image), reps)
    adj_img.copy_(2 * (image - target) / (img * img))
    shape.stats(reset=True)
    shape.render_forward(cam, light, pix, image=image)
    seg_fwd = shape.stats(reset=True)["segments"]
    ms_bwd = leg(lambda: shape.render_backward(cam, light, adj_img, pix), reps)
    shape.stats(reset=True)
    ms_allreduce = leg(shape.allreduce_grad, reps) if world > 1 else 0.0
    ms_adam = leg(lambda: (opt.step(), params.update()), reps)
    params["heights"] = h
    params.update()
    opt.reset()

    for _ in range(max(a.warmup, 3)):
        step()
    torch.cuda.synchronize()
    loss0 = float(loss.item())
    sampler = sampler_cls(local)
    if rank == 0:
        sampler.start(); time.sleep(0.3)
    run_step, launch = step, "eager"
    if os.environ.get("HF_BENCH_GRAPH", "1") != "0":
        try:
            run_step = hfb200.capture(step); launch = "cuda-graph"
        except Exception as exc:      # noqa: BLE001
            print(f"bench.py: capture refused ({type(exc).__name__}: {exc}); eager launches", file=sys.stderr, flush=True)
            torch.cuda.synchronize()
    for _ in range(2):
        run_step()
    params["heights"] = h           # the timed steps start from the initial field (Adam roughens it step by step: walks get longer)
    params.update()
    opt.reset()
    ms_step = timed(run_step, a.steps) / a.steps
    torch.cuda.synchronize()
    loss1 = float(loss.item())
    grad_checksum = float(shape.grad.double().abs().sum().item())
    collective = f"ncclAllReduce(grad {hs * hs * 4 / 2**20:.0f} MiB fp32, sum) per step, Adam replicated" if world > 1 else "none (1 GPU)"
    nccl_step, peer = {"ms_per_step": ms_step, "launch": launch}, None
    if world > 1 and os.environ.get("HF_BENCH_COLLECTIVE", "auto") != "nccl":
        # the same step over peer memory (include/hf_b200.h): partial gradient slabs to their owners by NVLink stores (zero lines are not
        # sent), the owner sums them, applies Adam to ITS slab and stores the new heights into every rank's grid. Checked against
        # the NCCL step on the first step from the same field before it may become the headline.
        try:
            def one_step(fn):
                params["heights"] = h; params.update(); opt.reset()
                fn(); torch.cuda.synchronize()
                return shape.heights.clone(), shape.grad.clone()
            h_nccl, g_nccl = one_step(step)
            handles = [None] * world
            dist.all_gather_object(handles, shape.peer_export())
            shape.peer_attach(rank, world, handles)

            def step_peer():
                shape.peer_zero_grad()
                loss.zero_()
                shape.render_step_l2(cam, light, target, pix, image=image, loss=loss)
                opt.step(peer=True)                      # reduce-scatter -> sharded Adam -> all-gather of the heights
                params.update()
            h_peer, _ = one_step(step_peer)
            big = g_nccl.abs() > 1e-3 * g_nccl.abs().max()       # Adam's first step moves a texel by lr whatever the gradient's size: compare
            dev_h = (h_peer - h_nccl).abs()[big].max() if bool(big.any()) else torch.zeros((), device=dev)      # where it is above rounding noise
            dist.all_reduce(dev_h, op=dist.ReduceOp.MAX)
            del h_nccl, g_nccl, h_peer, big
            run_peer, peer_launch = step_peer, "eager"
            if os.environ.get("HF_BENCH_GRAPH", "1") != "0":
                run_peer = hfb200.capture(step_peer); peer_launch = "cuda-graph"
            for _ in range(2):
                run_peer()
            params["heights"] = h; params.update(); opt.reset()
            ms_peer = timed(run_peer, a.steps) / a.steps
            peer = {"ms_per_step": ms_peer, "launch": peer_launch, "heights_max_abs_diff_vs_nccl_step": float(dev_h), "lr": 1e-4,
                    "what": "zero_grad -> path-traced L2 step -> hf_peer_adam_step (partial slabs to their owners over NVLink, owner sums + Adam on its slab, "
                            "new heights stored into every rank's grid) -> pyramid rebuild"}
            if float(dev_h) < 1e-5 and ms_peer < ms_step:
                ms_step, launch = ms_peer, peer_launch
                collective = ("reduce-scatter -> sharded Adam -> all-gather of the heights as kernels over peer memory (hf_peer_adam_step); "
                              "the ncclAllReduce + replicated Adam step is timed beside it (extras.nccl_allreduce_step)")
        except Exception as exc:      # noqa: BLE001
            print(f"bench.py: peer-memory path unavailable ({type(exc).__name__}: {exc}); NCCL allreduce only", file=sys.stderr, flush=True)
            peer = None
    clocks = sampler.stop() if rank == 0 else None

    # traversal counters of the PRIMARY rays (canonical traversal, counting kernel) on a 1/64 pixel sample: B_trav per segment of the model
    sub = torch.arange(0, img * img, 64, dtype=torch.int32, device=dev) if pix is None else pix[::64].contiguous()
    ray = shape.render_rays(cam, sub)
    pi = hfb200.PreliminaryIntersection3f(len(ray), dev, shape)
    shape.stats(reset=True)
    shape.ray_intersect_preliminary(ray, counted=True, out=pi)
    st = shape.stats(reset=True)
    n_node, n_leaf, hit_frac = st["nodes"] / st["rays"], st["leaves"] / st["rays"], st["hits"] / st["rays"]
    b_trav = 8.0 * n_node + 16.0 * n_leaf
    seg = torch.tensor([float(seg_fwd)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(seg)
    seg_total = float(seg.item())
    del ray, pi

    e2e = None
    if world == 1 and not a.no_e2e:
        pin = lambda x: x.detach().to("cpu").contiguous().pin_memory()
        h_t = pin(target); h_img = torch.empty(img, img, dtype=torch.float32).pin_memory()
        params["heights"] = h; params.update(); opt.reset()

        def host_step():
            shape.render_step_l2_host(cam, light, h_t, h_image=h_img)
            opt.step(); params.update()
            torch.cuda.current_stream().synchronize()
        host_step()
        ms_e = timed(host_step, reps) / reps
        e2e = {"value": n_total / (ms_e * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h_t.numel() * 4, "d2h_bytes_per_step": h_img.numel() * 4 + 8,
               "ms_per_step": ms_e, "api": "hf_render_step_l2_host (pinned host target in, image + loss out) + hf_adam_step + hf_parameters_changed"}

    cpu = None
    if world == 1 and rank == 0 and not a.no_cpu:
        cpu = c4_cpu(a, h.cpu().numpy(), cam, light, 1, 0)
    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return
    peak, peak_src = _peak()
    seg_local = float(seg_fwd)
    bytes_fwd = seg_local * (b_trav + 16.0) + npix_local * 4.0
    line = {
        "metric": "Mrays/s primary rays of a depth-3 path-traced inverse-rendering step (16384^2 heights, 2048^2 image x 8 spp)",
        "value": n_total / (ms_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": c4_workload(a), "rays_per_step": n_total, "segments_per_step": seg_total * 2,
                   "segments_what": "camera + bounce + visibility segments, traced once by the forward kernel and once more by the replayed adjoint",
                   "sharding": f"{a.tile}x{a.tile}-pixel image tiles, round-robin", "launch": launch,
                   "collective": collective,
                   "l2": f"heights {hs * hs * 4 / 2**30:.2f} GiB + quantised pyramid {hs * hs * 4 * 4 / 3 / 2**30:.2f} GiB exceed the 126 MB L2; no explicit flush",
                   "hit_fraction_primary": hit_frac},
        "clocks": clocks, "gpu_launches": 6 * a.steps,
        "kernels_per_step": ["hf_render_path_kernel<fwd>", "hf_l2_adjoint_kernel", "hf_render_path_kernel<bwd>", "hf_adam_prep_kernel", "hf_adam_kernel", "hf_pyramid_* (4)"],
        "roofline": {"bound": "hbm", "kernel": "hf_render_path_kernel<false>", "kernel_ms": ms_fwd, "achieved": bytes_fwd / (ms_fwd * 1e-3) / 1e9, "peak": peak,
                     "unit": "GB/s", "frac": bytes_fwd / (ms_fwd * 1e-3) / 1e9 / peak, "peak_source": peak_src, "algorithmic_bytes": bytes_fwd,
                     "model": f"segments x ({b_trav:.1f} B traversal [8 B x {n_node:.2f} nodes + 16 B x {n_leaf:.2f} leaves, canonical traversal counted on the primary rays] "
                              f"+ 16 B hit cell) + 4 B per pixel; {seg_local / (npix_local * spp):.2f} segments per camera sample",
                     "traffic": _traffic("hf_render_path_kernel<false>@C4", npix_local * spp), "traffic_unit": "bytes per launch (ncu dram__bytes_read + write, profiles/traffic.json)",
                     "backward": {"kernel": "hf_render_path_kernel<true>", "kernel_ms": ms_bwd,
                                  "achieved": (seg_local * (b_trav + 16.0 + 12.0)) / (ms_bwd * 1e-3) / 1e9,
                                  "frac": (seg_local * (b_trav + 16.0 + 12.0)) / (ms_bwd * 1e-3) / 1e9 / peak}},
        "extras": {"forward_kernel_ms": ms_fwd, "adjoint_kernel_ms": ms_bwd, "allreduce_ms": ms_allreduce, "adam_and_rebuild_ms": ms_adam,
                   "nccl_allreduce_step": nccl_step if world > 1 else None, "peer_adam_step": peer,
                   "loss_first_step_local": loss0, "loss_last_step_local": loss1, "grad_abs_sum": grad_checksum,
                   "mean_nodes_per_primary_ray": n_node, "mean_leaves_per_primary_ray": n_leaf},
    }
    if e2e is not None:
        line["e2e"] = e2e
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def c4_cpu(a, h_np, cam, light, steps, warmup):
    """The oracle's path tracer + PRB adjoint on all host cores, on a bounded pixel sample of the same step."""
    from oracle import hf_oracle as O
    O.build()
    O.set_threads(os.cpu_count() or 1)
    hs, img, spp = c4_sizes(a)
    c = O.OracleHeightfield(h_np, 0.15)
    c._pyramid()
    npix = img * img
    want = max(256, min(npix, (a.cpu_sample // 16) // spp))           # ~1/16 of bench.py's ray sample: three segments + visibility rays per sample
    pixels = np.arange(0, npix, max(1, npix // want), dtype=np.uint32)[:want]
    adj = np.linspace(-1, 1, npix).astype(np.float32)

    def one():
        t0 = time.perf_counter()
        _, samples = O.render_path_forward(c, cam, light, pixels)
        O.render_path_backward(c, cam, light, adj, pixels, samples=samples)
        return time.perf_counter() - t0
    for _ in range(warmup):
        one()
    ts = [one() for _ in range(max(1, steps))]
    n = pixels.size * spp
    return {"value": n / min(ts) / 1e6, "unit": "Mrays/s", "cores": O.get_threads(), "kind": "port",
            "sample": f"{n} camera samples ({pixels.size} pixels, every k-th, all {spp} spp) through oracle render_path_forward + render_path_backward, best of {len(ts)}",
            "ms": 1e3 * min(ts)}


def reference_c4(a):
    """--impl reference --config C4: the CPU arm alone (rank 0)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import hfb200
    from hfb200 import workloads as wl
    hs, img, spp = c4_sizes(a)
    cm = wl.CONFIGS["C4"]["camera"]
    h = wl.fbm_heights(hs, octaves=10, seed=0xC4).numpy()
    cam = hfb200.perspective_camera(cm["origin"], cm["target"], cm["up"], cm["fov"], img, img, spp, seed=0)
    light = hfb200.directional_light(C4_LIGHT["towards"], irradiance=C4_LIGHT["irradiance"], albedo=C4_LIGHT["albedo"],
                                     shadow_rays=C4_LIGHT["shadow_rays"], max_depth=C4_LIGHT["max_depth"])
    cpu = c4_cpu(a, h, cam, light, max(1, a.steps), min(a.warmup, 1))
    line = {"impl": "reference", "metric": "Mrays/s primary rays of a depth-3 path-traced inverse-rendering step (16384^2 heights, 2048^2 image x 8 spp)",
            "value": cpu["value"], "unit": "Mrays/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": cpu["ms"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": c4_workload(a), "sample": cpu["sample"]},
            "cpu_baseline": cpu, "e2e": {"value": cpu["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# =====================================================================================================================
# C5
# =====================================================================================================================
C5_SIZES = (512, 1024, 2048, 4096, 8192, 16384)
C5_COUNTS = tuple(1 << k for k in (20, 22, 24, 26, 28, 30))
C5_CHUNK = 1 << 26


def run_c5(a, sampler_cls):
    import hfb200
    from hfb200 import workloads as wl
    rank, world, local, dev, dist = _env(None)
    ev, barrier, timed = _timers(dev, world, dist)
    sizes = [int(x) for x in os.environ.get("HF_C5_SIZES", ",".join(map(str, C5_SIZES))).split(",")]
    counts = [1 << int(x) for x in os.environ.get("HF_C5_LOG2", "20,22,24,26,28,30").split(",")]
    chunk = min(C5_CHUNK, 1 << int(os.environ.get("HF_C5_CHUNK_LOG2", "26")))
    fields = ("t", "p", "n", "uv")
    cells = []
    e2e = None
    sampler = sampler_cls(local)
    if rank == 0:
        sampler.start()
    t_wall = time.perf_counter()
    comm_id = None
    for hs in sizes:
        h = wl.fbm_heights(hs, octaves=8, seed=0xC5, device=dev)
        shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
        del h
        shape.set_ray_coherence(os.environ.get("HF_C5_COHERENT", "0") == "1")      # hemisphere rays: unrelated -> the lane-pool kernel
        if world > 1:
            ids = [None] * world
            dist.all_gather_object(ids, hfb200.comm_unique_id() if rank == 0 else None)
            shape.comm_init(rank, world, ids[0])
            shape.zero_grad(); shape.allreduce_grad(); torch.cuda.synchronize()      # communicator set-up (lazy in NCCL) outside every timed region
        nmax = min(chunk, max(counts))
        si = hfb200.SurfaceInteraction3f(nmax, dev, shape, fields)
        pi = hfb200.PreliminaryIntersection3f(nmax, dev, shape)
        for n in counts:
            nchunks = (n + chunk - 1) // chunk
            mine = list(range(rank, nchunks, world))
            ms_i = ms_fb = 0.0
            hits = 0.0
            shape.zero_grad()
            for ci in mine:
                start, m = ci * chunk, min(chunk, n - ci * chunk)
                o, d = wl.hemisphere_rays(m, seed=5, device=dev, start=start)
                ray = hfb200.Ray3f(o, d, device=dev)
                adj = shape.prepare_adjoints(wl.hashed_adjoints(m, seed=1, device=dev, ids=torch.arange(start, start + m, dtype=torch.int64, device=dev)), m)
                si_v, pi_v = (si, pi) if m == nmax else (hfb200.SurfaceInteraction3f(m, dev, shape, fields), hfb200.PreliminaryIntersection3f(m, dev, shape))
                shape.ray_intersect_preliminary(ray, out=pi_v)              # warm-up (first touch of this chunk's pages)
                e0, e1, e2, e3 = ev(), ev(), ev(), ev()
                e0.record(); shape.ray_intersect_preliminary(ray, out=pi_v); e1.record()
                e2.record(); shape.ray_intersect(ray, out=(si_v, pi_v)); shape.backward(ray, pi_v, adj); e3.record()
                torch.cuda.synchronize()
                ms_i += e0.elapsed_time(e1); ms_fb += e2.elapsed_time(e3)
                hits += float(pi_v.is_valid().sum().item())
                del o, d, ray, adj
            ms_ar = 0.0
            if world > 1:
                barrier()                                      # every rank has finished its chunks: time the collective alone
                ms_ar = timed(shape.allreduce_grad, 1)
            t = torch.tensor([ms_i, ms_fb, hits], dtype=torch.float64, device=dev)
            if world > 1:
                tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
                tsum = t.clone(); dist.all_reduce(tsum)
                ms_i, ms_fb, hits = float(tmax[0]), float(tmax[1]) + ms_ar, float(tsum[2])
            cells.append({"heights": hs, "rays": n, "chunks": nchunks, "intersect_ms": ms_i, "intersect_mrays": n / (ms_i * 1e-3) / 1e6 if ms_i else None,
                          "fwd_bwd_ms": ms_fb, "fwd_bwd_mrays": n / (ms_fb * 1e-3) / 1e6 if ms_fb else None, "allreduce_ms": ms_ar, "hit_fraction": hits / n})
        if hs == 4096 and world == 1 and not a.no_e2e:
            e2e = c5_e2e(shape, dev, timed)
        # traversal counters of the model on 2^20 rays of this size
        o, d = wl.hemisphere_rays(1 << 20, seed=5, device=dev)
        ray = hfb200.Ray3f(o, d, device=dev)
        p1 = hfb200.PreliminaryIntersection3f(1 << 20, dev, shape)
        shape.stats(reset=True); shape.ray_intersect_preliminary(ray, counted=True, out=p1); st = shape.stats(reset=True)
        for c in cells:
            if c["heights"] == hs:
                c["mean_nodes_per_ray"], c["mean_leaves_per_ray"] = st["nodes"] / st["rays"], st["leaves"] / st["rays"]
        del shape, si, pi, ray, p1, o, d
        torch.cuda.empty_cache()
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return
    peak, peak_src = _peak()
    for c in cells:
        b = 8.0 * c["mean_nodes_per_ray"] + 16.0 * c["mean_leaves_per_ray"]
        c["intersect_frac"] = c["rays"] * (44.0 + b) / (c["intersect_ms"] * 1e-3) / 1e9 / peak / world
        c["fwd_bwd_frac"] = c["rays"] * (76.0 + b + c["hit_fraction"] * 108.0 + (1 - c["hit_fraction"]) * 4.0) / (c["fwd_bwd_ms"] * 1e-3) / 1e9 / peak / world
    tot_rays = sum(c["rays"] for c in cells)
    tot_ms = sum(c["fwd_bwd_ms"] for c in cells)
    head = [c for c in cells if c["heights"] == 4096 and c["rays"] == max(counts)] or cells[-1:]
    cpu = None
    if world == 1 and not a.no_cpu:
        cpu = c5_cpu(a, 1)
    line = {"metric": "Mrays/s fwd+bwd over the C5 sweep (incoherent hemisphere rays, 2^20..2^30 rays x 512^2..16384^2 heights)",
            "value": tot_rays / (tot_ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": 1, "warmup": 1, "ms_per_step": tot_ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C5: ray-count sweep " + "/".join(f"2^{int(np.log2(n))}" for n in counts) + " rays x heightfield sizes " + "/".join(map(str, sizes)) +
                                   " (fBm seed 0xC5), intersect-only and fwd+bwd, hemisphere origins (radius 3, seed 5)",
                       "ray_coherence_hint": "coherent (lock-step kernel)" if os.environ.get("HF_C5_COHERENT", "0") == "1" else "incoherent (hf_set_ray_coherence(0): lane-pool kernel)",
                       "chunk": chunk, "sharding": "chunks of <= 2^26 rays dealt round-robin; gradient grids summed by one allreduce per cell",
                       "l2": "every chunk's ray / result / adjoint SoA (>= 180 MB at 2^20 rays) exceeds L2 from 2^22 rays on; the 2^20 cells fit and say so by their rates"},
            "clocks": clocks, "gpu_launches": sum(3 * c["chunks"] for c in cells),
            "roofline": {"bound": "hbm", "kernel": "hf_intersect_pool_kernel<FUSED_SI> + hf_backward_kernel (whole fwd+bwd cell)", "achieved": head[0]["fwd_bwd_frac"] * peak, "peak": peak,
                         "unit": "GB/s", "frac": head[0]["fwd_bwd_frac"], "peak_source": peak_src, "traffic": None,
                         "model": "rays x (76 B fwd I/O + 8 B x nodes + 16 B x leaves [canonical traversal counted on 2^20 rays of the size] + 108 B per hit / 4 B per miss of the backward pass); "
                                  f"headline cell: {head[0]['heights']}^2 x {head[0]['rays']} rays"},
            "cells": cells, "wall_s": time.perf_counter() - t_wall}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if e2e is not None:
        line["e2e"] = e2e
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def c5_e2e(shape, dev, timed):
    """One cell end to end through the host-buffer entry point: 2^24 hemisphere rays on the 4096^2 field, pinned host SoA in
    (rays + adjoints), host SoA out (preliminary record, t / p / n / uv) + the gradient grid."""
    from hfb200 import workloads as wl
    n = 1 << 24
    o, d = wl.hemisphere_rays(n, seed=5, device=dev)
    adj = wl.hashed_adjoints(n, seed=1, device=dev)
    pin = lambda x: x.detach().to("cpu").contiguous().pin_memory()
    hr = {"ox": pin(o[0]), "oy": pin(o[1]), "oz": pin(o[2]), "dx": pin(d[0]), "dy": pin(d[1]), "dz": pin(d[2]),
          "maxt": torch.full((n,), float("inf")).pin_memory()}
    ha = {"t": pin(adj["t"]), "px": pin(adj["p"][0]), "py": pin(adj["p"][1]), "pz": pin(adj["p"][2]), "nx": pin(adj["n"][0]), "ny": pin(adj["n"][1]),
          "nz": pin(adj["n"][2]), "uvx": pin(adj["uv"][0]), "uvy": pin(adj["uv"][1])}
    hs = {k: torch.empty(n, dtype=torch.float32).pin_memory() for k in ("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy")}
    hp = {"t": torch.empty(n).pin_memory(), "u": torch.empty(n).pin_memory(), "v": torch.empty(n).pin_memory(),
          "prim_index": torch.empty(n, dtype=torch.int32).pin_memory()}
    hg = torch.empty(shape.H, shape.W, dtype=torch.float32).pin_memory()
    del o, d, adj

    def step():
        shape.zero_grad()
        torch.cuda.current_stream().synchronize()
        shape.forward_backward_host(hr, ha, hs, hg, n, h_prelim=hp)
    step()
    ms = timed(step, 3) / 3
    return {"value": n / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": n * 4 * 16, "d2h_bytes_per_step": n * 4 * 13 + hg.numel() * 4, "ms_per_step": ms,
            "api": "hf_forward_backward_host, one cell: 2^24 rays x 4096^2 heights"}


def c5_cpu(a, steps):
    from oracle import hf_oracle as O
    from hfb200 import workloads as wl
    O.build()
    O.set_threads(os.cpu_count() or 1)
    n = 1 << 21
    h = wl.fbm_heights(4096, octaves=8, seed=0xC5).numpy()
    c = O.OracleHeightfield(h, 0.3)
    c._pyramid()
    o, d = wl.hemisphere_rays(n, seed=5)
    adj = wl.hashed_adjoints(n, seed=1)
    r = O.Rays(o.numpy().T, d.numpy().T)
    adj_np = {"t": adj["t"].numpy(), "px": adj["p"][0].numpy(), "py": adj["p"][1].numpy(), "pz": adj["p"][2].numpy(), "nx": adj["n"][0].numpy(),
              "ny": adj["n"][1].numpy(), "nz": adj["n"][2].numpy(), "uvx": adj["uv"][0].numpy(), "uvy": adj["uv"][1].numpy()}

    def one():
        t0 = time.perf_counter()
        pi = c.ray_intersect_preliminary(r)
        c.compute_surface_interaction(r, pi, ray_flags=O.RF_ALL, fields=("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy"))
        c.backward(r, pi, adj_np, ray_flags=O.RF_ALL, literal=False)
        return time.perf_counter() - t0
    one()
    ts = [one() for _ in range(max(1, steps))]
    return {"value": n / min(ts) / 1e6, "unit": "Mrays/s", "cores": O.get_threads(), "kind": "port",
            "sample": f"{n} hemisphere rays on the 4096^2 field, forward + surface interaction + backward, best of {len(ts)}", "ms": 1e3 * min(ts)}


def reference_c5(a):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cpu = c5_cpu(a, max(1, a.steps))
    line = {"impl": "reference", "metric": "Mrays/s fwd+bwd over the C5 sweep (incoherent hemisphere rays, 2^20..2^30 rays x 512^2..16384^2 heights)",
            "value": cpu["value"], "unit": "Mrays/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": cpu["ms"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": "C5 (CPU arm: one cell)", "sample": cpu["sample"]},
            "cpu_baseline": cpu, "e2e": {"value": cpu["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)
