#!/usr/bin/env python
"""bench.py -- headline benchmark of the differentiable heightfield path.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run)
    python bench.py --impl reference ...                       (CPU arm: the oracle port on the host cores)

Metric (BASELINE.json): Mrays/s for forward + backward (intersection + surface interaction, then the adjoint
w.r.t. the height texels) on the C3 workload: 4096^2 fBm heights, 1024^2 image x 64 spp = 67,108,864 primary rays
per step. A step = zero_grad -> hf_ray_intersect (fused preliminary + surface interaction) -> hf_backward
[-> one NCCL allreduce of the gradient grid when N > 1]. Rays are sharded by 16x16-pixel image tile across ranks, so
total work is fixed as N grows ("strong" scaling, as BASELINE.json configs[2] states: one step at 1/2/4/8 B200).

`value`   : device-resident inputs, CUDA-event timed, max over ranks.
`e2e`     : the same step through the C ABI's host-buffer entry point (hf_forward_backward_host) with pinned HOST
            arrays: H2D of rays + adjoints and D2H of the results + gradient are inside the timed region.
`roofline`: the dominant kernel (hf_intersect_kernel<FUSED_SI>) against the measured HBM copy bandwidth, with the
            ALGORITHMIC byte model of SURVEY.md 8(d): 76 B/ray of I/O + 8 B per pyramid node + 16 B per leaf cell
            visited by the canonical traversal (counted on this batch by the counting variant of the kernel, which
            tests/ pin to the oracle's counters).
`cpu_baseline`: the oracle port (CPU restatement of the reference's llvm_ad_rgb arithmetic; the reference plugin
            itself is absent from the snapshot and the reference cannot be built here -- SURVEY.md F1/F2) on all
            host cores, on a bounded sample of the same rays.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "Mrays/s fwd+bwd heightfield intersect+grad (4096^2 heights, 1024^2 image x 64 spp)"
B_IO_FWD = 76          # ray 28 + prelim 16 + si (p 12, n 12, uv 8)              SURVEY.md 8(d)
B_BWD_HIT = 92 + 16    # a HIT: ray 28 + prelim 16 + adjoints 36 + 3 atomics 12, + re-fetch of the hit cell 16
B_BWD_MISS = 4         # a MISS: the kernel reads t and returns


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(kernel, rays):
    """dram bytes per launch from the committed ncu capture (profiles/traffic.json), scaled to this launch's ray count."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[kernel]
        return float(t["dram_bytes"]) * rays / float(t["rays"])
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--heights", type=int, default=4096, help="heightfield side (C3: 4096)")
    ap.add_argument("--image", type=int, default=1024, help="image side (C3: 1024)")
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--tile", type=int, default=16, help="image tile (pixels per side) of the deal over ranks; 16: the eight shards of C3 differ by 3.5 %% in kernel time, 32: 5.7 %%, 64: 16 %%")
    ap.add_argument("--cpu-sample", type=int, default=1 << 23, help="rays in the bounded CPU sample")
    ap.add_argument("--config", default="C3", choices=["C3", "C4", "C5"],
                    help="C3 (default): the headline metric of BASELINE.json; C4 / C5: configs[3] / configs[4] as the file states them (bench_configs.py)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def workload_name(a):
    if (a.heights, a.image, a.spp) == (4096, 1024, 64):
        return "C3: 4096^2 fBm heightfield (8 octaves, seed 0xC3), 1024^2 image x 64 spp primary rays, fwd+bwd"
    return f"reduced: {a.heights}^2 fBm heightfield, {a.image}^2 image x {a.spp} spp primary rays, fwd+bwd"


def adj_to_c(adj):
    return {"t": adj["t"], "px": adj["p"][0], "py": adj["p"][1], "pz": adj["p"][2], "nx": adj["n"][0], "ny": adj["n"][1],
            "nz": adj["n"][2], "uvx": adj["uv"][0], "uvy": adj["uv"][1]}


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores (cpu_baseline leg and --impl reference)
# ---------------------------------------------------------------------------------------------------------
def cpu_sample(a, heights_np, n_sample):
    """A bounded, strided sample of the step's rays (whole pixels: all spp of every k-th pixel)."""
    from hfb200 import workloads as wl
    cam = wl.CONFIGS["C3"]["camera"]
    npix = a.image * a.image
    want_pix = max(1, min(npix, n_sample // a.spp))
    stride = max(1, npix // want_pix)
    pix = torch.arange(0, npix, stride, dtype=torch.int64)[:want_pix]
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], a.image, a.image, a.spp, seed=0, pixels=pix)
    adj = wl.hashed_adjoints(o.shape[1], seed=1, ids=wl.global_ray_ids(pix, a.spp))
    return o, d, adj


def cpu_step(O, c, r, adj_np):
    """One forward + backward pass of the oracle over the sample; returns seconds."""
    t0 = time.perf_counter()
    pi = c.ray_intersect_preliminary(r)
    c.compute_surface_interaction(r, pi, ray_flags=O.RF_ALL, fields=("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy"))
    c.backward(r, pi, adj_np, ray_flags=O.RF_ALL, literal=False)
    return time.perf_counter() - t0


def run_cpu(a, heights_np, steps, warmup):
    from oracle import hf_oracle as O
    O.build()
    cores = os.cpu_count() or 1
    O.set_threads(cores)
    o, d, adj = cpu_sample(a, heights_np, a.cpu_sample)
    c = O.OracleHeightfield(heights_np, 0.3)
    t0 = time.perf_counter(); c._pyramid(); t_pyr = time.perf_counter() - t0
    r = O.Rays(o.numpy().T, d.numpy().T)
    adj_np = {k: v.numpy() for k, v in adj_to_c(adj).items()}
    for _ in range(warmup):
        cpu_step(O, c, r, adj_np)
    times = [cpu_step(O, c, r, adj_np) for _ in range(steps)]
    n = r.n
    return {"mrays": n / (sum(times) / len(times)) / 1e6, "best_mrays": n / min(times) / 1e6, "n": n, "cores": cores,
            "threads": O.get_threads(), "ms_per_step": 1e3 * sum(times) / len(times), "pyramid_ms": 1e3 * t_pyr}


def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from hfb200 import workloads as wl
    h = wl.fbm_heights(a.heights, octaves=8, seed=0xC3).numpy()
    res = run_cpu(a, h, max(1, a.steps), max(1, min(a.warmup, 2)))
    sample = f"{res['n']} rays = all {a.spp} spp of every k-th pixel of the {a.image}^2 image (1/{max(1, a.image * a.image * a.spp // res['n'])} of the step)"
    line = {"impl": "reference", "metric": METRIC, "value": res["mrays"], "unit": "Mrays/s", "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(a)},
            "cpu_baseline": {"value": res["mrays"], "unit": "Mrays/s", "cores": res["threads"], "kind": "port", "sample": sample,
                             "note": "CPU restatement of the reference llvm_ad_rgb arithmetic (oracle/hf_oracle.c, OpenMP); the reference "
                                     "plugin is absent from the snapshot and Mitsuba cannot be built here (SURVEY.md F1/F2)"},
            "e2e": {"value": res["mrays"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------
def sharded_gradient_check(rank, world, dev, tile):
    """N-rank path against the 1-rank path on a reduced configuration (512^2 heights, 256^2 image x 16 spp), inside the run: every
    rank traces its image tiles, the grids are summed by hf_allreduce_grad over the SAME communicator kind the timed step uses, and
    rank 0 compares with the grid it gets by tracing all rays alone. Returns the relative L2 difference (rank 0) or None."""
    import torch.distributed as dist
    import hfb200
    from hfb200 import workloads as wl
    hs, img, spp = 512, 256, 16
    cam = wl.CONFIGS["C3"]["camera"]
    h = wl.fbm_heights(hs, octaves=6, seed=0xC3, device=dev)

    def grad_of(shape, pix):
        o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], img, img, spp, seed=0, device=dev, pixels=pix)
        ray = hfb200.Ray3f(o, d)
        adj = wl.hashed_adjoints(o.shape[1], seed=1, device=dev, ids=wl.global_ray_ids(pix, spp))
        si, pi = shape.ray_intersect(ray, return_prelim=True, fields=("t", "p", "n", "uv"))
        shape.zero_grad()
        shape.backward(ray, pi, adj)

    sh = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
    ids = [None] * world
    dist.all_gather_object(ids, hfb200.comm_unique_id() if rank == 0 else None)
    sh.comm_init(rank, world, ids[0])
    grad_of(sh, wl.tile_pixels(img, img, tile, rank, world, device=dev))
    sh.allreduce_grad()
    torch.cuda.synchronize()
    g_n = sh.grad.double().clone()
    out = None
    if rank == 0:
        one = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
        grad_of(one, torch.arange(img * img, dtype=torch.int64, device=dev))
        torch.cuda.synchronize()
        g_1 = one.grad.double()
        out = float((g_n - g_1).norm() / g_1.norm())
        del one
    del sh
    dist.barrier()
    return out


def main_ours(a):
    import torch.distributed as dist
    import hfb200
    from hfb200 import workloads as wl

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the heightfield path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    # ---- workload (synthetic, identical on every rank; rays sharded by image tile) ----
    h = wl.fbm_heights(a.heights, octaves=8, seed=0xC3, device=dev)
    shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
    # the step traces the same rays every iteration, as an optimisation does: the launches learn which 32-ray groups are far above the
    # mean and start with those (include/hf_b200.h hf_set_adaptive_schedule; HF_BENCH_ADAPTIVE=0: the fixed schedule of round 1)
    adaptive = os.environ.get("HF_BENCH_ADAPTIVE", "1") != "0"
    shape.set_adaptive_schedule(adaptive)
    cam = wl.CONFIGS["C3"]["camera"]
    # dev knob: HF_BENCH_SHARD_OF=N on ONE GPU times the rays rank 0 of an N-GPU run would get (kernel scaling without the box)
    shard_of = int(os.environ.get("HF_BENCH_SHARD_OF", "0")) if world == 1 else 0
    shard_rank = int(os.environ.get("HF_BENCH_SHARD_RANK", "0")) if shard_of > 1 else rank       # ... or of any other rank of that run
    pix = wl.tile_pixels(a.image, a.image, a.tile, shard_rank, shard_of if shard_of > 1 else world, device=dev)
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], a.image, a.image, a.spp, seed=0, device=dev, pixels=pix)
    n_local = o.shape[1]
    n_total = a.image * a.image * a.spp if shard_of <= 1 else n_local
    ray = hfb200.Ray3f(o, d)
    # adjoints keyed by the GLOBAL ray id (pixel * spp + sample): every N solves the same problem, so the gradient grids (and their
    # checksums in `extras`) of the 1 / 2 / 4 / 8-GPU runs are comparable
    adj = wl.hashed_adjoints(n_local, seed=1, device=dev, ids=wl.global_ray_ids(pix, a.spp))
    adj_c = shape.prepare_adjoints(adj, n_local)
    fields = ("t", "p", "n", "uv")
    si = hfb200.SurfaceInteraction3f(n_local, dev, shape, fields)
    pi = hfb200.PreliminaryIntersection3f(n_local, dev, shape)
    del pix
    torch.cuda.empty_cache()

    if world > 1:
        ids = [None] * world
        dist.all_gather_object(ids, hfb200.comm_unique_id() if rank == 0 else None)
        shape.comm_init(rank, world, ids[0])

    shard_check = sharded_gradient_check(rank, world, dev, a.tile) if world > 1 else None

    ev = lambda: torch.cuda.Event(enable_timing=True)
    k_fwd, k_bwd = [], []

    zside = torch.cuda.Stream(device=dev)
    overlap_zero = os.environ.get("HF_BENCH_ZERO_OVERLAP", "1") != "0"

    def step(record=False):
        if record or not overlap_zero:
            shape.zero_grad()
        else:
            # the gradient grid is only needed by the backward kernel: clear it on a second stream beside the forward kernel
            # (it fits into the SMs the forward kernel's closing phase leaves idle); fork / join, so it is captured with the step
            cur = torch.cuda.current_stream()
            zside.wait_stream(cur)
            with torch.cuda.stream(zside):
                shape.zero_grad()
        if record:
            e0, e1, e2 = ev(), ev(), ev()
            e0.record()
        shape.ray_intersect(ray, out=(si, pi))
        if record:
            e1.record()
        elif overlap_zero:
            torch.cuda.current_stream().wait_stream(zside)
        shape.backward(ray, pi, adj_c)
        if record:
            e2.record()
            k_fwd.append((e0, e1)); k_bwd.append((e1, e2))
        if world > 1:
            shape.allreduce_grad()

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

    for _ in range(max(a.warmup, 3)):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    timed(lambda: step(True), a.steps)              # eager pass: per-kernel CUDA events (roofline legs below)
    # headline: the same step replayed from ONE CUDA graph (hfb200.capture): at N = 8 the step is ~1 ms of four launches and
    # the host's launch calls would otherwise show up in it. Falls back to eager launches if capture is refused.
    run_step, launch = (lambda: step(False)), "eager"
    if os.environ.get("HF_BENCH_GRAPH", "1") != "0":
        try:
            run_step = hfb200.capture(lambda: step(False))
            launch = "cuda-graph"
        except Exception as exc:      # noqa: BLE001
            print(f"bench.py: CUDA graph capture refused ({type(exc).__name__}: {exc}); eager launches", file=sys.stderr, flush=True)
            torch.cuda.synchronize()
    for _ in range(3):
        run_step()
    ms_total = timed(run_step, a.steps)
    ms_step = ms_total / a.steps
    sched_info = shape.adaptive_schedule_info("rays")
    adaptive_used = adaptive and sched_info[1] > 0          # (the library takes batches of <= 320 groups per resident warp only)
    ms_allreduce = timed(shape.allreduce_grad, 10) / 10 if world > 1 else 0.0     # the collective alone (64 MiB fp32 at C3)
    collective = "none (1 GPU)" if world == 1 else "ncclAllReduce(grad 64 MiB fp32, sum) per step"
    peer = None
    state = {"peer": False}
    if world > 1 and os.environ.get("HF_BENCH_COLLECTIVE", "auto") != "nccl":
        # The fused alternative (include/hf_b200.h "multi-GPU over peer memory"): every gradient scatter goes to the rank that owns the
        # texel, over NVLink, inside the backward kernel; what is left of the collective is an all-gather by peer stores. Verified in
        # this run against the NCCL step above on the full batch before it may become the headline.
        try:
            run_step(); torch.cuda.synchronize()
            g_nccl = shape.grad.clone()
            handles = [None] * world
            dist.all_gather_object(handles, shape.peer_export())
            shape.peer_attach(rank, world, handles)

            def step_peer():
                cur = torch.cuda.current_stream()
                zside.wait_stream(cur)
                with torch.cuda.stream(zside):              # gradient grid cleared beside the forward kernel
                    shape.peer_zero_grad()
                shape.ray_intersect(ray, out=(si, pi))
                cur.wait_stream(zside)
                shape.backward(ray, pi, adj_c)
                shape.peer_reduce_grad()                    # partial slabs -> owners (peer stores) | barrier | owner sums, stores the sum to every rank | barrier
            step_peer(); torch.cuda.synchronize()
            rel = (shape.grad.double() - g_nccl.double()).norm() / g_nccl.double().norm()
            if world > 1:
                dist.all_reduce(rel, op=dist.ReduceOp.MAX)
            del g_nccl
            run_peer, peer_launch = step_peer, "eager"
            if os.environ.get("HF_BENCH_GRAPH", "1") != "0":
                run_peer = hfb200.capture(step_peer); peer_launch = "cuda-graph"
            for _ in range(3):
                run_peer()
            ms_peer = timed(run_peer, a.steps) / a.steps
            torch.cuda.synchronize()
            state["checksum"] = float(shape.grad.double().abs().sum().item())     # (the leg below re-reduces the reduced grid: timing only)
            reduce_ms = timed(shape.peer_reduce_grad, 10) / 10
            run_peer()
            peer = {"ms_per_step": ms_peer, "mrays": n_total / (ms_peer * 1e-3) / 1e6, "launch": peer_launch, "grad_rel_l2_vs_nccl_step": float(rel),
                    "reduce_ms": reduce_ms,
                    "what": "zero_grad (beside the forward kernel) -> intersect/SI -> backward -> hf_peer_reduce_grad: partial slabs stored into their owners' "
                            "staging areas over NVLink, barrier kernel, owners sum in rank order and store the sum into every rank's grid, barrier kernel"}
            if os.environ.get("HF_BENCH_PEER_ATOMICS", "0") == "1":        # the rejected variant, for the record: scatters go to the owner with remote atomics
                shape.peer_set_mode(0)
                for _ in range(3):
                    step_peer()
                peer["remote_atomics_ms_per_step"] = timed(step_peer, a.steps) / a.steps
                shape.peer_set_mode(1)
            state["peer"] = True
        except Exception as exc:      # noqa: BLE001
            print(f"bench.py: peer-memory path unavailable ({type(exc).__name__}: {exc}); NCCL allreduce only", file=sys.stderr, flush=True)
            peer = None
    nccl_step = {"ms_per_step": ms_step, "mrays": n_total / (ms_step * 1e-3) / 1e6, "launch": launch, "allreduce_ms": ms_allreduce,
                 "what": "zero_grad -> intersect/SI -> backward -> ONE ncclAllReduce(grad, fp32, sum) on NCCL-registered memory"}
    if peer is not None and peer["grad_rel_l2_vs_nccl_step"] < 1e-4 and peer["ms_per_step"] < ms_step and os.environ.get("HF_BENCH_COLLECTIVE", "auto") != "nccl-headline":
        ms_step, launch = peer["ms_per_step"], peer["launch"]
        state["peer_headline"] = True
        collective = ("reduce-scatter + all-gather written as two kernels over peer memory (NVLink stores into IPC-mapped grids, device-side barriers); "
                      "the single-ncclAllReduce step is timed beside it (extras.nccl_allreduce_step)")
    clocks = sampler.stop() if rank == 0 else None
    value = n_total / (ms_step * 1e-3) / 1e6
    ms_fwd = float(np.mean([x.elapsed_time(y) for x, y in k_fwd]))
    ms_bwd = float(np.mean([x.elapsed_time(y) for x, y in k_bwd]))
    grad_checksum = state.get("checksum") or float(shape.grad.double().abs().sum().item())
    # how evenly the image tiles spread the work: the step ends with a collective, i.e. with the slowest rank
    kt = torch.tensor([ms_fwd, ms_bwd], dtype=torch.float64, device=dev)
    kt_min, kt_max = kt.clone(), kt.clone()
    if world > 1:
        dist.all_reduce(kt_min, op=dist.ReduceOp.MIN); dist.all_reduce(kt_max, op=dist.ReduceOp.MAX)
    rank_spread = {"fwd_ms_min": float(kt_min[0]), "fwd_ms_max": float(kt_max[0]), "bwd_ms_min": float(kt_min[1]), "bwd_ms_max": float(kt_max[1])}
    hits = float(pi.is_valid().float().mean().item())

    # ---- secondary modes (extras; not the headline) ----
    pi2 = hfb200.PreliminaryIntersection3f(n_local, dev, shape)
    ms_isect = timed(lambda: shape.ray_intersect_preliminary(ray, out=pi2), a.steps) / a.steps
    params = hfb200.traverse(shape)
    opt = hfb200.Adam(lr=1e-3, params=params)

    def full_step():
        if state["peer"]:                               # reduce-scatter -> sharded Adam -> all-gather of the heights, over peer memory
            cur = torch.cuda.current_stream()
            zside.wait_stream(cur)
            with torch.cuda.stream(zside):
                shape.peer_zero_grad()
            shape.ray_intersect(ray, out=(si, pi))
            cur.wait_stream(zside)
            shape.backward(ray, pi, adj_c)
            opt.step(peer=True)
        else:
            step()
            opt.step()
        params.update()

    full_step()                                     # eager once: builds the Adam bias-correction table for these betas
    run_full, full_launch = full_step, "eager"
    if os.environ.get("HF_BENCH_GRAPH", "1") != "0":
        try:
            run_full = hfb200.capture(full_step)    # Adam's step counter lives on the device: every replay is a correct next step
            full_launch = "cuda-graph"
        except Exception as exc:      # noqa: BLE001
            print(f"bench.py: full-step capture refused ({type(exc).__name__}: {exc}); eager launches", file=sys.stderr, flush=True)
            torch.cuda.synchronize()
    for _ in range(2):
        run_full()
    params["heights"] = h           # every timed run starts from the initial field (Adam roughens it step by step: walks get longer)
    params.update()
    opt.reset()
    ms_full = timed(run_full, a.steps) / a.steps
    params["heights"] = h           # restore the field for the legs below
    params.update()

    # ---- the optimisation step of BASELINE.json configs[2] through the fused render entry:
    #      zero_grad -> hf_render_step_l2 (in-register camera rays, direct illumination, box filter, L2 adjoint, backward)
    #      [-> allreduce] -> Adam -> pyramid rebuild. No ray / interaction record touches HBM.
    rs = run_render_step(a, shape, h, params, opt, rank, world, dev, timed, ev, state["peer"])      # (its e2e legs run at N = 1 only)
    params["heights"] = h
    params.update()

    # ---- traversal counters for the algorithmic byte model (outside every timed region) ----
    shape.stats(reset=True)
    shape.ray_intersect_preliminary(ray, counted=True, out=pi2)
    st = shape.stats(reset=True)
    cnt = torch.tensor([st["rays"], st["nodes"], st["leaves"], st["hits"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(cnt)
    rays_c, nodes_c, leaves_c, hits_c = [float(x) for x in cnt.tolist()]
    n_node, n_leaf = nodes_c / rays_c, leaves_c / rays_c
    b_trav = 8.0 * n_node + 16.0 * n_leaf
    del pi2

    # ---- e2e: host buffers through hf_forward_backward_host ----
    e2e = None
    if state["peer"]:
        shape.peer_detach()             # the host-buffer leg below sums with NCCL (its kernels must scatter locally again)
        state["peer"] = False
    if not a.no_e2e:
        e2e = run_e2e(a, shape, ray, adj, n_local, n_total, world, dev, timed)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    bpr_fwd = B_IO_FWD + b_trav
    ach_fwd = n_local * bpr_fwd / (ms_fwd * 1e-3) / 1e9
    bpr_bwd = hits * B_BWD_HIT + (1.0 - hits) * B_BWD_MISS          # a miss costs the t read only: charging it a full record inflated round 1's line
    ach_bwd = n_local * bpr_bwd / (ms_bwd * 1e-3) / 1e9
    tr_fwd = measured_traffic("hf_intersect_fast_kernel<FUSED_SI>", n_local)
    tr_bwd = measured_traffic("hf_backward_kernel", n_local)
    line = {
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a), "rays_per_step": n_total, "rays_per_gpu": n_local, "sharding": f"{a.tile}x{a.tile}-pixel image tiles, round-robin", "launch": launch,
                   "zero_grad": "on a second stream beside the forward kernel" if overlap_zero else "in stream order",
                   "collective": collective,
                   "l2": "inputs exceed L2 (ray + adjoint + output SoA > 5 GB per step vs 126 MB L2); no explicit flush",
                   "si_fields": "t,p,n,uv", "hit_fraction": hits,
                   "schedule": "adaptive (groups far above the mean cost of the previous launch go first)" if adaptive_used else
                               ("fixed (adaptive schedule enabled; the library keeps the fixed one above 320 groups per resident warp)" if adaptive else "fixed")},
        "clocks": clocks,
        "gpu_launches": ((3 if adaptive_used else 2) + (4 if state.get("peer_headline") else 0)) * a.steps,
        "kernels_per_step": ["hf_intersect_fast_kernel<FUSED_SI>"] + (["hf_pick_heavy_kernel"] if adaptive_used else []) + ["hf_backward_kernel"]
                            + (["hf_peer_send_partials_kernel", "hf_peer_barrier_kernel", "hf_peer_reduce_gather_kernel", "hf_peer_barrier_kernel"] if state.get("peer_headline") else []),
        "roofline": {"bound": "hbm", "kernel": "hf_intersect_fast_kernel<FUSED_SI>", "achieved": ach_fwd, "peak": peak, "unit": "GB/s",
                     "frac": ach_fwd / peak, "traffic": tr_fwd, "traffic_unit": "bytes per launch (ncu dram__bytes_read + write, profiles/traffic.json)",
                     "what_frac_means": "ALGORITHMIC bytes of SURVEY.md 8(d) per second over the measured HBM copy bandwidth -- the contract's figure, NOT DRAM utilisation: "
                                        "the traversal bytes of the model are served by L1 / L2 (see dram_frac) and the kernel is bound by instruction issue (limiter)",
                     "dram_gbs": (tr_fwd / (ms_fwd * 1e-3) / 1e9) if tr_fwd else None, "dram_frac": (tr_fwd / (ms_fwd * 1e-3) / 1e9 / peak) if tr_fwd else None,
                     "limiter": "alu/issue: ncu issue slots 69 %, ALU pipe 60 %, 22.7 of 32 lanes active, DRAM 17 % (profiles/r02_ncu_full_summary.txt)",
                     "algorithmic_bytes": n_local * bpr_fwd, "peak_source": peak_src, "bytes_per_ray": bpr_fwd,
                     "model": f"76 B I/O + 8 B x {n_node:.2f} nodes + 16 B x {n_leaf:.2f} leaves (canonical traversal, counted on this batch)",
                     "io_only_frac": n_local * B_IO_FWD / (ms_fwd * 1e-3) / 1e9 / peak, "kernel_ms": ms_fwd,
                     "backward": {"kernel": "hf_backward_kernel", "kernel_ms": ms_bwd, "achieved": ach_bwd, "frac": ach_bwd / peak,
                                  "bytes_per_ray": bpr_bwd, "model": f"hit {B_BWD_HIT} B (ray 28, prelim 16, adjoints 36, atomics 12, cell re-fetch 16) x {hits:.3f} + miss {B_BWD_MISS} B",
                                  "traffic": tr_bwd, "dram_gbs": (tr_bwd / (ms_bwd * 1e-3) / 1e9) if tr_bwd else None,
                                  "dram_frac": (tr_bwd / (ms_bwd * 1e-3) / 1e9 / peak) if tr_bwd else None,
                                  "limiter": "memory latency, then HBM: ncu issue slots 64 %, DRAM throughput 51 %, 1.6 eligible warps per cycle (profiles/r02_ncu_full_summary.txt)"}},
        "extras": {"intersect_only_mrays": n_total / (ms_isect * 1e-3) / 1e6, "intersect_only_ms": ms_isect,
                   "full_step_mrays": n_total / (ms_full * 1e-3) / 1e6, "full_step_ms": ms_full,
                   "full_step": "zero_grad + intersect/SI + backward + allreduce + Adam + pyramid rebuild", "full_step_launch": full_launch,
                   "allreduce_ms": ms_allreduce, "nccl_allreduce_step": nccl_step if world > 1 else None, "peer_scatter_step": peer,
                   "kernel_ms_over_ranks": rank_spread,
                   "grad_rel_l2_vs_single_rank": shard_check,
                   "grad_rel_l2_vs_single_rank_what": "N-rank tile shards + hf_allreduce_grad vs one rank tracing every ray, 512^2 heights, 256^2 x 16 spp, in this run",
                   "grad_abs_sum": grad_checksum, "mean_nodes_per_ray": n_node, "mean_leaves_per_ray": n_leaf,
                   "adaptive_schedule": dict(zip(("listed_groups", "groups"), sched_info)) if adaptive else None},
    }
    # SURVEY.md 8(d) "full step" byte model: rays x (B_trav + 16 + 12) + texels x (4 + 28 + 4 + 8/3) + pixels x 36
    texels, pixels_img = float(a.heights) ** 2, float(a.image) ** 2
    step_bytes = n_total * (b_trav + 28.0) + texels * (4 + 28 + 4 + 8.0 / 3.0) + pixels_img * 36.0
    kern_bytes = n_local * (b_trav + 28.0) + (pixels_img / world) * 12.0
    rs["roofline"] = {"bound": "hbm", "kernel": "hf_render_kernel<STEP_L2>", "kernel_ms": rs.pop("kernel_ms"),
                      "achieved": kern_bytes / (rs["kernel_ms_"] * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                      "frac": kern_bytes / (rs["kernel_ms_"] * 1e-3) / 1e9 / peak,
                      "model": f"rays x ({b_trav:.1f} B traversal + 16 B hit cell + 12 B gradient atomics) + 12 B per pixel"}
    rs.pop("kernel_ms_")
    rs["step_model_gbs"] = step_bytes / (rs["ms_per_step"] * 1e-3) / 1e9
    rs["step_model_frac"] = rs["step_model_gbs"] / (world * peak)       # aggregate bytes/s of the N ranks over N x one GPU's peak
    line["extras"]["render_step"] = rs
    line["config"]["step_ms"] = {"fwd_bwd": ms_step, "full_step": ms_full, "render_step": rs.get("ms_per_step"),
                                 "what": "fwd_bwd = the headline; full_step = zero_grad + intersect/SI + backward + collective + Adam + pyramid rebuild "
                                         "on materialised rays; render_step = the fused inverse-rendering step (hf_render_step_l2 + collective + Adam + rebuild)"}
    if world > 1:
        check = {"grad_rel_l2_vs_single_rank": shard_check,
                 "grad_rel_l2_vs_nccl_step": peer["grad_rel_l2_vs_nccl_step"] if peer is not None else None, "tolerance": 1e-4}
        line["config"]["multi_gpu_check"] = check
        line["nccl_step_value"] = nccl_step["mrays"]          # the literal "single ncclAllReduce" step, whichever step is the headline
        bad = [k for k in ("grad_rel_l2_vs_single_rank", "grad_rel_l2_vs_nccl_step") if check[k] is not None and not (check[k] <= 1e-4)]
    else:
        bad = []
    if e2e is not None:
        line["e2e"] = e2e
    if world == 1 and not a.no_cpu:
        res = run_cpu(a, h.cpu().numpy(), 2, 1)
        line["cpu_baseline"] = {"value": res["best_mrays"], "unit": "Mrays/s", "cores": res["threads"], "kind": "port",
                                "sample": f"{res['n']} rays (all {a.spp} spp of every k-th pixel), best of 2 after 1 warm-up",
                                "ms": res["ms_per_step"]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if bad:
        print(f"bench.py: multi-GPU gradient check failed: {bad} above 1e-4 ({line['config']['multi_gpu_check']})", file=sys.stderr, flush=True)
        sys.exit(3)


def run_render_step(a, shape, h, params, opt, rank, world, dev, timed, ev, peer=False):
    import hfb200
    from hfb200 import workloads as wl
    cam3 = wl.CONFIGS["C3"]["camera"]
    cam = hfb200.perspective_camera(cam3["origin"], cam3["target"], cam3["up"], cam3["fov"], a.image, a.image, a.spp, seed=0)
    light = hfb200.directional_light((0.3, 0.2, 1.0))
    shard_of = int(os.environ.get("HF_BENCH_SHARD_OF", "0")) if world == 1 else 0      # dev knob, see main_ours
    shard_rank = int(os.environ.get("HF_BENCH_SHARD_RANK", "0")) if shard_of > 1 else rank
    pix = None if world == 1 and shard_of <= 1 else \
        wl.tile_pixels(a.image, a.image, a.tile, shard_rank, shard_of if shard_of > 1 else world, device=dev).to(torch.int32)
    # target image: the same scene over another fBm field (seed 0x7A); every rank renders the pixels it owns
    tgt_shape = hfb200.load_dict({"type": "heightfield", "heights": wl.fbm_heights(a.heights, octaves=8, seed=0x7A, device=dev),
                                  "max_height": 0.3}, device=dev)
    target = tgt_shape.render_forward(cam, light, pix)
    del tgt_shape
    torch.cuda.empty_cache()
    image = torch.zeros_like(target)
    loss = torch.zeros(1, dtype=torch.float64, device=dev)
    k_times = []

    def step(record=False):
        if peer:
            shape.peer_zero_grad()
        else:
            shape.zero_grad()
        loss.zero_()
        if record:
            e0, e1 = ev(), ev()
            e0.record()
        shape.render_step_l2(cam, light, target, pix, image=image, loss=loss)
        if record:
            e1.record(); k_times.append((e0, e1))
        if peer:                                # reduce-scatter by peer stores -> sharded Adam -> heights stored into every rank's grid
            opt.step(peer=True)
        else:
            if world > 1:
                shape.allreduce_grad()
            opt.step()
        params.update()

    def kernel_only(record=True):          # the fused kernel alone on the initial field (heights do not move)
        if peer:
            shape.peer_zero_grad()
        else:
            shape.zero_grad()
        e0, e1 = ev(), ev()
        e0.record()
        shape.render_step_l2(cam, light, target, pix, image=image, loss=loss)
        e1.record()
        if record:
            k_times.append((e0, e1))

    params["heights"] = h
    params.update()
    for _ in range(3):
        kernel_only(False)
    timed(kernel_only, a.steps)
    kms = float(np.mean([x.elapsed_time(y) for x, y in k_times]))
    # real optimisation steps from the initial state (the field roughens as Adam moves it, walks get longer), replayed from ONE CUDA
    # graph (eager ctypes launches added up to 25 % of host noise to this number in round 1)
    opt.reset()
    for _ in range(3):
        step()
    params["heights"] = h
    params.update()
    opt.reset()
    step(); torch.cuda.synchronize(); loss0 = float(loss.item())
    run_step, launch = step, "eager"
    if os.environ.get("HF_BENCH_GRAPH", "1") != "0":
        try:
            run_step = hfb200.capture(step)
            launch = "cuda-graph"
        except Exception as exc:      # noqa: BLE001
            print(f"bench.py: render-step capture refused ({type(exc).__name__}: {exc}); eager launches", file=sys.stderr, flush=True)
            torch.cuda.synchronize()
    ms = timed(run_step, a.steps) / a.steps
    torch.cuda.synchronize()
    loss1 = float(loss.item())
    n_total = a.image * a.image * a.spp
    out = {"what": "zero_grad + hf_render_step_l2 (camera rays, direct illumination, box filter, L2 adjoint, backward in ONE kernel) + " +
                   ("reduce-scatter by peer stores + sharded Adam storing the heights into every rank's grid over NVLink" if peer else "allreduce + Adam") + " + pyramid rebuild",
           "collective": "peer memory: reduce-scatter -> sharded Adam -> all-gather of the heights" if peer else ("ncclAllReduce" if world > 1 else "none"), "ms_per_step": ms, "mrays": n_total / (ms * 1e-3) / 1e6, "launch": launch,
           "kernel_ms": kms, "kernel_ms_": kms, "kernel_mrays": n_total / world / (kms * 1e-3) / 1e6 * world,
           "loss_first_step_local": loss0, "loss_last_step_local": loss1, "gpu_launches_per_step": 6,
           "adaptive_schedule": dict(zip(("listed_pixels", "pixels"), shape.adaptive_schedule_info("pixels")))}
    if world == 1 and not a.no_e2e:
        out["e2e"] = run_render_e2e(a, shape, h, params, opt, cam, light, target, dev, timed, ms)
    return out


def run_render_e2e(a, shape, h, params, opt, cam, light, target, dev, timed, device_ms):
    """The optimisation step end to end through the host-buffer entry point hf_render_step_l2_host, two callers:
    (a) the optimiser lives on the HOST (the reference's Python Adam over llvm arrays): heights + target in, gradient grid + loss out;
    (b) the optimiser is the library's (hf_adam_step on the device): target in, image + loss out, then Adam + pyramid rebuild."""
    pin = lambda x: x.detach().to("cpu").contiguous().pin_memory()
    h_t, h_h = pin(target), pin(h)
    h_g = torch.empty(shape.H, shape.W, dtype=torch.float32).pin_memory()
    h_img = torch.empty(a.image, a.image, dtype=torch.float32).pin_memory()
    steps = max(2, min(a.steps, 5))

    def step_host_opt():
        shape.render_step_l2_host(cam, light, h_t, h_heights=h_h, h_grad=h_g)

    def step_dev_opt():
        shape.render_step_l2_host(cam, light, h_t, h_image=h_img)
        opt.step()
        params.update()
        torch.cuda.current_stream().synchronize()

    params["heights"] = h
    params.update()
    torch.cuda.synchronize()
    step_host_opt()
    ms_a = timed(step_host_opt, steps) / steps
    opt.reset()
    step_dev_opt()
    ms_b = timed(step_dev_opt, steps) / steps
    n_total = a.image * a.image * a.spp
    tb, hb = h_t.numel() * 4, h_h.numel() * 4
    return {"host_optimiser": {"value": n_total / (ms_a * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms_a, "h2d_bytes_per_step": tb + hb,
                               "d2h_bytes_per_step": hb + 8, "what": "heights + target in, gradient grid + loss out (hf_render_step_l2_host)"},
            "device_optimiser": {"value": n_total / (ms_b * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms_b, "h2d_bytes_per_step": tb,
                                 "d2h_bytes_per_step": tb + 8, "what": "target in, image + loss out, then hf_adam_step + pyramid rebuild on the device"},
            "device_step_ms": device_ms}


def run_e2e(a, shape, ray, adj, n_local, n_total, world, dev, timed):
    """The step through the reference-facing host-buffer call: every step copies the rays + adjoints host->device
    from pinned memory and the preliminary intersection, t/p/n/uv and the gradient grid device->host."""
    pin = lambda x: x.detach().to("cpu").contiguous().pin_memory()
    hr = {"ox": pin(ray.o[0]), "oy": pin(ray.o[1]), "oz": pin(ray.o[2]), "dx": pin(ray.d[0]), "dy": pin(ray.d[1]), "dz": pin(ray.d[2]),
          "maxt": pin(ray.maxt)}
    ha = {k: pin(v) for k, v in adj_to_c(adj).items()}
    names = ("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy")
    hs = {k: torch.empty(n_local, dtype=torch.float32).pin_memory() for k in names}
    hp = {"t": torch.empty(n_local).pin_memory(), "u": torch.empty(n_local).pin_memory(), "v": torch.empty(n_local).pin_memory(),
          "prim_index": torch.empty(n_local, dtype=torch.int32).pin_memory()}
    hg = torch.empty(shape.H, shape.W, dtype=torch.float32).pin_memory()
    h2d = n_local * 4 * (len(hr) + len(ha))
    d2h = n_local * 4 * (len(hs) + len(hp)) + hg.numel() * 4

    def step():
        shape.zero_grad()
        torch.cuda.current_stream().synchronize()
        if world > 1:
            # gradient must be summed before it is read back: forward+backward, allreduce, then the D2H of the grid
            shape.forward_backward_host(hr, ha, hs, None, n_local, h_prelim=hp)
            shape.allreduce_grad()
            hg.copy_(shape.grad, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        else:
            shape.forward_backward_host(hr, ha, hs, hg, n_local, h_prelim=hp)

    steps = max(2, min(a.steps, 5))
    step()
    ms = timed(step, steps) / steps
    return {"value": n_total / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "ms_per_step": ms, "steps": steps, "api": "hf_forward_backward_host (pinned host SoA in, host SoA + gradient grid out)",
            "result_checksum": float(hg.double().abs().sum().item())}


def main():
    a = parse_args()
    if a.config != "C3":
        import bench_configs as bc
        if a.impl == "reference":
            (bc.reference_c4 if a.config == "C4" else bc.reference_c5)(a)
        else:
            (bc.run_c4 if a.config == "C4" else bc.run_c5)(a, ClockSampler)
        return
    if a.impl == "reference":
        main_reference(a)
    else:
        main_ours(a)


if __name__ == "__main__":
    main()
