#!/usr/bin/env python
"""Measure the parity-test configurations of BASELINE.json on one GPU (C1, C2, C4 primary rays, C5 sweep) and print one
JSON line per case. bench.py measures the headline (C3); this fills the other rows of BASELINE.md section 5.

    python tools/run_configs.py [--cases C1,C2,C4,C5] [--reps 5] > gpurun_out/configs.jsonl
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import hfb200
from hfb200 import workloads as wl

PEAK = 6545.6
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps


def shape_modes(name, shape, o, d, reps, extra=None):
    """intersect-only, forward (+SI), backward on a device ray batch; counters for the byte model."""
    n = o.shape[1]
    ray = hfb200.Ray3f(o, d)
    pi = hfb200.PreliminaryIntersection3f(n, o.device, shape)
    si = hfb200.SurfaceInteraction3f(n, o.device, shape, ("t", "p", "n", "uv"))
    adj = shape.prepare_adjoints(wl.hashed_adjoints(n, seed=1, device=o.device), n)
    ms_i = timed(lambda: shape.ray_intersect_preliminary(ray, out=pi), reps)
    ms_f = timed(lambda: shape.ray_intersect(ray, out=(si, pi)), reps)
    ms_b = timed(lambda: (shape.zero_grad(), shape.backward(ray, pi, adj)), reps)
    shape.stats(reset=True)
    shape.ray_intersect_preliminary(ray, counted=True, out=pi)
    st = shape.stats(reset=True)
    nn, nl, hits = st["nodes"] / st["rays"], st["leaves"] / st["rays"], st["hits"] / st["rays"]
    btrav = 8 * nn + 16 * nl
    out = {"case": name, "rays": n, "hit_fraction": hits, "nodes_per_ray": nn, "leaves_per_ray": nl, "b_trav": btrav,
           "intersect_ms": ms_i, "intersect_mrays": n / ms_i / 1e3, "intersect_frac": n * (44 + btrav) / (ms_i * 1e-3) / 1e9 / PEAK,
           "forward_si_ms": ms_f, "forward_si_mrays": n / ms_f / 1e3, "forward_si_frac": n * (76 + btrav) / (ms_f * 1e-3) / 1e9 / PEAK,
           "backward_ms": ms_b, "backward_frac": n * (80 + 28 * hits) / (ms_b * 1e-3) / 1e9 / PEAK,
           "fwd_bwd_mrays": n / (ms_f + ms_b) / 1e3, "peak_gbs": PEAK}
    if extra:
        out.update(extra)
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="C1,C2,C4,C5")
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    cases = a.cases.split(",")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)

    if "C1" in cases:
        cfg = wl.CONFIGS["C1"]
        shape = hfb200.load_dict({"type": "heightfield", "heights": wl.make_heights(cfg["heights"], dev), "max_height": cfg["max_height"]})
        o, d = wl.config_rays("C1", device=dev, tile=0)
        shape_modes("C1 256^2 sine, 256^2 x 1 spp primary", shape, o, d, a.reps)
        del shape

    if "C2" in cases:
        cfg = wl.CONFIGS["C2"]
        cam0 = cfg["camera"]
        shape = hfb200.load_dict({"type": "heightfield", "heights": wl.make_heights(cfg["heights"], dev), "max_height": cfg["max_height"]})
        cam = hfb200.perspective_camera(cam0["origin"], cam0["target"], cam0["up"], cam0["fov"], 512, 512, 16, seed=2)
        for shadow in (False, True):
            light = hfb200.directional_light((0.3, 0.2, 1.0), shadow_rays=shadow)
            img = torch.zeros((512, 512), device=dev)
            ms = timed(lambda: shape.render_forward(cam, light, image=img), a.reps)
            n = 512 * 512 * 16
            _, prims = shape.render_forward(cam, light, return_prims=True)
            lit = float((prims != -1).float().mean())
            print(json.dumps({"case": "C2 1024^2 fBm, 512^2 x 16 spp direct illumination, forward" + (" + shadow rays" if shadow else ""),
                              "primary_rays": n, "shadow_rays": int(n * lit) if shadow else 0, "lit_fraction": lit, "render_forward_ms": ms,
                              "primary_mrays": n / ms / 1e3, "mean_pixel": float(img.mean())}), flush=True)
        o, d = wl.config_rays("C2", device=dev, tile=32, seed=2)
        shape_modes("C2 1024^2 fBm, 512^2 x 16 spp primary (ray-array API)", shape, o, d, a.reps)
        del shape

    if "C4" in cases:
        cfg = wl.CONFIGS["C4"]
        cam0 = cfg["camera"]
        h = wl.make_heights(cfg["heights"], dev)
        shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": cfg["max_height"]})
        del h
        torch.cuda.empty_cache()
        o, d = wl.config_rays("C4", device=dev, tile=32, seed=4)
        shape_modes("C4 16384^2 fBm (1 GiB), grazing camera, 2048^2 x 8 spp primary", shape, o, d, a.reps)
        del o, d
        cam = hfb200.perspective_camera(cam0["origin"], cam0["target"], cam0["up"], cam0["fov"], 2048, 2048, 8, seed=4)
        img = torch.zeros((2048, 2048), device=dev)
        for shadow in (False, True):
            light = hfb200.directional_light((0.3, 0.2, 1.0), shadow_rays=shadow)
            ms = timed(lambda: shape.render_forward(cam, light, image=img), a.reps)
            print(json.dumps({"case": "C4 render_forward" + (" with shadow rays (depth 1)" if shadow else ", primary rays only"),
                              "primary_rays": 2048 * 2048 * 8, "render_forward_ms": ms, "primary_mrays": 2048 * 2048 * 8 / ms / 1e3}), flush=True)
        del shape, img
        torch.cuda.empty_cache()

    if "C5" in cases:
        for size in (512, 2048, 4096, 16384):
            shape = hfb200.load_dict({"type": "heightfield", "heights": wl.fbm_heights(size, octaves=8, seed=0xC5, device=dev), "max_height": 0.3})
            for logn in (20, 24, 26):
                o, d = wl.hemisphere_rays(1 << logn, seed=5, device=dev)
                shape_modes(f"C5 {size}^2 fBm, 2^{logn} incoherent hemisphere rays", shape, o, d, max(2, a.reps if logn < 26 else 3))
                del o, d
            del shape
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
