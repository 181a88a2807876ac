"""Dev tool (under gpurun): time hf_ray_intersect on a slice of rank 0's 1/8 shard (groups g0..g1) to see what the slow region costs alone."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import hfb200
from hfb200 import workloads as wl
dev = torch.device("cuda", 0)
h = wl.fbm_heights(4096, octaves=8, seed=0xC3, device=dev)
shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
cam = wl.CONFIGS["C3"]["camera"]
pix = wl.tile_pixels(1024, 1024, 32, 0, 8, device=dev)
o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, device=dev, pixels=pix)
def run(g0, g1):
    oo = o[:, g0 * 32:g1 * 32].contiguous(); dd = d[:, g0 * 32:g1 * 32].contiguous()
    n = oo.shape[1]; ray = hfb200.Ray3f(oo, dd)
    si = hfb200.SurfaceInteraction3f(n, dev, shape, ("t", "p", "n", "uv")); pi = hfb200.PreliminaryIntersection3f(n, dev, shape)
    for _ in range(3): shape.ray_intersect(ray, out=(si, pi))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(5): shape.ray_intersect(ray, out=(si, pi))
    e1.record(); torch.cuda.synchronize()
    print(f"groups {g0}..{g1} ({n} rays, hit {float(pi.is_valid().float().mean()):.2f}): {e0.elapsed_time(e1) / 5 * 1e3:.1f} us per launch")
for a, b in [(78000, 82000), (79500, 80000), (79500, 79628), (144000, 148000), (144600, 144608), (100000, 104000), (200000, 204000), (100000, 100128)]:
    run(a, b)
