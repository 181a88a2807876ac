"""Dev tool (under gpurun, debug build): per-warp timeline of hf_intersect_fast_kernel on rank 0's shard of an N-way split.
HF_B200_LIB=.../libhf_b200_dbg.so python tools/dbg_timeline.py [shard_of]"""
import sys, os, ctypes as C
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import hfb200
from hfb200 import workloads as wl
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda", 0)
h = wl.fbm_heights(4096, octaves=8, seed=0xC3, device=dev)
shape = hfb200.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
cam = wl.CONFIGS["C3"]["camera"]
pix = wl.tile_pixels(1024, 1024, 32, 0, world, device=dev)
o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, device=dev, pixels=pix)
n = o.shape[1]; ray = hfb200.Ray3f(o, d)
si = hfb200.SurfaceInteraction3f(n, dev, shape, ("t", "p", "n", "uv")); pi = hfb200.PreliminaryIntersection3f(n, dev, shape)
lib = C.CDLL(os.environ["HF_B200_LIB"])
slow_sets = []
for rep in range(3):
    shape.ray_intersect(ray, out=(si, pi))
    torch.cuda.synchronize()
    buf = np.zeros(8192 * 6, np.uint64)
    assert lib.hf_debug_read(buf.ctypes.data_as(C.c_void_p)) == 0
    bb = buf.reshape(-1, 6); bb = bb[bb[:, 0] > 0]
    k = np.argsort(bb[:, 3])[-12:]
    slow_sets.append(sorted(int(x) for x in bb[k, 4]))
    print("rep", rep, "slowest fetch ids:", slow_sets[-1], "durations us:", [round(float(x) / 1e3) for x in bb[k, 3]])
valid = pi.is_valid().cpu().numpy()
ngroups = (n + 31) // 32
fetch_tail = min(ngroups, 592 * 8 * 4 * 2); nbig = (ngroups - fetch_tail) // 4
for tfetch in slow_sets[-1][-6:]:
    g0 = tfetch * 4 if tfetch < nbig else nbig * 4 + (tfetch - nbig); ng = 4 if tfetch < nbig else 1
    v = valid[g0 * 32:(g0 + ng) * 32]
    print(f"  fetch {tfetch}: groups {g0}..{g0 + ng - 1}, hit lanes {int(v.sum())}/{v.size}, pixel index {g0 // 2}")
b = buf.reshape(-1, 6); b = b[b[:, 0] > 0]
t0 = b[:, 0].min(); end = (b[:, 1] - t0) / 1e3; start = (b[:, 0] - t0) / 1e3
print(f"warps {len(b)}, rays {n}; start spread {start.max():.1f} us; end: min {end.min():.1f} median {np.median(end):.1f} p90 {np.percentile(end, 90):.1f} p99 {np.percentile(end, 99):.1f} max {end.max():.1f} us")
print("fetches per warp: min %d median %d max %d" % (b[:, 2].min(), np.median(b[:, 2]), b[:, 2].max()))
mx = b[:, 3] / 1e3
print(f"longest single fetch per warp: median {np.median(mx):.1f} p99 {np.percentile(mx, 99):.1f} max {mx.max():.1f} us")
k = np.argsort(end)[-8:]
for i in k: print(f"  warp {i}: end {end[i]:.1f} us, fetches {b[i,2]}, longest fetch {mx[i]:.1f} us (its #{b[i,4]}), last counter value {b[i,5]}")
hist, edges = np.histogram(end, bins=12)
print("end-time histogram:", [(round(float(e)), int(c)) for e, c in zip(edges[:-1], hist)])
