"""Dev tool: bundle-entry (hf_beam_enter) statistics on the C3 sample, host build. Not part of the product."""
import ctypes as C, os, subprocess, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import hf_oracle as O
import hfb200
from hfb200 import workloads as wl

def build(extra=()):
    so = "/tmp/trav_beam.so"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", *extra, "-o", so,
                           os.path.join(ROOT, "tests/trav_host.cpp"), "-L" + os.path.join(ROOT, "oracle/_build"),
                           "-lhf_oracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle/_build")])
    L = C.CDLL(so)
    L.trav_host_intersect_beam.argtypes = [C.POINTER(O._Shape), C.POINTER(O._Rays), C.c_size_t, C.POINTER(O._Prelim), O._u32p, O._u32p, O._u32p, O._f32p]
    return L

def run(L, hf, rays):
    out = O.Prelim(rays.n); it = np.zeros(rays.n, np.uint32); nc = np.zeros(rays.n, np.uint32)
    ng = (rays.n + 31) // 32
    bi = np.zeros(ng, np.uint32); te = np.zeros(ng, np.float32)
    s, r, o = hf._shape(), rays.c(), out.c()
    L.trav_host_intersect_beam(C.byref(s), C.byref(r), rays.n, C.byref(o), it.ctypes.data_as(O._u32p), nc.ctypes.data_as(O._u32p),
                               bi.ctypes.data_as(O._u32p), te.ctypes.data_as(O._f32p))
    return out, it, nc, bi, te

if __name__ == "__main__":
    n = int(os.environ.get("HF_SIM_N", "4096"))
    L = build(sys.argv[1:])
    h = wl.fbm_heights(n, octaves=8, seed=0xC3).numpy()
    hf = O.OracleHeightfield(h, 0.3)
    cam = wl.CONFIGS["C3"]["camera"]
    pix = torch.arange(0, 1024 * 1024, 257, dtype=torch.int64)
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, pixels=pix)
    r = O.Rays(o.numpy().T, d.numpy().T)
    ref = hf.ray_intersect_preliminary(r)
    out, it, nc, bi, te = run(L, hf, r)
    same = (out.t.view(np.uint32) == ref.t.view(np.uint32)) & (out.prim_index == ref.prim_index)
    hit = ref.valid(); hitw = hit.reshape(-1, 32).any(1)
    print(f"mismatches {(~same).sum()} of {r.n}; hit {hit.mean():.3f}")
    print(f"bundle iterations: mean {bi.mean():.1f} (hit bundles {bi[hitw].mean():.1f}), max {bi.max()}")
    print(f"per-ray walk iterations: hit rays {it[hit].mean():.2f}, all {it.mean():.2f}; warp-max hit warps {it.reshape(-1,32).max(1)[hitw].mean():.2f}; "
          f"cand rounds/hit warp {nc.reshape(-1,32).max(1)[hitw].mean():.2f}")
    if "-DHF_WALK_HIST" in sys.argv:
        L.trav_host_hist.restype = C.POINTER(C.c_ulonglong)
        hist = np.ctypeslib.as_array(L.trav_host_hist(), shape=(3, 32)).astype(np.float64)
        nh = hit.sum()
        print("per hit ray: level | descend | step | found")
        for l in range(10): print(f"{l:3d}  {hist[0,l]/nh:6.2f} {hist[1,l]/nh:6.2f} {hist[2,l]/nh:6.2f}")
