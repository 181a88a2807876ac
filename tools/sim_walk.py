"""Dev tool: compile the host instantiation of the production walk with given HF_START_LEVEL / HF_POP_LEVELS,
check it against the oracle and print per-warp iteration statistics on a C3-like sample. Not part of the product."""
import ctypes as C, os, subprocess, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import hf_oracle as O
import hfb200
from hfb200 import workloads as wl
from util import adversarial_rays, random_heights, random_rays

def build(start, pop):
    so = f"/tmp/trav_{start}_{pop}.so"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", f"-DHF_START_LEVEL={start}",
                           f"-DHF_POP_LEVELS={pop}", "-o", so, os.path.join(ROOT, "tests/trav_host.cpp"), "-L" + os.path.join(ROOT, "oracle/_build"),
                           "-lhf_oracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle/_build")])
    L = C.CDLL(so)
    L.trav_host_intersect.argtypes = [C.POINTER(O._Shape), C.POINTER(O._Rays), C.c_size_t, C.POINTER(O._Prelim), O._u32p, O._u32p, O._u32p]
    return L

def run(L, hf, rays):
    out = O.Prelim(rays.n); it = np.zeros(rays.n, np.uint32); nc = np.zeros(rays.n, np.uint32)
    run.rounds = np.zeros((rays.n, 8), np.uint32)
    s, r, o = hf._shape(), rays.c(), out.c()
    L.trav_host_intersect(C.byref(s), C.byref(r), rays.n, C.byref(o), it.ctypes.data_as(O._u32p), nc.ctypes.data_as(O._u32p),
                          run.rounds.ctypes.data_as(O._u32p))
    return out, it, nc

def check(L):
    bad = 0
    for (H, W, kind) in [(2, 2, "noise"), (3, 2, "noise"), (5, 7, "smooth"), (17, 17, "smooth"), (33, 20, "noise"), (64, 65, "smooth"), (12, 12, "flat"), (300, 257, "smooth")]:
        rng = np.random.default_rng(H * 131 + W)
        h = random_heights(rng, H, W, kind)
        hf = O.OracleHeightfield(h, 0.7)
        o, d = random_rays(rng, 200000); ao, ad = adversarial_rays(W, H, 0.7)
        o, d = np.concatenate([o, ao]), np.concatenate([d, ad])
        maxt = np.where(rng.uniform(size=o.shape[0]) < 0.15, rng.uniform(0.5, 4, o.shape[0]), np.inf).astype(np.float32)
        r = O.Rays(o, d, maxt)
        a = hf.ray_intersect_preliminary(r)
        b, _, _ = run(L, hf, r)
        same = (a.t.view(np.uint32) == b.t.view(np.uint32)) & (a.prim_index == b.prim_index) & (a.u.view(np.uint32) == b.u.view(np.uint32))
        bad += int((~same).sum())
        if (~same).any():
            i = np.nonzero(~same)[0][:5]; print("  mismatch", H, W, i, a.t[i], b.t[i], a.prim_index[i], b.prim_index[i])
    return bad

if __name__ == "__main__":
    n = int(os.environ.get("HF_SIM_N", "2048"))
    h = wl.fbm_heights(n, octaves=8, seed=0xC3).numpy()
    hf = O.OracleHeightfield(h, 0.3)
    cam = wl.CONFIGS["C3"]["camera"]
    pix = torch.arange(0, 1024 * 1024, 257, dtype=torch.int64)
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, pixels=pix)
    r = O.Rays(o.numpy().T, d.numpy().T)
    ref, nn, nl = hf.ray_intersect_preliminary(r, counters=True)
    print(f"field {n}^2, {r.n} rays, hit {ref.valid().mean():.3f}; canonical nodes/ray {nn.mean():.2f} leaves/ray {nl.mean():.2f}; per hit-ray iterations {(nn+nl)[ref.valid()].mean():.2f}")
    for arg in sys.argv[1:]:
        s, p = (int(x) for x in arg.split(","))
        L = build(s, p)
        bad = check(L)
        out, it, nc = run(L, hf, r)
        same = (out.t.view(np.uint32) == ref.t.view(np.uint32)) & (out.prim_index == ref.prim_index)
        w_it = it.reshape(-1, 32).max(1); w_nc = nc.reshape(-1, 32).max(1)
        hitw = ref.valid().reshape(-1, 32).any(1)
        rd = run.rounds.reshape(-1, 32, 8)
        lockstep = rd.max(1).sum(1)            # sum over rounds of the slowest lane of the round
        lanes = it.reshape(-1, 32).sum(1)
        print(f"   per hit warp: lock-step rounds {lockstep[hitw].mean():.2f} iterations vs slowest lane {w_it[hitw].mean():.2f}; "
              f"lane utilisation {lanes[hitw].sum() / (32.0 * lockstep[hitw].sum()):.3f}")
        print(f"start {s:2d} pop {p:2d}: mismatches(random/adv) {bad}, C3 mismatches {(~same).sum()}; iters/ray {it.mean():.2f} (hit rays {it[ref.valid()].mean():.2f}); "
              f"warp-max iters: all {w_it.mean():.2f}, hit warps {w_it[hitw].mean():.2f}; cand rounds/hit warp {w_nc[hitw].mean():.2f} (per hit ray {nc[ref.valid()].mean():.2f})")
