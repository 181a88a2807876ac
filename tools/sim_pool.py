"""Dev tool: instruction-count model of two warp schedules on the C3 sample (host build of the production walk):
  A. today's: one ray per lane per 32-ray group, lock-step rounds (walk until every working lane has a candidate, test them together);
  B. a lane pool ("while-while"): a warp owns P rays; a lane whose ray is finished takes the next ray of the pool at once; a round
     lasts at most K walk iterations, then the lanes that hold a candidate are tested together.
Costs in warp instructions: C_IT per lock-step walk iteration, C_MT per primitive round, C_INIT per (re)fill, C_OUT per 32 results."""
import sys, os, subprocess, ctypes as C, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import hf_oracle as O
from hfb200 import workloads as wl
R = 96
subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", f"-DHF_HOST_ROUNDS={R}", "-o", "/tmp/trav_tail.so",
                       os.path.join(ROOT, "tests/trav_host.cpp"), "-L" + os.path.join(ROOT, "oracle/_build"), "-lhf_oracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle/_build")])
L = C.CDLL('/tmp/trav_tail.so')
L.trav_host_intersect.argtypes = [C.POINTER(O._Shape), C.POINTER(O._Rays), C.c_size_t, C.POINTER(O._Prelim), O._u32p, O._u32p, O._u32p]
n = 4096
h = wl.fbm_heights(n, octaves=8, seed=0xC3).numpy()
hf = O.OracleHeightfield(h, 0.3)
cam = wl.CONFIGS["C3"]["camera"]
# whole pixels (64 spp = 2 groups), pairs of neighbouring pixels so that a pool of 128 rays is what a 4-group fetch would be
if os.environ.get("HF_SIM_C5"):
    o, d = wl.hemisphere_rays(1 << 17, seed=5)
else:
    base = torch.arange(0, 1024 * 1024 - 1, 122, dtype=torch.int64)
    pix = torch.stack([base, base + 1], 1).reshape(-1)
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, pixels=pix)
r = O.Rays(o.numpy().T, d.numpy().T)
out = O.Prelim(r.n); it = np.zeros(r.n, np.uint32); nc = np.zeros(r.n, np.uint32); rounds = np.zeros((r.n, R), np.uint32)
s, rr, oo = hf._shape(), r.c(), out.c()
L.trav_host_intersect(C.byref(s), C.byref(rr), r.n, C.byref(oo), it.ctypes.data_as(O._u32p), nc.ctypes.data_as(O._u32p), rounds.ctypes.data_as(O._u32p))
hit = out.valid()
C_IT, C_MT, C_INIT, C_OUT = 69.0, 130.0, 150.0, 300.0
G = r.n // 32
rd = rounds.reshape(G, 32, R).astype(np.int64)
act = rd > 0
costA = (rd.max(1) * C_IT + act.any(1) * C_MT).sum(1) + C_INIT + C_OUT
print(f"A: {costA.mean():.0f} warp instructions per 32 rays (measured on the GPU: 1510); lock-step iterations {rd.max(1).sum(1).mean():.1f}, rounds {act.any(1).sum(1).mean():.2f}")

def simB(P, K, refill_min):
    tot = 0.0; iters_tot = 0; mt_tot = 0; fill_tot = 0
    npool = r.n // P
    for p in range(0, npool, 7):           # a sample of the pools
        seq = rounds[p * P:(p + 1) * P]; ncs = nc[p * P:(p + 1) * P]; hh = hit[p * P:(p + 1) * P]
        nxt = 0
        ray = [-1] * 32; k = [0] * 32; rem = [0] * 32
        cost = 0.0
        while True:
            idle = [l for l in range(32) if ray[l] < 0]
            if nxt < P and len(idle) >= refill_min or all(x < 0 for x in ray):
                if nxt >= P and all(x < 0 for x in ray): break
                filled = False
                for l in idle:
                    if nxt >= P: break
                    ray[l] = nxt; k[l] = 0; rem[l] = int(seq[nxt, 0]); nxt += 1; filled = True
                    if rem[l] == 0: ray[l] = -1      # inactive ray / missed the slab: costs only the fill
                if filled: cost += C_INIT; fill_tot += 1
                continue
            live = [l for l in range(32) if ray[l] >= 0]
            if not live:
                if nxt >= P: break
                refill_min = 1; continue
            step = min(K, max(rem[l] for l in live))
            cost += step * C_IT; iters_tot += step
            cand = False
            for l in live:
                if rem[l] <= step:
                    # event: candidate (round index k < nc) or end of ray
                    j = ray[l]
                    if k[l] < ncs[j]:
                        cand = True
                        k[l] += 1
                        if k[l] == ncs[j] and hh[j]: ray[l] = -1            # that candidate was the hit
                        else:
                            rem[l] = int(seq[j, k[l]]) if k[l] < R else 0
                            if rem[l] == 0: ray[l] = -1
                    else:
                        ray[l] = -1
                else:
                    rem[l] -= step
            if cand: cost += C_MT; mt_tot += 1
        cost += C_OUT * P / 32
        tot += cost
    npl = len(range(0, npool, 7))
    return tot / npl / (P / 32), iters_tot / npl / (P / 32), mt_tot / npl / (P / 32), fill_tot / npl / (P / 32)

for P in (256,):
    for K in (4, 8, 1000):
        for rm in (1, 8, 16):
            c, i_, m_, f_ = simB(P, K, rm)
            print(f"B: pool {P} K {K} refill when >= {rm} idle: {c:.0f} per 32 rays (iterations {i_:.1f}, primitive rounds {m_:.2f}, fills {f_:.2f})")
