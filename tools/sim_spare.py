"""Dev tool: lock-step cost model of a "spare candidate" schedule on the C3 sample (host build of the production walk).
  A. today's: walk until every working lane holds ONE candidate, test them together, repeat;
  B. a lane that holds a candidate keeps walking (it would idle otherwise) and queues up to Q candidates; after a failed primitive
     test it has the next candidate ready instead of forcing a new lock-step walk phase on the warp.
Costs in warp instructions: C_IT per lock-step walk iteration, C_MT per primitive round."""
import sys, os, subprocess, ctypes as C, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import hf_oracle as O
from hfb200 import workloads as wl
R = 32
subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", f"-DHF_HOST_ROUNDS={R}", "-o", "/tmp/trav_spare.so",
                       os.path.join(ROOT, "tests/trav_host.cpp"), "-L" + os.path.join(ROOT, "oracle/_build"), "-lhf_oracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle/_build")])
L = C.CDLL('/tmp/trav_spare.so')
L.trav_host_intersect.argtypes = [C.POINTER(O._Shape), C.POINTER(O._Rays), C.c_size_t, C.POINTER(O._Prelim), O._u32p, O._u32p, O._u32p]
n = int(os.environ.get("HF_SIM_N", "4096"))
h = wl.fbm_heights(n, octaves=8, seed=0xC3).numpy()
hf = O.OracleHeightfield(h, 0.3)
cam = wl.CONFIGS["C3"]["camera"]
pix = torch.arange(0, 1024 * 1024, 509, dtype=torch.int64)
o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, pixels=pix)
r = O.Rays(o.numpy().T, d.numpy().T)
out = O.Prelim(r.n); it = np.zeros(r.n, np.uint32); nc = np.zeros(r.n, np.uint32); rounds = np.zeros((r.n, R), np.uint32)
s, rr, oo = hf._shape(), r.c(), out.c()
L.trav_host_intersect(C.byref(s), C.byref(rr), r.n, C.byref(oo), it.ctypes.data_as(O._u32p), nc.ctypes.data_as(O._u32p), rounds.ctypes.data_as(O._u32p))
hit = out.valid()
C_IT, C_MT = 69.0, 130.0
G = r.n // 32

def sim(Q):
    tot_it = 0; tot_mt = 0
    for g in range(G):
        sl = slice(g * 32, g * 32 + 32)
        seq = rounds[sl].astype(np.int64); ncs = nc[sl].astype(np.int64); hh = hit[sl]
        k = np.zeros(32, np.int64); rem = seq[:, 0].copy(); q = np.zeros(32, np.int64); tested = np.zeros(32, np.int64)
        walk_done = rem == 0
        resolved = walk_done.copy()
        while not resolved.all():
            # walk phase
            while True:
                need = ~resolved & (q == 0) & ~walk_done
                if not need.any(): break
                adv = ~resolved & ~walk_done & (q < Q)
                step = int(rem[adv & need].min())             # jump to the next event among the lanes the warp waits for ...
                step = min(step, int(rem[adv].min()))         # ... or any advancing lane
                tot_it += step
                rem[adv] -= step
                done = adv & (rem == 0)
                for l in np.nonzero(done)[0]:
                    if k[l] < ncs[l]:
                        q[l] += 1
                        if hh[l] and k[l] == ncs[l] - 1: walk_done[l] = True      # nothing recorded past the hit
                    else:
                        walk_done[l] = True
                    k[l] += 1
                    if not walk_done[l]:
                        rem[l] = seq[l, k[l]] if k[l] < R else 1
                        if rem[l] == 0: walk_done[l] = True
            # primitive round
            have = ~resolved & (q > 0)
            if have.any():
                tot_mt += 1
                for l in np.nonzero(have)[0]:
                    if hh[l] and tested[l] == ncs[l] - 1: resolved[l] = True
                    tested[l] += 1; q[l] -= 1
            resolved |= (~resolved & walk_done & (q == 0))
    return tot_it / G, tot_mt / G

for Q in (1, 2, 3):
    a, b = sim(Q)
    print(f"Q={Q}: lock-step iterations {a:.2f}, primitive rounds {b:.2f} per 32 rays -> {a * C_IT + b * C_MT:.0f} warp instructions (walk + primitive)")
