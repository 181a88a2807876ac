"""Dev tool: lock-step cost model of 32-ray groups on the C3 sample: how much of a group's time is spent with few lanes active
(what a segment-parallel finish of the last rays could win). Host build of the production walk."""
import sys, os, subprocess, ctypes as C, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
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
step = int(os.environ.get("PIXSTEP", "61"))
pix = torch.arange(0, 1024*1024, step, dtype=torch.int64)
o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, pixels=pix)
r = O.Rays(o.numpy().T, d.numpy().T)
out = O.Prelim(r.n); it = np.zeros(r.n, np.uint32); nc = np.zeros(r.n, np.uint32); rounds = np.zeros((r.n, R), np.uint32)
s, rr, oo = hf._shape(), r.c(), out.c()
L.trav_host_intersect(C.byref(s), C.byref(rr), r.n, C.byref(oo), it.ctypes.data_as(O._u32p), nc.ctypes.data_as(O._u32p), rounds.ctypes.data_as(O._u32p))
G = r.n // 32
rd = rounds.reshape(G, 32, R).astype(np.float64); ncg = nc.reshape(G, 32)
# lane is active in round k if it recorded iterations there (k <= nc: the round that ends the ray included)
act = rd > 0
MT = float(os.environ.get("MT_COST", "2.5"))       # cost of a primitive round in units of one loop iteration
cost_round = rd.max(1) + MT * (act.any(1))         # (G, R)
nact = act.sum(1)                                  # (G, R)
tot = cost_round.sum(1)
for thr in (1, 2, 4, 8):
    few = (cost_round * ((nact <= thr) & (nact > 0))).sum(1)
    print(f"lanes <= {thr}: share of all group time {few.sum()/tot.sum():.3f}; in the 1% slowest groups {few[np.argsort(tot)[-G//100:]].sum()/np.sort(tot)[-G//100:].sum():.3f}")
print("groups", G, "mean cost", tot.mean(), "max", tot.max(), "p99", np.percentile(tot, 99), "max rounds", ncg.max())
w = np.argsort(tot)[-5:]
for g in w:
    print("group", g, "cost", tot[g], "rounds", int(act.any(0)[g].sum()) if False else int((nact[g] > 0).sum()), "active lanes per round", nact[g][:40])
