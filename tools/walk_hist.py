"""Dev tool: per-level histogram of walk iterations (descend / step / found) per hit ray on the C3 sample."""
import sys, os, subprocess, ctypes as C, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import sim_walk as S
from oracle import hf_oracle as O
from hfb200 import workloads as wl
extra = sys.argv[1:]
subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", "-DHF_WALK_HIST", *extra, "-o", "/tmp/trav_hist.so",
                       os.path.join(ROOT, "tests/trav_host.cpp"), "-L" + os.path.join(ROOT, "oracle/_build"), "-lhf_oracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle/_build")])
L = C.CDLL('/tmp/trav_hist.so')
L.trav_host_intersect.argtypes = [C.POINTER(O._Shape), C.POINTER(O._Rays), C.c_size_t, C.POINTER(O._Prelim), O._u32p, O._u32p, O._u32p]
L.trav_host_hist.restype = C.POINTER(C.c_ulonglong)
n = int(os.environ.get("HF_SIM_N", "4096"))
h = wl.fbm_heights(n, octaves=8, seed=0xC3).numpy()
hf = O.OracleHeightfield(h, 0.3)
cam = wl.CONFIGS["C3"]["camera"]
pix = torch.arange(0, 1024*1024, 257, dtype=torch.int64)
o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, pixels=pix)
r = O.Rays(o.numpy().T, d.numpy().T)
out, it, nc = S.run(L, hf, r)
ref = hf.ray_intersect_preliminary(r)
print("mismatch", int(((out.t.view(np.uint32) != ref.t.view(np.uint32)) | (out.prim_index != ref.prim_index)).sum()))
hist = np.ctypeslib.as_array(L.trav_host_hist(), shape=(3,32)).astype(np.float64)
nh = out.valid().sum()
print("per hit ray: level | descend | step | found")
for l in range(13): print(f"{l:3d}  {hist[0,l]/nh:6.2f} {hist[1,l]/nh:6.2f} {hist[2,l]/nh:6.2f}")
print("totals", hist.sum(1)/nh, "iters/hit ray", it[out.valid()].mean())
rd = S.run.rounds.reshape(-1, 32, 8); hitw = out.valid().reshape(-1, 32).any(1)
print("lock-step iterations per hit warp", rd.max(1).sum(1)[hitw].mean(), "MT rounds", nc.reshape(-1,32).max(1)[hitw].mean())
