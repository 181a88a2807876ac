"""Dev tool: group the source page of an .ncu-rep by code region (warp instructions, lane utilisation) for one kernel id."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"] + sys.argv[2:], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
sec = None; hdr = None; agg = collections.defaultdict(lambda: [0, 0]); text = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": sec = r[1].split('/')[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or r[0] == '': continue
    d = dict(zip(hdr, r))
    try:
        k = (sec, int(r[0])); agg[k][0] += int(d.get("Instructions Executed") or 0); agg[k][1] += int(d.get("Thread Instructions Executed") or 0)
        text[k] = r[1]
    except ValueError:
        pass
def region(f, l):
    if f.startswith("hf_traverse"):
        if 213 <= l <= 310: return "walk loop"
        if 108 <= l <= 128: return "walk int helpers"
        if 129 <= l <= 212: return "walk init"
        if l >= 320: return "beam march"
    if f.startswith("hf_device"):
        if 100 <= l <= 156: return "cell_intersect/MT"
        return "hf_device other (vec, xform, rcp, rsqrt, vtx)"
    if f.startswith("hf_walk_common"): return "walk_common (staging, scatter)"
    if f.startswith("hf_render"): return "hf_render.cu"
    if f.startswith("hf_kernels"): return "hf_kernels.cu"
    return "intrinsics/" + f[:18]
g = collections.defaultdict(lambda: [0, 0])
for (f, l), v in agg.items():
    r = region(f, l); g[r][0] += v[0]; g[r][1] += v[1]
tot = sum(v[0] for v in g.values())
print("total warp inst", tot)
for r, v in sorted(g.items(), key=lambda kv: -kv[1][0]):
    print(f"{r:48s} {v[0]/tot*100:5.1f}%  thr/inst {v[1]/max(v[0],1):5.1f}")
