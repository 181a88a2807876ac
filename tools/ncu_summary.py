"""Dev tool: print the metrics we care about from an .ncu-rep (raw page) and the hottest source lines."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__issue_inst0.avg.pct_of_peak_sustained_active"]
for vals in rows[2:]:
    print("== kernel:", vals[hdr.index("Kernel Name")][:80])
    for w in WANT:
        if w in hdr:
            i = hdr.index(w); print(f"  {w:70s} {vals[i]:>18s} {units[i]}")
    for i, h in enumerate(hdr):
        if "warp_issue_stalled" in h and h.endswith("per_warp_active.pct") and float(vals[i] or 0) > 3:
            print(f"  {h:70s} {vals[i]:>18s}")
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    sec = None; hdr = None; agg = collections.defaultdict(lambda: [0, 0, 0]); text = {}
    for r in rows:
        if not r: continue
        if r[0] == "File Path": sec = r[1].split('/')[-1]; continue
        if r[0] == "Function Name": continue
        if r[0] == "Line No": hdr = r; continue
        if hdr is None or r[0] == '': continue
        d = dict(zip(hdr, r))
        try:
            k = (sec, int(r[0])); agg[k][0] += int(d.get("Instructions Executed") or 0); agg[k][1] += int(d.get("Thread Instructions Executed") or 0); agg[k][2] += int(d.get("# Samples") or 0)
            text[k] = r[1]
        except ValueError:
            pass
    tot = sum(v[0] for v in agg.values()); tots = sum(v[2] for v in agg.values())
    print("total warp inst", tot, "samples", tots)
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:int(sys.argv[2])]:
        print(f"{k[0][:14]}:{k[1]:4d} inst {v[0]/tot*100:5.1f}% thr/inst {v[1]/max(v[0],1):5.1f} samp {v[2]/tots*100:5.1f}% | {text[k][:100]}")
