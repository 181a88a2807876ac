#!/usr/bin/env python
"""SASS of the three hot kernels, with the loops marked and instruction counts per region, into profiles/<prefix>_sass_*.txt.

    python tools/sass_listing.py [prefix=r03]

A loop is a backward branch (BRA to a lower address); its body is the address range [target, branch]. For every loop the header of
the listing gives the number of instructions inside it and their split over issue classes (fp32 = FADD/FMUL/FFMA, alu = integer /
compare / select / min-max / shift / logic / convert, which share the half-rate ALU pipe on sm_100, lsu = loads / stores / atomics,
sfu = MUFU, ctl = branches / barriers / votes / shuffles), so that the claims in profiles/*_notes.md ("N instructions in the walk loop, M on
the ALU pipe") can be checked against the tree and compared between rounds. Needs only cuobjdump (no GPU)."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mitsuba3-differentiable-heightfield-rendering_b200", "lib", "libhf_b200.so")
KERNELS = {
    "fast": ("_Z24hf_intersect_fast_kernelILi4ELb0EEv8HfParams12HfLaunchInfo7hf_rays9hf_prelim5hf_siPhjmyyPy", "hf_intersect_fast_kernel<FUSED_SI_LEAN, false>"),
    "backward": ("_Z18hf_backward_kernelILb0ELb1EEv8HfParams7hf_rays9hf_prelim9hf_si_adjjPfm", "hf_backward_kernel<false, true>"),
    "render": ("_Z16hf_render_kernelILi2ELb0ELb0EEv8HfParams12HfLaunchInfo12HfRenderArgsPfPjPKfS6_S3_PdPy", "hf_render_kernel<STEP_L2, false, false>"),
}
CLASSES = (("fp32", r"^(FADD|FMUL|FFMA|FADD2|FMUL2|FFMA2)\b"), ("sfu", r"^MUFU"),
           ("lsu", r"^(LDG|STG|LDS|STS|LDC|LDCU|LD|ST|RED|REDG|ATOM|ATOMG|ATOMS|LDL|STL|UBLKCP|SYNCS|LDSM)\b"),
           ("ctl", r"^(BRA|BSSY|BSYNC|EXIT|WARPSYNC|BAR|VOTE|VOTEU|SHFL|MATCH|REDUX|NANOSLEEP|RET|CALL|BREAK|YIELD|ELECT|MEMBAR|ERRBAR|CCTL|NOP|DEPBAR)\b"),
           ("uniform", r"^(U[A-Z0-9]+|R2UR|S2UR|S2R|CS2R|R2P|P2R)\b"))


def classify(op):
    for name, rx in CLASSES:
        if re.match(rx, op):
            return name
    return "alu"


def main():
    prefix = sys.argv[1] if len(sys.argv) > 1 else "r03"
    for short, (mangled, pretty) in KERNELS.items():
        txt = subprocess.run(["cuobjdump", "-sass", "-fun", mangled, LIB], capture_output=True, text=True, check=True).stdout
        ins = []            # (addr, opcode, text)
        for line in txt.splitlines():
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
            if not m:
                continue
            body = re.sub(r"^@!?U?P\d+\s+", "", m.group(2).strip())
            ins.append((int(m.group(1), 16), body.split()[0].split(".")[0], m.group(2).strip()))
        loops = []
        for addr, op, text in ins:
            if op == "BRA":
                t = re.search(r"0x([0-9a-f]+)", text)
                if t and int(t.group(1), 16) <= addr:
                    loops.append((int(t.group(1), 16), addr))
        loops.sort(key=lambda r: (r[0], -r[1]))
        out = [f"# {pretty}  ({mangled})", f"# cuobjdump -sass of lib/libhf_b200.so; {len(ins)} instructions", "#"]
        tot = {}
        for _, op, _t in ins:
            tot[classify(op)] = tot.get(classify(op), 0) + 1
        out.append("# whole kernel: " + ", ".join(f"{k} {v}" for k, v in sorted(tot.items())))
        for k, (lo, hi) in enumerate(loops):
            body = [i for i in ins if lo <= i[0] <= hi]
            cnt = {}
            for _, op, _t in body:
                cnt[classify(op)] = cnt.get(classify(op), 0) + 1
            inner = [j for j, (l2, h2) in enumerate(loops) if j != k and lo <= l2 and h2 <= hi]
            out.append(f"# loop L{k}: 0x{lo:04x} .. 0x{hi:04x}, {len(body)} instructions ({', '.join(f'{a} {b}' for a, b in sorted(cnt.items()))})"
                       + (f"; contains {', '.join('L%d' % j for j in inner)}" if inner else ""))
        out.append("#")
        begins, ends = {}, {}
        for k, (lo, hi) in enumerate(loops):
            begins.setdefault(lo, []).append(k); ends.setdefault(hi, []).append(k)
        for addr, op, text in ins:
            for k in begins.get(addr, []):
                out.append(f"        >>> loop L{k} begins")
            out.append(f"/*{addr:04x}*/ {text}")
            for k in ends.get(addr, []):
                out.append(f"        <<< loop L{k} ends")
        path = os.path.join(ROOT, "profiles", f"{prefix}_sass_{short}.txt")
        with open(path, "w") as f:
            f.write("\n".join(out) + "\n")
        print(path, len(ins), "instructions,", len(loops), "loops")


if __name__ == "__main__":
    main()
