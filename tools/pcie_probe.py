"""Dev tool (under gpurun): what the box's PCIe link gives pinned host <-> device copies, one direction alone and both at once
(the ceiling of the host-buffer e2e number in bench.py)."""
import torch
n = 1 << 30
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    for s in (s1, s2): torch.cuda.current_stream().wait_stream(s)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
def h2d():
    with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
def both(): h2d(); d2h()
for name, fn, b in (("h2d", h2d, n), ("d2h", d2h, n), ("both", both, 2 * n)):
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    ms = timed(fn)
    print(f"{name}: {b / ms / 1e6:.1f} GB/s ({ms:.1f} ms per GiB{' each way' if name == 'both' else ''})")
