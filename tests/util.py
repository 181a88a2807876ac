"""Shared helpers for the tests: random / adversarial ray batches and comparison utilities."""
import numpy as np


def random_heights(rng, H, W, kind="smooth"):
    if kind == "flat":
        return np.full((H, W), 0.5, np.float32)
    if kind == "noise":
        return rng.uniform(0, 1, (H, W)).astype(np.float32)
    y, x = np.mgrid[0:H, 0:W]
    h = 0.5 + 0.25 * np.sin(2 * np.pi * 3 * x / max(W - 1, 1)) * np.cos(2 * np.pi * 2 * y / max(H - 1, 1))
    h += 0.1 * rng.uniform(-1, 1, (H, W))
    return h.astype(np.float32)


def random_rays(rng, n, zlo=-0.2, zhi=1.2, spread=1.6):
    """Origins in a box around the field, aimed at random points inside/near it."""
    o = np.stack([rng.uniform(-spread, spread, n), rng.uniform(-spread, spread, n), rng.uniform(zlo, zhi + 1.5, n)], 1)
    tgt = np.stack([rng.uniform(-1.1, 1.1, n), rng.uniform(-1.1, 1.1, n), rng.uniform(zlo, zhi, n)], 1)
    d = tgt - o
    d *= rng.uniform(0.3, 2.0, (n, 1))          # not normalised
    return o.astype(np.float32), d.astype(np.float32)


def adversarial_rays(W, H, max_height=1.0):
    """Axis-parallel, grazing, corner-crossing, starting-inside and degenerate rays."""
    xs = [(2 * i - (W - 1)) / (W - 1) for i in range(W)]
    ys = [(2 * j - (H - 1)) / (H - 1) for j in range(H)]
    o, d = [], []
    for x in (xs[0], xs[W // 2], xs[-1], 0.123, -0.77):
        for y in (ys[0], ys[H // 2], ys[-1], 0.311, -0.5):
            o.append((x, y, 5.0)); d.append((0, 0, -1))          # straight down (vertices / edges / interior)
            o.append((x, y, -5.0)); d.append((0, 0, 1))          # straight up (back side)
    for y in (ys[H // 3], 0.2718, ys[0], ys[-1]):
        for z in (0.1, 0.5 * max_height, 0.9 * max_height, max_height):
            o.append((-3, y, z)); d.append((1, 0, 0))            # d.y == 0, grazing along x
            o.append((3, y, z)); d.append((-1, 0, 0))
            o.append((y, -3, z)); d.append((0, 1, 0))
            o.append((y, 3, z)); d.append((0, -2, 0))
    for s in (1, -1):
        o.append((-2 * s, -2 * s, 0.5 * max_height)); d.append((s, s, 0))          # exactly along the cell diagonals
        o.append((-2 * s, 2 * s, 0.4 * max_height)); d.append((s, -s, 0.001))
        o.append((-2 * s, -2 * s, 2.0)); d.append((s, s, -0.7))
    o.append((0.01, 0.02, 0.5 * max_height)); d.append((0.3, 0.2, 0.1))               # starts inside the bbox
    o.append((0.01, 0.02, 0.5 * max_height)); d.append((0.3, 0.2, -0.4))
    o.append((0, 0, 3)); d.append((0, 0, 0))                                           # degenerate direction
    o.append((5, 5, 5)); d.append((1, 1, 1))                                           # points away
    o.append((0, 0, 3)); d.append((0, 0, 1))                                           # above, going up
    o.append((0.3, -0.2, 3)); d.append((1e-8, 0, -1))                                  # nearly vertical
    o.append((-1.5, 0.1, 0.05)); d.append((1, 0.001, 1e-4))                            # very shallow
    return np.array(o, np.float32), np.array(d, np.float32)


def ulp_diff(a, b):
    """Distance in units of last place between fp32 arrays (inf == inf -> 0)."""
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    ai = a.view(np.int32).astype(np.int64); bi = b.view(np.int32).astype(np.int64)
    ai = np.where(ai < 0, np.int64(-2**31) - ai, ai); bi = np.where(bi < 0, np.int64(-2**31) - bi, bi)
    return np.abs(ai - bi)
