"""Synthetic workloads of SURVEY.md 8(d): procedural heightfields, pinhole camera rays, hashed
adjoints. Pure tensor plumbing (torch, any device) -- nothing here is on the timed path.

Everything is a deterministic function of integer counters through a lowbias32-style hash, so a
test can regenerate the same inputs on the CPU for the oracle and on the GPU for the kernels.
"""
from __future__ import annotations

import math

import torch

M32 = 0xFFFFFFFF


def hash32(x: torch.Tensor) -> torch.Tensor:
    """lowbias32 on int64 tensors holding uint32 values."""
    x = x & M32
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & M32
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & M32
    x = x ^ (x >> 16)
    return x


def hash_uniform(seed: int, idx: torch.Tensor, channel: int = 0) -> torch.Tensor:
    """U[0,1) with 24 bits, fp32, from (seed, idx, channel)."""
    h = hash32(idx.to(torch.int64) ^ hash32(torch.full_like(idx, (seed * 0x9E3779B1 + channel * 0x85EBCA6B) & M32, dtype=torch.int64)))
    return (h >> 8).to(torch.float32) * (1.0 / 16777216.0)


# ---- heightfields ---------------------------------------------------------------------------
def sine_heights(n: int = 256, device="cpu") -> torch.Tensor:
    """C1: h = 0.5 + 0.25 sin(2 pi 4u) cos(2 pi 4v), u = i/(n-1), v = j/(n-1); fp64 then rounded."""
    u = torch.arange(n, dtype=torch.float64, device=device) / (n - 1)
    h = 0.5 + 0.25 * torch.sin(2 * math.pi * 4 * u)[None, :] * torch.cos(2 * math.pi * 4 * u)[:, None]
    return h.to(torch.float32).contiguous()


def fbm_heights(n: int, octaves: int = 8, seed: int = 0xC3, base_freq: float = 4.0, device="cpu",
                rows_per_chunk: int = 1024) -> torch.Tensor:
    """Value-noise fBm (lacunarity 2, gain 0.5), normalised to [0, 1]. (n, n) fp32."""
    out = torch.empty((n, n), dtype=torch.float32, device=device)
    xs = torch.arange(n, dtype=torch.float32, device=device) / (n - 1)
    for r0 in range(0, n, rows_per_chunk):
        r1 = min(n, r0 + rows_per_chunk)
        ys = torch.arange(r0, r1, dtype=torch.float32, device=device) / (n - 1)
        acc = torch.zeros((r1 - r0, n), dtype=torch.float32, device=device)
        amp, freq = 0.5, base_freq
        for o in range(octaves):
            fx, fy = xs * freq, ys * freq
            ix, iy = fx.floor(), fy.floor()
            tx, ty = fx - ix, fy - iy
            sx, sy = tx * tx * (3 - 2 * tx), ty * ty * (3 - 2 * ty)
            ixl, iyl = ix.to(torch.int64), iy.to(torch.int64)

            def lat(jx, jy):
                key = (jx[None, :] * 73856093) ^ (jy[:, None] * 19349663) ^ ((o + 1) * 83492791)
                return hash_uniform(seed, key, o)

            v00, v10 = lat(ixl, iyl), lat(ixl + 1, iyl)
            v01, v11 = lat(ixl, iyl + 1), lat(ixl + 1, iyl + 1)
            a = v00 + (v10 - v00) * sx[None, :]
            b = v01 + (v11 - v01) * sx[None, :]
            acc += amp * (a + (b - a) * sy[:, None])
            amp *= 0.5
            freq *= 2.0
        out[r0:r1] = acc
    lo, hi = out.min(), out.max()
    return ((out - lo) / (hi - lo)).contiguous()


# ---- cameras ----------------------------------------------------------------------------------
def _normalize(v):
    n = math.sqrt(sum(c * c for c in v))
    return tuple(c / n for c in v)


def _cross(a, b):
    return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])


def camera_rays(origin, target, up, fov_deg, width, height, spp, seed=0, device="cpu", pixels=None,
                jitter=True):
    """Pinhole primary rays (restating what src/sensors/perspective.cpp:198-234 produces: origin at the
    camera position, normalised direction, maxt = inf). Ray index = pixel * spp + sample as in
    src/render/integrator.cpp:251-265. `pixels`: optional int64 tensor of pixel ids (y*width + x) for a shard.
    Returns (o (3,N), d (3,N)) fp32."""
    fwd = _normalize(tuple(t - o for t, o in zip(target, origin)))
    right = _normalize(_cross(fwd, up))
    upv = _cross(right, fwd)
    if pixels is None:
        pixels = torch.arange(width * height, dtype=torch.int64, device=device)
    pixels = pixels.to(device)
    pid = pixels.repeat_interleave(spp)
    sid = torch.arange(spp, dtype=torch.int64, device=device).repeat(pixels.numel())
    px, py = (pid % width).to(torch.float32), (pid // width).to(torch.float32)
    if jitter and spp > 1:
        key = pid * spp + sid
        jx, jy = hash_uniform(seed, key, 0), hash_uniform(seed, key, 1)
    else:
        jx = jy = torch.full_like(px, 0.5)
    tanh = math.tan(math.radians(fov_deg) * 0.5)
    sx = (2.0 * (px + jx) / width - 1.0) * tanh
    sy = (1.0 - 2.0 * (py + jy) / height) * tanh * (height / width)
    d = torch.stack([fwd[k] + sx * right[k] + sy * upv[k] for k in range(3)])
    d = d / d.norm(dim=0, keepdim=True)
    o = torch.tensor(origin, dtype=torch.float32, device=device)[:, None].expand(3, d.shape[1]).contiguous()
    return o, d.contiguous()


def tile_pixels(width, height, tile, rank, world, device="cpu", skew=3):
    """Image tiles (tile x tile pixels) dealt round-robin to ranks; pixels of a rank ordered tile by tile, row-major
    inside a tile. The deal runs over a SKEWED tile order -- row r of tiles is rotated by skew * r columns -- because a plain
    row-major deal gives rank k the same tile COLUMNS in every row whenever the tiles per row divide by the number of ranks (32 tiles
    per row, 8 ranks: vertical stripes), and the cost of a pixel depends on its column: measured at 8 ranks on C3, the slowest rank
    was 18 % behind the fastest (profiles/r02_notes.md). Counts per rank still differ by at most one tile."""
    tx, ty = (width + tile - 1) // tile, (height + tile - 1) // tile
    slots = torch.arange(rank, tx * ty, world, dtype=torch.int64, device=device)       # positions in the skewed order
    t_y = slots // tx
    t_x = (slots % tx - skew * t_y) % tx                                                # slot (r, c) holds tile column (c - skew r) mod tx
    loc = torch.arange(tile * tile, dtype=torch.int64, device=device)
    lx, ly = loc % tile, loc // tile
    x = (t_x[:, None] * tile + lx[None, :]).reshape(-1)
    y = (t_y[:, None] * tile + ly[None, :]).reshape(-1)
    ok = (x < width) & (y < height)
    return (y * width + x)[ok]


def hemisphere_rays(n, seed=5, radius=3.0, device="cpu", start=0):
    """C5: origin uniform on the upper hemisphere of `radius`, target uniform on [-1,1]^2 x {0}. Rays start .. start + n - 1 of the
    stream (a chunk of a larger batch sees the same rays as the whole batch would)."""
    idx = torch.arange(start, start + n, dtype=torch.int64, device=device)
    u1, u2 = hash_uniform(seed, idx, 0), hash_uniform(seed, idx, 1)
    z = u1
    r = torch.sqrt(torch.clamp(1 - z * z, min=0))
    phi = 2 * math.pi * u2
    o = torch.stack([radius * r * torch.cos(phi), radius * r * torch.sin(phi), radius * z])
    tgt = torch.stack([2 * hash_uniform(seed, idx, 2) - 1, 2 * hash_uniform(seed, idx, 3) - 1, torch.zeros_like(z)])
    d = tgt - o
    d = d / d.norm(dim=0, keepdim=True)
    return o.contiguous(), d.contiguous()


def global_ray_ids(pixels, spp):
    """Ray id = pixel * spp + sample (src/render/integrator.cpp:251-265) for a list of pixel ids, in the order camera_rays emits them."""
    pixels = pixels.to(torch.int64)
    return pixels.repeat_interleave(spp) * spp + torch.arange(spp, dtype=torch.int64, device=pixels.device).repeat(pixels.numel())


def hashed_adjoints(n, seed=1, device="cpu", ids=None):
    """Upstream adjoints dt, dp, dn, duv ~ U[-1,1] from hash(seed, ray, channel). `ids`: the GLOBAL id of every
    ray (global_ray_ids) -- a shard of an image then sees exactly the adjoints the single-rank run gives those rays."""
    idx = torch.arange(n, dtype=torch.int64, device=device) if ids is None else ids.to(device=device, dtype=torch.int64)
    assert idx.numel() == n
    ch = lambda c: 2 * hash_uniform(seed, idx, c) - 1
    return {"t": ch(0), "p": torch.stack([ch(1), ch(2), ch(3)]), "n": torch.stack([ch(4), ch(5), ch(6)]),
            "uv": torch.stack([ch(7), ch(8)])}


# ---- named configurations (BASELINE.json configs / SURVEY.md 8(d)) ------------------------------
CONFIGS = {
    "C1": dict(heights=("sine", 256), max_height=0.25, camera=dict(origin=(0, 0, 3), target=(0, 0, 0), up=(0, 1, 0), fov=40.0),
               image=256, spp=1),
    "C2": dict(heights=("fbm", 1024, 6, 0xC2), max_height=0.3,
               camera=dict(origin=(0, -2.5, 1.8), target=(0, 0, 0), up=(0, 0, 1), fov=40.0), image=512, spp=16),
    "C3": dict(heights=("fbm", 4096, 8, 0xC3), max_height=0.3,
               camera=dict(origin=(0, -2.5, 1.8), target=(0, 0, 0), up=(0, 0, 1), fov=40.0), image=1024, spp=64),
    "C4": dict(heights=("fbm", 16384, 10, 0xC4), max_height=0.15,
               camera=dict(origin=(-1.4, 0, 0.25), target=(1, 0, 0.05), up=(0, 0, 1), fov=35.0), image=2048, spp=8),
}


def make_heights(spec, device="cpu"):
    if spec[0] == "sine":
        return sine_heights(spec[1], device=device)
    return fbm_heights(spec[1], octaves=spec[2], seed=spec[3], device=device)


def config_rays(name, device="cpu", seed=0, rank=0, world=1, tile=32, image=None, spp=None):
    cfg = CONFIGS[name]
    cam = cfg["camera"]
    image = image or cfg["image"]
    spp = spp or cfg["spp"]
    pixels = tile_pixels(image, image, tile, rank, world, device=device) if world > 1 or tile else None
    return camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], image, image, spp, seed=seed,
                       device=device, pixels=pixels, jitter=spp > 1)
