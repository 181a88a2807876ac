"""The production walk (csrc/hf_traverse.cuh) instantiated on the HOST and checked bit-for-bit against the oracle's canonical DDA (which test_oracle.py pins to brute force):
conservativeness of the warp-uniform hierarchical DDA and of the bundle march (hf_beam_enter) on random, adversarial
and camera rays -- without a GPU. The GPU parity tests then pin the device build of the same header to the oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from util import adversarial_rays, random_heights, random_rays

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def trav(oracle, tmp_path_factory):
    so = str(tmp_path_factory.mktemp("trav") / "trav_host.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", "-o", so,
                           os.path.join(ROOT, "tests", "trav_host.cpp"), "-L" + os.path.join(ROOT, "oracle", "_build"),
                           "-lhf_oracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle", "_build")])
    O = oracle
    L = C.CDLL(so)
    L.trav_host_intersect.argtypes = [C.POINTER(O._Shape), C.POINTER(O._Rays), C.c_size_t, C.POINTER(O._Prelim), O._u32p, O._u32p, O._u32p]
    L.trav_host_intersect_beam.argtypes = [C.POINTER(O._Shape), C.POINTER(O._Rays), C.c_size_t, C.POINTER(O._Prelim), O._u32p, O._u32p,
                                           O._u32p, O._f32p]
    return L


def _run(L, O, hf, rays, beam):
    out = O.Prelim(rays.n)
    s, r, o = hf._shape(), rays.c(), out.c()
    if beam:
        te = np.zeros((rays.n + 31) // 32, np.float32)
        L.trav_host_intersect_beam(C.byref(s), C.byref(r), rays.n, C.byref(o), None, None, None, te.ctypes.data_as(O._f32p))
        return out, te
    L.trav_host_intersect(C.byref(s), C.byref(r), rays.n, C.byref(o), None, None, None)
    return out, None


def _same(a, b):
    return (a.t.view(np.uint32) == b.t.view(np.uint32)) & (a.prim_index == b.prim_index) & (a.u.view(np.uint32) == b.u.view(np.uint32)) & \
           (a.v.view(np.uint32) == b.v.view(np.uint32))


@pytest.mark.parametrize("H,W,kind", [(2, 2, "noise"), (3, 2, "noise"), (5, 7, "smooth"), (17, 17, "smooth"), (33, 20, "noise"),
                                      (64, 65, "smooth"), (12, 12, "flat"), (300, 257, "smooth")])
def test_walk_equals_canonical_dda(oracle, trav, H, W, kind):
    O = oracle
    rng = np.random.default_rng(H * 131 + W)
    hf = O.OracleHeightfield(random_heights(rng, H, W, kind), 0.7)
    n = 60000 if H * W > 5000 else 200000
    o, d = random_rays(rng, n)
    ao, ad = adversarial_rays(W, H, 0.7)
    o, d = np.concatenate([o, ao]), np.concatenate([d, ad])
    maxt = np.where(rng.uniform(size=o.shape[0]) < 0.15, rng.uniform(0.5, 4, o.shape[0]), np.inf).astype(np.float32)
    r = O.Rays(o, d, maxt)
    ref = hf.ray_intersect_preliminary(r)          # canonical DDA; test_oracle.py pins it to brute force away from ties
    got, _ = _run(trav, O, hf, r, beam=False)
    assert _same(ref, got).all()
    got, _ = _run(trav, O, hf, r, beam=True)      # random origins: no bundles form, same path through the other entry
    assert _same(ref, got).all()


AFFINE = np.array([[1.5 * np.cos(0.4), -np.sin(0.4), 0.1, 0.3], [1.5 * np.sin(0.4), np.cos(0.4), 0.0, -0.2],
                   [0.0, 0.2, 2.0, 0.5]], np.float32)


@pytest.mark.parametrize("H,W,kind", [(17, 17, "smooth"), (64, 65, "smooth"), (100, 517, "noise")])
def test_walk_under_an_affine_transform(oracle, trav, H, W, kind):
    """Axis-parallel adversarial rays pushed through to_world come back to object space almost axis-parallel, a hair off the
    cell boundaries, travelling in either direction: the mirrored walk must still see the canonical cells."""
    O = oracle
    rng = np.random.default_rng(H * 131 + W)
    hf = O.OracleHeightfield(random_heights(rng, H, W, kind), 0.7, to_world=AFFINE)
    o, d = random_rays(rng, 20000)
    ao, ad = adversarial_rays(W, H, 0.7)
    o, d = np.concatenate([o, ao, ao[:, [1, 0, 2]] * np.float32([1, -1, 1])]), np.concatenate([d, ad, ad[:, [1, 0, 2]] * np.float32([1, -1, 1])])
    o = (o @ AFFINE[:, :3].T + AFFINE[:, 3]).astype(np.float32); d = (d @ AFFINE[:, :3].T).astype(np.float32)
    r = O.Rays(o, d, np.full(o.shape[0], np.inf, np.float32))
    ref = hf.ray_intersect_preliminary(r)
    got, _ = _run(trav, O, hf, r, beam=False)
    assert _same(ref, got).all()


@pytest.mark.parametrize("H,W", [(17, 17), (33, 20), (130, 257)])
def test_walk_next_to_cell_boundaries(oracle, trav, H, W):
    """Rays that run a few ulp beside a row / column boundary with a tiny transverse drift of either sign, in all four
    travel directions, starting outside and inside the grid."""
    O = oracle
    rng = np.random.default_rng(H * 17 + W)
    hf = O.OracleHeightfield(random_heights(rng, H, W, "smooth"), 0.7)
    n = 40000
    xs = (2 * rng.integers(0, W, n) - (W - 1)) / (W - 1)
    ys = (2 * rng.integers(0, H, n) - (H - 1)) / (H - 1)
    off = rng.integers(-8, 9, n) * np.float32(2.0 ** -24) * rng.choice([1.0, 0.25, 4.0], n)
    drift = rng.choice([0.0, 1e-9, -1e-9, 1e-7, -1e-7, 1e-5, -1e-5], n)
    along = rng.choice([1.0, -1.0], n) * rng.uniform(0.5, 2.0, n)
    start = np.where(rng.uniform(size=n) < 0.5, -3.0 * np.sign(along), rng.uniform(-1, 1, n))
    z0 = rng.uniform(0.0, 0.8, n); dz = rng.choice([0.0, -0.05, 0.05, -0.3], n)
    horiz = rng.uniform(size=n) < 0.5                       # run along x beside a row boundary, or along y beside a column boundary
    o = np.where(horiz[:, None], np.stack([start, ys + off, z0], 1), np.stack([xs + off, start, z0], 1)).astype(np.float32)
    d = np.where(horiz[:, None], np.stack([along, drift, dz], 1), np.stack([drift, along, dz], 1)).astype(np.float32)
    r = O.Rays(o, d, np.full(n, np.inf, np.float32))
    ref = hf.ray_intersect_preliminary(r)
    got, _ = _run(trav, O, hf, r, beam=False)
    bad = ~_same(ref, got)
    # a ray that runs exactly along a shared cell edge hits (or just misses) the two cells at u, v = 0 +- rounding: which of the
    # two the closest hit lands in is a cell-edge tie (excluded by the parity definition); everything else is exact
    def on_edge(p):
        return p.valid() & ((np.abs(p.u) < 2e-3) | (np.abs(p.v) < 2e-3) | (np.abs(1 - p.u - p.v) < 2e-3))    # the rays run <= 2e-6 beside an edge of a >= 8e-3 wide cell
    tie = bad & (on_edge(ref) | on_edge(got))
    both = tie & ref.valid() & got.valid()
    close = np.isclose(ref.t, got.t, rtol=1e-4)
    # where the two differ by more, the closer one is itself an edge hit the other traversal's column did not register
    assert (close | ((got.t < ref.t) & on_edge(got)) | ((ref.t < got.t) & on_edge(ref)))[both].all()
    bad &= ~tie
    assert not bad.any(), f"{int(bad.sum())} mismatches, first at ray {int(np.nonzero(bad)[0][0])}: o={o[bad][0]} d={d[bad][0]}"
    assert tie.sum() < 0.002 * n


def _bundle_rays(rng, n_bundles, origin_fn, spread):
    """n_bundles x 32 rays; every bundle has one origin and directions within `spread` of a random aim point."""
    o = np.repeat(origin_fn(n_bundles), 32, axis=0)
    aim = np.stack([rng.uniform(-1.2, 1.2, n_bundles), rng.uniform(-1.2, 1.2, n_bundles), rng.uniform(0.0, 0.6, n_bundles)], 1)
    tgt = np.repeat(aim, 32, axis=0) + rng.uniform(-1, 1, (n_bundles * 32, 3)) * np.repeat(spread(n_bundles), 32, axis=0)[:, None]
    d = tgt - o
    d *= np.repeat(rng.uniform(0.2, 3.0, (n_bundles, 1)), 32, axis=0)
    return o.astype(np.float32), d.astype(np.float32)


@pytest.mark.parametrize("H,W,kind,mh", [(64, 65, "smooth", 0.7), (33, 20, "noise", 0.7), (257, 300, "smooth", 0.3), (12, 12, "flat", 1.0),
                                          (2, 2, "noise", 1.0), (513, 513, "noise", 0.05)])
def test_bundle_march_is_conservative(oracle, trav, H, W, kind, mh):
    """Bundles from above, from below, from inside the box, grazing and wide: the bundled walk must give the canonical answer."""
    O = oracle
    rng = np.random.default_rng(H * 7 + W)
    hf = O.OracleHeightfield(random_heights(rng, H, W, kind), mh)
    nb = 1500 if H * W > 50000 else 4000
    sets = [
        (lambda k: np.stack([rng.uniform(-2, 2, k), rng.uniform(-2, 2, k), rng.uniform(0.8, 3, k)], 1), lambda k: rng.uniform(1e-4, 0.05, k)),   # above
        (lambda k: np.stack([rng.uniform(-2, 2, k), rng.uniform(-2, 2, k), rng.uniform(-2, -0.1, k)], 1), lambda k: rng.uniform(1e-4, 0.05, k)),  # below
        (lambda k: np.stack([rng.uniform(-1, 1, k), rng.uniform(-1, 1, k), rng.uniform(0, mh, k)], 1), lambda k: rng.uniform(1e-3, 0.3, k)),      # inside
        (lambda k: np.stack([rng.uniform(-3, -1.2, k), rng.uniform(-1, 1, k), rng.uniform(0.3 * mh, 1.1 * mh, k)], 1), lambda k: rng.uniform(1e-4, 0.02, k)),  # grazing
        (lambda k: np.stack([rng.uniform(-2, 2, k), rng.uniform(-2, 2, k), rng.uniform(0.5, 3, k)], 1), lambda k: rng.uniform(0.3, 1.5, k)),      # wide
    ]
    skipped = 0
    for origin_fn, spread in sets:
        o, d = _bundle_rays(rng, nb, origin_fn, spread)
        maxt = np.where(rng.uniform(size=o.shape[0]) < 0.1, rng.uniform(0.5, 4, o.shape[0]), np.inf).astype(np.float32)
        active = (rng.uniform(size=o.shape[0]) < 0.97).astype(np.uint8)
        r = O.Rays(o, d, maxt, active)
        ref = hf.ray_intersect_preliminary(r)
        got, te = _run(trav, O, hf, r, beam=True)
        bad = ~_same(ref, got)
        assert not bad.any(), f"{int(bad.sum())} mismatches, first at ray {int(np.nonzero(bad)[0][0])}"
        skipped += int((te > 0).sum())
    assert skipped > 0 or H * W <= 4          # the march actually advanced some bundles


def test_bundle_march_on_camera_rays(oracle, trav):
    import torch
    import hfb200
    from hfb200 import workloads as wl
    O = oracle
    h = wl.fbm_heights(1024, octaves=8, seed=0xC3).numpy()
    hf = O.OracleHeightfield(h, 0.3)
    cam = wl.CONFIGS["C3"]["camera"]
    pix = torch.arange(0, 1024 * 1024, 523, dtype=torch.int64)
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 1024, 1024, 64, seed=0, pixels=pix)
    r = O.Rays(o.numpy().T, d.numpy().T)
    ref = hf.ray_intersect_preliminary(r)
    got, te = _run(trav, O, hf, r, beam=True)
    assert _same(ref, got).all()
    hitb = ref.valid().reshape(-1, 32).any(1)
    assert (te[hitb] > 0).mean() > 0.9     # nearly every hit bundle skipped part of its way


@pytest.mark.parametrize("case", ["offset", "flat0", "flat_high", "negative", "tiny", "far_origin", "steep"])
def test_walk_code_space_extremes(oracle, trav, case):
    """The walk compares z in CODE units (hf_traverse.cuh, HF_CODE_SPACE): fields whose z range is tiny next to their magnitude, flat
    fields (degenerate code step), negative max_height, denormal-small heights, ray origins far outside the code range. The closest
    hit must still be the canonical one, bit for bit."""
    O = oracle
    rng = np.random.default_rng(hash(case) % 1000 + 5)
    H, W, mh = 40, 33, 0.7
    h = random_heights(rng, H, W, "smooth")
    if case == "offset":
        h, mh = (0.9999 + 1e-4 * h).astype(np.float32), 100.0          # z in [99.99, 100]: range / magnitude = 1e-4
    elif case == "flat0":
        h = np.zeros((H, W), np.float32)
    elif case == "flat_high":
        h, mh = np.full((H, W), 1.0, np.float32), 1000.0
    elif case == "negative":
        mh = -0.7
    elif case == "tiny":
        h = (h * 1e-30).astype(np.float32)
    elif case == "steep":
        h = rng.uniform(0, 1, (H, W)).astype(np.float32); mh = 50.0
    hf = O.OracleHeightfield(h, mh)
    zc = float(np.mean(h) * mh)
    n = 60000
    o, d = random_rays(rng, n)
    o[:, 2] = o[:, 2] * (abs(mh) if case in ("steep",) else 1.0) + (zc if case in ("offset", "flat_high") else 0.0)
    tgt = np.stack([rng.uniform(-1.1, 1.1, n), rng.uniform(-1.1, 1.1, n), zc + rng.uniform(-0.3, 0.3, n) * (abs(mh) if case == "steep" else 1.0)], 1)
    if case == "far_origin":
        o = (o * np.float32([1, 1, 3e4])).astype(np.float32)              # |z| up to 1e5: 10^10 codes away
    d = (tgt - o).astype(np.float32)
    ao, ad = adversarial_rays(W, H, mh)
    ao[:, 2] += np.float32(zc if case in ("offset", "flat_high") else 0.0)
    o, d = np.concatenate([o.astype(np.float32), ao]), np.concatenate([d, ad])
    r = O.Rays(o, d, np.full(o.shape[0], np.inf, np.float32))
    ref = hf.ray_intersect_preliminary(r)
    got, _ = _run(trav, O, hf, r, beam=False)
    assert ref.valid().sum() > 1000
    assert _same(ref, got).all()


def test_prim_index_division_by_multiplication(tmp_path):
    """csrc/hf_divmagic.h: floor(n / d) == (n * m) >> (31 + l) for every 31-bit numerator; checked on the divisors heightfields
    produce (W - 1) and on the numerators around every multiple of d."""
    src = tmp_path / "dm.cpp"
    src.write_text('#include "%s"\n#include <stdio.h>\nint main(){ unsigned ds[] = {1,2,3,4,5,7,15,16,17,63,255,256,1023,4095,4096,8191,16383,16384,65535,'
                   '1000003,8388606,8388607}; unsigned long long bad = 0, tot = 0; for (unsigned d : ds) { HfDivMagic m = hf_divmagic(d);'
                   ' for (unsigned long long q = 0; q * d < 0x7fffffffull; q += (0x7fffffffull / d) / 4099 + 1) for (long long e = -2; e <= 2; ++e) {'
                   ' long long n = (long long) (q * d) + e; if (n < 0 || n > 0x7fffffffll) continue; ++tot; if (hf_divide((unsigned) n, m.m, m.s) != (unsigned) n / d) ++bad; }'
                   ' for (unsigned long long n = 0x7fffffffull - 70000; n <= 0x7fffffffull; ++n) { ++tot; if (hf_divide((unsigned) n, m.m, m.s) != (unsigned) n / d) ++bad; } }'
                   ' printf("%%llu %%llu\\n", bad, tot); return bad != 0; }\n' % os.path.join(ROOT, "mitsuba3-differentiable-heightfield-rendering_b200", "csrc", "hf_divmagic.h"))
    exe = str(tmp_path / "dm")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, str(src)])
    out = subprocess.check_output([exe]).decode().split()
    assert out[0] == "0" and int(out[1]) > 1000000
