"""GPU parity tests: the sm_100a kernels, called through the C ABI (include/hf_b200.h via the ctypes host mirror),
against the CPU oracle on identical seeded inputs.

Tolerances (BASELINE.json north star): hit/miss mask and prim_index bit-exact away from cell-edge ties; t and
normals within 1e-5 relative; height gradients within 1e-4 relative. Because the kernels use the oracle's op order
(explicit fmaf, IEEE div/sqrt, no contraction) the forward quantities are in fact expected to be BIT-IDENTICAL and
the tests assert that; only the gradient (fp32 atomics in arbitrary order vs fp64 accumulation) needs a tolerance.

The heightfield level is "parity unpinned" (the plugin is absent from the reference snapshot); the
triangle level is pinned by the reference's known answers (test_mesh.py test10/15/16), repeated here on the GPU.
"""
import numpy as np
import pytest
import torch

from util import adversarial_rays, random_heights, random_rays

pytestmark = pytest.mark.gpu
f32 = np.float32


@pytest.fixture(scope="module")
def hf():
    import hfb200
    return hfb200


def _mk(hf, oracle, h, mh=1.0, to_world=None, flip=False):
    props = {"type": "heightfield", "heights": torch.from_numpy(np.ascontiguousarray(h)), "max_height": mh,
             "flip_normals": flip}
    if to_world is not None:
        props["to_world"] = to_world
    g = hf.load_dict(props)
    c = oracle.OracleHeightfield(h, mh, to_world=to_world, flip_normals=flip)
    return g, c


def _rays(hf, oracle, o, d, maxt=None, active=None):
    ray = hf.Ray3f(torch.from_numpy(o.T.copy()), torch.from_numpy(d.T.copy()),
                   None if maxt is None else torch.from_numpy(maxt))
    return ray, oracle.Rays(o, d, maxt, active)


def _np(t):
    return t.detach().cpu().numpy()


def _same_bits(a, b):
    a = np.asarray(a, f32); b = np.asarray(b, f32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


SI_MAP = {"p": ("px", "py", "pz"), "n": ("nx", "ny", "nz"), "uv": ("uvx", "uvy"), "dp_du": ("dpdu_x", "dpdu_y", "dpdu_z"),
          "dp_dv": ("dpdv_x", "dpdv_y", "dpdv_z"), "sh_n": ("shn_x", "shn_y", "shn_z"), "sh_s": ("shs_x", "shs_y", "shs_z"),
          "sh_t": ("sht_x", "sht_y", "sht_z"), "wi": ("wi_x", "wi_y", "wi_z")}


def _si_dict(si):
    out = {"t": _np(si.t)}
    if getattr(si, "boundary_test", None) is not None:
        out["boundary_test"] = _np(si.boundary_test)
    for k, names in SI_MAP.items():
        a = _np(getattr(si, k))
        for i, nm in enumerate(names):
            out[nm] = a[i]
    return out


def _adj_np(adj):
    out = {}
    for k, v in adj.items():
        v = np.asarray(v, f32)
        if k == "t":
            out["t"] = v
        else:
            for i, nm in enumerate(SI_MAP[k]):
                out[nm] = v[i]
    return out


# ---- acceleration data -------------------------------------------------------------------------------
@pytest.mark.parametrize("H,W", [(2, 2), (3, 2), (5, 7), (37, 50), (64, 65), (129, 200), (257, 1025)])
def test_pyramid_bit_exact(hf, oracle, H, W):
    rng = np.random.default_rng(H * 7 + W)
    h = rng.uniform(-1, 2, (H, W)).astype(f32)
    g, c = _mk(hf, oracle, h, -0.8)
    lv_g, lv_c = g.pyramid_levels(), c.pyramid_levels()
    assert len(lv_g) == len(lv_c)
    for a, b in zip(lv_g, lv_c):
        assert tuple(a.shape) == b.shape
        assert np.array_equal(_np(a), b)
    assert np.array_equal(g.bbox(), c.bbox())
    assert g.primitive_count() == c.primitive_count() == 2 * (H - 1) * (W - 1)


@pytest.mark.parametrize("H,W", [(2, 2), (3, 2), (5, 7), (17, 33), (64, 65), (100, 517), (700, 300)])
def test_quantised_pyramid_is_conservative(hf, oracle, H, W):
    """csrc/hf_qpyr.h: every node of the padded, quantised pyramid the walk reads decodes to bounds that contain the exact
    (min, max) of its cells; padding nodes hold the empty code; the layout is the documented power-of-two one."""
    rng = np.random.default_rng(H * 11 + W)
    h = (rng.uniform(-1, 2, (H, W)) * rng.choice([1.0, 1e-3, 50.0])).astype(f32)
    mh = -0.8
    g, _ = _mk(hf, oracle, h, mh)
    levels, qs, zb = g.qpyramid_levels()
    NX, NY = W - 1, H - 1
    Mx, My = 1 << int(np.ceil(np.log2(max(NX, 1)))), 1 << int(np.ceil(np.log2(max(NY, 1))))
    assert len(levels) - 1 == max(int(np.log2(Mx)), int(np.log2(My)), 1)
    z = (f32(mh) * h).astype(f32)
    cmin = np.minimum(np.minimum(z[:-1, :-1], z[:-1, 1:]), np.minimum(z[1:, :-1], z[1:, 1:]))
    cmax = np.maximum(np.maximum(z[:-1, :-1], z[:-1, 1:]), np.maximum(z[1:, :-1], z[1:, 1:]))
    pmin = np.full((My, Mx), np.inf, f32); pmax = np.full((My, Mx), -np.inf, f32)
    pmin[:NY, :NX] = cmin; pmax[:NY, :NX] = cmax
    qs32, zb32 = f32(qs), f32(zb)
    for l, t in enumerate(levels):
        q = _np(t).view(np.uint32)
        rows, pitch = max(My >> l, 1), max(Mx >> l, 1)
        assert q.shape == (rows, pitch)
        fy, fx = My // rows, Mx // pitch
        emin = pmin.reshape(rows, fy, pitch, fx).min(axis=(1, 3)); emax = pmax.reshape(rows, fy, pitch, fx).max(axis=(1, 3))
        lo = (q & 0xFFFF).astype(np.float64); hi = (q >> 16).astype(np.float64)
        # fma(code, qs, zb) in fp32: the product code * qs is exact in fp64, one rounding to fp32 after the add
        dlo = (lo * np.float64(qs32) + np.float64(zb32)).astype(f32); dhi = (hi * np.float64(qs32) + np.float64(zb32)).astype(f32)
        real = np.isfinite(emin)
        assert (q[~real] == 0x0000FFFF).all()
        assert (dlo[real] <= emin[real]).all() and (dhi[real] >= emax[real]).all()
        if l == 0:            # quantisation costs at most a few codes per bound
            assert (emin[real] - dlo[real]).max() <= 4.5 * qs32 + 1e-30 and (dhi[real] - emax[real]).max() <= 4.5 * qs32 + 1e-30


def test_capture_replays_the_step(hf, oracle):
    """hfb200.capture: zero_grad + ray_intersect + backward recorded into one CUDA graph give the eager result on every replay."""
    rng = np.random.default_rng(5)
    h = random_heights(rng, 65, 65, "smooth")
    g, _ = _mk(hf, oracle, h, 0.7)
    o, d = random_rays(rng, 50000)
    ray = hf.Ray3f(torch.from_numpy(o.T.copy()), torch.from_numpy(d.T.copy()))
    n = o.shape[0]
    adj = g.prepare_adjoints({"t": torch.ones(n), "n": torch.ones(3, n)}, n)
    si = hf.SurfaceInteraction3f(n, g.device, g, ("t", "p", "n", "uv")); pi = hf.PreliminaryIntersection3f(n, g.device, g)

    def step():
        g.zero_grad()
        g.ray_intersect(ray, out=(si, pi))
        g.backward(ray, pi, adj)

    step()
    torch.cuda.synchronize()
    ref_t, ref_g = _np(pi.t).copy(), _np(g.grad).astype(np.float64).copy()
    replay = hf.capture(step)
    for _ in range(3):
        pi.t.zero_()
        replay()
        torch.cuda.synchronize()
        assert np.array_equal(_np(pi.t).view(np.uint32), ref_t.view(np.uint32))
        assert np.linalg.norm(_np(g.grad).astype(np.float64) - ref_g) <= 1e-5 * np.linalg.norm(ref_g)


# ---- the reference's known answers, on the GPU ---------------------------------------------------------
def test_gpu_matches_drjit_heightfield_golden(hf, oracle):
    """DIRECT link GPU <-> reference-compiled arithmetic: tests/golden/drjit_heightfield.npz holds whole-heightfield closest hits
    evaluated by the reference's own Dr.Jit headers over every triangle (oracle/make_golden_heightfield.py). The CUDA walk must give
    the same mask / prim_index / t / u / v bit for bit away from cell-edge ties. (The per-triangle vectors of drjit_triangle.npz are
    arbitrary triangles, which no heightfield cell can express: that link is transitive, through the oracle.)"""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "drjit_heightfield.npz"))
    for name in ("a", "b", "c"):
        c = {k[len(name) + 1:]: g[k] for k in g.files if k.startswith(name + "_")}
        shape = hf.load_dict({"type": "heightfield", "heights": torch.from_numpy(c["h"]), "max_height": float(c["mh"]),
                              "to_world": c["to_world"], "to_object": c["to_object"]})
        ray = hf.Ray3f(torch.from_numpy(c["o"].T.copy()), torch.from_numpy(c["d"].T.copy()), torch.from_numpy(c["maxt"]))
        pi = shape.ray_intersect_preliminary(ray)
        t, u, v = _np(pi.t), _np(pi.prim_uv[0]), _np(pi.prim_uv[1])
        prim = _np(pi.prim_index).view(np.uint32)
        rv, gv = np.isfinite(c["t"]), np.isfinite(t)
        diff = (rv != gv) | (gv & ((prim != c["prim"]) | ~_same_bits(t, c["t"]) | ~_same_bits(u, c["u"]) | ~_same_bits(v, c["v"])))
        edge = lambda uu, vv: np.minimum.reduce([np.abs(uu), np.abs(vv), np.abs(1 - uu - vv)])
        near = (rv & (edge(c["u"], c["v"]) < 1e-5)) | (gv & (edge(u, v) < 1e-5))
        with np.errstate(invalid="ignore"):
            t_tie = rv & gv & (np.abs(c["t"] - t) <= 4 * np.spacing(np.abs(c["t"])))
        bad = diff & ~near & ~t_tie
        assert bad.sum() == 0, f"case {name}: {int(bad.sum())} rays differ from the reference arithmetic away from ties"
        assert diff[:2500].mean() < 2e-3 and gv.sum() > 1000
        assert np.array_equal(_np(shape.ray_test(ray)), gv)


def test_test10_known_answers(hf, oracle):
    """src/render/tests/test_mesh.py:257-292 on the flat 2x2 field (== rectangle.obj)."""
    g, _ = _mk(hf, oracle, np.zeros((2, 2), f32))
    ray = hf.Ray3f(torch.tensor([[-0.3, 0.3], [-0.3, 0.3], [-10.0, -10.0]]), torch.tensor([[0.0, 0.0], [0.0, 0.0], [1.0, 1.0]]))
    pi = g.ray_intersect_preliminary(ray)
    assert np.allclose(_np(pi.t), 10)
    assert _np(pi.prim_index).tolist() == [0, 1]
    assert np.allclose(_np(pi.prim_uv).T, [[0.35, 0.3], [0.3, 0.35]])
    si = pi.compute_surface_interaction(ray)
    assert np.allclose(_np(si.p).T, [[-0.3, -0.3, 0], [0.3, 0.3, 0]], atol=1e-6)
    assert np.allclose(_np(si.uv).T, [[0.35, 0.35], [0.65, 0.65]])
    assert np.allclose(_np(si.dp_du).T, [[2, 0, 0]] * 2) and np.allclose(_np(si.dp_dv).T, [[0, 2, 0]] * 2)
    assert np.allclose(_np(si.n).T, [[0, 0, 1]] * 2)
    assert g.ray_test(ray).tolist() == [True, True]


def test_test16_known_answers(hf, oracle):
    """test_mesh.py:536-671 (see tests/test_oracle.py::test_test16_params_backward for the vertex mapping)."""
    g, _ = _mk(hf, oracle, np.zeros((2, 2), f32))
    ray = hf.Ray3f(torch.tensor([[0.99999], [0.99999], [-10.0]]), torch.tensor([[0.0], [0.0], [1.0]]))
    pi = g.ray_intersect_preliminary(ray)

    def grad(adj):
        g.zero_grad()
        g.backward(ray, pi, adj)
        return _np(g.grad).ravel().copy()

    z1, o1 = torch.zeros(3, 1), torch.ones(1)
    assert np.allclose(grad({"t": o1}), [0, 0, 0, 1], atol=1e-5)
    assert np.allclose(grad({"p": torch.tensor([[0.0], [0.0], [1.0]])}), [0, 0, 0, 1], atol=1e-5)
    assert np.allclose(grad({"n": torch.tensor([[1.0], [0.0], [0.0]])}), [0, 0, 0.5, -0.5], atol=1e-5)
    assert np.allclose(grad({"n": torch.tensor([[0.0], [1.0], [0.0]])}), [0, 0.5, 0, -0.5], atol=1e-5)
    assert np.allclose(grad({"dp_du": torch.tensor([[0.0], [0.0], [1.0]])}), [0, 0, -1, 1], atol=1e-5)
    assert np.allclose(grad({"dp_dv": torch.tensor([[0.0], [0.0], [1.0]])}), [0, -1, 0, 1], atol=1e-5)
    assert np.all(grad({"p": z1}) == 0)


# ---- forward parity -----------------------------------------------------------------------------------
CASES = [(2, 2, "noise", None), (3, 2, "noise", None), (5, 7, "smooth", None), (17, 17, "smooth", "affine"),
         (33, 20, "noise", None), (64, 65, "smooth", "affine"), (12, 12, "flat", None), (256, 256, "smooth", None),
         (100, 517, "noise", "affine")]
AFFINE = np.array([[1.5 * np.cos(0.4), -np.sin(0.4), 0.1, 0.3], [1.5 * np.sin(0.4), np.cos(0.4), 0.0, -0.2],
                   [0.0, 0.2, 2.0, 0.5]], f32)


@pytest.mark.parametrize("H,W,kind,xf", CASES)
def test_forward_bit_exact_vs_oracle(hf, oracle, H, W, kind, xf):
    rng = np.random.default_rng(H * 131 + W)
    h = random_heights(rng, H, W, kind)
    tw = AFFINE if xf else None
    g, c = _mk(hf, oracle, h, 0.7, to_world=tw, flip=(H % 2 == 1))
    o, d = random_rays(rng, 20000)
    ao, ad = adversarial_rays(W, H, 0.7)
    o, d = np.concatenate([o, ao]), np.concatenate([d, ad])
    if tw is not None:
        o = (o @ tw[:, :3].T + tw[:, 3]).astype(f32); d = (d @ tw[:, :3].T).astype(f32)
    n = o.shape[0]
    maxt = np.where(rng.uniform(size=n) < 0.15, rng.uniform(0.5, 4, n), np.inf).astype(f32)
    active = rng.uniform(size=n) < 0.9
    ray, cr = _rays(hf, oracle, o, d, maxt, active)
    act_t = torch.from_numpy(active)

    pc = c.ray_intersect_preliminary(cr)
    pg = g.ray_intersect_preliminary(ray, active=act_t)
    t, u, v, prim = _np(pg.t), _np(pg.prim_uv[0]), _np(pg.prim_uv[1]), _np(pg.prim_index).view(np.uint32)
    same = _same_bits(t, pc.t) & _same_bits(u, pc.u) & _same_bits(v, pc.v) & (prim == pc.prim_index)
    assert same.all(), f"{(~same).sum()} of {n} preliminary intersections differ from the oracle DDA: {np.nonzero(~same)[0][:8]}"
    assert not np.isfinite(t[~active]).any()
    assert np.array_equal(_np(g.ray_test(ray, active=act_t)), pc.valid())

    # against brute force: identical away from cell-edge ties (the north-star contract)
    if H * W <= 70 * 70:
        pb = c.ray_intersect_preliminary(cr, brute=True)
        both = np.isfinite(t) & pb.valid()
        diff = (np.isfinite(t) != pb.valid()) | (both & ((prim != pb.prim_index) | (t != pb.t)))
        edge = lambda uu, vv: np.minimum.reduce([np.abs(uu), np.abs(vv), np.abs(1 - uu - vv)])
        near = (np.isfinite(t) & (edge(u, v) < 1e-5)) | (pb.valid() & (edge(pb.u, pb.v) < 1e-5))
        with np.errstate(invalid="ignore"):
            tie = both & (np.abs(t - pb.t) <= 4 * np.spacing(np.abs(pb.t)))
        assert (diff & ~near & ~tie).sum() == 0

    every = hf.SurfaceInteraction3f.ALL_FIELDS          # all 28 members, boundary_test included (zero unless its flag is set)
    for flags in (oracle.RF_ALL, oracle.RF_UV, oracle.RF_ALL | oracle.RF_FOLLOWSHAPE, oracle.RF_ALL | oracle.RF_BOUNDARYTEST):
        sc = c.compute_surface_interaction(cr, pc, ray_flags=flags)
        sg = _si_dict(g.compute_surface_interaction(ray, pg, ray_flags=flags, active=act_t, fields=every))
        for k in sc:
            ok = _same_bits(sg[k], sc[k])
            assert ok.all(), f"si.{k} differs in {(~ok).sum()} lanes (flags={flags:#x}); first {np.nonzero(~ok)[0][:5]}"
    # fused ray_intersect == preliminary + compute_surface_interaction
    bt = hf.RayFlags.All | hf.RayFlags.BoundaryTest
    sf, pf = g.ray_intersect(ray, ray_flags=bt, active=act_t, return_prelim=True, fields=every)
    sfd, s2 = _si_dict(sf), _si_dict(g.compute_surface_interaction(ray, pg, ray_flags=bt, active=act_t, fields=every))
    for k in s2:
        assert _same_bits(sfd[k], s2[k]).all(), k
    assert torch.equal(pf.t, pg.t) and torch.equal(pf.prim_index, pg.prim_index)


def test_c1_config_parity(hf, oracle):
    """BASELINE.json configs[0]: 256^2 sine heightfield, 256^2 image, 1 spp primary rays + d/d(heights)."""
    from hfb200 import workloads as wl
    h = wl.sine_heights(256)
    cam = wl.CONFIGS["C1"]["camera"]
    o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], 256, 256, 1, jitter=False)
    n = o.shape[1]
    g, c = _mk(hf, oracle, h.numpy(), 0.25)
    ray = hf.Ray3f(o, d)
    cr = oracle.Rays(o.numpy().T, d.numpy().T)
    pc = c.ray_intersect_preliminary(cr)
    si, pg = g.ray_intersect(ray, return_prelim=True)
    assert _same_bits(_np(pg.t), pc.t).all() and np.array_equal(_np(pg.prim_index).view(np.uint32), pc.prim_index)
    assert 0.85 < pc.valid().mean() < 1.0
    adj = wl.hashed_adjoints(n, seed=1)
    g.zero_grad(); g.backward(ray, pg, adj)
    ref = c.backward(cr, pc, _adj_np({k: v.numpy() for k, v in adj.items()}), literal=True)
    got = _np(g.grad).astype(np.float64)
    assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)
    big = np.abs(ref) > 1e-3 * np.abs(ref).max()
    assert np.max(np.abs(got - ref)[big] / np.abs(ref)[big]) < 1e-4
    # d(image)/d(heights) for Lambert shading max(0, n.l): the image adjoint of a unit image gradient
    l = np.array([0.3, 0.2, 1.0]); l /= np.linalg.norm(l)
    nrm = _np(si.n)
    lit = (nrm * l[:, None].astype(f32)).sum(0) > 0
    adj_n = torch.from_numpy((l[:, None] * lit[None, :]).astype(f32))
    g.zero_grad(); g.backward(ray, pg, {"n": adj_n})
    ref = c.backward(cr, pc, _adj_np({"n": adj_n.numpy()}), literal=True)
    got = _np(g.grad).astype(np.float64)
    assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)


# ---- backward parity ----------------------------------------------------------------------------------
@pytest.mark.parametrize("H,W,flags_extra,xf", [(9, 11, 0, None), (40, 33, 0, "affine"), (40, 33, 0x80, "affine"),
                                                (128, 128, 0, None), (6, 7, 0, "affine")])
def test_backward_vs_oracle_fp64(hf, oracle, H, W, flags_extra, xf):
    rng = np.random.default_rng(H + 1000 * W + flags_extra)
    h = random_heights(rng, H, W)
    tw = AFFINE if xf else None
    g, c = _mk(hf, oracle, h, 0.6, to_world=tw, flip=(W % 2 == 1))
    # ray density of the real workloads (tens of rays per texel): the per-texel criterion below is about fp32
    # atomic accumulation, whose error grows with the number of cancelling terms per texel
    o, d = random_rays(rng, min(50000, 60 * H * W))
    if tw is not None:
        o = (o @ tw[:, :3].T + tw[:, 3]).astype(f32); d = (d @ tw[:, :3].T).astype(f32)
    n = o.shape[0]
    ray, cr = _rays(hf, oracle, o, d)
    pg = g.ray_intersect_preliminary(ray)
    pc = c.ray_intersect_preliminary(cr)
    adj = {"t": rng.normal(size=n).astype(f32), "p": rng.normal(size=(3, n)).astype(f32), "n": rng.normal(size=(3, n)).astype(f32),
           "uv": rng.normal(size=(2, n)).astype(f32), "dp_du": rng.normal(size=(3, n)).astype(f32),
           "dp_dv": rng.normal(size=(3, n)).astype(f32)}
    flags = oracle.RF_ALL | flags_extra
    g.zero_grad()
    g.backward(ray, pg, {k: torch.from_numpy(v) for k, v in adj.items()}, ray_flags=flags)
    got = _np(g.grad).astype(np.float64)
    for literal in (True, False):
        ref = c.backward(cr, pc, _adj_np(adj), ray_flags=flags, literal=literal)
        assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref), (literal, np.linalg.norm(got - ref) / np.linalg.norm(ref))
    big = np.abs(ref) > 1e-3 * np.abs(ref).max()
    assert np.max(np.abs(got - ref)[big] / np.abs(ref)[big]) < 1e-4
    # DetachShape: no gradient at all (mesh.cpp:713-717)
    g.zero_grad()
    g.backward(ray, pg, {"t": torch.from_numpy(adj["t"])}, ray_flags=oracle.RF_ALL | oracle.RF_DETACHSHAPE)
    assert float(g.grad.abs().max()) == 0.0
    # gradient accumulates across calls, zero_grad resets
    g.backward(ray, pg, {"t": torch.from_numpy(adj["t"])})
    g1 = g.grad.clone()
    g.backward(ray, pg, {"t": torch.from_numpy(adj["t"])})
    assert torch.allclose(g.grad, 2 * g1, rtol=1e-5, atol=1e-6 * float(g1.abs().max()))


def test_backward_heavy_contention(hf, oracle):
    """64 spp onto a coarse field: every warp hits 1-2 cells (the warp-aggregation path)."""
    from hfb200 import workloads as wl
    rng = np.random.default_rng(4)
    h = random_heights(rng, 17, 17)
    g, c = _mk(hf, oracle, h, 0.3)
    o, d = wl.camera_rays((0, -2.5, 1.8), (0, 0, 0), (0, 0, 1), 40.0, 64, 64, 64, seed=3)
    n = o.shape[1]
    ray = hf.Ray3f(o, d)
    cr = oracle.Rays(o.numpy().T, d.numpy().T)
    pg = g.ray_intersect_preliminary(ray); pc = c.ray_intersect_preliminary(cr)
    assert _same_bits(_np(pg.t), pc.t).all()
    adj = wl.hashed_adjoints(n, seed=9)
    g.zero_grad(); g.backward(ray, pg, adj)
    ref = c.backward(cr, pc, _adj_np({k: v.numpy() for k, v in adj.items()}), literal=False)
    got = _np(g.grad).astype(np.float64)
    assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)


# ---- optimiser + parameter updates ----------------------------------------------------------------------
def test_adam_bit_exact_and_traverse(hf, oracle):
    rng = np.random.default_rng(12)
    H, W = 33, 47
    h = random_heights(rng, H, W)
    g, c = _mk(hf, oracle, h, 0.5)
    params = hf.traverse(g)
    assert set(params) == {"heights", "max_height", "to_world"}
    opt = hf.Adam(lr=1e-2, params=params)
    p = h.copy(); m = np.zeros_like(p); v = np.zeros_like(p)
    for t in range(1, 5):
        gr = rng.normal(size=(H, W)).astype(f32); gr[::3, ::2] = 0
        g.grad.copy_(torch.from_numpy(gr))
        opt.mask_updates = (t % 2 == 0)
        opt.step()
        keys = params.update()
        assert keys == ["heights"]
        oracle.adam_step(p, gr, m, v, lr=1e-2, t=t, mask_updates=(t % 2 == 0))
        assert np.array_equal(_np(g.heights), p)
    # the acceleration data follows the update
    c2 = oracle.OracleHeightfield(p, 0.5)
    for a, b in zip(g.pyramid_levels(), c2.pyramid_levels()):
        assert np.array_equal(_np(a), b)
    # parameter update through SceneParameters (util.py:185-232)
    params["heights"] = torch.from_numpy(h)
    params["max_height"] = 0.25
    params.update()
    c3 = oracle.OracleHeightfield(h, 0.25)
    assert np.array_equal(g.bbox(), c3.bbox())
    with pytest.raises(hf.HeightfieldError):
        params["heights"] = torch.zeros(3, 3)


def test_adam_bias_correction_beyond_the_table(hf, oracle):
    """optimizers.py:266-269 keeps counting t for good. The library tabulates lr_scale(t) for t < 65536 and used to stay on the last
    entry afterwards, which is wrong for betas close to 1 (beta2 = 0.9999: 0.99929 at the end of the table, 0.99954 at t = 70000,
    1 in the limit). 69 999 steps with a zero gradient leave p, m, v untouched (0 / (0 + eps)) and bring the device counter to
    t = 69 999; the next step, with a real gradient, must be the oracle's step at t = 70 000. Also: a second pair of betas gets a
    table of its own and the first pair still works afterwards (tables are never rewritten under a captured graph)."""
    rng = np.random.default_rng(5)
    H, W = 8, 12
    h = random_heights(rng, H, W)
    g, _ = _mk(hf, oracle, h, 0.5)
    g.grad.zero_()
    for _ in range(69999):
        g.adam_step(1e-2, beta_2=0.9999)
    torch.cuda.synchronize()
    assert np.array_equal(_np(g.heights), h)
    gr = rng.normal(size=(H, W)).astype(f32)
    g.grad.copy_(torch.from_numpy(gr))
    g.adam_step(1e-2, beta_2=0.9999)
    p = h.copy(); m = np.zeros_like(p); v = np.zeros_like(p)
    oracle.adam_step(p, gr, m, v, lr=1e-2, beta2=0.9999, t=70000)
    got = _np(g.heights)
    step, want = got - h, p - h
    assert np.abs(step).min() > 1e-3                                            # every texel moved
    assert np.allclose(step, want, rtol=2e-6, atol=0), np.abs(step / want - 1).max()   # the saturated table entry is 2.5e-4 off
    # other betas, then the first pair again: t = 70001 and 70002
    g.adam_step(1e-2)
    oracle.adam_step(p, gr, m, v, lr=1e-2, t=70001)
    g.adam_step(1e-2, beta_2=0.9999)
    oracle.adam_step(p, gr, m, v, lr=1e-2, beta2=0.9999, t=70002)
    assert np.allclose(_np(g.heights), p, rtol=0, atol=1e-6)


def test_boundary_test_flag(hf, oracle):
    """RayFlags::BoundaryTest (mesh.cpp:845-888 for a mesh without vertex normals; interaction.h:497-498): the reference's known
    answers of src/render/tests/test_mesh.py:928-957 on the flat 2 x 2 field (the rectangle [-1, 1]^2 as two triangles), then the
    GPU against the oracle bit for bit on a rough field under an affine transform, through the fused and the two-call path."""
    BT = hf.RayFlags.BoundaryTest
    flat, _ = _mk(hf, oracle, np.zeros((2, 2), f32))
    o = np.array([[2, 0, 1], [0.9999, 0.9999, 1], [0.99999, 0.0, 1], [0.9, 0.0, 1], [-1 / 3, -1 / 3, 1]], f32)
    d = np.tile(np.array([[0, 0, -1]], f32), (5, 1))
    ray, _ = _rays(hf, oracle, o, d)
    si = flat.ray_intersect(ray, ray_flags=BT, fields=("t", "boundary_test"))
    b, t = _np(si.boundary_test), _np(si.t)
    assert np.isinf(t[0]) and b[0] > 1e6               # test22: no intersection
    assert b[1] < 1e-3 and b[2] < 1e-4 and b[3] > 1e-1 # test22: near a corner, near an edge, far from the silhouette
    assert abs(b[4] - 1.0) < 1e-5                      # barycentre of the lower-left triangle
    si0 = flat.ray_intersect(ray, fields=("t", "boundary_test"))
    assert not _np(si0.boundary_test).any()            # flag not set: the field of a zero-initialised record

    rng = np.random.default_rng(8)
    h = random_heights(rng, 41, 57, "noise")
    g, c = _mk(hf, oracle, h, 0.6, to_world=AFFINE)
    o, d = random_rays(rng, 50000)
    ray, cr = _rays(hf, oracle, o, d)
    pc = c.ray_intersect_preliminary(cr)
    want = c.compute_surface_interaction(cr, pc, ray_flags=0x2 | 0x4 | 0x8 | 0x40)["boundary_test"]
    hit = pc.valid()
    assert hit.sum() > 5000 and (want[~hit] == f32(1e8)).all() and (want[hit] >= 0).all() and (want[hit] <= 1.0001).all()
    si, pi = g.ray_intersect(ray, ray_flags=hf.RayFlags.All | BT, return_prelim=True, fields=hf.SurfaceInteraction3f.ALL_FIELDS)
    assert _same_bits(_np(si.boundary_test), want).all()
    si2 = g.compute_surface_interaction(ray, pi, ray_flags=hf.RayFlags.All | BT, fields=("boundary_test",))
    assert _same_bits(_np(si2.boundary_test), want).all()


def test_error_behaviour(hf, oracle):
    with pytest.raises(hf.HeightfieldError):
        hf.load_dict({"type": "sphere"})
    with pytest.raises(hf.HeightfieldError):
        hf.load_dict({"type": "heightfield"})
    with pytest.raises(hf.HeightfieldError):
        hf.load_dict({"type": "heightfield", "heights": torch.zeros(1, 5)})
    with pytest.raises(hf.HeightfieldError):
        hf.load_dict({"type": "heightfield", "heights": torch.zeros(4, 4), "bogus": 1})
    g, _ = _mk(hf, oracle, np.zeros((4, 4), f32))
    ray = hf.Ray3f(torch.zeros(3, 4), torch.ones(3, 4))
    pi = g.ray_intersect_preliminary(ray)
    with pytest.raises(hf.HeightfieldError, match="DetachShape"):
        g.compute_surface_interaction(ray, pi, ray_flags=0x80 | 0x100)
    n0 = hf.Ray3f(torch.zeros(3, 0), torch.zeros(3, 0))
    assert g.ray_intersect_preliminary(n0).t.numel() == 0       # empty batch


# ---- host-buffer entry points (the llvm_ad_rgb caller's arrays live in host memory) -------------------------
def test_host_api_matches_device_api(hf, oracle):
    from hfb200 import workloads as wl
    rng = np.random.default_rng(21)
    H, W = 200, 150
    h = random_heights(rng, H, W)
    g, c = _mk(hf, oracle, h, 0.4)
    n = (1 << 21) + 12345            # more than one staging chunk, ragged tail
    o, d = wl.hemisphere_rays(n, seed=5)
    adj = wl.hashed_adjoints(n, seed=2)
    ray = hf.Ray3f(o, d)
    si, pi = g.ray_intersect(ray, return_prelim=True)
    g.zero_grad(); g.backward(ray, pi, adj)
    grad_dev = g.grad.clone()

    pin = lambda x: x.contiguous().pin_memory()
    hr = {"ox": pin(o[0]), "oy": pin(o[1]), "oz": pin(o[2]), "dx": pin(d[0]), "dy": pin(d[1]), "dz": pin(d[2]),
          "maxt": pin(torch.full((n,), float("inf")))}
    ha = {"t": pin(adj["t"]), "px": pin(adj["p"][0]), "py": pin(adj["p"][1]), "pz": pin(adj["p"][2]),
          "nx": pin(adj["n"][0]), "ny": pin(adj["n"][1]), "nz": pin(adj["n"][2]), "uvx": pin(adj["uv"][0]), "uvy": pin(adj["uv"][1])}
    hs = {k: torch.empty(n).pin_memory() for k in ("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy")}
    hp = {"t": torch.empty(n).pin_memory(), "u": torch.empty(n).pin_memory(), "v": torch.empty(n).pin_memory(),
          "prim_index": torch.empty(n, dtype=torch.int32).pin_memory()}
    hg = torch.empty(H, W).pin_memory()
    g.zero_grad()
    torch.cuda.synchronize()
    g.forward_backward_host(hr, ha, hs, hg, n, h_prelim=hp)
    assert torch.equal(hp["t"], pi.t.cpu()) and torch.equal(hp["prim_index"], pi.prim_index.cpu())
    assert torch.equal(hs["px"], si.p[0].cpu()) and torch.equal(hs["nz"], si.n[2].cpu()) and torch.equal(hs["uvy"], si.uv[1].cpu())
    assert torch.allclose(hg, grad_dev.cpu(), rtol=1e-4, atol=1e-5 * float(grad_dev.abs().max()))
    # oracle spot check on a subset
    idx = rng.choice(n, 20000, replace=False)
    cr = oracle.Rays(o.numpy().T[idx], d.numpy().T[idx])
    pc = c.ray_intersect_preliminary(cr)
    assert _same_bits(hp["t"].numpy()[idx], pc.t).all()


def test_host_api_waits_for_a_rebuild_on_a_side_stream(hf, oracle):
    """include/hf_b200.h, ordering contract of the *_host entry points: new heights and the pyramid rebuild are queued on a non-blocking
    stream of the caller behind ~100 ms of other work; the host entry that follows must walk the NEW field (its internal streams
    used to wait for the legacy stream only and read a stale pyramid), and a zero_grad queued the same way must not wipe what the
    host entry scatters."""
    from hfb200 import workloads as wl
    rng = np.random.default_rng(77)
    H, W = 300, 260
    h_old, h_new = random_heights(rng, H, W), random_heights(rng, H, W, kind="noise")
    g, _ = _mk(hf, oracle, h_old, 0.5)
    ref, _ = _mk(hf, oracle, h_new, 0.5)
    n = 200000
    o, d = wl.hemisphere_rays(n, seed=9)
    adj = wl.hashed_adjoints(n, seed=4)
    ray = hf.Ray3f(o, d)
    si, pi = ref.ray_intersect(ray, return_prelim=True)
    ref.zero_grad(); ref.backward(ray, pi, adj)
    torch.cuda.synchronize()
    pin = lambda x: x.contiguous().pin_memory()
    hr = {"ox": pin(o[0]), "oy": pin(o[1]), "oz": pin(o[2]), "dx": pin(d[0]), "dy": pin(d[1]), "dz": pin(d[2]),
          "maxt": pin(torch.full((n,), float("inf")))}
    ha = {"t": pin(adj["t"]), "px": pin(adj["p"][0]), "py": pin(adj["p"][1]), "pz": pin(adj["p"][2]),
          "nx": pin(adj["n"][0]), "ny": pin(adj["n"][1]), "nz": pin(adj["n"][2]), "uvx": pin(adj["uv"][0]), "uvy": pin(adj["uv"][1])}
    hs = {k: torch.empty(n).pin_memory() for k in ("t", "px", "nz")}
    hp = {"t": torch.empty(n).pin_memory(), "u": torch.empty(n).pin_memory(), "v": torch.empty(n).pin_memory(),
          "prim_index": torch.empty(n, dtype=torch.int32).pin_memory()}
    hg = torch.empty(H, W).pin_memory()
    new_dev = torch.from_numpy(h_new).cuda()
    side = torch.cuda.Stream()                      # torch's side streams are non-blocking
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        torch.cuda._sleep(200_000_000)              # ~0.1 s at 1.9 GHz: the host entry below is called long before this ends
        g.heights.copy_(new_dev)
        g.parameters_changed()
        g.zero_grad()
    g.forward_backward_host(hr, ha, hs, hg, n, h_prelim=hp)
    assert torch.equal(hp["t"], pi.t.cpu()) and torch.equal(hp["prim_index"], pi.prim_index.cpu())
    assert torch.equal(hs["px"], si.p[0].cpu()) and torch.equal(hs["nz"], si.n[2].cpu())
    gref = ref.grad.cpu()
    assert float(gref.abs().sum()) > 0
    assert torch.allclose(hg, gref, rtol=1e-4, atol=1e-5 * float(gref.abs().max()))
    torch.cuda.synchronize()


def test_host_api_every_si_field_and_empty_batch(hf, oracle):
    """hf_ray_intersect_host / hf_forward_backward_host with ALL 28 surface-interaction members requested (the staging slab is
    derived from sizeof(hf_si): wi_z used to be left without a slot), compared bit for bit with the device API; then the n == 0 call
    that only reads the gradient grid back (null rays allowed)."""
    import hfb200
    from hfb200 import workloads as wl
    rng = np.random.default_rng(33)
    H, W = 90, 120
    g, c = _mk(hf, oracle, random_heights(rng, H, W), 0.5, to_world=AFFINE)
    n = 70001
    o, d = wl.hemisphere_rays(n, seed=9)
    ray = hf.Ray3f(o, d)
    every = hf.RayFlags.All | hf.RayFlags.BoundaryTest
    si, pi = g.ray_intersect(ray, ray_flags=every, return_prelim=True, fields=hf.SurfaceInteraction3f.ALL_FIELDS)     # every field
    want = _si_dict(si)
    pin = lambda x: x.contiguous().pin_memory()
    hr = {"ox": pin(o[0]), "oy": pin(o[1]), "oz": pin(o[2]), "dx": pin(d[0]), "dy": pin(d[1]), "dz": pin(d[2]),
          "maxt": pin(torch.full((n,), float("inf")))}
    hs = {k: torch.full((n,), float("nan")).pin_memory() for k in hfb200.SI_FIELDS}
    hp = {"t": torch.empty(n).pin_memory(), "u": torch.empty(n).pin_memory(), "v": torch.empty(n).pin_memory(),
          "prim_index": torch.empty(n, dtype=torch.int32).pin_memory()}
    torch.cuda.synchronize()
    g.ray_intersect_host(hr, hs, n, ray_flags=every, h_prelim=hp)
    assert len(hfb200.SI_FIELDS) == 28
    for k in hfb200.SI_FIELDS:
        assert not torch.isnan(hs[k]).any(), f"host API left {k} unwritten"
        assert _same_bits(hs[k].numpy(), want[k]).all(), k
    assert torch.equal(hp["prim_index"], pi.prim_index.cpu())
    # forward + backward through the host entry with every field as well
    adj = wl.hashed_adjoints(n, seed=4)
    ha = {"t": pin(adj["t"]), "px": pin(adj["p"][0]), "py": pin(adj["p"][1]), "pz": pin(adj["p"][2]),
          "nx": pin(adj["n"][0]), "ny": pin(adj["n"][1]), "nz": pin(adj["n"][2]), "uvx": pin(adj["uv"][0]), "uvy": pin(adj["uv"][1])}
    g.zero_grad(); g.backward(ray, pi, adj)
    torch.cuda.synchronize()
    grad_dev = g.grad.clone().cpu()
    for k in hs:
        hs[k].fill_(float("nan"))
    hg = torch.empty(H, W).pin_memory()
    g.zero_grad()
    torch.cuda.synchronize()
    g.forward_backward_host(hr, ha, hs, hg, n, ray_flags=every, h_prelim=hp)
    for k in hfb200.SI_FIELDS:
        assert _same_bits(hs[k].numpy(), want[k]).all(), k
    assert torch.allclose(hg, grad_dev, rtol=1e-4, atol=1e-5 * float(grad_dev.abs().max()))
    # n == 0: nothing to trace, the call only copies the gradient grid out (rays may be null)
    hg2 = torch.zeros(H, W).pin_memory()
    g.forward_backward_host({}, {}, {}, hg2, 0)
    assert torch.equal(hg2, hg)
    g.forward_backward_host({}, {}, {}, None, 0)                         # and the fully empty call is a no-op


# ---- full-size properties (sizes the oracle cannot finish: size-independent invariants) -------------------------
def test_full_size_invariants(hf, oracle):
    """4096^2 heights (BASELINE.json C3 field), 4 M primary rays: p = o + t d, hits inside the bbox, ray_test ==
    is_valid, backward linear in the adjoints, a sampled subset bit-exact against the oracle, sum rule for dt."""
    from hfb200 import workloads as wl
    dev = "cuda"
    h = wl.fbm_heights(4096, octaves=8, seed=0xC3, device=dev)
    g = hf.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3})
    o, d = wl.config_rays("C3", device=dev, image=256, spp=64, tile=32)
    n = o.shape[1]
    ray = hf.Ray3f(o, d)
    si, pi = g.ray_intersect(ray, return_prelim=True)
    ok = pi.is_valid()
    assert 0.3 < float(ok.float().mean()) <= 1.0
    assert torch.equal(g.ray_test(ray), ok)
    onray = o + pi.t * d
    err = ((si.p - onray).abs().max(0).values)[ok]
    assert float(err.max()) < 2e-5 * 4
    bb = g.bbox()
    p_ok = si.p[:, ok]
    assert bool((p_ok >= torch.tensor(bb[:3], device=dev)[:, None] - 1e-4).all()) and bool((p_ok <= torch.tensor(bb[3:], device=dev)[:, None] + 1e-4).all())
    assert float((si.n[:, ok].norm(dim=0) - 1).abs().max()) < 1e-5
    prim = pi.prim_index[ok].to(torch.int64) & 0xFFFFFFFF
    assert int(prim.max()) < g.primitive_count()
    # linearity + accumulation
    adj = wl.hashed_adjoints(n, seed=1, device=dev)
    g.zero_grad(); g.backward(ray, pi, adj); g1 = g.grad.clone()
    g.zero_grad(); g.backward(ray, pi, {k: 2 * v for k, v in adj.items()}); g2 = g.grad.clone()
    assert torch.allclose(g2, 2 * g1, rtol=1e-4, atol=1e-5 * float(g1.abs().max()))
    # sampled subset against the oracle (rays are independent units)
    idx = torch.randperm(n, device=dev)[:30000]
    hc = h.cpu().numpy()
    c = oracle.OracleHeightfield(hc, 0.3)
    cr = oracle.Rays(o[:, idx].T.cpu().numpy(), d[:, idx].T.cpu().numpy())
    pc = c.ray_intersect_preliminary(cr)
    assert _same_bits(_np(pi.t[idx]), pc.t).all()
    assert np.array_equal(_np(pi.prim_index[idx]).view(np.uint32), pc.prim_index)
    # whole-batch gradient against the oracle's fp64 accumulation
    cr_all = oracle.Rays(o.T.cpu().numpy(), d.T.cpu().numpy())
    pc_all = oracle.Prelim(t=_np(pi.t), u=_np(pi.prim_uv[0]), v=_np(pi.prim_uv[1]), prim_index=_np(pi.prim_index).view(np.uint32))
    ref = c.backward(cr_all, pc_all, _adj_np({k: v.cpu().numpy() for k, v in adj.items()}), literal=False)
    got = _np(g1).astype(np.float64)
    assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)


# ---- bundled kernel (hf_intersect_beam_kernel): identical answers to the plain kernel and the oracle -------------
def test_bundled_kernel_bit_identical(hf, oracle, monkeypatch):
    """HF_B200_BUNDLES=1 routes ray_intersect / ray_intersect_preliminary through the bundle march. Camera rays (bundles form),
    random rays (no bundles), ragged sizes, an active mask and a transformed shape: bit-identical to the plain kernel."""
    from hfb200 import workloads as wl
    dev = "cuda"
    h = wl.fbm_heights(1024, octaves=8, seed=0xC3, device=dev)
    tw = np.array([[0.9, 0.1, 0.0, 0.05], [-0.1, 0.9, 0.05, -0.02], [0.0, 0.02, 1.1, 0.01]], np.float32)
    for to_world in (None, tw):
        props = {"type": "heightfield", "heights": h, "max_height": 0.3}
        if to_world is not None:
            props["to_world"] = to_world
        monkeypatch.setenv("HF_B200_BUNDLES", "0")
        plain = hf.load_dict(dict(props))
        monkeypatch.setenv("HF_B200_BUNDLES", "1")
        bund = hf.load_dict(dict(props))
        o, d = wl.config_rays("C3", device=dev, image=96, spp=64, tile=32)
        rng = np.random.default_rng(5)
        ro, rd = random_rays(rng, 50001)
        o = torch.cat([o, torch.from_numpy(ro.T.copy()).to(dev)], 1)[:, :-7]      # ragged: not a multiple of 32
        d = torch.cat([d, torch.from_numpy(rd.T.copy()).to(dev)], 1)[:, :-7]
        n = o.shape[1]
        active = torch.from_numpy((rng.uniform(size=n) < 0.95)).to(dev)
        ray = hf.Ray3f(o, d)
        for act in (None, active):
            sa, pa = plain.ray_intersect(ray, return_prelim=True, active=act)
            sb, pb = bund.ray_intersect(ray, return_prelim=True, active=act)
            assert torch.equal(pa.t.view(torch.int32), pb.t.view(torch.int32))
            assert torch.equal(pa.prim_index, pb.prim_index)
            assert torch.equal(pa.prim_uv.view(torch.int32), pb.prim_uv.view(torch.int32))
            assert torch.equal(sa.p.view(torch.int32), sb.p.view(torch.int32))
            assert torch.equal(sa.n.view(torch.int32), sb.n.view(torch.int32))
            pc = bund.ray_intersect_preliminary(ray, active=act)
            assert torch.equal(pa.t.view(torch.int32), pc.t.view(torch.int32))
        assert 0.3 < float(pa.is_valid().float().mean()) < 1.0


# ---- lane-pool kernel for incoherent batches (hf_set_ray_coherence(0)): identical answers to the lock-step kernel and the oracle ----
@pytest.mark.parametrize("H,W,xf,n", [(64, 65, None, 100003), (200, 150, "affine", 70001), (5, 7, None, 33), (1024, 1024, None, 1 << 20)])
def test_pool_kernel_bit_identical(hf, oracle, H, W, xf, n):
    """Every lane of a warp on its own ray, refilled from a pool, results queued in shared memory and written by the whole warp: the
    same bits as the 32-rays-in-lock-step kernel for the preliminary record, every surface-interaction field and ray_test -- on random
    incoherent rays, ragged sizes (pools, queues and the tail), an active mask, finite maxt and a transformed shape."""
    from hfb200 import workloads as wl
    rng = np.random.default_rng(H * 7 + W + n)
    h = random_heights(rng, H, W, "smooth") if H < 1000 else wl.fbm_heights(H, octaves=8, seed=0xC5).numpy()
    tw = AFFINE if xf else None
    g, c = _mk(hf, oracle, h, 0.4, to_world=tw)
    if H < 1000:
        o, d = random_rays(rng, n)
        ao, ad = adversarial_rays(W, H, 0.4)
        o, d = np.concatenate([o, ao]), np.concatenate([d, ad])
    else:
        ot, dt = wl.hemisphere_rays(n, seed=5)
        o, d = ot.numpy().T.copy(), dt.numpy().T.copy()
    if tw is not None:
        o = (o @ tw[:, :3].T + tw[:, 3]).astype(f32); d = (d @ tw[:, :3].T).astype(f32)
    m = o.shape[0]
    maxt = np.where(rng.uniform(size=m) < 0.15, rng.uniform(0.5, 4, m), np.inf).astype(f32)
    active = rng.uniform(size=m) < 0.9
    ray, cr = _rays(hf, oracle, o, d, maxt, active)
    act_t = torch.from_numpy(active)
    want_si, want_pi = g.ray_intersect(ray, active=act_t, return_prelim=True)
    want_hit = g.ray_test(ray, active=act_t)
    want_lean = g.ray_intersect(ray, active=act_t, fields=("t", "p", "n", "uv"))
    g.set_ray_coherence(False)
    try:
        si, pi = g.ray_intersect(ray, active=act_t, return_prelim=True)
        pi2 = g.ray_intersect_preliminary(ray, active=act_t)
        hitp = g.ray_test(ray, active=act_t)
        lean = g.ray_intersect(ray, active=act_t, fields=("t", "p", "n", "uv"))
    finally:
        g.set_ray_coherence(True)
    for a, b in ((pi.t, want_pi.t), (pi.prim_uv, want_pi.prim_uv), (pi2.t, want_pi.t), (pi2.prim_uv, want_pi.prim_uv)):
        assert _same_bits(_np(a), _np(b)).all()
    assert torch.equal(pi.prim_index, want_pi.prim_index) and torch.equal(pi2.prim_index, want_pi.prim_index)
    assert torch.equal(hitp, want_hit)
    ws, gs = _si_dict(want_si), _si_dict(si)
    for k in ws:
        assert _same_bits(gs[k], ws[k]).all(), k
    for k in ("t", "p", "n", "uv"):
        assert _same_bits(_np(getattr(lean, k)), _np(getattr(want_lean, k))).all(), k
    assert 0.05 < float(want_pi.is_valid().float().mean()) < 1.0
    if m <= 200000:                                   # and the oracle, on the smaller cases
        pc = c.ray_intersect_preliminary(cr)
        assert _same_bits(_np(pi.t), pc.t).all()


# ---- adaptive schedule (hf_set_adaptive_schedule): the order in which 32-ray groups are handed out must not show in any result ----
def test_adaptive_schedule_bit_identical(hf):
    """Launch k + 1 starts with the groups launch k found far above the mean cost and skips them in the natural order: every ray is
    traced exactly once whatever the list holds -- plain launches, a CUDA graph replay, a smaller and a larger batch in between
    (lists made for another batch size), the preliminary entry point, and with the hint switched off again."""
    from hfb200 import workloads as wl
    dev = torch.device("cuda", 0)
    h = wl.fbm_heights(1024, octaves=8, seed=0xC2, device=dev)
    g = hf.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
    cam = wl.CONFIGS["C3"]["camera"]                      # grazing enough for slope-skimming groups

    def batch(img, spp):
        o, d = wl.camera_rays(cam["origin"], cam["target"], cam["up"], cam["fov"], img, img, spp, seed=2, device=dev)
        return hf.Ray3f(o, d, device=dev)

    ray, small, large = batch(256, 16), batch(192, 16), batch(320, 16)         # 2^20 rays (32768 groups), 18432 and 51200 groups
    want = {id(r): g.ray_intersect(r, return_prelim=True) for r in (ray, small, large)}

    def same(r, got):
        (wsi, wpi), (si, pi) = want[id(r)], got
        assert torch.equal(pi.prim_index, wpi.prim_index)
        assert torch.equal(pi.t.view(torch.int32), wpi.t.view(torch.int32)) and torch.equal(pi.prim_uv.view(torch.int32), wpi.prim_uv.view(torch.int32))
        ws, gs = _si_dict(wsi), _si_dict(si)
        for k in ws:
            assert _same_bits(gs[k], ws[k]).all(), k

    g.set_adaptive_schedule(True)
    assert g.adaptive_schedule_info("rays") == (0, 0)
    listed = []
    for _ in range(3):
        same(ray, g.ray_intersect(ray, return_prelim=True))
        listed.append(g.adaptive_schedule_info("rays"))
    assert all(grp == 32768 for _, grp in listed) and listed[-1][0] > 0, listed        # something was learnt ...
    assert listed[-1][0] < 32768 // 8, listed                                           # ... and it is a short list
    pi = g.ray_intersect_preliminary(ray)
    assert torch.equal(pi.prim_index, want[id(ray)][1].prim_index) and torch.equal(pi.t.view(torch.int32), want[id(ray)][1].t.view(torch.int32))
    # a CUDA graph of the launch (the list is read and rewritten on the device: replays keep learning)
    si_g = hf.SurfaceInteraction3f(len(ray), dev, g, ("t", "p", "n", "uv"))
    pi_g = hf.PreliminaryIntersection3f(len(ray), dev, g)
    replay = hf.capture(lambda: g.ray_intersect(ray, out=(si_g, pi_g)))
    for _ in range(3):
        si_g.t.fill_(-1); pi_g.prim_index.zero_()
        replay()
        torch.cuda.synchronize()
        assert torch.equal(pi_g.prim_index, want[id(ray)][1].prim_index)
        assert torch.equal(si_g.t.view(torch.int32), want[id(ray)][0].t.view(torch.int32))
        assert torch.equal(si_g.n.view(torch.int32), want[id(ray)][0].n.view(torch.int32))
    # other batch sizes in between: the list of the last batch does not fit them (ids beyond a smaller batch, a larger buffer needed)
    for r in (small, ray, large, ray, small, small, large, large, ray):
        same(r, g.ray_intersect(r, return_prelim=True))
    replay()                                                  # recorded against the buffers that `large` has outgrown: still exact
    torch.cuda.synchronize()
    assert torch.equal(pi_g.prim_index, want[id(ray)][1].prim_index)
    g.set_adaptive_schedule(False)
    same(ray, g.ray_intersect(ray, return_prelim=True))
    assert g.adaptive_schedule_info("rays") == (0, 0)
