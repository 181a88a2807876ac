"""GPU parity AT THE SIZES of BASELINE.json's configurations: the scenes are the configs' own (fields, cameras,
sample counts), the GPU renders / traces them in full, and the oracle checks a seeded subset of the very same pixels or rays --
pixels and rays are independent units, and the fused kernels are shown to give a pixel the same bits whether it is rendered alone or
inside the full image. Tolerances as everywhere: hit triangles and forward quantities bit-exact, pixel means 1e-6 (1e-5 for paths),
gradients 1e-4 relative L2 against the oracle's fp64 accumulation.

    C2  1024^2 fBm, 512^2 x 16 spp, direct illumination WITH shadow rays, forward
    C4  grazing camera, depth-3 paths with gradients: 2048^2 (the largest field the oracle's path tracer turns around in under a
        minute on a sample) and, for the 16384^2 field itself (1 GiB of heights), a ray subset against the oracle's DDA + invariants
    C5  incoherent hemisphere rays on a 4096^2 field, intersect + surface interaction + backward
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
f32 = np.float32


@pytest.fixture(scope="module")
def hf():
    import hfb200
    return hfb200


def _bits(a, b):
    a = np.asarray(a, f32); b = np.asarray(b, f32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def _adj_c(adj):
    return {"t": adj["t"], "px": adj["p"][0], "py": adj["p"][1], "pz": adj["p"][2], "nx": adj["n"][0], "ny": adj["n"][1],
            "nz": adj["n"][2], "uvx": adj["uv"][0], "uvy": adj["uv"][1]}


def test_c2_direct_illumination_at_size(hf, oracle):
    """configs[1]: 1024^2 procedural heightfield, 512^2 image, 16 spp, direct illumination with shadow rays, forward only."""
    from hfb200 import workloads as wl
    cfg = wl.CONFIGS["C2"]
    h = wl.make_heights(cfg["heights"], device="cuda")
    g = hf.load_dict({"type": "heightfield", "heights": h, "max_height": cfg["max_height"]})
    c = oracle.OracleHeightfield(h.cpu().numpy(), cfg["max_height"])
    cm = cfg["camera"]
    cam = hf.perspective_camera(cm["origin"], cm["target"], cm["up"], cm["fov"], cfg["image"], cfg["image"], cfg["spp"], seed=2)
    light = hf.directional_light((0.8, 0.2, 0.25), irradiance=2.0, albedo=0.8, shadow_rays=True)      # a low sun: the fBm ridges cast shadows
    npix = cfg["image"] ** 2
    full, prims_full = g.render_forward(cam, light, return_prims=True)                    # 4 194 304 samples + as many shadow rays
    rng = np.random.default_rng(0xC2)
    sub = np.sort(rng.permutation(npix)[:1500]).astype(np.uint32)                         # 24 000 samples
    img, prims = g.render_forward(cam, light, sub, return_prims=True)
    ref_img, value, rprim, _ = oracle.render_forward(c, cam, light, sub)
    assert np.array_equal(prims.cpu().numpy().view(np.uint32).reshape(rprim.shape), rprim)
    got = img.cpu().numpy().reshape(-1).astype(np.float64)
    assert np.allclose(got[sub], ref_img[sub], rtol=1e-6, atol=1e-7)
    # the full-size launch gives those pixels the same bits (and the same hit triangles) as the subset launch
    assert torch.equal(full.reshape(-1)[torch.from_numpy(sub.astype(np.int64)).cuda()].view(torch.int32),
                       img.reshape(-1)[torch.from_numpy(sub.astype(np.int64)).cuda()].view(torch.int32))
    pf = prims_full.cpu().numpy().view(np.uint32).reshape(npix, cfg["spp"])
    assert np.array_equal(pf[sub], rprim)
    lit = rprim != 0xFFFFFFFF
    assert 0.1 < lit.mean() < 0.99
    _, _, p_noshadow, _ = oracle.render_forward(c, cam, dict(light, shadow_rays=False), sub)
    assert (p_noshadow != 0xFFFFFFFF).sum() > lit.sum()                                   # shadow rays do remove light at this size


@pytest.mark.parametrize("shadow", [True, False])
def test_c4_shaped_paths_with_gradients(hf, oracle, shadow):
    """configs[3] in shape: fBm 10 octaves seed 0xC4, max_height 0.15, the grazing camera, 8 spp, depth-3 paths with gradients --
    on a 2048^2 field so that the oracle's path tracer finishes a 2 000-pixel sample in seconds."""
    from hfb200 import workloads as wl
    cfg = wl.CONFIGS["C4"]
    n_h, image = 2048, 512
    h = wl.fbm_heights(n_h, octaves=10, seed=0xC4, device="cuda")
    g = hf.load_dict({"type": "heightfield", "heights": h, "max_height": cfg["max_height"]})
    c = oracle.OracleHeightfield(h.cpu().numpy(), cfg["max_height"])
    cm = cfg["camera"]
    cam = hf.perspective_camera(cm["origin"], cm["target"], cm["up"], cm["fov"], image, image, cfg["spp"], seed=4)
    light = hf.directional_light((0.3, 0.2, 1.0), irradiance=2.0, albedo=0.8, shadow_rays=shadow, max_depth=3)
    npix = image * image
    rng = np.random.default_rng(0xC4)
    sub = np.sort(rng.permutation(npix)[:2000]).astype(np.uint32)
    adj = rng.uniform(-1, 1, npix).astype(f32)
    full = g.render_forward(cam, light)
    img = g.render_forward(cam, light, sub)
    ref_img, samples = oracle.render_path_forward(c, cam, light, sub)
    got = img.cpu().numpy().reshape(-1).astype(np.float64)
    assert np.allclose(got[sub], ref_img[sub], rtol=1e-5, atol=1e-7)
    idx = torch.from_numpy(sub.astype(np.int64)).cuda()
    assert torch.equal(full.reshape(-1)[idx].view(torch.int32), img.reshape(-1)[idx].view(torch.int32))
    value, vprim, _ = samples
    assert (vprim[:, 0] != 0xFFFFFFFF).mean() > 0.3 and (vprim[:, 2] != 0xFFFFFFFF).sum() > 20       # grazing rays hit, paths reach depth 3
    ref = oracle.render_path_backward(c, cam, light, adj, sub, samples=samples)
    g.zero_grad(); g.render_backward(cam, light, torch.from_numpy(adj), sub)
    ga = g.grad.cpu().numpy().astype(np.float64)
    assert np.linalg.norm(ref) > 0 and np.linalg.norm(ga - ref) <= 1e-4 * np.linalg.norm(ref)
    # the one-call L2 step (what bench.py --config C4 times) equals forward -> loss adjoint -> backward on the same pixels
    target = torch.zeros_like(full)
    loss = torch.zeros(1, dtype=torch.float64, device="cuda")
    g.zero_grad(); g.render_step_l2(cam, light, target, sub, loss=loss)
    a2 = np.zeros(npix); a2[sub] = 2 * ref_img[sub] / npix
    ref2 = oracle.render_path_backward(c, cam, light, a2, sub, samples=samples)
    g2 = g.grad.cpu().numpy().astype(np.float64)
    assert np.linalg.norm(g2 - ref2) <= 1e-4 * np.linalg.norm(ref2)
    assert abs(float(loss) - float((ref_img[sub] ** 2).sum() / npix)) <= 1e-5 * float(loss)


def test_c4_field_itself_16384(hf, oracle):
    """The 16384^2 field of configs[3] (1 GiB of heights; texels and pyramid nodes no longer fit L2): grazing primary rays of the C4
    camera, a 20 000-ray sample against the oracle's DDA (bit-exact t / prim_index / p / n), and size-independent invariants on the
    whole batch: p = o + t d, ray_test == is_valid, hits inside the bbox, backward linear in the adjoints."""
    from hfb200 import workloads as wl
    cfg = wl.CONFIGS["C4"]
    h = wl.make_heights(cfg["heights"], device="cuda")
    g = hf.load_dict({"type": "heightfield", "heights": h, "max_height": cfg["max_height"]})
    cm = cfg["camera"]
    cam = hf.perspective_camera(cm["origin"], cm["target"], cm["up"], cm["fov"], 512, 512, cfg["spp"], seed=4)
    ray = g.render_rays(cam)                                   # 2 097 152 grazing rays
    n = len(ray)
    si, pi = g.ray_intersect(ray, return_prelim=True, fields=("t", "p", "n", "uv"))
    ok = pi.is_valid()
    assert 0.3 < float(ok.float().mean()) <= 1.0
    assert torch.equal(g.ray_test(ray), ok)
    onray = ray.o + pi.t * ray.d
    assert float(((si.p - onray).abs().max(0).values)[ok].max()) < 1e-4
    bb = g.bbox()
    p_ok = si.p[:, ok]
    assert bool((p_ok >= torch.tensor(bb[:3], device="cuda")[:, None] - 1e-4).all()) and bool((p_ok <= torch.tensor(bb[3:], device="cuda")[:, None] + 1e-4).all())
    adj = wl.hashed_adjoints(n, seed=1, device="cuda")
    g.zero_grad(); g.backward(ray, pi, adj); g1 = g.grad.clone()
    g.zero_grad(); g.backward(ray, pi, {k: -3 * v for k, v in adj.items()})
    assert torch.allclose(g.grad, -3 * g1, rtol=1e-4, atol=1e-5 * float(g1.abs().max()))
    del g1
    # the oracle on the same 1 GiB field: a sample of the rays, bit for bit
    c = oracle.OracleHeightfield(h.cpu().numpy(), cfg["max_height"])
    idx = torch.randperm(n, device="cuda")[:20000]
    cr = oracle.Rays(ray.o[:, idx].T.cpu().numpy(), ray.d[:, idx].T.cpu().numpy())
    pc = c.ray_intersect_preliminary(cr)
    assert _bits(pi.t[idx].cpu().numpy(), pc.t).all()
    v = pc.valid()
    assert np.array_equal(pi.prim_index[idx].cpu().numpy().view(np.uint32)[v], pc.prim_index[v])
    sc = c.compute_surface_interaction(cr, pc, fields=("px", "py", "pz", "nx", "ny", "nz"))
    for k, a in (("px", si.p[0]), ("py", si.p[1]), ("pz", si.p[2]), ("nx", si.n[0]), ("ny", si.n[1]), ("nz", si.n[2])):
        assert _bits(a[idx].cpu().numpy()[v], sc[k][v]).all(), k
    # gradient of that sample against fp64
    sub_ray = hf.Ray3f(ray.o[:, idx].contiguous(), ray.d[:, idx].contiguous())
    sub_pi = g.ray_intersect_preliminary(sub_ray)
    sub_adj = {k: (t[idx] if t.dim() == 1 else t[:, idx]).contiguous() for k, t in adj.items()}
    g.zero_grad(); g.backward(sub_ray, sub_pi, sub_adj)
    ref = c.backward(cr, pc, {k: x.cpu().numpy() for k, x in _adj_c(sub_adj).items()}, literal=False)
    got = g.grad.cpu().numpy().astype(np.float64)
    assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)


def test_c5_incoherent_rays_on_4096(hf, oracle):
    """configs[4]: origins uniform on the upper hemisphere of radius 3, targets uniform on the field (seed 5), 4096^2 fBm seed 0xC5:
    2^22 rays on the GPU; a 30 000-ray subset bit-exact against the oracle (preliminary record + t, p, n, uv), the gradient of a
    2^18-ray subset against the oracle's fp64 accumulation."""
    from hfb200 import workloads as wl
    h = wl.fbm_heights(4096, octaves=8, seed=0xC5, device="cuda")
    g = hf.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3})
    n = 1 << 22
    o, d = wl.hemisphere_rays(n, seed=5, device="cuda")
    ray = hf.Ray3f(o, d)
    si, pi = g.ray_intersect(ray, return_prelim=True, fields=("t", "p", "n", "uv"))
    ok = pi.is_valid()
    assert float(ok.float().mean()) > 0.6
    assert torch.equal(g.ray_test(ray), ok)
    c = oracle.OracleHeightfield(h.cpu().numpy(), 0.3)
    idx = torch.randperm(n, device="cuda")[:30000]
    cr = oracle.Rays(o[:, idx].T.cpu().numpy(), d[:, idx].T.cpu().numpy())
    pc = c.ray_intersect_preliminary(cr)
    v = pc.valid()
    assert _bits(pi.t[idx].cpu().numpy(), pc.t).all()
    assert np.array_equal(pi.prim_index[idx].cpu().numpy().view(np.uint32)[v], pc.prim_index[v])
    assert _bits(pi.prim_uv[0][idx].cpu().numpy()[v], pc.u[v]).all() and _bits(pi.prim_uv[1][idx].cpu().numpy()[v], pc.v[v]).all()
    sc = c.compute_surface_interaction(cr, pc, fields=("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy"))
    for k, a in (("px", si.p[0]), ("py", si.p[1]), ("pz", si.p[2]), ("nx", si.n[0]), ("ny", si.n[1]), ("nz", si.n[2]),
                 ("uvx", si.uv[0]), ("uvy", si.uv[1])):
        assert _bits(a[idx].cpu().numpy()[v], sc[k][v]).all(), k
    m = 1 << 18
    adj = wl.hashed_adjoints(m, seed=3, device="cuda")
    sub_ray = hf.Ray3f(o[:, :m].contiguous(), d[:, :m].contiguous())
    sub_pi = g.ray_intersect_preliminary(sub_ray)
    g.zero_grad(); g.backward(sub_ray, sub_pi, adj)
    cr2 = oracle.Rays(o[:, :m].T.cpu().numpy(), d[:, :m].T.cpu().numpy())
    pc2 = oracle.Prelim(t=sub_pi.t.cpu().numpy(), u=sub_pi.prim_uv[0].cpu().numpy(), v=sub_pi.prim_uv[1].cpu().numpy(),
                        prim_index=sub_pi.prim_index.cpu().numpy().view(np.uint32))
    ref = c.backward(cr2, pc2, {k: x.cpu().numpy() for k, x in _adj_c(adj).items()}, literal=False)
    got = g.grad.cpu().numpy().astype(np.float64)
    assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)
