"""GPU parity of the render step on the heightfield alone (include/hf_b200.h "render step"):
in-register camera rays, direct illumination, box-filter image, PRB-style adjoint and the fused L2 step, against the
oracle's restatement (oracle/hf_oracle.c hfo_camera_rays / hfo_render_samples + hfo_backward).

Tolerances: generated rays, per-sample hit triangles and visibility bit-exact; pixel means within 1e-6 relative (fp32 sum
of <= 256 samples vs fp64); height gradients within 1e-4 relative L2 of the oracle's fp64 accumulation.
"""
import numpy as np
import pytest
import torch

from util import random_heights

pytestmark = pytest.mark.gpu
f32 = np.float32


@pytest.fixture(scope="module")
def hf():
    import hfb200
    return hfb200


TW = np.array([[0.9, 0.1, 0.0, 0.05], [-0.1, 0.9, 0.05, -0.02], [0.0, 0.02, 1.1, 0.01]], f32)


def _scene(hf, oracle, H=96, W=80, mh=0.3, to_world=None, flip=False, kind="smooth", seed=3):
    rng = np.random.default_rng(seed)
    h = random_heights(rng, H, W, kind)
    props = {"type": "heightfield", "heights": torch.from_numpy(h), "max_height": mh, "flip_normals": flip}
    if to_world is not None:
        props["to_world"] = to_world
    return hf.load_dict(props), oracle.OracleHeightfield(h, mh, to_world=to_world, flip_normals=flip)


def _cam(hf, w=48, h=40, spp=16, seed=7, origin=(0.3, -2.5, 1.8)):
    return hf.perspective_camera(origin, (0, 0, 0), (0, 0, 1), 40.0, w, h, spp, seed)


def test_generated_rays_bit_exact(hf, oracle):
    g, c = _scene(hf, oracle)
    for spp in (1, 5, 64):
        cam = _cam(hf, spp=spp)
        pix = np.array([0, 7, 48 * 40 - 1, 100, 101, 999], np.uint32)
        for px in (None, pix):
            ray = g.render_rays(cam, px)
            ref = oracle.camera_rays(cam, px)
            for a, b in ((ray.o[0], ref.ox), (ray.o[1], ref.oy), (ray.o[2], ref.oz), (ray.d[0], ref.dx), (ray.d[1], ref.dy),
                         (ray.d[2], ref.dz), (ray.maxt, ref.maxt)):
                assert np.array_equal(a.cpu().numpy().view(np.uint32), np.asarray(b, f32).view(np.uint32))


@pytest.mark.parametrize("spp,shadow,xf,flip", [(1, False, None, False), (16, False, None, False), (64, True, None, False),
                                                (100, True, "tw", False), (33, False, "tw", True), (256, False, None, False),
                                                (5, True, "tw", False), (12, False, None, False), (8, True, None, True)])
def test_forward_image_and_prims(hf, oracle, spp, shadow, xf, flip):
    g, c = _scene(hf, oracle, to_world=TW if xf else None, flip=flip, kind="noise" if shadow else "smooth", mh=0.5 if shadow else 0.3)
    cam = _cam(hf, spp=spp, origin=(0.3, -2.5, -1.8) if flip else (0.3, -2.5, 1.8))     # flipped normals face a camera below the sheet
    ldir = (0.6, 0.2, -0.5) if flip else (0.6, 0.2, 0.5)
    light = hf.directional_light(ldir, irradiance=3.0, albedo=0.8, shadow_rays=shadow)
    rng = np.random.default_rng(1)
    some = rng.permutation(48 * 40)[:777].astype(np.uint32)            # ragged: 777 pixels, not a multiple of 32
    for px in (None, some):
        img, prims = g.render_forward(cam, light, px, return_prims=True)
        ref_img, value, rprim, _ = oracle.render_forward(c, cam, light, px)
        assert np.array_equal(prims.cpu().numpy().view(np.uint32).reshape(rprim.shape), rprim)
        got = img.cpu().numpy().reshape(-1).astype(np.float64)
        assert np.allclose(got, ref_img, rtol=1e-6, atol=1e-7)
        assert (rprim != 0xFFFFFFFF).mean() > 0.03
        if shadow:     # the visibility ray does remove light somewhere
            _, v2, p2, _ = oracle.render_forward(c, cam, dict(light, shadow_rays=False), px)
            assert (p2 != 0xFFFFFFFF).sum() > (rprim != 0xFFFFFFFF).sum()


@pytest.mark.parametrize("spp,shadow,xf", [(16, False, None), (64, True, "tw"), (40, False, "tw"), (12, True, None), (2, False, "tw")])
def test_backward_matches_oracle_fp64(hf, oracle, spp, shadow, xf):
    g, c = _scene(hf, oracle, to_world=TW if xf else None)
    cam = _cam(hf, spp=spp)
    light = hf.directional_light((0.6, 0.2, 0.5), irradiance=2.0, albedo=0.7, shadow_rays=shadow)
    rng = np.random.default_rng(2)
    adj = rng.uniform(-1, 1, 48 * 40).astype(f32)
    px = rng.permutation(48 * 40)[:1500].astype(np.uint32)
    for pixels in (None, px):
        ref = oracle.render_backward(c, cam, light, adj, pixels)
        nref = np.linalg.norm(ref)
        assert nref > 0
        # (a) re-tracing backward
        g.zero_grad(); g.render_backward(cam, light, torch.from_numpy(adj), pixels)
        ga = g.grad.cpu().numpy().astype(np.float64)
        assert np.linalg.norm(ga - ref) <= 1e-4 * nref
        # (b) backward from the forward pass's hit triangles
        _, prims = g.render_forward(cam, light, pixels, return_prims=True)
        g.zero_grad(); g.render_backward(cam, light, torch.from_numpy(adj), pixels, prims=prims)
        gb = g.grad.cpu().numpy().astype(np.float64)
        assert np.linalg.norm(gb - ref) <= 1e-4 * nref
        big = np.abs(ref) > 1e-6 * np.abs(ref).max()
        assert np.allclose(ga[big], ref[big], rtol=2e-3) and np.allclose(gb[big], ref[big], rtol=2e-3)


@pytest.mark.parametrize("spp", [64, 8, 3])          # below 32 spp several pixels share a warp round
def test_fused_l2_step_equals_forward_loss_backward(hf, oracle, spp):
    g, c = _scene(hf, oracle)
    cam = _cam(hf, spp=spp)
    light = hf.directional_light((0.6, 0.2, 0.5))
    npix = 48 * 40
    g2, _ = _scene(hf, oracle, seed=11)                      # another field renders the target
    target = g2.render_forward(cam, light)
    rng = np.random.default_rng(4)
    px = rng.permutation(npix)[:1111].astype(np.uint32)
    for pixels in (None, px):
        ids = np.arange(npix) if pixels is None else pixels.astype(np.int64)
        img = g.render_forward(cam, light, pixels)
        diff = (img - target).reshape(-1)
        adj = torch.zeros(npix, device=img.device)
        adj[torch.from_numpy(ids).to(img.device)] = 2 * diff[torch.from_numpy(ids).to(img.device)] / npix
        g.zero_grad(); g.render_backward(cam, light, adj, pixels)
        want = g.grad.clone()
        want_loss = float((diff[torch.from_numpy(ids).to(img.device)].double() ** 2).sum() / npix)
        loss = torch.zeros(1, dtype=torch.float64, device=img.device)
        img2 = torch.zeros_like(img)
        g.zero_grad(); g.render_step_l2(cam, light, target, pixels, image=img2, loss=loss)
        assert torch.equal(img2.view(torch.int32), img.view(torch.int32))          # deterministic pixel means
        assert abs(float(loss) - want_loss) <= 1e-6 * max(want_loss, 1e-12)
        err = float((g.grad - want).double().norm() / want.double().norm())
        assert err < 1e-5
        # and against the oracle end to end
        ref_img, *_ = oracle.render_forward(c, cam, light, pixels)
        tgt = target.cpu().numpy().reshape(-1).astype(np.float64)
        adj_ref = np.zeros(npix); adj_ref[ids] = 2 * (ref_img[ids] - tgt[ids]) / npix
        ref = oracle.render_backward(c, cam, light, adj_ref, pixels)
        got = g.grad.cpu().numpy().astype(np.float64)
        assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)


def test_render_equals_the_shape_api_on_the_same_rays(hf, oracle):
    """The fused path is the shape API composed: render_rays -> ray_intersect -> Lambert in torch -> backward(adjoint of n)."""
    g, c = _scene(hf, oracle, to_world=TW)
    cam = _cam(hf, spp=32)
    light = hf.directional_light((0.6, 0.2, 0.5), irradiance=1.5, albedo=0.9)
    ray = g.render_rays(cam)
    si, pi = g.ray_intersect(ray, return_prelim=True)
    l = torch.tensor(light["direction"], device=si.n.device)[:, None]
    cosl = (si.n * l).sum(0)
    lit = pi.is_valid() & (cosl > 0) & ((si.n * -ray.d).sum(0) > 0)          # one-sided Lambertian (diffuse.cpp:127-139)
    scale = np.float32(0.9 / np.pi * 1.5)
    value = torch.where(lit, float(scale) * cosl, torch.zeros_like(cosl))
    img_ref = value.reshape(-1, 32).double().mean(1)
    img = g.render_forward(cam, light)
    assert torch.allclose(img.reshape(-1).double(), img_ref, rtol=1e-5, atol=1e-7)
    adj_img = torch.linspace(-1, 1, 48 * 40, device=img.device)
    a_v = (adj_img.repeat_interleave(32) / 32 * float(scale)) * lit
    g.zero_grad(); g.backward(ray, pi, {"n": a_v[None, :] * l})
    want = g.grad.clone()
    g.zero_grad(); g.render_backward(cam, light, adj_img)
    assert float((g.grad - want).norm() / want.norm()) < 1e-5


def test_render_error_behaviour(hf, oracle):
    g, _ = _scene(hf, oracle)
    light = hf.directional_light((0, 0, 1))
    cam = _cam(hf, spp=512)
    target = torch.zeros((40, 48), device="cuda")
    g.render_forward(cam, light)                                     # forward takes any spp
    with pytest.raises(hf.HeightfieldError, match="spp"):
        g.render_step_l2(cam, light, target)
    bad = dict(_cam(hf), right=[1.0, 0.2, 0.0])
    with pytest.raises(hf.HeightfieldError, match="orthonormal"):
        g.render_forward(bad, light)
    with pytest.raises(hf.HeightfieldError, match="unit"):
        g.render_forward(_cam(hf), dict(light, direction=[0.0, 0.0, 2.0]))
    with pytest.raises(hf.HeightfieldError, match="npix"):
        import ctypes as C
        c, l = hf._camera_struct(_cam(hf)), hf._light_struct(light)
        hf._check(hf._load().hf_render_forward(g._ctx, C.byref(c), C.byref(l), None, 5, C.c_void_p(target.data_ptr()), None, None))


# ---- path tracing -------------------------------------------------------------------------------
@pytest.mark.parametrize("depth,shadow,xf,spp", [(2, False, None, 16), (3, True, None, 32), (3, True, "tw", 40), (4, False, None, 8),
                                                 (2, True, None, 12), (3, False, "tw", 1)])
def test_path_forward_and_backward(hf, oracle, depth, shadow, xf, spp):
    """Radiance per pixel within 1e-5 relative of the oracle (fp32 sums; the bounce directions go through the same double-precision
    sincos on both sides), height gradient within 1e-4 relative L2 of the oracle's fp64 PRB adjoint."""
    g, c = _scene(hf, oracle, to_world=TW if xf else None, kind="noise", mh=0.5)
    cam = _cam(hf, spp=spp)
    light = hf.directional_light((0.6, 0.2, 0.5), irradiance=2.0, albedo=0.8, shadow_rays=shadow, max_depth=depth)
    rng = np.random.default_rng(9)
    adj = rng.uniform(-1, 1, 48 * 40).astype(f32)
    px = rng.permutation(48 * 40)[:901].astype(np.uint32)
    for pixels in (None, px):
        img = g.render_forward(cam, light, pixels)
        ref_img, samples = oracle.render_path_forward(c, cam, light, pixels)
        got = img.cpu().numpy().reshape(-1).astype(np.float64)
        assert np.allclose(got, ref_img, rtol=1e-5, atol=1e-7)
        value, vprim, _ = samples
        assert (vprim[:, depth - 1] != 0xFFFFFFFF).sum() > 50                      # paths do reach the last vertex
        d1, *_ = oracle.render_forward(c, cam, dict(light, max_depth=1), pixels)
        assert ref_img.sum() > 1.02 * d1.sum()                                     # and the bounces carry light
        ref = oracle.render_path_backward(c, cam, light, adj, pixels, samples=samples)
        g.zero_grad(); g.render_backward(cam, light, torch.from_numpy(adj), pixels)
        ga = g.grad.cpu().numpy().astype(np.float64)
        assert np.linalg.norm(ga - ref) <= 1e-4 * np.linalg.norm(ref)


def test_path_l2_step_and_errors(hf, oracle):
    g, c = _scene(hf, oracle, kind="noise", mh=0.5)
    cam = _cam(hf, spp=32)
    light = hf.directional_light((0.6, 0.2, 0.5), albedo=0.8, shadow_rays=True, max_depth=3)
    npix = 48 * 40
    g2, _ = _scene(hf, oracle, kind="noise", mh=0.5, seed=12)
    target = g2.render_forward(cam, light)
    img = g.render_forward(cam, light)
    adj = 2 * (img - target).reshape(-1) / npix
    g.zero_grad(); g.render_backward(cam, light, adj)
    want = g.grad.clone()
    loss = torch.zeros(1, dtype=torch.float64, device=img.device)
    img2 = torch.zeros_like(img)
    g.zero_grad(); g.render_step_l2(cam, light, target, image=img2, loss=loss)
    assert torch.equal(img2.view(torch.int32), img.view(torch.int32))
    assert abs(float(loss) - float(((img - target).double() ** 2).mean())) <= 1e-9 + 1e-6 * float(loss)
    assert float((g.grad - want).double().norm() / want.double().norm()) < 1e-5
    with pytest.raises(hf.HeightfieldError, match="max_depth"):
        g.render_forward(cam, dict(light, max_depth=7))
    with pytest.raises(hf.HeightfieldError, match="replayed"):
        g.render_forward(cam, light, return_prims=True)


def test_host_render_step_matches_device_step(hf, oracle):
    """hf_render_step_l2_host: heights + target + pixel list from HOST memory in, image / gradient grid / loss out -- the same bits
    as the device entry point on the same inputs, with and without a new set of heights."""
    g, _ = _scene(hf, oracle)
    g2, _ = _scene(hf, oracle, seed=11)
    cam = _cam(hf, spp=64)
    light = hf.directional_light((0.6, 0.2, 0.5), shadow_rays=True)
    target = g2.render_forward(cam, light)
    npix = 48 * 40
    rng = np.random.default_rng(6)
    for pixels in (None, rng.permutation(npix)[:999].astype(np.int32)):
        loss = torch.zeros(1, dtype=torch.float64, device="cuda")
        img = torch.zeros_like(target)
        g.zero_grad(); g.render_step_l2(cam, light, target, pixels, image=img, loss=loss)
        torch.cuda.synchronize()
        want_grad, want_img, want_loss = g.grad.clone().cpu(), img.cpu(), float(loss)
        h_t = target.cpu().pin_memory()
        h_img = torch.full((40, 48), -1.0).pin_memory(); h_grad = torch.empty(g.H, g.W).pin_memory()
        h_h = g.heights.cpu().pin_memory()
        h_px = None if pixels is None else torch.from_numpy(pixels).pin_memory()
        for heights_in in (None, h_h):
            got_loss = g.render_step_l2_host(cam, light, h_t, h_heights=heights_in, h_image=h_img, h_grad=h_grad, h_pixels=h_px)
            assert torch.equal(h_img.view(torch.int32), want_img.view(torch.int32))
            assert abs(got_loss - want_loss) <= 1e-12 + 1e-9 * abs(want_loss)
            assert float((h_grad - want_grad).double().norm() / want_grad.double().norm()) < 1e-5


def test_render_step_under_the_adaptive_schedule(hf):
    """hf_set_adaptive_schedule for pixel lists (unit of work = a pixel with >= 32 samples): forward image and hit triangles identical, the
    fused L2 step's image identical and its gradient equal up to the order of the atomic additions, launch after launch (the list
    of expensive pixels changes between them) and through a pixel list."""
    from hfb200 import workloads as wl
    dev = torch.device("cuda", 0)
    h = wl.fbm_heights(1024, octaves=8, seed=0xC2, device=dev)
    g = hf.load_dict({"type": "heightfield", "heights": h, "max_height": 0.3}, device=dev)
    c3 = wl.CONFIGS["C3"]["camera"]
    cam = hf.perspective_camera(c3["origin"], c3["target"], c3["up"], c3["fov"], 320, 256, 64, seed=4)       # 81920 pixels
    light = hf.directional_light((0.3, 0.2, 1.0), shadow_rays=True)
    target = torch.zeros(256, 320, device=dev)
    rng = np.random.default_rng(9)
    some = rng.permutation(320 * 256)[:60000].astype(np.int32)

    def step(pixels):
        img = torch.full((256, 320), -1.0, device=dev)
        loss = torch.zeros(1, dtype=torch.float64, device=dev)
        g.zero_grad(); g.render_step_l2(cam, light, target, pixels, image=img, loss=loss)
        torch.cuda.synchronize()
        return img.clone(), g.grad.double().clone(), float(loss)

    want = {k: step(p) for k, p in (("all", None), ("some", some))}
    want_img, want_prim = g.render_forward(cam, light, return_prims=True)
    g.set_adaptive_schedule(True)
    for it in range(3):
        for k, p in (("all", None), ("all", None), ("some", some), ("all", None)):
            img, grad, loss = step(p)
            assert torch.equal(img.view(torch.int32), want[k][0].view(torch.int32)), (it, k)
            assert float((grad - want[k][1]).norm() / want[k][1].norm()) < 1e-5
            assert abs(loss - want[k][2]) <= 1e-9 * abs(want[k][2])
        got_img, got_prim = g.render_forward(cam, light, return_prims=True)
        assert torch.equal(got_img.view(torch.int32), want_img.view(torch.int32)) and torch.equal(got_prim, want_prim)
    listed, units = g.adaptive_schedule_info("pixels")
    print("adaptive schedule, pixels:", listed, "of", units)
    assert units == 320 * 256 and listed < units // 4, (listed, units)
    g.set_adaptive_schedule(False)
