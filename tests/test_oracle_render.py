"""CPU checks of the oracle's render-step restatement (hfo_camera_rays / hfo_render_samples / render_backward): the
definition the GPU render kernels are compared with (tests/test_gpu_render.py) must itself be right."""
import numpy as np

from util import random_heights

f32 = np.float32


def _cam(spp=8, w=24, h=20, seed=3):
    import hfb200          # host-side helper only (no GPU needed): the fp32 camera frame both sides consume
    return hfb200.perspective_camera((0.2, -2.5, 1.8), (0, 0, 0), (0, 0, 1), 40.0, w, h, spp, seed)


def test_camera_rays_are_a_pinhole(oracle):
    cam = _cam(spp=4)
    r = oracle.camera_rays(cam)
    n = 24 * 20 * 4
    assert r.n == n and np.all(np.isinf(r.maxt))
    d = np.stack([r.dx, r.dy, r.dz], 1).astype(np.float64)
    assert np.allclose(np.linalg.norm(d, axis=1), 1, atol=2e-7)
    assert np.all(r.ox == f32(cam["origin"][0])) and np.all(r.oz == f32(cam["origin"][2]))
    # film coordinates recovered from the direction fall inside the sample's own pixel (perspective.cpp:212-231)
    R, U, F = (np.asarray(cam[k], np.float64) for k in ("right", "up", "forward"))
    sx, sy = (d @ R) / (d @ F), (d @ U) / (d @ F)
    tx = cam["tan_half_fov"]; ty = tx * 20 / 24
    fx, fy = (sx / tx + 1) / 2 * 24, (1 - sy / ty) / 2 * 20
    pid = np.repeat(np.arange(24 * 20), 4)
    assert np.all(np.floor(fx + 1e-4).astype(int) >= pid % 24 - 0) and np.all(fx <= pid % 24 + 1 + 1e-4) and np.all(fx >= pid % 24 - 1e-4)
    assert np.all(fy <= pid // 24 + 1 + 1e-4) and np.all(fy >= pid // 24 - 1e-4)
    # spp == 1: pixel centres; a pixel list selects the same rays
    c1 = oracle.camera_rays(_cam(spp=1), np.array([5, 77], np.uint32))
    call = oracle.camera_rays(_cam(spp=1))
    assert c1.dx[0] == call.dx[5] and c1.dz[1] == call.dz[77]


def test_render_samples_agree_with_the_shape_level_oracle(oracle):
    rng = np.random.default_rng(0)
    hf = oracle.OracleHeightfield(random_heights(rng, 40, 36, "smooth"), 0.4)
    cam = _cam(spp=8)
    light = {"direction": [float(f32(x)) for x in np.array([0.6, 0.2, 0.5]) / np.linalg.norm([0.6, 0.2, 0.5])], "irradiance": 2.0, "albedo": 0.5}
    value, prim, pi = oracle.render_samples(hf, cam, light)
    rays = oracle.camera_rays(cam)
    ref = hf.ray_intersect_preliminary(rays)
    assert np.array_equal(ref.t.view(np.uint32), pi.t.view(np.uint32)) and np.array_equal(ref.prim_index, pi.prim_index)
    si = hf.compute_surface_interaction(rays, ref)
    cosl = (np.stack([si["nx"], si["ny"], si["nz"]], 1).astype(np.float64) @ np.asarray(light["direction"], np.float64))
    want = np.where(ref.valid() & (cosl > 0), 0.5 / np.pi * 2.0 * cosl, 0.0)
    assert np.allclose(value.reshape(-1), want, rtol=2e-6, atol=1e-7)
    assert np.array_equal(prim.reshape(-1) != 0xFFFFFFFF, value.reshape(-1) > 0)
    # a visibility ray can only remove light, and it does on a rough field lit at a grazing angle
    rough = oracle.OracleHeightfield(random_heights(rng, 40, 36, "noise"), 0.6)
    lg = dict(light, direction=[float(f32(x)) for x in np.array([0.9, 0.1, 0.2]) / np.linalg.norm([0.9, 0.1, 0.2])])
    v0, p0, _ = oracle.render_samples(rough, cam, lg)
    v1, p1, _ = oracle.render_samples(rough, cam, dict(lg, shadow_rays=True))
    assert np.all((v1 == v0) | (v1 == 0)) and (v1 == 0).sum() > (v0 == 0).sum()


def test_render_backward_is_the_derivative_with_hit_triangles_held_fixed(oracle):
    """PRB without reparameterisation: visibility / hit position detached, the value moves through the facet normal only."""
    rng = np.random.default_rng(1)
    h = random_heights(rng, 24, 22, "smooth")
    mh = 0.4
    hf = oracle.OracleHeightfield(h, mh)
    cam = _cam(spp=4, w=16, h=12)
    ld = np.array([0.5, 0.3, 0.6]); ld /= np.linalg.norm(ld)
    light = {"direction": [float(f32(x)) for x in ld], "irradiance": 1.3, "albedo": 0.9}
    adj = rng.uniform(-1, 1, 16 * 12)
    samples = oracle.render_samples(hf, cam, light)
    g = oracle.render_backward(hf, cam, light, adj, samples=samples)
    value, prim, _ = samples
    scale = 0.9 / np.pi * 1.3
    W = 22

    def objective(hh):          # fp64, facet normals of the FIXED triangles
        tot = 0.0
        a_s = np.repeat(adj, 4) / 4
        for i in np.nonzero(prim.reshape(-1) != 0xFFFFFFFF)[0]:
            p = int(prim.reshape(-1)[i]); cell, k = p >> 1, p & 1
            cy, cx = divmod(cell, W - 1)
            X = lambda ii: (2 * ii - (W - 1)) / (W - 1); Y = lambda jj: (2 * jj - 23) / 23
            P = lambda ii, jj: np.array([X(ii), Y(jj), mh * hh[jj, ii]])
            p0 = P(cx + 1, cy)
            p1, p2 = (P(cx, cy + 1), P(cx, cy)) if k == 0 else (P(cx + 1, cy + 1), P(cx, cy + 1))
            n = np.cross(p1 - p0, p2 - p0); n /= np.linalg.norm(n)
            tot += a_s[i] * scale * float(n @ np.asarray(light["direction"], np.float64))
        return tot

    h64 = h.astype(np.float64)
    idx = np.argsort(-np.abs(g).reshape(-1))[:6]
    for flat in idx:
        j, i = divmod(int(flat), W)
        e = 1e-4
        hp, hm = h64.copy(), h64.copy(); hp[j, i] += e; hm[j, i] -= e
        fd = (objective(hp) - objective(hm)) / (2 * e)
        assert abs(fd - g[j, i]) <= 2e-4 * max(abs(fd), 1e-6), (fd, g[j, i])
