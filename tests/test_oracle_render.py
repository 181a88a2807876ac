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
    d64 = np.stack([rays.dx, rays.dy, rays.dz], 1).astype(np.float64)
    front = (np.stack([si["nx"], si["ny"], si["nz"]], 1).astype(np.float64) * -d64).sum(1) > 0      # one-sided BSDF, diffuse.cpp:127-139
    want = np.where(ref.valid() & (cosl > 0) & front, 0.5 / np.pi * 2.0 * cosl, 0.0)
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


def test_path_depth1_is_direct_and_path_adjoint_matches_fd(oracle):
    """max_depth = 1 reproduces the direct integrator; for max_depth = 3 the PRB adjoint equals central differences of the
    radiance with the path geometry (triangles, bounce directions, visibility) held fixed -- prb.py's detached sampling."""
    import hfb200
    rng = np.random.default_rng(5)
    h = random_heights(rng, 20, 18, "noise")
    mh, W, H = 0.5, 18, 20
    hf = oracle.OracleHeightfield(h, mh)
    cam = _cam(spp=4, w=12, h=10)
    l1 = hfb200.directional_light((0.5, 0.3, 0.6), irradiance=1.3, albedo=0.9, shadow_rays=True, max_depth=1)
    img_d, *_ = oracle.render_forward(hf, cam, l1)
    img_p, _ = oracle.render_path_forward(hf, cam, l1)
    assert np.array_equal(img_d, img_p)
    light = dict(l1, max_depth=3)
    adj = rng.uniform(-1, 1, 12 * 10)
    value, vprim, vadj = oracle.render_path_samples(hf, cam, light)
    assert (vprim[:, 1] != 0xFFFFFFFF).sum() > 20 and (vprim[:, 2] != 0xFFFFFFFF).sum() > 5
    g = oracle.render_path_backward(hf, cam, light, adj, samples=(value, vprim, vadj))
    scale, alb = 0.9 / np.pi * 1.3, 0.9
    ld = np.asarray(light["direction"], np.float64)

    def normal(hh, prim):
        cell, k = int(prim) >> 1, int(prim) & 1
        cy, cx = divmod(cell, W - 1)
        P = lambda ii, jj: np.array([(2 * ii - (W - 1)) / (W - 1), (2 * jj - (H - 1)) / (H - 1), mh * hh[jj, ii]])
        p0 = P(cx + 1, cy)
        p1, p2 = (P(cx, cy + 1), P(cx, cy)) if k == 0 else (P(cx + 1, cy + 1), P(cx, cy + 1))
        n = np.cross(p1 - p0, p2 - p0)
        return n / np.linalg.norm(n)

    # radiance of the FIXED paths as an explicit fp64 function of the facet normals (detached: triangles, wo_k, visibility):
    # L = sum_k alb^k lit_k scale (n_k . l) prod_{j<k} (n_j . wo_j) / (n_j . wo_j)_primal
    info = np.zeros((vprim.shape[0], 4, 5))
    oracle.render_path_samples(hf, cam, light, info=info)
    h64 = h.astype(np.float64)
    a_s = np.repeat(adj, 4) / 4
    paths = [(i, [int(p) for p in vprim[i] if p != 0xFFFFFFFF]) for i in np.nonzero(vprim[:, 0] != 0xFFFFFFFF)[0]]

    def objective(hh):
        tot = 0.0
        for i, verts in paths:
            w, L = 1.0, 0.0
            for k, p in enumerate(verts):
                n = normal(hh, p)
                lit, wo, cwo = info[i, k, 0], info[i, k, 1:4], info[i, k, 4]
                L += w * (alb ** k) * lit * scale * float(n @ ld)
                if cwo > 0:
                    w *= float(n @ wo) / cwo
            tot += a_s[i] * L
        return tot

    # the primal agrees with the kernel-order fp32 radiance
    L0 = objective(h64)
    assert abs(L0 - float((a_s * value.reshape(-1)).sum())) <= 1e-5 * abs(L0) + 1e-7
    idx = np.argsort(-np.abs(g).reshape(-1))[:6]
    for flat in idx:
        j, i = divmod(int(flat), W)
        e = 1e-5
        hp, hm = h64.copy(), h64.copy(); hp[j, i] += e; hm[j, i] -= e
        fd = (objective(hp) - objective(hm)) / (2 * e)
        assert abs(fd - g[j, i]) <= 5e-4 * max(abs(fd), 1e-6), (fd, g[j, i])
