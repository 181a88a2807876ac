"""Generate tests/golden/drjit_triangle.npz from the reference's OWN Dr.Jit header arithmetic
(oracle/_ref/libdrjit_ref.so, built by `make -C oracle ref` from /root/reference/ext/drjit).

Run here (where /root/reference exists); the vectors are committed so that the GPU box, which has no
/root/reference, can still pin the oracle and the kernels to the reference's op order.

    python oracle/make_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import hf_oracle as O  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "drjit_triangle.npz")


def main():
    R = O.ref_lib()
    if R is None:
        O.build()
        R = O.ref_lib()
    assert R is not None, "oracle/_ref/libdrjit_ref.so missing (needs /root/reference)"
    rng = np.random.default_rng(20261104)
    n = 4096
    f32 = np.float32
    # triangles of heightfield-cell shape (small xy extent, arbitrary z) plus fully random ones
    p0 = rng.uniform(-1, 1, (n, 3)).astype(f32)
    scale = np.where(np.arange(n)[:, None] % 2 == 0, 2.0 / 255.0, 1.0).astype(f32)
    p1 = (p0 + rng.uniform(-1, 1, (n, 3)).astype(f32) * scale).astype(f32)
    p2 = (p0 + rng.uniform(-1, 1, (n, 3)).astype(f32) * scale).astype(f32)
    # rays aimed at a random barycentric point (so most hit), some jittered off
    b = rng.dirichlet([1, 1, 1], n).astype(f32)
    tgt = (p0 * b[:, :1] + p1 * b[:, 1:2] + p2 * b[:, 2:3]).astype(f32)
    tgt += (rng.uniform(-1, 1, (n, 3)) * (rng.uniform(0, 1, (n, 1)) < 0.3) * 0.01).astype(f32)
    o = (tgt + rng.normal(0, 1, (n, 3)) * rng.uniform(0.5, 5, (n, 1))).astype(f32)
    d = (tgt - o).astype(f32)
    d = (d * rng.uniform(0.2, 2.0, (n, 1))).astype(f32)      # directions are not normalised in general
    maxt = np.where(rng.uniform(0, 1, n) < 0.2, rng.uniform(0, 2, n), np.inf).astype(f32)
    # a few exact edge cases: axis-parallel rays, degenerate triangle, ray through a vertex
    o[:8] = [[-0.3, -0.3, -10]] * 8; d[:8] = [[0, 0, 1]] * 8
    p0[:8] = [[1, -1, 0]] * 8; p1[:8] = [[-1, 1, 0]] * 8; p2[:8] = [[-1, -1, 0]] * 8; maxt[:8] = np.inf
    o[1] = [0.3, 0.3, -10]; p1[1] = [1, 1, 0]; p2[1] = [-1, 1, 0]
    o[2] = [1, -1, -10]                 # through vertex p0
    o[3] = [0, 0, -10]                  # on the diagonal edge
    p2[4] = p1[4]                       # degenerate
    d[5] = [1, 0, 0]                    # parallel to the plane
    maxt[6] = 9.999                     # maxt just short
    maxt[7] = 10.0                      # maxt exactly at the hit

    t = np.empty(n, f32); u = np.empty(n, f32); v = np.empty(n, f32); hit = np.empty(n, np.uint8)
    tt, uu, vv = C.c_float(), C.c_float(), C.c_float()
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    for i in range(n):
        oi, di, a0, a1, a2 = (np.ascontiguousarray(x[i]) for x in (o, d, p0, p1, p2))
        hit[i] = R.ref_moeller_trumbore(fp(oi), fp(di), C.c_float(maxt[i]), fp(a0), fp(a1), fp(a2),
                                        C.byref(tt), C.byref(uu), C.byref(vv))
        t[i], u[i], v[i] = tt.value, uu.value, vv.value

    # interpolation + face normal (mesh.cpp:738-757)
    b1 = rng.uniform(0, 1, n).astype(f32); b2 = (rng.uniform(0, 1, n) * (1 - b1)).astype(f32)
    pi = np.empty((n, 3), f32); ni = np.empty((n, 3), f32)
    for i in range(n):
        a0, a1, a2 = (np.ascontiguousarray(x[i]) for x in (p0, p1, p2))
        po, no = np.empty(3, f32), np.empty(3, f32)
        R.ref_interpolate_and_normal(fp(a0), fp(a1), fp(a2), C.c_float(b1[i]), C.c_float(b2[i]), fp(po), fp(no))
        pi[i], ni[i] = po, no

    # affine transforms (transform.h:104-139)
    m = rng.normal(0, 1, (64, 12)).astype(f32)
    q = rng.normal(0, 2, (64, 3)).astype(f32)
    xp = np.empty((64, 3), f32); xv = np.empty((64, 3), f32)
    for i in range(64):
        mi, qi = np.ascontiguousarray(m[i]), np.ascontiguousarray(q[i])
        a, bb = np.empty(3, f32), np.empty(3, f32)
        R.ref_transform_affine_point(fp(mi), fp(qi), fp(a)); R.ref_transform_vector(fp(mi), fp(qi), fp(bb))
        xp[i], xv[i] = a, bb

    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, o=o, d=d, maxt=maxt, p0=p0, p1=p1, p2=p2, t=t, u=u, v=v, hit=hit,
                        b1=b1, b2=b2, p_interp=pi, n_face=ni, m=m, q=q, xform_point=xp, xform_vector=xv)
    print(f"wrote {OUT}: {n} triangle cases ({int(hit.sum())} hits), 64 transforms")


if __name__ == "__main__":
    main()
