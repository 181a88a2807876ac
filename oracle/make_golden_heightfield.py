"""Generate tests/golden/drjit_heightfield.npz: closest hits of whole (small) heightfields evaluated THROUGH the reference's own
Dr.Jit header arithmetic (oracle/_ref/libdrjit_ref.so :: ref_heightfield_brute_force -- every triangle of the DESIGN.md section 2 mesh
through the Moeller-Trumbore sequence of include/mitsuba/render/mesh.h:357-380, rays moved to object space by transform_affine).

Run here (where /root/reference exists); the vectors are committed so that the GPU box, which has no /root/reference, can pin the
oracle's brute force, its DDA and the CUDA kernels to reference-compiled arithmetic at the level of whole heightfields.

    python oracle/make_golden_heightfield.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import hf_oracle as O  # noqa: E402
from util import adversarial_rays, random_heights, random_rays  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "drjit_heightfield.npz")
AFFINE = np.array([[1.5 * np.cos(0.4), -np.sin(0.4), 0.1, 0.3], [1.5 * np.sin(0.4), np.cos(0.4), 0.0, -0.2], [0.0, 0.2, 2.0, 0.5]], np.float32)


def main():
    O.build()
    assert O.ref_lib() is not None, "oracle/_ref/libdrjit_ref.so missing (needs /root/reference): make -C oracle ref"
    out = {}
    cases = [("a", 33, 33, "smooth", 0.7, None), ("b", 21, 33, "noise", 0.4, AFFINE), ("c", 2, 2, "flat", 1.0, None)]
    for name, H, W, kind, mh, tw in cases:
        rng = np.random.default_rng(20261105 + H * 100 + W)
        h = random_heights(rng, H, W, kind)
        o, d = random_rays(rng, 2500)
        ao, ad = adversarial_rays(W, H, mh)
        o, d = np.concatenate([o, ao]), np.concatenate([d, ad])
        if tw is not None:
            o = (o @ tw[:, :3].T + tw[:, 3]).astype(np.float32); d = (d @ tw[:, :3].T).astype(np.float32)
        maxt = np.where(rng.uniform(size=o.shape[0]) < 0.15, rng.uniform(0.5, 4, o.shape[0]), np.inf).astype(np.float32)
        hf = O.OracleHeightfield(h, mh, to_world=tw)           # only for the fp64-inverted to_object the product also uses
        to_object = np.asarray(hf.to_object, np.float32).reshape(-1)
        t, u, v, prim = O.ref_heightfield_brute_force(h, mh, to_object, o, d, maxt)
        out.update({f"{name}_h": h, f"{name}_mh": np.float32(mh), f"{name}_to_world": np.eye(4, dtype=np.float32)[:3] if tw is None else tw,
                    f"{name}_to_object": to_object.reshape(3, 4), f"{name}_o": o, f"{name}_d": d, f"{name}_maxt": maxt,
                    f"{name}_t": t, f"{name}_u": u, f"{name}_v": v, f"{name}_prim": prim})
        print(f"case {name}: {H}x{W} {kind}, {o.shape[0]} rays, {int(np.isfinite(t).sum())} hits")
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
