"""hf_ray_intersect_preliminary_static: the preliminary intersection for callers that cannot allocate device memory.

Some users of this library are not allowed to call cudaMalloc: a plugin whose host application owns the device-memory budget
through its own pool, a latency-bound loop that cannot afford the device-wide synchronisation an allocation implies, a process
whose memory was partitioned up front. They cannot build a context (hf_create allocates the grids), so the library offers one
entry point that works on the caller's own heights / rays / result arrays and keeps its acceleration data in a fixed 9 MiB array
inside the library image (include/hf_b200.h). This file checks, through the C ABI, that it returns what the context-based entry
point returns -- bit for bit, the tolerance every forward test of this repository uses -- on two calls in a row over DIFFERENT
fields: the array is shared and keeps the last call's contents, so the second call is the one that would see leftovers.

The file sorts last on purpose (pytest -x stops at the first failure; a failure here must not hide the other tests) and does not
depend on anything the other tests leave behind: it passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from util import adversarial_rays, random_heights, random_rays

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bits(t):
    return t.detach().cpu().numpy().view(np.uint32)


def _compare(hf, h, o, d, mh, to_world=None, maxt=None, active=None):
    """One call of the context-free entry point against the context-based one on identical inputs. Returns the number of hits."""
    ht = torch.from_numpy(np.ascontiguousarray(h))
    props = {"type": "heightfield", "heights": ht, "max_height": mh}
    if to_world is not None:
        props["to_world"] = to_world
    shape = hf.load_dict(props)
    ray = hf.Ray3f(torch.from_numpy(o.T.copy()), torch.from_numpy(d.T.copy()), None if maxt is None else torch.from_numpy(maxt))
    act = None if active is None else torch.from_numpy(active)
    want = shape.ray_intersect_preliminary(ray, active=act)

    mine = ht.cuda()                                   # the caller's own device copy of the heights
    out = hf.PreliminaryIntersection3f(len(ray), mine.device)
    for f in (out.t, out.prim_uv, out.prim_index):     # poison: every element must be written by the call
        f.view(torch.int32).fill_(0x7fc00123)
    got = hf.ray_intersect_preliminary_static(mine, ray, max_height=mh, to_world=to_world, active=act, out=out)
    torch.cuda.synchronize()

    assert np.array_equal(_bits(got.t), _bits(want.t)), "t differs from hf_ray_intersect_preliminary"
    hit = np.isfinite(want.t.cpu().numpy())
    assert np.array_equal(_bits(got.prim_index)[hit], _bits(want.prim_index)[hit])
    assert np.array_equal(_bits(got.prim_uv)[:, hit], _bits(want.prim_uv)[:, hit])
    assert torch.equal(mine.cpu(), ht), "the caller's heights are an input: they must come back untouched"
    return int(hit.sum())


@pytest.mark.gpu
def test_static_entry_matches_the_context_entry_on_two_calls_in_a_row():
    import hfb200 as hf
    rng = np.random.default_rng(20260311)
    # call 1: a mid-sized field under an affine transform, random + adversarial rays
    h1 = random_heights(rng, 193, 257)
    tw = np.array([[1.5, 0.2, 0.0, 0.3], [-0.1, 1.2, 0.1, -0.2], [0.0, 0.3, 0.8, 0.1]], np.float32)
    o1, d1 = random_rays(rng, 60000)
    oa, da = adversarial_rays(257, 193, 0.7)
    n1 = _compare(hf, h1, np.concatenate([o1, oa]), np.concatenate([d1, da]), 0.7, to_world=tw)
    assert n1 > 10000
    # call 2, straight after it: a SMALLER field with other heights (whatever call 1 left beyond this field's nodes, and in the
    # work counter, is still in the array), finite maxt and an active mask
    h2 = random_heights(rng, 48, 64, kind="noise")
    o2, d2 = random_rays(rng, 40000)
    maxt = rng.uniform(0.2, 3.0, 40000).astype(np.float32)
    active = (rng.uniform(0, 1, 40000) > 0.25).astype(np.uint8)
    n2 = _compare(hf, h2, o2, d2, 1.0, maxt=maxt, active=active)
    assert n2 > 2000
    # call 3: the largest field the array holds (1025 x 1025 texels), then call 1's field again
    h3 = random_heights(rng, 1025, 1025)
    o3, d3 = random_rays(rng, 100000)
    assert _compare(hf, h3, o3, d3, 0.4) > 20000
    assert _compare(hf, h1, o1, d1, 0.7, to_world=tw) > 10000


@pytest.mark.gpu
def test_static_entry_empty_batch_and_oversized_field_on_the_gpu():
    import hfb200 as hf
    ray = hf.Ray3f(torch.zeros(3, 0), torch.zeros(3, 0))
    pi = hf.ray_intersect_preliminary_static(torch.zeros(8, 8, device="cuda"), ray)
    assert pi.t.numel() == 0
    big = torch.zeros(1025, 2049, device="cuda")
    ray = hf.Ray3f(torch.tensor([[0.0], [0.0], [2.0]]), torch.tensor([[0.0], [0.0], [-1.0]]))
    with pytest.raises(hf.HeightfieldError, match="static scratch"):
        hf.ray_intersect_preliminary_static(big, ray)


def test_static_entry_rejects_what_does_not_fit_before_any_device_work():
    """Runs without a GPU: the size and argument checks come before the first CUDA call."""
    import hfb200
    if not os.path.exists(hfb200.lib_path()):
        import __graft_entry__
        __graft_entry__.build()
    L = hfb200._load()
    d = hfb200._Desc()
    d.max_height = 1.0
    d.to_world[:] = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]
    fake = C.c_void_p(256)                              # never dereferenced: the call must fail first
    rays, out = hfb200._Rays(), hfb200._Prelim()
    for w, h in ((2049, 1025), (4096, 4096), (2, 3000000), (1 << 23, 1 << 23)):
        d.width, d.height = w, h
        assert L.hf_ray_intersect_preliminary_static(C.byref(d), fake, C.byref(rays), C.byref(out), 1, None) == 1
        msg = L.hf_last_error().decode()
        assert "static scratch" in msg and str(9 << 20) in msg, msg
    d.width, d.height = 1, 7
    assert L.hf_ray_intersect_preliminary_static(C.byref(d), fake, C.byref(rays), C.byref(out), 1, None) == 1
    d.width, d.height = 64, 64
    assert L.hf_ray_intersect_preliminary_static(None, fake, C.byref(rays), C.byref(out), 1, None) == 1
    assert L.hf_ray_intersect_preliminary_static(C.byref(d), None, C.byref(rays), C.byref(out), 1, None) == 1
    assert L.hf_ray_intersect_preliminary_static(C.byref(d), fake, C.byref(rays), C.byref(out), 1, None) == 1   # null ray components
    assert b"null component" in L.hf_last_error()
    assert hfb200.STATIC_SCRATCH_BYTES == 9 << 20
