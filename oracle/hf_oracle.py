"""ctypes front end of the CPU oracle (oracle/hf_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the cpu_baseline /
--impl reference legs of bench.py. The product package never imports this module.

Parity status: triangle primitive pinned by the reference's known-answer tests
(src/render/tests/test_mesh.py:257-292, 458-531, 536-671) and by oracle/_ref (Dr.Jit headers);
heightfield level "parity unpinned" (plugin absent from the reference snapshot).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libhf_oracle.so")
_REF_PATH = os.path.join(_HERE, "_ref", "libdrjit_ref.so")

RF_MINIMAL, RF_UV, RF_DPDUV, RF_SHADINGFRAME = 0x1, 0x2, 0x4, 0x8
RF_BOUNDARYTEST = 0x40
RF_FOLLOWSHAPE, RF_DETACHSHAPE = 0x80, 0x100
RF_ALL = RF_UV | RF_DPDUV | RF_SHADINGFRAME

_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)
_u8p = C.POINTER(C.c_uint8)
_f64p = C.POINTER(C.c_double)


class _Shape(C.Structure):
    _fields_ = [("W", C.c_int32), ("H", C.c_int32), ("max_height", C.c_float), ("flip_normals", C.c_int32),
                ("to_world", C.c_float * 12), ("to_object", C.c_float * 12), ("heights", _f32p)]


class _Rays(C.Structure):
    _fields_ = [(k, _f32p) for k in ("ox", "oy", "oz", "dx", "dy", "dz", "maxt")] + [("active", _u8p)]


class _Prelim(C.Structure):
    _fields_ = [("t", _f32p), ("u", _f32p), ("v", _f32p), ("prim_index", _u32p)]


SI_FIELDS = ("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy",
             "dpdu_x", "dpdu_y", "dpdu_z", "dpdv_x", "dpdv_y", "dpdv_z",
             "shn_x", "shn_y", "shn_z", "shs_x", "shs_y", "shs_z", "sht_x", "sht_y", "sht_z",
             "wi_x", "wi_y", "wi_z", "boundary_test")
ADJ_FIELDS = ("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy",
              "dpdu_x", "dpdu_y", "dpdu_z", "dpdv_x", "dpdv_y", "dpdv_z")


class _SI(C.Structure):
    _fields_ = [(k, _f32p) for k in SI_FIELDS]


class _Adj(C.Structure):
    _fields_ = [(k, _f32p) for k in ADJ_FIELDS]


class _Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3), ("forward", C.c_float * 3),
                ("tan_half_fov", C.c_float), ("width", C.c_uint32), ("height", C.c_uint32), ("spp", C.c_uint32), ("seed", C.c_uint32)]


class _Light(C.Structure):
    _fields_ = [("direction", C.c_float * 3), ("irradiance", C.c_float), ("albedo", C.c_float), ("shadow_rays", C.c_int32),
                ("max_depth", C.c_int32)]


def build(force: bool = False) -> None:
    """Compile the oracle (and, where /root/reference exists, oracle/_ref)."""
    if force or not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "hf_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "all"])
    if os.path.isdir("/root/reference/ext/drjit/include") and (force or not os.path.exists(_REF_PATH)):
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.hfo_moeller_trumbore.restype = C.c_int
        L.hfo_moeller_trumbore.argtypes = [_f32p, _f32p, C.c_float, _f32p, _f32p, _f32p, _f32p, _f32p, _f32p]
        L.hfo_interpolate_and_normal.argtypes = [_f32p, _f32p, _f32p, C.c_float, C.c_float, _f32p, _f32p]
        L.hfo_xform_point.argtypes = [_f32p, _f32p, _f32p]
        L.hfo_xform_vector.argtypes = [_f32p, _f32p, _f32p]
        L.hfo_vertex_x.restype = C.c_float
        L.hfo_vertex_x.argtypes = [C.c_int32, C.c_int32]
        L.hfo_pyramid_build.restype = C.c_void_p
        L.hfo_pyramid_build.argtypes = [C.POINTER(_Shape)]
        L.hfo_pyramid_free.argtypes = [C.c_void_p]
        L.hfo_pyramid_levels.restype = C.c_int
        L.hfo_pyramid_levels.argtypes = [C.c_void_p]
        L.hfo_pyramid_level_dims.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.hfo_pyramid_level_data.restype = _f32p
        L.hfo_pyramid_level_data.argtypes = [C.c_void_p, C.c_int]
        L.hfo_bbox.argtypes = [C.POINTER(_Shape), C.c_void_p, _f32p]
        L.hfo_intersect_brute.argtypes = [C.POINTER(_Shape), C.POINTER(_Rays), C.c_size_t, C.POINTER(_Prelim)]
        L.hfo_intersect_dda.argtypes = [C.POINTER(_Shape), C.c_void_p, C.POINTER(_Rays), C.c_size_t,
                                        C.POINTER(_Prelim), _u32p, _u32p]
        L.hfo_ray_test.argtypes = [C.POINTER(_Shape), C.c_void_p, C.POINTER(_Rays), C.c_size_t, _u8p]
        L.hfo_surface_interaction.argtypes = [C.POINTER(_Shape), C.POINTER(_Rays), C.POINTER(_Prelim),
                                              C.c_uint32, C.c_size_t, C.POINTER(_SI)]
        L.hfo_backward.argtypes = [C.POINTER(_Shape), C.POINTER(_Rays), C.POINTER(_Prelim), C.POINTER(_Adj),
                                   C.c_uint32, C.c_size_t, _f64p, C.c_int]
        L.hfo_adam_step.argtypes = [_f32p, _f32p, _f32p, _f32p, C.c_size_t, C.c_double, C.c_double, C.c_double,
                                    C.c_double, C.c_int, C.c_int]
        L.hfo_camera_rays.argtypes = [C.POINTER(_Camera), _u32p, C.c_size_t] + [_f32p] * 7
        L.hfo_render_samples.argtypes = [C.POINTER(_Shape), C.c_void_p, C.POINTER(_Camera), C.POINTER(_Light), _u32p, C.c_size_t,
                                         _f32p, _u32p, C.POINTER(_Prelim)]
        L.hfo_render_path_samples.argtypes = [C.POINTER(_Shape), C.c_void_p, C.POINTER(_Camera), C.POINTER(_Light), _u32p, C.c_size_t,
                                              _f32p, _u32p, _f64p, _f64p]
        L.hfo_set_threads.argtypes = [C.c_int]
        L.hfo_get_threads.restype = C.c_int
        _lib = L
    return _lib


_ref = None


def ref_lib():
    """The reference's own Dr.Jit header arithmetic (oracle/_ref), or None if it was never built."""
    global _ref
    if _ref is None and os.path.exists(_REF_PATH):
        R = C.CDLL(_REF_PATH)
        R.ref_moeller_trumbore.restype = C.c_int
        R.ref_moeller_trumbore.argtypes = [_f32p, _f32p, C.c_float, _f32p, _f32p, _f32p, _f32p, _f32p, _f32p]
        R.ref_interpolate_and_normal.argtypes = [_f32p, _f32p, _f32p, C.c_float, C.c_float, _f32p, _f32p]
        R.ref_transform_affine_point.argtypes = [_f32p, _f32p, _f32p]
        R.ref_transform_vector.argtypes = [_f32p, _f32p, _f32p]
        if hasattr(R, "ref_heightfield_brute_force"):
            R.ref_heightfield_brute_force.argtypes = [_f32p, C.c_int, C.c_int, C.c_float, _f32p, _f32p, _f32p, _f32p, C.c_size_t,
                                                      _f32p, _f32p, _f32p, _u32p]
        _ref = R
    return _ref


def _fp(a):
    return a.ctypes.data_as(_f32p) if a is not None else None


def ref_heightfield_brute_force(heights, max_height, to_object, o, d, maxt):
    """Closest hit of every ray over ALL triangles of the heightfield, evaluated by the reference's own Dr.Jit header arithmetic
    (oracle/drjit_ref.cpp). heights (H, W); to_object 12 floats; o, d (N, 3); returns (t, u, v, prim_index)."""
    R = ref_lib()
    assert R is not None and hasattr(R, "ref_heightfield_brute_force"), "oracle/_ref is not built (needs /root/reference)"
    h = _f32(heights); oo = _f32(o); dd = _f32(d); mt = _f32(maxt); to = _f32(np.asarray(to_object).reshape(-1)[:12])
    n = oo.shape[0]
    t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
    prim = np.empty(n, np.uint32)
    R.ref_heightfield_brute_force(_fp(h), h.shape[1], h.shape[0], float(max_height), _fp(to), _fp(oo), _fp(dd), _fp(mt), n,
                                  _fp(t), _fp(u), _fp(v), prim.ctypes.data_as(_u32p))
    return t, u, v, prim


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def moeller_trumbore(o, d, maxt, p0, p1, p2, use_ref=False):
    """One triangle test; returns (hit, t, u, v). use_ref=True runs the Dr.Jit-header build."""
    o, d, p0, p1, p2 = (_f32(x) for x in (o, d, p0, p1, p2))
    t, u, v = C.c_float(), C.c_float(), C.c_float()
    fn = ref_lib().ref_moeller_trumbore if use_ref else lib().hfo_moeller_trumbore
    hit = fn(_fp(o), _fp(d), C.c_float(maxt), _fp(p0), _fp(p1), _fp(p2), C.byref(t), C.byref(u), C.byref(v))
    return bool(hit), np.float32(t.value), np.float32(u.value), np.float32(v.value)


def invert_affine(to_world):
    """fp64 inverse of a 3x4 affine, rounded to fp32 (to_object definition)."""
    m = np.eye(4, dtype=np.float64)
    m[:3, :4] = np.asarray(to_world, dtype=np.float64).reshape(3, 4)
    return np.linalg.inv(m)[:3, :4].astype(np.float32)


class Rays:
    """SoA ray batch on the host."""

    def __init__(self, o, d, maxt=None, active=None):
        o = _f32(o).reshape(-1, 3)
        d = _f32(d).reshape(-1, 3)
        self.n = o.shape[0]
        self.ox, self.oy, self.oz = (np.ascontiguousarray(o[:, k]) for k in range(3))
        self.dx, self.dy, self.dz = (np.ascontiguousarray(d[:, k]) for k in range(3))
        self.maxt = np.full(self.n, np.inf, np.float32) if maxt is None else \
            np.ascontiguousarray(np.broadcast_to(_f32(maxt), (self.n,)))
        self.active = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)

    @classmethod
    def from_soa(cls, ox, oy, oz, dx, dy, dz, maxt, active=None):
        r = cls.__new__(cls)
        r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, r.maxt = (_f32(x) for x in (ox, oy, oz, dx, dy, dz, maxt))
        r.n = r.ox.shape[0]
        r.active = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
        return r

    def c(self):
        return _Rays(_fp(self.ox), _fp(self.oy), _fp(self.oz), _fp(self.dx), _fp(self.dy), _fp(self.dz),
                     _fp(self.maxt), self.active.ctypes.data_as(_u8p) if self.active is not None else None)


class Prelim:
    def __init__(self, n=None, t=None, u=None, v=None, prim_index=None):
        if t is None:
            self.t = np.empty(n, np.float32); self.u = np.empty(n, np.float32)
            self.v = np.empty(n, np.float32); self.prim_index = np.empty(n, np.uint32)
        else:
            self.t, self.u, self.v = _f32(t), _f32(u), _f32(v)
            self.prim_index = np.ascontiguousarray(prim_index, dtype=np.uint32)

    def c(self):
        return _Prelim(_fp(self.t), _fp(self.u), _fp(self.v), self.prim_index.ctypes.data_as(_u32p))

    def valid(self):
        return np.isfinite(self.t)


class OracleHeightfield:
    """CPU oracle of the `heightfield` shape (API mirrors include/mitsuba/render/shape.h:137-183, 253)."""

    def __init__(self, heights, max_height=1.0, to_world=None, flip_normals=False, to_object=None):
        self.heights = np.ascontiguousarray(heights, dtype=np.float32)
        assert self.heights.ndim == 2 and min(self.heights.shape) >= 2
        self.H, self.W = self.heights.shape
        self.max_height = float(max_height)
        self.to_world = np.eye(4, dtype=np.float32)[:3, :4].copy() if to_world is None else \
            _f32(to_world).reshape(-1)[:12].reshape(3, 4).copy()
        self.to_object = invert_affine(self.to_world) if to_object is None else _f32(to_object).reshape(3, 4)
        self.flip_normals = bool(flip_normals)
        self._pyr = None

    # -- plumbing ---------------------------------------------------------
    def _shape(self):
        s = _Shape()
        s.W, s.H, s.max_height, s.flip_normals = self.W, self.H, self.max_height, int(self.flip_normals)
        s.to_world[:] = self.to_world.reshape(-1).tolist()
        s.to_object[:] = self.to_object.reshape(-1).tolist()
        s.heights = _fp(self.heights)
        return s

    def _pyramid(self):
        if self._pyr is None:
            self._pyr = lib().hfo_pyramid_build(C.byref(self._shape()))
        return self._pyr

    def parameters_changed(self):
        if self._pyr is not None:
            lib().hfo_pyramid_free(self._pyr)
            self._pyr = None

    def __del__(self):
        try:
            self.parameters_changed()
        except Exception:
            pass

    # -- API --------------------------------------------------------------
    def primitive_count(self):
        return 2 * (self.W - 1) * (self.H - 1)

    def bbox(self):
        out = np.empty(6, np.float32)
        lib().hfo_bbox(C.byref(self._shape()), self._pyramid(), _fp(out))
        return out

    def pyramid_levels(self):
        """List of (min,max) arrays of shape (h, w, 2) for levels 1..Ltop."""
        p = self._pyramid()
        out = []
        for l in range(1, lib().hfo_pyramid_levels(p) + 1):
            w, h = C.c_int(), C.c_int()
            lib().hfo_pyramid_level_dims(p, l, C.byref(w), C.byref(h))
            ptr = lib().hfo_pyramid_level_data(p, l)
            out.append(np.ctypeslib.as_array(ptr, shape=(h.value, w.value, 2)).copy())
        return out

    def ray_intersect_preliminary(self, rays: Rays, brute=False, counters=False):
        out = Prelim(rays.n)
        s, r, o = self._shape(), rays.c(), out.c()
        if brute:
            lib().hfo_intersect_brute(C.byref(s), C.byref(r), rays.n, C.byref(o))
            return out
        nn = np.zeros(rays.n, np.uint32) if counters else None
        nl = np.zeros(rays.n, np.uint32) if counters else None
        lib().hfo_intersect_dda(C.byref(s), self._pyramid(), C.byref(r), rays.n, C.byref(o),
                                nn.ctypes.data_as(_u32p) if counters else None,
                                nl.ctypes.data_as(_u32p) if counters else None)
        return (out, nn, nl) if counters else out

    def ray_test(self, rays: Rays):
        hit = np.empty(rays.n, np.uint8)
        s, r = self._shape(), rays.c()
        lib().hfo_ray_test(C.byref(s), self._pyramid(), C.byref(r), rays.n, hit.ctypes.data_as(_u8p))
        return hit.astype(bool)

    def compute_surface_interaction(self, rays: Rays, pi: Prelim, ray_flags=RF_ALL, fields=SI_FIELDS):
        out = {k: np.empty(rays.n, np.float32) for k in fields}
        si = _SI(**{k: _fp(out[k]) for k in fields})
        s, r, p = self._shape(), rays.c(), pi.c()
        lib().hfo_surface_interaction(C.byref(s), C.byref(r), C.byref(p), ray_flags, rays.n, C.byref(si))
        return out

    def backward(self, rays: Rays, pi: Prelim, adj: dict, ray_flags=RF_ALL, literal=True, grad=None):
        """adj: dict of upstream adjoints keyed like ADJ_FIELDS (missing = 0). Returns float64 (H, W)."""
        grad = np.zeros((self.H, self.W), np.float64) if grad is None else grad
        keep = {k: _f32(v) for k, v in adj.items()}
        a = _Adj(**{k: _fp(keep[k]) for k in keep})
        s, r, p = self._shape(), rays.c(), pi.c()
        lib().hfo_backward(C.byref(s), C.byref(r), C.byref(p), C.byref(a), ray_flags, rays.n,
                           grad.ctypes.data_as(_f64p), int(literal))
        return grad


def adam_step(p, g, m, v, lr, beta1=0.9, beta2=0.999, eps=1e-8, t=1, mask_updates=False):
    """In place on contiguous fp32 arrays (src/python/python/ad/optimizers.py:263-302)."""
    for a in (p, g, m, v):
        assert a.dtype == np.float32 and a.flags.c_contiguous
    lib().hfo_adam_step(_fp(p), _fp(g), _fp(m), _fp(v), p.size, lr, beta1, beta2, eps, int(t), int(mask_updates))


def set_threads(n):
    lib().hfo_set_threads(int(n))


def get_threads():
    return lib().hfo_get_threads()


# ---- render step on the heightfield alone -----------------------------------------------
def make_camera(cam: dict) -> _Camera:
    """cam: dict with origin/right/up/forward (3 floats each), tan_half_fov, width, height, spp, seed -- the fp32 values the
    product's perspective_camera() helper produced (the oracle does not recompute the frame)."""
    c = _Camera()
    for k in ("origin", "right", "up", "forward"):
        getattr(c, k)[:] = [float(x) for x in cam[k]]
    c.tan_half_fov = float(cam["tan_half_fov"])
    c.width, c.height, c.spp, c.seed = int(cam["width"]), int(cam["height"]), int(cam["spp"]), int(cam.get("seed", 0))
    return c


def make_light(light: dict) -> _Light:
    l = _Light()
    l.direction[:] = [float(x) for x in light["direction"]]
    l.irradiance, l.albedo, l.shadow_rays = float(light.get("irradiance", 1.0)), float(light.get("albedo", 1.0)), int(bool(light.get("shadow_rays", False)))
    l.max_depth = int(light.get("max_depth", 1))
    return l


def _pix(pixels):
    if pixels is None:
        return None, None
    a = np.ascontiguousarray(pixels, np.uint32)
    return a, a.ctypes.data_as(_u32p)


def camera_rays(cam: dict, pixels=None) -> "Rays":
    c = make_camera(cam)
    a, ap = _pix(pixels)
    npix = c.width * c.height if a is None else a.size
    n = npix * c.spp
    f = [np.empty(n, np.float32) for _ in range(7)]
    lib().hfo_camera_rays(C.byref(c), ap, npix, *[_fp(x) for x in f])
    return Rays.from_soa(*f)


def render_samples(hf: "OracleHeightfield", cam: dict, light: dict, pixels=None):
    """-> (value (npix, spp) fp32, lit-triangle ids (npix, spp) uint32, Prelim of every sample)"""
    c, l = make_camera(cam), make_light(light)
    a, ap = _pix(pixels)
    npix = c.width * c.height if a is None else a.size
    n = npix * c.spp
    value, prim, pi = np.empty(n, np.float32), np.empty(n, np.uint32), Prelim(n)
    s, pc = hf._shape(), pi.c()
    lib().hfo_render_samples(C.byref(s), hf._pyramid(), C.byref(c), C.byref(l), ap, npix, _fp(value),
                             prim.ctypes.data_as(_u32p), C.byref(pc))
    return value.reshape(npix, c.spp), prim.reshape(npix, c.spp), pi


def render_forward(hf, cam, light, pixels=None):
    """Box-filter image in fp64, indexed by pixel id (zeros where not rendered), plus the per-sample data."""
    value, prim, pi = render_samples(hf, cam, light, pixels)
    img = np.zeros(int(cam["width"]) * int(cam["height"]), np.float64)
    ids = np.arange(img.size) if pixels is None else np.asarray(pixels, np.int64)
    img[ids] = value.astype(np.float64).mean(1)
    return img, value, prim, pi


def render_backward(hf, cam, light, adj_image, pixels=None, samples=None):
    """d(loss)/d(heights) in fp64 for a per-pixel adjoint image: each lit sample passes adj / spp * scale * l to the adjoint of
    its normal (hfo_backward), visibility and hit position detached."""
    value, prim, pi = samples if samples is not None else render_samples(hf, cam, light, pixels)
    npix, spp = value.shape
    ids = np.arange(npix) if pixels is None else np.asarray(pixels, np.int64)
    lit = (prim != 0xFFFFFFFF).reshape(-1)
    scale = np.float32(float(light.get("albedo", 1.0)) / np.pi * float(light.get("irradiance", 1.0)))
    a_v = (np.repeat(np.asarray(adj_image, np.float64)[ids], spp) / spp * float(scale)) * lit
    rays = camera_rays(cam, pixels)
    pi2 = Prelim(t=np.where(lit, pi.t, np.float32(np.inf)).astype(np.float32), u=pi.u, v=pi.v, prim_index=pi.prim_index)
    ldir = [float(x) for x in light["direction"]]
    adj = {"nx": (a_v * ldir[0]).astype(np.float32), "ny": (a_v * ldir[1]).astype(np.float32), "nz": (a_v * ldir[2]).astype(np.float32)}
    return hf.backward(rays, pi2, adj, literal=False)


def render_path_samples(hf, cam, light, pixels=None, info=None):
    """Path tracing (light['max_depth'] in 2..4): -> (value (npix, spp) fp32, vertex triangles (npix*spp, 4) uint32,
    d radiance / d n_k (npix*spp, 4, 3) float64)."""
    c, l = make_camera(cam), make_light(light)
    a, ap = _pix(pixels)
    npix = c.width * c.height if a is None else a.size
    n = npix * c.spp
    value, vprim, vadj = np.empty(n, np.float32), np.empty((n, 4), np.uint32), np.empty((n, 4, 3), np.float64)
    s = hf._shape()
    lib().hfo_render_path_samples(C.byref(s), hf._pyramid(), C.byref(c), C.byref(l), ap, npix, _fp(value),
                                  vprim.ctypes.data_as(_u32p), vadj.ctypes.data_as(_f64p),
                                  info.ctypes.data_as(_f64p) if info is not None else None)
    return value.reshape(npix, c.spp), vprim, vadj


def render_path_forward(hf, cam, light, pixels=None):
    value, vprim, vadj = render_path_samples(hf, cam, light, pixels)
    img = np.zeros(int(cam["width"]) * int(cam["height"]), np.float64)
    ids = np.arange(img.size) if pixels is None else np.asarray(pixels, np.int64)
    img[ids] = value.astype(np.float64).mean(1)
    return img, (value, vprim, vadj)


def render_path_backward(hf, cam, light, adj_image, pixels=None, samples=None):
    """fp64 gradient: every path vertex passes a_v * g_k to the adjoint of its facet normal (hfo_backward, n adjoint only)."""
    value, vprim, vadj = samples if samples is not None else render_path_samples(hf, cam, light, pixels)
    npix, spp = value.shape
    n = npix * spp
    ids = np.arange(npix) if pixels is None else np.asarray(pixels, np.int64)
    a_v = np.repeat(np.asarray(adj_image, np.float64)[ids], spp) / spp
    grad = np.zeros((hf.H, hf.W), np.float64)
    dummy = Rays(np.tile(np.array([[0, 0, 5]], np.float32), (n, 1)), np.tile(np.array([[0, 0, -1]], np.float32), (n, 1)))
    for k in range(4):
        ok = vprim[:, k] != 0xFFFFFFFF
        if not ok.any():
            continue
        pi = Prelim(t=np.where(ok, np.float32(1), np.float32(np.inf)).astype(np.float32), u=np.full(n, 0.3, np.float32),
                    v=np.full(n, 0.3, np.float32), prim_index=np.where(ok, vprim[:, k], 0).astype(np.uint32))
        a = a_v[:, None] * vadj[:, k, :] * ok[:, None]
        # split the adjoint into fp32 head + tail so that the fp32 interface of hfo_backward loses nothing
        for part in (a.astype(np.float32), (a - a.astype(np.float32).astype(np.float64)).astype(np.float32)):
            hf.backward(dummy, pi, {"nx": part[:, 0].copy(), "ny": part[:, 1].copy(), "nz": part[:, 2].copy()}, literal=False, grad=grad)
    return grad
