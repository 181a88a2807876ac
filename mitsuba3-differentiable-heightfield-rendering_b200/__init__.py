"""B200-native differentiable `heightfield` shape: host-side mirror of the Mitsuba 3 Shape plugin
interface (include/mitsuba/render/shape.h:137-197, 253, 432, 533-550) over the C ABI of
include/hf_b200.h.

Everything that computes runs in lib/libhf_b200.so (hand-written sm_100a kernels). PyTorch is used
only for device memory, streams and torch.distributed plumbing. There is no CPU fallback: importing
works anywhere, but constructing a Heightfield without the compiled library or without a CUDA
device raises.

Reference-facing names (citations into /root/reference):
    load_dict({'type': 'heightfield', ...})      src/core/plugin.cpp:174-200
    Heightfield.ray_intersect_preliminary        shape.h:137
    Heightfield.ray_test                         shape.h:153
    Heightfield.compute_surface_interaction      shape.h:179-183
    Heightfield.ray_intersect                    shape.h:197
    Heightfield.bbox / primitive_count           shape.h:253 / 432
    traverse(shape) -> SceneParameters           src/python/python/util.py:292, 185-232
    Adam                                         src/python/python/ad/optimizers.py:203-309
"""
from __future__ import annotations

import ctypes as C
import enum
import os

import numpy as np
import torch

__all__ = ["RayFlags", "Ray3f", "PreliminaryIntersection3f", "SurfaceInteraction3f", "Heightfield",
           "load_dict", "traverse", "SceneParameters", "Adam", "lib_path", "HeightfieldError", "perspective_camera",
           "directional_light"]

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.environ.get("HF_B200_LIB") or os.path.join(_HERE, "lib", "libhf_b200.so")   # override: kernel experiments only
_f32p = C.POINTER(C.c_float)


def lib_path() -> str:
    return _LIB


class HeightfieldError(RuntimeError):
    pass


class RayFlags(enum.IntFlag):
    """Bit-identical to include/mitsuba/render/interaction.h:19-69."""
    Empty = 0x0
    Minimal = 0x1
    UV = 0x2
    dPdUV = 0x4
    ShadingFrame = 0x8
    dNGdUV = 0x10
    dNSdUV = 0x20
    BoundaryTest = 0x40
    FollowShape = 0x80
    DetachShape = 0x100
    All = 0x2 | 0x4 | 0x8
    AllNonDifferentiable = 0x2 | 0x4 | 0x8 | 0x100


# ---- ctypes mirror of include/hf_b200.h ------------------------------------------------------
class _Desc(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("max_height", C.c_float),
                ("to_world", C.c_float * 12), ("to_object", C.c_float * 12), ("has_to_object", C.c_int32),
                ("flip_normals", C.c_int32), ("device", C.c_int32)]


class _Rays(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("ox", "oy", "oz", "dx", "dy", "dz", "maxt", "active")]


class _Prelim(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("t", "u", "v", "prim_index")]


SI_FIELDS = ("t", "px", "py", "pz", "nx", "ny", "nz", "uvx", "uvy",
             "dpdu_x", "dpdu_y", "dpdu_z", "dpdv_x", "dpdv_y", "dpdv_z",
             "shn_x", "shn_y", "shn_z", "shs_x", "shs_y", "shs_z", "sht_x", "sht_y", "sht_z",
             "wi_x", "wi_y", "wi_z", "boundary_test")
ADJ_FIELDS = SI_FIELDS[:15]


class _SI(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in SI_FIELDS]


class _Adj(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ADJ_FIELDS]


class _Adam(C.Structure):
    _fields_ = [("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("epsilon", C.c_double),
                ("mask_updates", C.c_int32)]


class _Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3), ("forward", C.c_float * 3),
                ("tan_half_fov", C.c_float), ("width", C.c_uint32), ("height", C.c_uint32), ("spp", C.c_uint32), ("seed", C.c_uint32)]


class _Light(C.Structure):
    _fields_ = [("direction", C.c_float * 3), ("irradiance", C.c_float), ("albedo", C.c_float), ("shadow_rays", C.c_int32),
                ("max_depth", C.c_int32)]


class _Stats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("rays", "nodes", "leaves", "hits", "segments")]


_lib = None


def _load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB):
        raise HeightfieldError(
            f"compiled CUDA library missing: {_LIB}. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"(nvcc, sm_100a). There is no CPU fallback.")
    L = C.CDLL(_LIB)
    L.hf_last_error.restype = C.c_char_p
    vp, sz, u32 = C.c_void_p, C.c_size_t, C.c_uint32
    sig = {
        "hf_abi_version": [],
        "hf_create": [C.POINTER(vp), C.POINTER(_Desc)],
        "hf_destroy": [vp],
        "hf_set_heights": [vp, vp, C.c_int, vp],
        "hf_heights": [vp, C.POINTER(vp)],
        "hf_grad": [vp, C.POINTER(vp)],
        "hf_zero_grad": [vp, vp],
        "hf_set_to_world": [vp, _f32p, _f32p],
        "hf_set_max_height": [vp, C.c_float],
        "hf_parameters_changed": [vp, vp],
        "hf_parameters_grad_enabled": [vp, C.POINTER(C.c_int)],
        "hf_set_grad_enabled": [vp, C.c_int],
        "hf_set_ray_coherence": [vp, C.c_int],
        "hf_set_adaptive_schedule": [vp, C.c_int],
        "hf_adaptive_schedule_info": [vp, C.c_int, C.POINTER(u32), C.POINTER(u32)],
        "hf_bbox": [vp, _f32p],
        "hf_primitive_count": [vp, C.POINTER(u32)],
        "hf_size": [vp, C.POINTER(u32), C.POINTER(u32)],
        "hf_pyramid_level": [vp, u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), C.POINTER(vp)],
        "hf_qpyramid_level": [vp, u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), C.POINTER(vp), C.POINTER(C.c_float), C.POINTER(C.c_float)],
        "hf_ray_intersect_preliminary": [vp, C.POINTER(_Rays), C.POINTER(_Prelim), sz, vp],
        "hf_ray_intersect_preliminary_counted": [vp, C.POINTER(_Rays), C.POINTER(_Prelim), sz, vp],
        "hf_ray_intersect_preliminary_static": [C.POINTER(_Desc), vp, C.POINTER(_Rays), C.POINTER(_Prelim), sz, vp],
        "hf_stats_get": [vp, C.POINTER(_Stats), C.c_int],
        "hf_ray_test": [vp, C.POINTER(_Rays), vp, sz, vp],
        "hf_compute_surface_interaction": [vp, C.POINTER(_Rays), C.POINTER(_Prelim), u32, C.POINTER(_SI), sz, vp],
        "hf_ray_intersect": [vp, C.POINTER(_Rays), u32, C.POINTER(_Prelim), C.POINTER(_SI), sz, vp],
        "hf_backward": [vp, C.POINTER(_Rays), C.POINTER(_Prelim), C.POINTER(_Adj), u32, sz, vp],
        "hf_adam_step": [vp, C.POINTER(_Adam), vp],
        "hf_adam_reset": [vp, vp],
        "hf_comm_unique_id": [C.POINTER(C.c_uint8)],
        "hf_comm_init": [vp, C.c_int, C.c_int, C.POINTER(C.c_uint8)],
        "hf_comm_destroy": [vp],
        "hf_allreduce_grad": [vp, vp],
        "hf_render_rays": [vp, C.POINTER(_Camera), vp, sz, C.POINTER(_Rays), vp],
        "hf_render_forward": [vp, C.POINTER(_Camera), C.POINTER(_Light), vp, sz, vp, vp, vp],
        "hf_render_backward": [vp, C.POINTER(_Camera), C.POINTER(_Light), vp, sz, vp, vp, vp],
        "hf_render_step_l2": [vp, C.POINTER(_Camera), C.POINTER(_Light), vp, sz, vp, vp, vp, vp],
        "hf_host_alloc": [C.POINTER(vp), sz],
        "hf_host_free": [vp],
        "hf_ray_intersect_host": [vp, C.POINTER(_Rays), u32, C.POINTER(_Prelim), C.POINTER(_SI), sz],
        "hf_forward_backward_host": [vp, C.POINTER(_Rays), C.POINTER(_Adj), u32, C.POINTER(_Prelim), C.POINTER(_SI), vp, sz],
        "hf_render_step_l2_host": [vp, C.POINTER(_Camera), C.POINTER(_Light), vp, sz, vp, vp, vp, vp, C.POINTER(C.c_double)],
        "hf_peer_export": [vp, C.POINTER(C.c_uint8)],
        "hf_peer_attach": [vp, C.c_int, C.c_int, C.POINTER(C.c_uint8)],
        "hf_peer_detach": [vp],
        "hf_peer_barrier": [vp, vp],
        "hf_peer_zero_grad": [vp, vp],
        "hf_peer_reduce_grad": [vp, vp],
        "hf_peer_set_mode": [vp, C.c_int],
        "hf_peer_adam_step": [vp, C.POINTER(_Adam), vp],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise HeightfieldError(_load().hf_last_error().decode("utf-8", "replace"))


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def capture(fn, warmup=1):
    """Record the launches `fn` makes (every entry point here is asynchronous on the current stream, the NCCL allreduce included)
    into one CUDA graph and return a callable that replays it: a step of a few ~1 ms kernels is otherwise bound by the host's
    launch calls. `fn` must not allocate, synchronise or read results back; arguments are baked in (the buffers are re-used)."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(max(1, warmup)):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    def replay():
        g.replay()
    replay.graph = g              # keeps the graph (and its memory pool) alive as long as the callable
    return replay


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class _DevView:
    """Zero-copy torch view of a device buffer owned by the C library."""

    def __init__(self, ptr, shape, owner):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 3, "strides": None}
        self._owner = owner


# ---- records (interaction.h) --------------------------------------------------------------
def _soa(x, rows, name, device):
    t = torch.as_tensor(x, dtype=torch.float32, device=device)
    if t.dim() == 2 and t.shape[0] != rows and t.shape[1] == rows:
        t = t.t()
    if t.dim() != 2 or t.shape[0] != rows:
        raise HeightfieldError(f"{name}: expected shape ({rows}, N) or (N, {rows})")
    return t.contiguous()


class Ray3f:
    """SoA ray batch on the device: o, d are (3, N) tensors, maxt is (N,)."""

    def __init__(self, o, d, maxt=None, device="cuda"):
        self.o = _soa(o, 3, "o", device)
        self.d = _soa(d, 3, "d", device)
        n = self.o.shape[1]
        if maxt is None:
            self.maxt = torch.full((n,), float("inf"), dtype=torch.float32, device=self.o.device)
        else:
            self.maxt = torch.as_tensor(maxt, dtype=torch.float32, device=self.o.device).expand(n).contiguous()

    def __len__(self):
        return self.o.shape[1]

    def _c(self, active=None):
        return _Rays(_ptr(self.o[0]), _ptr(self.o[1]), _ptr(self.o[2]), _ptr(self.d[0]), _ptr(self.d[1]), _ptr(self.d[2]),
                     _ptr(self.maxt), _ptr(active))


class PreliminaryIntersection3f:
    """interaction.h:586-691; invalid <=> t == +inf (639-646)."""

    def __init__(self, n, device, shape=None):
        self.t = torch.empty(n, dtype=torch.float32, device=device)
        self.prim_uv = torch.empty((2, n), dtype=torch.float32, device=device)
        self.prim_index = torch.empty(n, dtype=torch.int32, device=device)   # uint32 bit pattern
        self.shape_index = 0xFFFFFFFF     # rectangle.cpp:222-223
        self.shape = shape

    def is_valid(self):
        return self.t != float("inf")

    def _c(self):
        return _Prelim(_ptr(self.t), _ptr(self.prim_uv[0]), _ptr(self.prim_uv[1]), _ptr(self.prim_index))

    def compute_surface_interaction(self, ray, ray_flags=RayFlags.All, active=None):
        """interaction.h:658-684"""
        return self.shape.compute_surface_interaction(ray, self, ray_flags, active=active)


class SurfaceInteraction3f:
    """World-space record (interaction.h). `fields` selects which members are materialised; the kernels skip the
    rest (null pointers in hf_si). Default: everything RayFlags.All produces."""
    ALL_FIELDS = ("t", "p", "n", "uv", "dp_du", "dp_dv", "sh_n", "sh_s", "sh_t", "wi", "boundary_test")
    DEFAULT_FIELDS = ALL_FIELDS[:-1]          # boundary_test only on request (RayFlags.BoundaryTest)
    _ROWS = {"p": 3, "n": 3, "uv": 2, "dp_du": 3, "dp_dv": 3, "sh_n": 3, "sh_s": 3, "sh_t": 3, "wi": 3}
    _CNAMES = {"p": "p", "n": "n", "dp_du": "dpdu_", "dp_dv": "dpdv_", "sh_n": "shn_", "sh_s": "shs_", "sh_t": "sht_", "wi": "wi_"}

    def __init__(self, n, device, shape=None, fields=None):
        fields = self.DEFAULT_FIELDS if fields is None else tuple(fields)
        unknown = set(fields) - set(self.ALL_FIELDS)
        if unknown:
            raise HeightfieldError(f"unknown surface interaction fields: {sorted(unknown)}")
        self.t = torch.empty(n, dtype=torch.float32, device=device) if "t" in fields else None
        for k, rows in self._ROWS.items():
            setattr(self, k, torch.empty((rows, n), dtype=torch.float32, device=device) if k in fields else None)
        self.boundary_test = torch.empty(n, dtype=torch.float32, device=device) if "boundary_test" in fields else None
        self.prim_index = None
        self.shape = shape

    def is_valid(self):
        return self.t != float("inf")

    def _c(self):
        f = {"t": self.t, "boundary_test": self.boundary_test}
        for key, cname in self._CNAMES.items():
            t = getattr(self, key)
            for k, ax in enumerate("xyz"):
                f[cname + ax] = t[k] if t is not None else None
        f["uvx"], f["uvy"] = (self.uv[0], self.uv[1]) if self.uv is not None else (None, None)
        return _SI(**{k: _ptr(f[k]) for k in SI_FIELDS})


def _adj_struct(adj, n, device):
    """adj: dict with any of t (N,), p (3,N), n (3,N), uv (2,N), dp_du (3,N), dp_dv (3,N)."""
    keep, f = [], {}
    def put(names, val, rows):
        t = torch.as_tensor(val, dtype=torch.float32, device=device)
        t = t.reshape(rows, n).contiguous() if rows > 1 else t.reshape(n).contiguous()
        keep.append(t)
        if rows == 1:
            f[names[0]] = t
        else:
            for k, nm in enumerate(names):
                f[nm] = t[k]
    for key, names in (("t", ("t",)), ("p", ("px", "py", "pz")), ("n", ("nx", "ny", "nz")), ("uv", ("uvx", "uvy")),
                       ("dp_du", ("dpdu_x", "dpdu_y", "dpdu_z")), ("dp_dv", ("dpdv_x", "dpdv_y", "dpdv_z"))):
        if key in adj and adj[key] is not None:
            put(names, adj[key], len(names))
    unknown = set(adj) - {"t", "p", "n", "uv", "dp_du", "dp_dv"}
    if unknown:
        raise HeightfieldError(f"unknown adjoint fields: {sorted(unknown)}")
    return _Adj(**{k: _ptr(f.get(k)) for k in ADJ_FIELDS}), keep


# ---- the shape ------------------------------------------------------------------------------
class Heightfield:
    """`heightfield` Shape plugin (definition: DESIGN.md).

    Properties (constructor dict, cf. Shape(props) src/render/shape.cpp:22-78 and
    src/textures/bitmap.cpp:124-143): `heights` ((H, W) array / tensor) XOR `filename` (.npy),
    `max_height` (default 1), `to_world` (3x4 or 4x4, default identity), `flip_normals` (default False).
    """

    def __init__(self, props: dict, device=None):
        props = dict(props)
        if props.pop("type", "heightfield") != "heightfield":
            raise HeightfieldError("this plugin only implements type='heightfield'")
        if ("heights" in props) == ("filename" in props):
            raise HeightfieldError("heightfield: specify exactly one of `heights` or `filename`")
        if not torch.cuda.is_available():
            raise HeightfieldError("no CUDA device: the heightfield plugin has no CPU fallback")
        L = _load()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        h = props.pop("heights", None)
        if h is None:
            from .formats import load_heights          # .npy / .pfm / .pgm (formats.py)
            h = load_heights(props.pop("filename"))
        h = torch.as_tensor(h, dtype=torch.float32)
        if h.dim() == 3 and h.shape[2] == 1:      # (H, W, 1) tensors as in bitmap.cpp:259-263
            h = h[:, :, 0]
        if h.dim() != 2:
            raise HeightfieldError("heightfield: `heights` must have shape (H, W)")
        self.H, self.W = int(h.shape[0]), int(h.shape[1])
        tw = props.pop("to_world", None)
        tw = np.eye(4, dtype=np.float32) if tw is None else np.asarray(
            tw.cpu() if isinstance(tw, torch.Tensor) else tw, dtype=np.float32)
        tw = tw.reshape(-1)[:12].reshape(3, 4)
        self._to_world = tw.copy()
        self._max_height = float(props.pop("max_height", 1.0))
        self.flip_normals = bool(props.pop("flip_normals", False))
        to_object = props.pop("to_object", None)
        props.pop("bsdf", None)            # children of the shape that this package does not model (shape.cpp:27-78)
        props.pop("id", None)
        if props:
            raise HeightfieldError(f"heightfield: unreferenced properties {sorted(props)}")   # Properties::mark_queried semantics

        d = _Desc()
        d.width, d.height, d.max_height = self.W, self.H, self._max_height
        d.to_world[:] = tw.reshape(-1).tolist()
        if to_object is not None:
            d.to_object[:] = np.asarray(to_object, dtype=np.float32).reshape(-1)[:12].tolist()
            d.has_to_object = 1
        d.flip_normals, d.device = int(self.flip_normals), dev.index or 0
        self._ctx = C.c_void_p()
        _check(L.hf_create(C.byref(self._ctx), C.byref(d)))
        hp, gp = C.c_void_p(), C.c_void_p()
        _check(L.hf_heights(self._ctx, C.byref(hp)))
        _check(L.hf_grad(self._ctx, C.byref(gp)))
        with torch.cuda.device(dev):
            self.heights = torch.as_tensor(_DevView(hp.value, (self.H, self.W), self), device=dev)
            self.grad = torch.as_tensor(_DevView(gp.value, (self.H, self.W), self), device=dev)
            self.heights.copy_(h.to(dev))
        self.parameters_changed()

    def __del__(self):
        try:
            if getattr(self, "_ctx", None) and _lib is not None:
                _lib.hf_destroy(self._ctx)
                self._ctx = None
        except Exception:
            pass

    # -- traverse()/parameters_changed (shape.h:533-550) ------------------------------------
    @property
    def max_height(self):
        return self._max_height

    @max_height.setter
    def max_height(self, v):
        self._max_height = float(v)
        _check(_load().hf_set_max_height(self._ctx, self._max_height))

    @property
    def to_world(self):
        return self._to_world

    @to_world.setter
    def to_world(self, m):
        m = np.asarray(m, dtype=np.float32).reshape(-1)[:12].reshape(3, 4).copy()
        arr = (C.c_float * 12)(*m.reshape(-1).tolist())
        _check(_load().hf_set_to_world(self._ctx, arr, None))
        self._to_world = m

    def traverse(self, callback=None):
        params = {"heights": self.heights, "max_height": self._max_height, "to_world": self._to_world}
        if callback is not None:
            for k, v in params.items():
                callback(k, v)
        return params

    def parameters_changed(self, keys=()):
        """Rebuild the acceleration data (min/max pyramid) after a parameter update."""
        with torch.cuda.device(self.device):
            _check(_load().hf_parameters_changed(self._ctx, _stream()))

    def parameters_grad_enabled(self):
        e = C.c_int()
        _check(_load().hf_parameters_grad_enabled(self._ctx, C.byref(e)))
        return bool(e.value)

    def set_grad_enabled(self, enabled=True):
        _check(_load().hf_set_grad_enabled(self._ctx, int(enabled)))

    def set_ray_coherence(self, coherent=True):
        """The `coherent` hint of Scene::ray_intersect (src/render/scene.cpp:141-150) for the ray-array entry points: False = the
        rays of a batch are unrelated (every lane of a warp keeps its own ray); results are bit-identical either way."""
        _check(_load().hf_set_ray_coherence(self._ctx, int(bool(coherent))))

    def set_adaptive_schedule(self, on=True):
        """include/hf_b200.h hf_set_adaptive_schedule: launches over batches of the same size (the rays or pixels of an optimisation's
        iterations) start with the units of work the previous launch found far more expensive than the mean. Results do not depend on it."""
        _check(_load().hf_set_adaptive_schedule(self._ctx, int(bool(on))))

    def adaptive_schedule_info(self, kind="rays"):
        """(listed, units): how many of the last launch's units of work (kind "rays": 32-ray groups of a ray batch; "pixels": pixels of a
        render step) the next launch will start with"""
        a, b = C.c_uint32(0), C.c_uint32(0)
        _check(_load().hf_adaptive_schedule_info(self._ctx, {"rays": 0, "pixels": 1}[kind], C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def zero_grad(self):
        with torch.cuda.device(self.device):
            _check(_load().hf_zero_grad(self._ctx, _stream()))

    # -- queries ----------------------------------------------------------------------------
    def bbox(self):
        out = (C.c_float * 6)()
        _check(_load().hf_bbox(self._ctx, out))
        return np.array(out[:], dtype=np.float32)

    def primitive_count(self):
        n = C.c_uint32()
        _check(_load().hf_primitive_count(self._ctx, C.byref(n)))
        return n.value

    def pyramid_levels(self):
        """Acceleration data as a list of (h, w, 2) (min,max) device tensors for levels 1..Ltop (zero-copy views)."""
        L = _load()
        nl = C.c_uint32()
        _check(L.hf_pyramid_level(self._ctx, 0, C.byref(nl), None, None, None))
        out = []
        for l in range(1, nl.value + 1):
            w, h, ptr = C.c_uint32(), C.c_uint32(), C.c_void_p()
            _check(L.hf_pyramid_level(self._ctx, l, None, C.byref(w), C.byref(h), C.byref(ptr)))
            with torch.cuda.device(self.device):
                out.append(torch.as_tensor(_DevView(ptr.value, (h.value, w.value, 2), self), device=self.device))
        return out

    def qpyramid_levels(self):
        """The quantised pyramid the production walk reads: ([(rows, pitch) int32 device tensors for levels 0..Ltop], qs, zb);
        node = (hi << 16) | lo, bound = fma(code, qs, zb) (csrc/hf_qpyr.h). Zero-copy views."""
        L = _load()
        nl, qs, zb = C.c_uint32(), C.c_float(), C.c_float()
        _check(L.hf_qpyramid_level(self._ctx, 0, C.byref(nl), None, None, None, C.byref(qs), C.byref(zb)))
        out = []
        for l in range(0, nl.value + 1):
            p, r, ptr = C.c_uint32(), C.c_uint32(), C.c_void_p()
            _check(L.hf_qpyramid_level(self._ctx, l, None, C.byref(p), C.byref(r), C.byref(ptr), None, None))
            view = _DevView(ptr.value, (r.value, p.value), self)
            view.__cuda_array_interface__["typestr"] = "<i4"
            with torch.cuda.device(self.device):
                out.append(torch.as_tensor(view, device=self.device))
        return out, qs.value, zb.value

    def _mask(self, active, n):
        if active is None:
            return None
        a = torch.as_tensor(active, device=self.device).to(torch.uint8).contiguous()
        if a.numel() != n:
            raise HeightfieldError("active mask has the wrong length")
        return a

    def ray_intersect_preliminary(self, ray: Ray3f, active=None, counted=False, out=None):
        n = len(ray)
        pi = out if out is not None else PreliminaryIntersection3f(n, self.device, self)
        a = self._mask(active, n)
        r, p = ray._c(a), pi._c()
        fn = _load().hf_ray_intersect_preliminary_counted if counted else _load().hf_ray_intersect_preliminary
        with torch.cuda.device(self.device):
            _check(fn(self._ctx, C.byref(r), C.byref(p), n, _stream()))
        return pi

    def ray_test(self, ray: Ray3f, active=None):
        n = len(ray)
        hit = torch.empty(n, dtype=torch.uint8, device=self.device)
        a = self._mask(active, n)
        r = ray._c(a)
        with torch.cuda.device(self.device):
            _check(_load().hf_ray_test(self._ctx, C.byref(r), _ptr(hit), n, _stream()))
        return hit.bool()

    def compute_surface_interaction(self, ray: Ray3f, pi: PreliminaryIntersection3f, ray_flags=RayFlags.All,
                                    recursion_depth=0, active=None, fields=None, out=None):
        n = len(ray)
        si = out if out is not None else SurfaceInteraction3f(n, self.device, self, fields)
        a = self._mask(active, n)
        r, p, s = ray._c(a), pi._c(), si._c()
        with torch.cuda.device(self.device):
            _check(_load().hf_compute_surface_interaction(self._ctx, C.byref(r), C.byref(p), int(ray_flags), C.byref(s), n, _stream()))
        si.prim_index = pi.prim_index
        return si

    def ray_intersect(self, ray: Ray3f, ray_flags=RayFlags.All, active=None, return_prelim=False, fields=None, out=None):
        """out = (si, pi) reuses preallocated records (no allocation on the hot path)."""
        n = len(ray)
        si, pi = out if out is not None else (SurfaceInteraction3f(n, self.device, self, fields),
                                              PreliminaryIntersection3f(n, self.device, self))
        a = self._mask(active, n)
        r, p, s = ray._c(a), pi._c(), si._c()
        with torch.cuda.device(self.device):
            _check(_load().hf_ray_intersect(self._ctx, C.byref(r), int(ray_flags), C.byref(p), C.byref(s), n, _stream()))
        si.prim_index = pi.prim_index
        return (si, pi) if return_prelim else si

    def backward(self, ray: Ray3f, pi: PreliminaryIntersection3f, adjoints, ray_flags=RayFlags.All, active=None):
        """Accumulate d(loss)/d(heights) into self.grad given upstream adjoints of si fields
        (reverse mode of compute_surface_interaction, src/render/mesh.cpp:695-789). `adjoints` is a dict
        (t (N,), p/n/dp_du/dp_dv (3,N), uv (2,N)) or the result of prepare_adjoints()."""
        n = len(ray)
        a = self._mask(active, n)
        adj, keep = adjoints if isinstance(adjoints, tuple) else _adj_struct(adjoints, n, self.device)
        r, p = ray._c(a), pi._c()
        with torch.cuda.device(self.device):
            _check(_load().hf_backward(self._ctx, C.byref(r), C.byref(p), C.byref(adj), int(ray_flags), n, _stream()))
        del keep

    def prepare_adjoints(self, adjoints: dict, n):
        """Validate / lay out an adjoint dict once so that backward() does no tensor work per call."""
        return _adj_struct(adjoints, n, self.device)

    def stats(self, reset=True):
        s = _Stats()
        _check(_load().hf_stats_get(self._ctx, C.byref(s), int(reset)))
        return {"rays": s.rays, "nodes": s.nodes, "leaves": s.leaves, "hits": s.hits, "segments": s.segments}

    # -- optimiser / multi-GPU --------------------------------------------------------------
    def adam_step(self, lr, beta_1=0.9, beta_2=0.999, epsilon=1e-8, mask_updates=False):
        cfg = _Adam(lr, beta_1, beta_2, epsilon, int(mask_updates))
        with torch.cuda.device(self.device):
            _check(_load().hf_adam_step(self._ctx, C.byref(cfg), _stream()))

    def adam_reset(self):
        with torch.cuda.device(self.device):
            _check(_load().hf_adam_reset(self._ctx, _stream()))

    def comm_init(self, rank, world_size, unique_id: bytes):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        with torch.cuda.device(self.device):
            _check(_load().hf_comm_init(self._ctx, rank, world_size, buf))
            # hf_comm_init may move the gradient grid into NCCL-registered memory: wrap the pointer again
            gp = C.c_void_p()
            _check(_load().hf_grad(self._ctx, C.byref(gp)))
            self.grad = torch.as_tensor(_DevView(gp.value, (self.H, self.W), self), device=self.device)

    def allreduce_grad(self):
        with torch.cuda.device(self.device):
            _check(_load().hf_allreduce_grad(self._ctx, _stream()))

    # -- multi-GPU over peer memory: scatters go to the texel's owner over NVLink (include/hf_b200.h "multi-GPU over peer memory") --
    PEER_HANDLE_BYTES = 256

    def _rewrap_grad(self):
        gp = C.c_void_p()
        _check(_load().hf_grad(self._ctx, C.byref(gp)))
        self.grad = torch.as_tensor(_DevView(gp.value, (self.H, self.W), self), device=self.device)

    def peer_export(self) -> bytes:
        """CUDA IPC handles of this rank's gradient grid, heights and barrier counter; ship them to every rank."""
        buf = (C.c_uint8 * self.PEER_HANDLE_BYTES)()
        with torch.cuda.device(self.device):
            _check(_load().hf_peer_export(self._ctx, buf))
            self._rewrap_grad()                # the grid leaves NCCL-registered memory if hf_comm_init had put it there
        return bytes(buf)

    def peer_attach(self, rank, world_size, handles):
        """handles: the peer_export() bytes of ranks 0 .. world_size - 1, in rank order."""
        blob = b"".join(handles)
        if len(blob) != world_size * self.PEER_HANDLE_BYTES:
            raise HeightfieldError("peer_attach: need one 256-byte handle block per rank")
        buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
        with torch.cuda.device(self.device):
            _check(_load().hf_peer_attach(self._ctx, rank, world_size, buf))
        self.peer_world = world_size

    def peer_detach(self):
        with torch.cuda.device(self.device):
            _check(_load().hf_peer_detach(self._ctx))
        self.peer_world = 1

    def peer_barrier(self):
        with torch.cuda.device(self.device):
            _check(_load().hf_peer_barrier(self._ctx, _stream()))

    def peer_zero_grad(self):
        with torch.cuda.device(self.device):
            _check(_load().hf_peer_zero_grad(self._ctx, _stream()))

    def peer_reduce_grad(self):
        """After the scatter kernels: every rank's gradient grid holds the sum over ranks (bit-identical everywhere)."""
        with torch.cuda.device(self.device):
            _check(_load().hf_peer_reduce_grad(self._ctx, _stream()))

    def peer_set_mode(self, mode):
        """1 (default): local scatter + reduce-scatter by peer stores; 0: scatters go to the owner with remote atomics."""
        _check(_load().hf_peer_set_mode(self._ctx, int(mode)))

    def peer_adam_step(self, lr, beta_1=0.9, beta_2=0.999, epsilon=1e-8, mask_updates=False):
        cfg = _Adam(lr, beta_1, beta_2, epsilon, int(mask_updates))
        with torch.cuda.device(self.device):
            _check(_load().hf_peer_adam_step(self._ctx, C.byref(cfg), _stream()))

    # -- render step on the heightfield alone (sensor + direct illumination + film + adjoint) ------
    def _pixels(self, cam, pixels):
        if pixels is None:
            return None, int(cam["width"]) * int(cam["height"])
        t = torch.as_tensor(pixels, device=self.device).to(torch.int32).contiguous()     # uint32 bit pattern
        return t, t.numel()

    def render_rays(self, cam: dict, pixels=None) -> "Ray3f":
        """The primary rays the render entry points generate (index = k * spp + sample for the k-th listed pixel)."""
        c = _camera_struct(cam)
        px, npix = self._pixels(cam, pixels)
        n = npix * c.spp
        ray = Ray3f(torch.empty((3, n), device=self.device), torch.empty((3, n), device=self.device),
                    torch.empty(n, device=self.device), device=self.device)
        r = ray._c()
        with torch.cuda.device(self.device):
            _check(_load().hf_render_rays(self._ctx, C.byref(c), _ptr(px), npix, C.byref(r), _stream()))
        return ray

    def render_forward(self, cam: dict, light: dict, pixels=None, image=None, return_prims=False):
        """Primal image (height, width) fp32; pixels not listed keep their previous value (zeros for a fresh image)."""
        c, l = _camera_struct(cam), _light_struct(light)
        px, npix = self._pixels(cam, pixels)
        if image is None:
            image = torch.zeros((c.height, c.width), dtype=torch.float32, device=self.device)
        prims = torch.empty(npix * c.spp, dtype=torch.int32, device=self.device) if return_prims else None
        with torch.cuda.device(self.device):
            _check(_load().hf_render_forward(self._ctx, C.byref(c), C.byref(l), _ptr(px), npix, _ptr(image), _ptr(prims), _stream()))
        return (image, prims) if return_prims else image

    def render_backward(self, cam: dict, light: dict, adj_image, pixels=None, prims=None):
        """self.grad += J^T adj_image (ADIntegrator.render_backward, common.py:784-970, for this one shape)."""
        c, l = _camera_struct(cam), _light_struct(light)
        px, npix = self._pixels(cam, pixels)
        adj = torch.as_tensor(adj_image, dtype=torch.float32, device=self.device).contiguous()
        if adj.numel() != c.width * c.height:
            raise HeightfieldError("adjoint image has the wrong size")
        with torch.cuda.device(self.device):
            _check(_load().hf_render_backward(self._ctx, C.byref(c), C.byref(l), _ptr(px), npix, _ptr(adj), _ptr(prims), _stream()))

    def render_step_l2(self, cam: dict, light: dict, target, pixels=None, image=None, loss=None):
        """Forward + L2 loss against `target` + backward in one kernel. `loss`: 1-element float64 device tensor that the
        call ADDS mean-squared-error contributions to (or None)."""
        c, l = _camera_struct(cam), _light_struct(light)
        px, npix = self._pixels(cam, pixels)
        if target.dtype != torch.float32 or not target.is_contiguous() or target.numel() != c.width * c.height:
            raise HeightfieldError("target must be a contiguous fp32 (height, width) device tensor")
        with torch.cuda.device(self.device):
            _check(_load().hf_render_step_l2(self._ctx, C.byref(c), C.byref(l), _ptr(px), npix, _ptr(target), _ptr(image), _ptr(loss), _stream()))
        return image

    # -- host-buffer path (llvm_ad_rgb keeps its arrays in host memory) ------------------------
    def forward_backward_host(self, h_rays: dict, h_adj: dict, h_si: dict, h_grad, n, ray_flags=RayFlags.All,
                              h_prelim: dict | None = None):
        """All arguments are dicts of 1-D HOST torch tensors (pinned for full PCIe speed), keyed by the C field
        names of hf_rays / hf_si_adj / hf_si / hf_prelim. h_grad: (H, W) host tensor or None."""
        r = _Rays(**{k: _ptr(h_rays.get(k)) for k in ("ox", "oy", "oz", "dx", "dy", "dz", "maxt", "active")})
        a = _Adj(**{k: _ptr(h_adj.get(k)) for k in ADJ_FIELDS})
        s = _SI(**{k: _ptr(h_si.get(k)) for k in SI_FIELDS})
        p = _Prelim(**{k: _ptr(h_prelim.get(k)) for k in ("t", "u", "v", "prim_index")}) if h_prelim else None
        with torch.cuda.device(self.device):
            _check(_load().hf_forward_backward_host(self._ctx, C.byref(r), C.byref(a), int(ray_flags),
                                                    C.byref(p) if p is not None else None, C.byref(s), _ptr(h_grad), n))

    def ray_intersect_host(self, h_rays: dict, h_si: dict, n, ray_flags=RayFlags.All, h_prelim: dict | None = None):
        r = _Rays(**{k: _ptr(h_rays.get(k)) for k in ("ox", "oy", "oz", "dx", "dy", "dz", "maxt", "active")})
        s = _SI(**{k: _ptr(h_si.get(k)) for k in SI_FIELDS})
        p = _Prelim(**{k: _ptr(h_prelim.get(k)) for k in ("t", "u", "v", "prim_index")}) if h_prelim else None
        with torch.cuda.device(self.device):
            _check(_load().hf_ray_intersect_host(self._ctx, C.byref(r), int(ray_flags),
                                                 C.byref(p) if p is not None else None, C.byref(s), n))

    def render_step_l2_host(self, cam: dict, light: dict, h_target, h_heights=None, h_image=None, h_grad=None, h_pixels=None):
        """The optimisation step with HOST tensors (pinned for full PCIe speed): heights (optional) + target in, image / gradient
        grid (optional) + loss out. Returns the loss (float)."""
        c, l = _camera_struct(cam), _light_struct(light)
        npix = c.width * c.height if h_pixels is None else h_pixels.numel()
        loss = C.c_double(0.0)
        with torch.cuda.device(self.device):
            _check(_load().hf_render_step_l2_host(self._ctx, C.byref(c), C.byref(l), _ptr(h_pixels), npix, _ptr(h_heights), _ptr(h_target),
                                                  _ptr(h_image), _ptr(h_grad), C.byref(loss)))
        return loss.value

    def __repr__(self):
        return (f"Heightfield[\n  size = {self.W}x{self.H},\n  max_height = {self._max_height},\n"
                f"  primitive_count = {self.primitive_count()},\n  flip_normals = {self.flip_normals}\n]")


STATIC_SCRATCH_BYTES = 9 << 20


def ray_intersect_preliminary_static(heights, ray: Ray3f, max_height=1.0, to_world=None, flip_normals=False, active=None, out=None):
    """`Heightfield.ray_intersect_preliminary` for a caller that may not allocate device memory: no shape object, no context.

    `heights` is the caller's own (H, W) float32 CUDA tensor; the library reads it in place and keeps its acceleration data in a
    fixed 9 MiB array that is part of the loaded library (include/hf_b200.h, hf_ray_intersect_preliminary_static), so the C call
    itself allocates nothing. Fields whose pyramids do not fit (beyond 1025 x 1025 texels) raise HeightfieldError. Pass `out`
    (a PreliminaryIntersection3f made ahead of time) to keep this wrapper from allocating the result through torch as well. Calls
    share the array: keep them on one stream, or order them with events. Same results, bit for bit, as the shape's method."""
    if not (isinstance(heights, torch.Tensor) and heights.is_cuda and heights.dtype == torch.float32 and heights.dim() == 2
            and heights.is_contiguous()):
        raise HeightfieldError("ray_intersect_preliminary_static: `heights` must be a contiguous (H, W) float32 CUDA tensor")
    L = _load()
    d = _Desc()
    d.height, d.width, d.max_height = int(heights.shape[0]), int(heights.shape[1]), float(max_height)
    tw = np.eye(4, dtype=np.float32) if to_world is None else np.asarray(
        to_world.cpu() if isinstance(to_world, torch.Tensor) else to_world, dtype=np.float32)
    d.to_world[:] = tw.reshape(-1)[:12].tolist()
    d.flip_normals, d.device = int(bool(flip_normals)), heights.device.index or 0
    n = len(ray)
    pi = out if out is not None else PreliminaryIntersection3f(n, heights.device)
    m = None
    if active is not None:
        m = torch.as_tensor(active, device=heights.device).to(torch.uint8).contiguous()
    with torch.cuda.device(heights.device):
        r, p = ray._c(m), pi._c()
        _check(L.hf_ray_intersect_preliminary_static(C.byref(d), _ptr(heights), C.byref(r), C.byref(p), n, _stream()))
    return pi


def perspective_camera(origin, target, up, fov, width, height, spp, seed=0) -> dict:
    """`perspective` sensor with to_world = look_at(origin, target, up) and a horizontal field of view `fov` in degrees
    (src/sensors/perspective.cpp:63-150): the orthonormal frame is computed in fp64 and rounded to fp32."""
    o, t, u = (np.asarray(x, np.float64) for x in (origin, target, up))
    f = t - o
    f /= np.linalg.norm(f)
    r = np.cross(f, u)
    r /= np.linalg.norm(r)
    upv = np.cross(r, f)
    f32 = lambda v: [float(np.float32(x)) for x in v]
    return {"origin": f32(o), "right": f32(r), "up": f32(upv), "forward": f32(f),
            "tan_half_fov": float(np.float32(np.tan(np.radians(fov) * 0.5))), "width": int(width), "height": int(height),
            "spp": int(spp), "seed": int(seed)}


def directional_light(towards, irradiance=1.0, albedo=1.0, shadow_rays=False, max_depth=1) -> dict:
    """`directional` emitter seen from the surface (unit vector towards the emitter) + the `diffuse` BSDF's reflectance."""
    d = np.asarray(towards, np.float64)
    d /= np.linalg.norm(d)
    return {"direction": [float(np.float32(x)) for x in d], "irradiance": float(irradiance), "albedo": float(albedo),
            "shadow_rays": bool(shadow_rays), "max_depth": int(max_depth)}


def _camera_struct(cam: dict) -> _Camera:
    c = _Camera()
    for k in ("origin", "right", "up", "forward"):
        getattr(c, k)[:] = [float(x) for x in cam[k]]
    c.tan_half_fov = float(cam["tan_half_fov"])
    c.width, c.height, c.spp, c.seed = int(cam["width"]), int(cam["height"]), int(cam["spp"]), int(cam.get("seed", 0))
    return c


def _light_struct(light: dict) -> _Light:
    l = _Light()
    l.direction[:] = [float(x) for x in light["direction"]]
    l.irradiance, l.albedo = float(light.get("irradiance", 1.0)), float(light.get("albedo", 1.0))
    l.shadow_rays = int(bool(light.get("shadow_rays", False)))
    l.max_depth = int(light.get("max_depth", 1))
    return l


def comm_unique_id() -> bytes:
    buf = (C.c_uint8 * 128)()
    _check(_load().hf_comm_unique_id(buf))
    return bytes(buf)


def load_dict(d: dict, device=None):
    """mi.load_dict for the one plugin this package implements (src/core/plugin.cpp:174-200)."""
    if d.get("type") != "heightfield":
        raise HeightfieldError(f"plugin '{d.get('type')}' is not provided by this package (only 'heightfield')")
    return Heightfield(d, device=device)


class SceneParameters(dict):
    """src/python/python/util.py:185-232: dictionary of exposed parameters + update()."""

    def __init__(self, shape: Heightfield, prefix=""):
        super().__init__()
        self._shape, self._prefix, self._dirty = shape, prefix, set()
        for k, v in shape.traverse().items():
            dict.__setitem__(self, prefix + k, v)

    def __setitem__(self, key, value):
        if key not in self:
            raise KeyError(key)
        base = key[len(self._prefix):]
        if base == "heights":
            value = torch.as_tensor(value, dtype=torch.float32, device=self._shape.device)
            if tuple(value.shape) != (self._shape.H, self._shape.W):     # cf. bitmap.cpp:272-287
                raise HeightfieldError("heights: size mismatch on update")
            self._shape.heights.copy_(value)
            value = self._shape.heights
        elif base == "max_height":
            self._shape.max_height = value
        elif base == "to_world":
            self._shape.to_world = value
            value = self._shape.to_world
        dict.__setitem__(self, key, value)
        self._dirty.add(base)

    def set_dirty(self, key):
        self._dirty.add(key[len(self._prefix):])

    def update(self, values=None):
        if values:
            for k, v in dict(values).items():
                self[k] = v
        keys = sorted(self._dirty)
        self._dirty.clear()
        if keys:
            self._shape.parameters_changed(keys)
        return keys


def traverse(shape: Heightfield, prefix=""):
    return SceneParameters(shape, prefix)


class Adam:
    """src/python/python/ad/optimizers.py:203-309 for the heightfield's `heights` parameter. The update runs
    in one fused kernel on the library's own buffers (state m, v live next to the heights)."""

    def __init__(self, lr, beta_1=0.9, beta_2=0.999, epsilon=1e-8, mask_updates=False, params: SceneParameters = None):
        assert 0 <= beta_1 < 1 and 0 <= beta_2 < 1 and lr > 0 and epsilon > 0
        self.lr, self.beta_1, self.beta_2, self.epsilon, self.mask_updates = lr, beta_1, beta_2, epsilon, mask_updates
        self.params = params
        if params is not None:
            params._shape.adam_reset()

    def step(self, peer=False):
        """peer=True (after Heightfield.peer_attach): reduce-scatter of the gradient, Adam on this rank's slab, updated heights stored
        into every rank's height grid (barriers included); params.update() next."""
        s = self.params._shape
        if peer:
            s.peer_adam_step(self.lr, self.beta_1, self.beta_2, self.epsilon, self.mask_updates)
        else:
            s.adam_step(self.lr, self.beta_1, self.beta_2, self.epsilon, self.mask_updates)
        self.params.set_dirty(self.params._prefix + "heights")

    def reset(self, key=None):
        self.params._shape.adam_reset()
