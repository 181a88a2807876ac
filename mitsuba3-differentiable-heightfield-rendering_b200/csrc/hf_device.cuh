// hf_device.cuh -- device-side arithmetic of the heightfield path, in the reference's op order.
//
// Parity target is the reference's llvm_ad_rgb lowering: no a*b+c
// contraction, single-rounding fma only where the reference source says fmadd/fmsub/fnmadd,
// IEEE division for rcp, sqrt(1/x) for rsqrt, denormals preserved. This translation unit is
// therefore compiled with -fmad=false -prec-div=true -prec-sqrt=true -ftz=false, and every fma
// below is an explicit fmaf().
//
// Citations are into /root/reference.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "hf_divmagic.h"

#define HF_MAX_LEVELS 26
#define HF_MAX_PEERS 8            // GPUs of one NVSwitch box
#define HF_INF __int_as_float(0x7f800000)

// 2^-20: relative slack on the ray's z interval during traversal (never culls a node that
// contains a hit because of rounding in the traversal arithmetic)
#define HF_Z_SLACK 9.5367431640625e-07f

struct HfParams {
    int W, H, NX, NY, Ltop;
    int flip_normals;
    float mh;                 // max_height
    float rw, rh;             // 1/(W-1), 1/(H-1)
    float x0, y0;             // X(0), Y(0)
    float sgx, sgy;           // object -> grid units: 0.5*(W-1), 0.5*(H-1)
    float to_world[12];       // row-major 3x4
    float to_object[12];
    const float *heights;     // (H, W) row-major
    const float2 *pyr;        // levels 1..Ltop, (min,max) of z = mh*h
    int lvl_off[HF_MAX_LEVELS];   // node offset of level l in pyr
    int lvl_w[HF_MAX_LEVELS];     // nodes per row at level l
    // quantised, power-of-two padded pyramid with a level 0 (hf_qpyr.h): what the production walk reads
    const uint32_t *qpyr;
    uint32_t qoff[HF_MAX_LEVELS]; // node offset of level l in qpyr; pitch of level l = max((xmask + 1) >> l, 1)
    int QLtop, xmask, ymask;      // top level; padded cell counts minus one
    uint32_t nx_magic;            // cell / NX without a divide (hf_divmagic.h); prim_index < 2^31 is checked by hf_create
    int32_t nx_shift;
    // Multi-GPU peer scatter (hf_peer_attach): the gradient grid is cut into peer_n slabs of slab_elems texels,
    // slab r is OWNED by rank r, and grad_peer[r] is rank r's grid mapped into this process over NVLink (CUDA IPC). A scatter then
    // goes straight to the owner (red.global.add.f32 on peer memory): the reduce-scatter of the gradient happens inside the kernels
    // that produce it, texel by texel, instead of a 64 MiB collective behind them. peer_n <= 1: the local grid only.
    float *grad_peer[HF_MAX_PEERS];
    uint32_t slab_magic;
    int32_t slab_shift;
    int32_t peer_n;
};

// gradient scatter of one value: the texel's owner takes it (see HfParams::grad_peer)
__device__ __forceinline__ void hf_grad_add(const HfParams &P, float *__restrict__ grad, uint32_t idx, float val) {
#ifndef HF_NO_PEER_ATOMICS          // dev builds: measure what the owner lookup costs the single-GPU kernels
    if (P.peer_n > 1) {
        // several GPUs add to the same texel of the owner's grid over NVLink: only system scope makes that atomic across devices
        // (atomicAdd is device scope, red.gpu)
        atomicAdd_system(P.grad_peer[hf_divide(idx, P.slab_magic, P.slab_shift)] + idx, val);
        return;
    }
#endif
    atomicAdd(grad + idx, val);
}

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 neg3(f3 a) { return mk3(-a.x, -a.y, -a.z); }

// ext/drjit/include/drjit/array_base.h:603-631
__device__ __forceinline__ float dot3(f3 a, f3 b) {
    float r = a.x * b.x;
    r = fmaf(a.y, b.y, r);
    r = fmaf(a.z, b.z, r);
    return r;
}
// ext/drjit/include/drjit/array_router.h:648-661: fmsub(a.yzx, b.zxy, a.zxy * b.yzx)
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
    return mk3(fmaf(a.y, b.z, -(a.z * b.y)),
               fmaf(a.z, b.x, -(a.x * b.z)),
               fmaf(a.x, b.y, -(a.y * b.x)));
}
// array_router.h:629-638
__device__ __forceinline__ float sqnorm3(f3 v) {
    float r = v.x * v.x;
    r = fmaf(v.y, v.y, r);
    r = fmaf(v.z, v.z, r);
    return r;
}
// ext/drjit/ext/drjit-core/src/op.cpp:1194-1203, 1231-1236
__device__ __forceinline__ float rcp_llvm(float x) { return __fdiv_rn(1.0f, x); }
__device__ __forceinline__ float rsqrt_llvm(float x) { return __fsqrt_rn(__fdiv_rn(1.0f, x)); }
// array_router.h:644-646
__device__ __forceinline__ f3 normalize3(f3 v) {
    float s = rsqrt_llvm(sqnorm3(v));
    return mk3(v.x * s, v.y * s, v.z * s);
}
// include/mitsuba/core/transform.h:104-111
__device__ __forceinline__ f3 xform_point(const float *m, f3 p) {
    float r0 = m[3], r1 = m[7], r2 = m[11];
    r0 = fmaf(m[0], p.x, r0); r1 = fmaf(m[4], p.x, r1); r2 = fmaf(m[8],  p.x, r2);
    r0 = fmaf(m[1], p.y, r0); r1 = fmaf(m[5], p.y, r1); r2 = fmaf(m[9],  p.y, r2);
    r0 = fmaf(m[2], p.z, r0); r1 = fmaf(m[6], p.z, r1); r2 = fmaf(m[10], p.z, r2);
    return mk3(r0, r1, r2);
}
// transform.h:129-139
__device__ __forceinline__ f3 xform_vector(const float *m, f3 v) {
    float r0 = m[0] * v.x, r1 = m[4] * v.x, r2 = m[8] * v.x;
    r0 = fmaf(m[1], v.y, r0); r1 = fmaf(m[5], v.y, r1); r2 = fmaf(m[9],  v.y, r2);
    r0 = fmaf(m[2], v.z, r0); r1 = fmaf(m[6], v.z, r1); r2 = fmaf(m[10], v.z, r2);
    return mk3(r0, r1, r2);
}
// transform.h:141-153: columns of the inverse transpose == rows of to_object
__device__ __forceinline__ f3 xform_normal(const float *to, f3 n) {
    float r0 = to[0] * n.x, r1 = to[1] * n.x, r2 = to[2] * n.x;
    r0 = fmaf(to[4], n.y, r0); r1 = fmaf(to[5], n.y, r1); r2 = fmaf(to[6],  n.y, r2);
    r0 = fmaf(to[8], n.z, r0); r1 = fmaf(to[9], n.z, r1); r2 = fmaf(to[10], n.z, r2);
    return mk3(r0, r1, r2);
}
// y = M_lin^T x (adjoint of xform_vector; plain fp32, not parity-critical)
__device__ __forceinline__ f3 lin_T(const float *m, f3 x) {
    return mk3(m[0] * x.x + m[4] * x.y + m[8] * x.z,
               m[1] * x.x + m[5] * x.y + m[9] * x.z,
               m[2] * x.x + m[6] * x.y + m[10] * x.z);
}
__device__ __forceinline__ f3 lin(const float *m, f3 x) {
    return mk3(m[0] * x.x + m[1] * x.y + m[2] * x.z,
               m[4] * x.x + m[5] * x.y + m[6] * x.z,
               m[8] * x.x + m[9] * x.y + m[10] * x.z);
}

// include/mitsuba/core/vector.h:116-136
__device__ __forceinline__ float mulsign_(float a, float b) {
    return __int_as_float(__float_as_int(a) ^ (__float_as_int(b) & 0x80000000));
}
__device__ __forceinline__ void coordinate_system(f3 n, f3 &s, f3 &t) {
    float sign = copysignf(1.0f, n.z);
    float a = -rcp_llvm(sign + n.z);
    float b = n.x * n.y * a;
    s = mk3(mulsign_((n.x * n.x) * a, n.z) + 1.0f, mulsign_(b, n.z), mulsign_(-n.x, n.z));
    t = mk3(b, fmaf(n.y, n.y * a, sign), -n.y);
}

// include/mitsuba/render/mesh.h:357-380
__device__ __forceinline__ bool moeller_trumbore(f3 o, f3 d, float maxt, f3 p0, f3 p1, f3 p2,
                                                 float &t, float &u, float &v) {
    f3 e1 = sub3(p1, p0), e2 = sub3(p2, p0);
    f3 pvec = cross3(d, e2);
    float inv_det = rcp_llvm(dot3(e1, pvec));
    f3 tvec = sub3(o, p0);
    u = dot3(tvec, pvec) * inv_det;
    bool active = (u >= 0.f) && (u <= 1.f);
    f3 qvec = cross3(tvec, e1);
    v = dot3(d, qvec) * inv_det;
    active = active && (v >= 0.f) && (u + v <= 1.f);
    t = dot3(e2, qvec) * inv_det;
    active = active && (t >= 0.f) && (t <= maxt);
    return active;
}

// Heightfield definition: X(i) = float(2i-(W-1)) * (1/float(W-1))
__device__ __forceinline__ float vtx(int i, int W, float rw) { return (float) (2 * i - (W - 1)) * rw; }
// texcoord of a vertex: fmadd(x, 0.5, 0.5) (src/shapes/rectangle.cpp:312-313)
__device__ __forceinline__ float tex(float x) { return fmaf(x, 0.5f, 0.5f); }

// Both triangles of cell (cx,cy) given its four corner z values; closest wins, tie -> smaller index.
// Triangle 0 = (p10, p01, p00), triangle 1 = (p10, p11, p01): diagonal and winding of
// resources/data/common/meshes/rectangle.obj faces "2 3 1" / "2 4 3".
__device__ __forceinline__ bool cell_intersect(const HfParams &P, int cx, int cy,
                                               float z00, float z10, float z01, float z11,
                                               f3 o, f3 d, float maxt,
                                               float &t, float &u, float &v, uint32_t &prim) {
    float x0 = vtx(cx, P.W, P.rw), x1 = vtx(cx + 1, P.W, P.rw);
    float y0 = vtx(cy, P.H, P.rh), y1 = vtx(cy + 1, P.H, P.rh);
    f3 p00 = mk3(x0, y0, z00), p10 = mk3(x1, y0, z10), p01 = mk3(x0, y1, z01), p11 = mk3(x1, y1, z11);
    float t0, u0, v0, t1, u1, v1;
    bool h0 = moeller_trumbore(o, d, maxt, p10, p01, p00, t0, u0, v0);
    bool h1 = moeller_trumbore(o, d, maxt, p10, p11, p01, t1, u1, v1);
    uint32_t base = 2u * ((uint32_t) cy * (uint32_t) P.NX + (uint32_t) cx);
    if (h1 && (!h0 || t1 < t0)) { t = t1; u = u1; v = v1; prim = base + 1u; return true; }
    if (h0) { t = t0; u = u0; v = v0; prim = base; return true; }
    return false;
}

// Vertices of triangle k of cell (cx,cy) + the texel index of each vertex
__device__ __forceinline__ void cell_triangle(const HfParams &P, uint32_t prim, f3 &p0, f3 &p1, f3 &p2,
                                              uint32_t &i0, uint32_t &i1, uint32_t &i2) {
    uint32_t cell = prim >> 1, k = prim & 1u;
    uint32_t cy = hf_divide(cell, P.nx_magic, P.nx_shift), cx = cell - cy * (uint32_t) P.NX;
#ifdef HF_DEBUG_BOUNDS
    if (cy >= (uint32_t) P.NY) __trap();       // checked build: a primitive index that is not a triangle of this field
#endif
    float x0 = vtx((int) cx, P.W, P.rw), x1 = vtx((int) cx + 1, P.W, P.rw);
    float y0 = vtx((int) cy, P.H, P.rh), y1 = vtx((int) cy + 1, P.H, P.rh);
    uint32_t b = cy * (uint32_t) P.W + cx;
    const float *h = P.heights;
    i0 = b + 1u;                                  // p10
    p0 = mk3(x1, y0, P.mh * __ldg(h + i0));
    if (k == 0) {
        i1 = b + (uint32_t) P.W; i2 = b;          // p01, p00
        p1 = mk3(x0, y1, P.mh * __ldg(h + i1));
        p2 = mk3(x0, y0, P.mh * __ldg(h + i2));
    } else {
        i1 = b + (uint32_t) P.W + 1u; i2 = b + (uint32_t) P.W;   // p11, p01
        p1 = mk3(x1, y1, P.mh * __ldg(h + i1));
        p2 = mk3(x0, y1, P.mh * __ldg(h + i2));
    }
}
