// drjit_ref.cpp -- the reference's OWN array arithmetic, compiled from the Dr.Jit headers where
// they lie under /root/reference/ext/drjit (header-only scalar mode; nothing is copied into this
// repo). Built by oracle/Makefile into oracle/_ref/libdrjit_ref.so. TEST INFRASTRUCTURE ONLY.
//
// It pins the op order of the restatement in hf_oracle.c (dot / cross / fmadd / squared_norm /
// transform_affine) bit-for-bit. rcp and rsqrt are written as 1/x and sqrt(1/x): that is the
// llvm_ad_rgb lowering (ext/drjit/ext/drjit-core/src/op.cpp:1194-1203, 1231-1236), not what
// dr::rcp does on a scalar/packet array.
//
// The vector type is dr::Array<dr::Packet<float, 4>, 3> -- an array OF arrays, exactly like the
// JIT variants' Array<LLVMArray<float>, 3> and like the plugin's `_packet` Embree callbacks
// (include/mitsuba/render/shape.h:594-617). That matters: for an array of arrays dot() is
// mul + fmadd + fmadd (array_base.h:618-622), whereas the flat scalar Array<float, 3> takes the
// SSE4.2 _mm_dp_ps path (packet_sse42.h:1675-1677), which rounds differently and is not what
// llvm_ad_rgb computes.
//
// The function body of ref_moeller_trumbore follows include/mitsuba/render/mesh.h:357-380 with
// Dr.Jit arrays; mesh.h itself needs the full Mitsuba build (config.h, variants) and cannot be
// included on its own.
#include <drjit/array.h>
#include <drjit/packet.h>
#include <cmath>
#include <cstddef>
#include <cstdint>

namespace dr = drjit;
using FP = dr::Packet<float, 4>;
using V3 = dr::Array<FP, 3>;

static inline V3 mk(float a, float b, float c) { V3 r; r[0] = FP(a); r[1] = FP(b); r[2] = FP(c); return r; }
static inline V3 ld(const float *p) { return mk(p[0], p[1], p[2]); }

extern "C" {

int ref_moeller_trumbore(const float *o_, const float *d_, float maxt,
                         const float *p0_, const float *p1_, const float *p2_,
                         float *t_, float *u_, float *v_) {
    V3 o = ld(o_), d = ld(d_), p0 = ld(p0_), p1 = ld(p1_), p2 = ld(p2_);
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 pvec = dr::cross(d, e2);
    FP inv_det = FP(1.0f) / dr::dot(e1, pvec);
    V3 tvec = o - p0;
    FP u = dr::dot(tvec, pvec) * inv_det;
    auto active = u >= 0.f && u <= 1.f;
    V3 qvec = dr::cross(tvec, e1);
    FP v = dr::dot(d, qvec) * inv_det;
    active &= v >= 0.f && u + v <= 1.f;
    FP t = dr::dot(e2, qvec) * inv_det;
    active &= t >= 0.f && t <= maxt;
    *t_ = t[0]; *u_ = u[0]; *v_ = v[0];
    return active[0] ? 1 : 0;
}

// src/render/mesh.cpp:738-757: p = fmadd(p0,b0, fmadd(p1,b1, p2*b2)); n = normalize(cross(dp0,dp1))
void ref_interpolate_and_normal(const float *p0_, const float *p1_, const float *p2_, float b1, float b2,
                                float *p_out, float *n_out) {
    V3 p0 = ld(p0_), p1 = ld(p1_), p2 = ld(p2_);
    FP b0 = FP(1.f) - FP(b1) - FP(b2);
    V3 dp0 = p1 - p0, dp1 = p2 - p0;
    V3 p = dr::fmadd(p0, b0, dr::fmadd(p1, FP(b1), p2 * FP(b2)));
    V3 c = dr::cross(dp0, dp1);
    V3 n = c * dr::sqrt(FP(1.0f) / dr::squared_norm(c));
    for (int i = 0; i < 3; ++i) { p_out[i] = p[i][0]; n_out[i] = n[i][0]; }
}

// include/mitsuba/core/transform.h:104-111 with a row-major 3x4 matrix m
void ref_transform_affine_point(const float *m, const float *p_, float *out) {
    V3 col0 = mk(m[0], m[4], m[8]), col1 = mk(m[1], m[5], m[9]), col2 = mk(m[2], m[6], m[10]), col3 = mk(m[3], m[7], m[11]);
    V3 r = col3;
    r = dr::fmadd(col0, FP(p_[0]), r);
    r = dr::fmadd(col1, FP(p_[1]), r);
    r = dr::fmadd(col2, FP(p_[2]), r);
    for (int i = 0; i < 3; ++i) out[i] = r[i][0];
}

// include/mitsuba/core/transform.h:129-139
void ref_transform_vector(const float *m, const float *v_, float *out) {
    V3 col0 = mk(m[0], m[4], m[8]), col1 = mk(m[1], m[5], m[9]), col2 = mk(m[2], m[6], m[10]);
    V3 r = col0;
    r *= FP(v_[0]);
    r = dr::fmadd(col1, FP(v_[1]), r);
    r = dr::fmadd(col2, FP(v_[2]), r);
    for (int i = 0; i < 3; ++i) out[i] = r[i][0];
}

// Whole-heightfield brute force THROUGH the arithmetic above: the mesh of DESIGN.md section 2 (vertex
// P(i,j) = (X(i), Y(j), max_height h[j W + i]), cell (cx,cy) -> triangle 0 = (p10, p01, p00), triangle 1 = (p10, p11, p01), the
// "procedural mesh" pattern of src/shapes/cube.cpp:106-163) is enumerated triangle by triangle, the ray is taken to object space by
// transform_affine (src/shapes/rectangle.cpp:206-214, t preserved) and every triangle goes through the Moeller-Trumbore sequence of
// include/mitsuba/render/mesh.h:357-380 in Dr.Jit arrays. Closest hit, tie: smaller prim_index. This ties hit mask / prim_index /
// t / uv of whole heightfields to reference-compiled code rather than to the C restatement (the heightfield LAYOUT stays this
// repo's definition: the plugin is absent from the snapshot).
void ref_heightfield_brute_force(const float *h, int W, int H, float max_height, const float *to_object,
                                 const float *o_, const float *d_, const float *maxt, size_t n,
                                 float *t_out, float *u_out, float *v_out, uint32_t *prim_out) {
    const float rw = 1.0f / (float) (W - 1), rh = 1.0f / (float) (H - 1);
    for (size_t r = 0; r < n; ++r) {
        float oo[3], dd[3];
        ref_transform_affine_point(to_object, o_ + 3 * r, oo);
        ref_transform_vector(to_object, d_ + 3 * r, dd);
        float best = INFINITY, bu = 0.f, bv = 0.f; uint32_t bp = 0;
        for (int cy = 0; cy < H - 1; ++cy)
            for (int cx = 0; cx < W - 1; ++cx) {
                const float x0 = (float) (2 * cx - (W - 1)) * rw, x1 = (float) (2 * (cx + 1) - (W - 1)) * rw;
                const float y0 = (float) (2 * cy - (H - 1)) * rh, y1 = (float) (2 * (cy + 1) - (H - 1)) * rh;
                const float p00[3] = { x0, y0, max_height * h[(size_t) cy * W + cx] }, p10[3] = { x1, y0, max_height * h[(size_t) cy * W + cx + 1] };
                const float p01[3] = { x0, y1, max_height * h[(size_t) (cy + 1) * W + cx] }, p11[3] = { x1, y1, max_height * h[(size_t) (cy + 1) * W + cx + 1] };
                float t, u, v;
                const uint32_t base = 2u * ((uint32_t) cy * (uint32_t) (W - 1) + (uint32_t) cx);
                if (ref_moeller_trumbore(oo, dd, maxt[r], p10, p01, p00, &t, &u, &v) && t < best) { best = t; bu = u; bv = v; bp = base; }
                if (ref_moeller_trumbore(oo, dd, maxt[r], p10, p11, p01, &t, &u, &v) && t < best) { best = t; bu = u; bv = v; bp = base + 1u; }
            }
        t_out[r] = best; u_out[r] = bu; v_out[r] = bv; prim_out[r] = bp;
    }
}

}  // extern "C"
