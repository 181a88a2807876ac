// hf_backward.cu -- reverse mode of compute_surface_interaction w.r.t. the height texels.
//
// Replaces, for this shape, the reverse-mode sweep the reference's AD tape runs over src/render/mesh.cpp:695-789 (barycentric
// interpolation, face normal, uv, dp_du / dp_dv, with t and prim_uv re-derived by Moeller-Trumbore, mesh.cpp:728-735) and the
// scatter_reduce(Add) that is the adjoint of the vertex gather (ext/drjit/src/autodiff/autodiff.cpp:1309-1328; the reference JIT's
// own intrinsic is reduce_add_f32, ext/drjit/ext/drjit-core/src/cuda_eval.cpp:805-888).
//
// Closed form: the hit stays on the ray (p = o + t d), so the adjoint of p acts through t; the adjoint of uv reaches the heights
// only through the Moeller-Trumbore barycentrics; the adjoints of n, dp_du, dp_dv act through the triangle's edges. Only the z of
// a vertex depends on a texel, so each hit ray contributes three values, one per vertex of its triangle.
//
// This translation unit is compiled with fused multiply-add contraction (csrc/Makefile; division and square root stay IEEE): its
// results are compared with the oracle's fp64 accumulation at 1e-4 relative (the north star's gradient tolerance, per texel and in
// L2), not bit for bit -- the reference's own scatter order is nondeterministic. The forward kernels keep the reference's op order
// (-fmad=false). Division and 1/sqrt stay at <= 1 ulp: the texel sums cancel heavily, the kernel is bound by memory latency, and the
// fast forms bought nothing measurable.
//
// Scatter: ONE match.any on the hit triangle per ray (the three texel indices are functions of prim_index), then one log-step
// segmented reduction inside each peer group that carries the three vertex values together, then three red.global.add.f32 per
// distinct triangle per warp. Primary rays of one pixel hit one to three triangles, so ~10x fewer atomics reach L2 than rays x 3.
#include "hf_kernels.h"

#ifndef HF_BLOCK
#define HF_BLOCK 256
#endif
#define FULL_MASK 0xffffffffu

struct HfBwdXf {            // the two 3x3 linear parts the adjoint needs, by value in registers / constant bank
    float tw[9];            // to_world linear part, row-major
    float to[9];            // to_object linear part, row-major
};

// 1 / sqrt(x) to ~1 ulp: the hardware approximation (2 ulp) + one Newton step. The texel sums cancel heavily (tens of signed terms two
// to three orders of magnitude above their sum), so per-term errors of 2 ulp already show at the 1e-4 per-texel tolerance.
__device__ __forceinline__ float bw_rsqrt(float x) {
    const float y = rsqrtf(x);
    return y * fmaf(-0.5f * x, y * y, 1.5f);
}

template <bool IDENT>
__device__ __forceinline__ f3 bw_lin_T(const float *m, f3 x) {      // M^T x
    if (IDENT) return x;
    return mk3(m[0] * x.x + m[3] * x.y + m[6] * x.z, m[1] * x.x + m[4] * x.y + m[7] * x.z, m[2] * x.x + m[5] * x.y + m[8] * x.z);
}
template <bool IDENT>
__device__ __forceinline__ f3 bw_lin(const float *m, f3 x) {        // M x
    if (IDENT) return x;
    return mk3(m[0] * x.x + m[1] * x.y + m[2] * x.z, m[3] * x.x + m[4] * x.y + m[5] * x.z, m[6] * x.x + m[7] * x.y + m[8] * x.z);
}

// GENERAL: FollowShape and / or adjoints of dp_du, dp_dv are present. IDENT: to_world is the identity (the plugin's default).
// The kernel is bound by memory latency (ncu: 0.9 eligible warps per cycle, DRAM 38 %), so what matters is how many of a ray's ~20
// independent loads are in flight together: with the default register budget (40) ptxas serialises them to save registers and the
// kernel time swings by 40 % with unrelated changes of the arithmetic. HF_BWD_MIN_CTAS resident CTAs per SM = the register budget.
#ifndef HF_BWD_MIN_CTAS
#define HF_BWD_MIN_CTAS 3              // measured (C3, ms): 3 -> 0.647, 4 -> 0.679, 5 -> 0.721, 6 -> 0.706
#endif
template <bool GENERAL, bool IDENT>
__global__ void __launch_bounds__(HF_BLOCK, HF_BWD_MIN_CTAS)
hf_backward_kernel(const __grid_constant__ HfParams P, const hf_rays R, const hf_prelim PI, const hf_si_adj A,
                   const uint32_t flags, float *__restrict__ grad, const size_t n) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const float *tw = P.to_world, *to = P.to_object;
    // 3x3 linear parts (row stride 4 in HfParams)
    const float TW[9] = { tw[0], tw[1], tw[2], tw[4], tw[5], tw[6], tw[8], tw[9], tw[10] };
    const float TO[9] = { to[0], to[1], to[2], to[4], to[5], to[6], to[8], to[9], to[10] };
    const float flip = P.flip_normals ? -1.f : 1.f;

    const size_t ntiles = (n + HF_BLOCK - 1) / HF_BLOCK;
    // The chain t -> prim_index -> texels is three dependent DRAM round trips per tile. t and prim_index of the NEXT tile are
    // fetched before this tile is processed (two registers), so a tile starts with both in hand: one round trip for the ray /
    // adjoint / texel loads. (A miss then costs 8 bytes instead of 4.)
    float t_next = HF_INF;
    uint32_t prim_next = 0u;
    if ((size_t) blockIdx.x < ntiles) {
        const size_t j = (size_t) blockIdx.x * HF_BLOCK + threadIdx.x;
        if (j < n) { t_next = __ldcs(PI.t + j); prim_next = __ldcs(PI.prim_index + j); }
    }
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const size_t i = tile * HF_BLOCK + threadIdx.x;
        const float t = t_next;
        uint32_t prim = prim_next;
        t_next = HF_INF;
        {
            const size_t j = (tile + gridDim.x) * HF_BLOCK + threadIdx.x;
            if (tile + gridDim.x < ntiles && j < n) { t_next = __ldcs(PI.t + j); prim_next = __ldcs(PI.prim_index + j); }
        }
        bool valid = i < n && (t != HF_INF);
        if (valid && R.active) valid = R.active[i] != 0;
        const unsigned act = __ballot_sync(FULL_MASK, valid);
        if (act == 0u) continue;                                // warp-uniform: sky / ground regions of the image
        uint32_t i0 = 0u, i1 = 0u, i2 = 0u;
        if (!valid) prim = 0u;
        float g0 = 0.f, g1 = 0.f, g2 = 0.f;
        if (valid) {
            // Every load of the ray first, in program order (asm volatile keeps ptxas from sinking them behind the arithmetic to save
            // registers): ~20 independent requests in flight per thread is what hides the DRAM latency this kernel is bound by.
            #define LDS_CS(dst, ptr) asm volatile("ld.global.cs.f32 %0, [%1];" : "=f"(dst) : "l"((ptr) + i))
            #define LDA(dst, ptr) do { dst = 0.f; if (ptr) LDS_CS(dst, ptr); } while (0)
            const uint32_t cell = prim >> 1, k = prim & 1u;
            const uint32_t cy = hf_divide(cell, P.nx_magic, P.nx_shift), cx = cell - cy * (uint32_t) P.NX;
#ifdef HF_DEBUG_BOUNDS
            if (cy >= (uint32_t) P.NY) __trap();                // checked build: a primitive index that is not a triangle of this field
#endif
            const uint32_t b = cy * (uint32_t) P.W + cx;
            const float *h = P.heights;
            i0 = b + 1u;
            i1 = b + (uint32_t) P.W + k;
            i2 = k ? b + (uint32_t) P.W : b;
            float h0, h1, h2;
            asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(h0) : "l"(h + i0));
            asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(h1) : "l"(h + i1));
            asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(h2) : "l"(h + i2));
            f3 dw, ow = mk3(0.f, 0.f, 0.f);
            LDS_CS(dw.x, R.dx); LDS_CS(dw.y, R.dy); LDS_CS(dw.z, R.dz);
            float u, v;
            LDS_CS(u, PI.u); LDS_CS(v, PI.v);
            float a_t, apx, apy, apz, anx, any_, anz, a_uvx, a_uvy;
            LDA(a_t, A.t); LDA(apx, A.px); LDA(apy, A.py); LDA(apz, A.pz); LDA(anx, A.nx); LDA(any_, A.ny); LDA(anz, A.nz); LDA(a_uvx, A.uvx); LDA(a_uvy, A.uvy);
            float adu[3] = { 0.f, 0.f, 0.f }, adv[3] = { 0.f, 0.f, 0.f };
            if (GENERAL) {                                      // only FollowShape reads the origin; dp_du / dp_dv adjoints are rare
                LDS_CS(ow.x, R.ox); LDS_CS(ow.y, R.oy); LDS_CS(ow.z, R.oz);
                LDA(adu[0], A.dpdu_x); LDA(adu[1], A.dpdu_y); LDA(adu[2], A.dpdu_z); LDA(adv[0], A.dpdv_x); LDA(adv[1], A.dpdv_y); LDA(adv[2], A.dpdv_z);
            }
            #undef LDA
            #undef LDS_CS
            const f3 a_pw = mk3(apx, apy, apz);
            const f3 a_nw = mk3(flip * anx, flip * any_, flip * anz);      // mesh.cpp:837-840

            // the hit triangle from prim_index: triangle 0 = (p10, p01, p00), triangle 1 = (p10, p11, p01)
            const float z0 = P.mh * h0, z1 = P.mh * h1, z2 = P.mh * h2;
            const float x0 = vtx((int) cx, P.W, P.rw), x1 = vtx((int) cx + 1, P.W, P.rw);
            const float y0 = vtx((int) cy, P.H, P.rh), y1 = vtx((int) cy + 1, P.H, P.rh);
            const f3 p0 = mk3(x1, y0, z0);
            const f3 p1 = mk3(k ? x1 : x0, y1, z1);
            const f3 p2 = mk3(x0, k ? y1 : y0, z2);
            const f3 e1 = sub3(p1, p0), e2 = sub3(p2, p0);
            const f3 c = mk3(e1.y * e2.z - e1.z * e2.y, e1.z * e2.x - e1.x * e2.z, e1.x * e2.y - e1.y * e2.x);

            // n_w = normalize(TO^T normalize(c)): adjoint of the facet normal w.r.t. the edges (only the z parts are needed below)
            const float inv_clen = bw_rsqrt(c.x * c.x + c.y * c.y + c.z * c.z);
            const f3 no = mk3(c.x * inv_clen, c.y * inv_clen, c.z * inv_clen);
            f3 a_no;
            if (IDENT) {
                a_no = a_nw;                                    // |no| = 1: the outer normalisation is the identity map's
            } else {
                const f3 m = bw_lin_T<false>(TO, no);
                const float inv_mlen = bw_rsqrt(m.x * m.x + m.y * m.y + m.z * m.z);
                const f3 nw = mk3(m.x * inv_mlen, m.y * inv_mlen, m.z * inv_mlen);
                const float dn = nw.x * a_nw.x + nw.y * a_nw.y + nw.z * a_nw.z;
                const f3 a_m = mk3((a_nw.x - nw.x * dn) * inv_mlen, (a_nw.y - nw.y * dn) * inv_mlen, (a_nw.z - nw.z * dn) * inv_mlen);
                a_no = bw_lin<false>(TO, a_m);
            }
            const float dno = no.x * a_no.x + no.y * a_no.y + no.z * a_no.z;
            const f3 a_c = mk3((a_no.x - no.x * dno) * inv_clen, (a_no.y - no.y * dno) * inv_clen, (a_no.z - no.z * dno) * inv_clen);
            float a_e1z = e2.x * a_c.y - e2.y * a_c.x;          // (e2 x a_c).z
            float a_e2z = a_c.x * e1.y - a_c.y * e1.x;          // (a_c x e1).z

            // d0 = uv(p1) - uv(p0), d1 = uv(p2) - uv(p0) from the fp32 vertex texcoords fmadd(x, 0.5, 0.5) the forward pass interpolates
            // (src/shapes/rectangle.cpp:312-313). NOT 0.5 * e: the rounded texcoords are the function being differentiated, and their
            // differences deviate from the exact half-edge by up to 1e-5 relative -- far above the 1e-4 per-texel tolerance once the
            // texel sums cancel.
#ifdef HF_BWD_HALF_EDGE          // dev builds only: the form that fails the per-texel tolerance, to measure what the exact form costs
            const float d0x = 0.5f * e1.x, d0y = 0.5f * e1.y, d1x = 0.5f * e2.x, d1y = 0.5f * e2.y;
#else
            const float d0x = tex(p1.x) - tex(p0.x), d0y = tex(p1.y) - tex(p0.y), d1x = tex(p2.x) - tex(p0.x), d1y = tex(p2.y) - tex(p0.y);
#endif
            if (GENERAL && (flags & HF_RAY_DPDUV)) {            // mesh.cpp:778-789
                const f3 a_du = bw_lin_T<IDENT>(TW, mk3(adu[0], adu[1], adu[2]));
                const f3 a_dv = bw_lin_T<IDENT>(TW, mk3(adv[0], adv[1], adv[2]));
                const float det = d0x * d1y - d0y * d1x;
                if (det != 0.f) {
                    const float r = 1.0f / det;
                    a_e1z += (a_du.z * d1y - a_dv.z * d1x) * r;
                    a_e2z += (a_dv.z * d0x - a_du.z * d0y) * r;
                }
            }
            f3 a_p = bw_lin_T<IDENT>(TW, a_pw);
            const float b1 = u, b2 = v, b0 = 1.f - u - v;
            float wz;       // z component of the adjoint that is distributed with the barycentrics
            if (GENERAL && (flags & HF_RAY_FOLLOWSHAPE)) {
                // mesh.cpp:728-729, 751-752: barycentrics detached, t = sqrt(|p-o|^2/|d|^2)
                const f3 po = mk3(p0.x * b0 + p1.x * b1 + p2.x * b2, p0.y * b0 + p1.y * b1 + p2.y * b2, p0.z * b0 + p1.z * b1 + p2.z * b2);
                const f3 pw = xform_point(P.to_world, po);
                const f3 rel = sub3(pw, ow);
                const float dd = dw.x * dw.x + dw.y * dw.y + dw.z * dw.z;
                const float tt = sqrtf((rel.x * rel.x + rel.y * rel.y + rel.z * rel.z) / dd);
                if (tt > 0.f) {
                    const float kk = a_t / (tt * dd);
                    const f3 extra = bw_lin_T<IDENT>(TW, mk3(rel.x * kk, rel.y * kk, rel.z * kk));
                    a_p = mk3(a_p.x + extra.x, a_p.y + extra.y, a_p.z + extra.z);
                }
                wz = a_p.z;
            } else {
                const f3 d = bw_lin<IDENT>(TO, dw);             // object-space direction (linear part of to_object)
                const float a_tt = a_t + (d.x * a_p.x + d.y * a_p.y + d.z * a_p.z);
                const float a_u = d0x * a_uvx + d0y * a_uvy;
                const float a_v = d1x * a_uvx + d1y * a_uvy;
                const f3 pvec = mk3(d.y * e2.z - d.z * e2.y, d.z * e2.x - d.x * e2.z, d.x * e2.y - d.y * e2.x);
                const float det = e1.x * pvec.x + e1.y * pvec.y + e1.z * pvec.z;
                const float de1z = d.x * e1.y - d.y * e1.x;
                wz = (a_v * de1z - a_tt * c.z - a_u * pvec.z) / det;           // IEEE division (see bw_rsqrt)
            }
            g0 = P.mh * (wz * b0 - (a_e1z + a_e2z));
            g1 = P.mh * (wz * b1 + a_e1z);
            g2 = P.mh * (wz * b2 + a_e2z);
        }

        // ---- warp-aggregated scatter. All lanes of the warp are here (act is warp-uniform).
        const uint32_t lead_prim = __shfl_sync(FULL_MASK, prim, __ffs((int) act) - 1);
        if (__all_sync(FULL_MASK, !valid || prim == lead_prim)) {
            // one triangle for the whole warp (the common case at >= 32 samples per pixel): a plain butterfly, invalid lanes hold 0
            #pragma unroll
            for (int s = 16; s > 0; s >>= 1) {
                g0 += __shfl_xor_sync(FULL_MASK, g0, s);
                g1 += __shfl_xor_sync(FULL_MASK, g1, s);
                g2 += __shfl_xor_sync(FULL_MASK, g2, s);
            }
            if (lane == (unsigned) (__ffs((int) act) - 1)) {
                hf_grad_add(P, grad, i0, g0); hf_grad_add(P, grad, i1, g1); hf_grad_add(P, grad, i2, g2);
            }
        } else if (valid) {
            unsigned peers = __match_any_sync(act, prim);
            const bool leader = (peers & lt) == 0u;
            unsigned rel = __popc(peers & lt);                     // my rank inside the group
            peers &= (0xfffffffeu << lane);                        // peers above me
            while (__any_sync(act, peers != 0u)) {
                const int next = __ffs((int) peers);               // 1-based lane of the next peer above, 0 = none
                const int src = next ? next - 1 : (int) lane;
                const float o0 = __shfl_sync(act, g0, src), o1 = __shfl_sync(act, g1, src), o2 = __shfl_sync(act, g2, src);
                if (next) { g0 += o0; g1 += o1; g2 += o2; }
                const unsigned retired = __ballot_sync(act, (rel & 1u) != 0u);   // odd ranks retire: drop them from every peer set
                peers &= ~retired;
                rel >>= 1;
            }
            if (leader) { hf_grad_add(P, grad, i0, g0); hf_grad_add(P, grad, i1, g1); hf_grad_add(P, grad, i2, g2); }
        }
    }
}

static inline bool hf_is_identity(const float m[12]) {
    static const float I[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
    for (int k = 0; k < 12; ++k) if (m[k] != I[k]) return false;
    return true;
}

cudaError_t hf_launch_backward(const HfParams &P, const HfLaunchInfo &li, const hf_rays &rays, const hf_prelim &pi,
                               const hf_si_adj &adj, uint32_t flags, float *grad, size_t n, cudaStream_t s) {
    if (n == 0 || (flags & HF_RAY_DETACHSHAPE)) return cudaSuccess;   // mesh.cpp:713-717
    const size_t ntiles = (n + HF_BLOCK - 1) / HF_BLOCK, cap = (size_t) li.num_sms * 8;
    const int grid = (int) (ntiles < cap ? ntiles : cap);
    const bool have_duv = (flags & HF_RAY_DPDUV) && (adj.dpdu_x || adj.dpdu_y || adj.dpdu_z || adj.dpdv_x || adj.dpdv_y || adj.dpdv_z);
    const bool general = (flags & HF_RAY_FOLLOWSHAPE) || have_duv;
    const bool ident = hf_is_identity(P.to_world) && hf_is_identity(P.to_object);
    if (general) {
        if (ident) hf_backward_kernel<true, true><<<grid, HF_BLOCK, 0, s>>>(P, rays, pi, adj, flags, grad, n);
        else hf_backward_kernel<true, false><<<grid, HF_BLOCK, 0, s>>>(P, rays, pi, adj, flags, grad, n);
    } else {
        if (ident) hf_backward_kernel<false, true><<<grid, HF_BLOCK, 0, s>>>(P, rays, pi, adj, flags, grad, n);
        else hf_backward_kernel<false, false><<<grid, HF_BLOCK, 0, s>>>(P, rays, pi, adj, flags, grad, n);
    }
    return cudaGetLastError();
}
