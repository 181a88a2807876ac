// hf_kernels.cu -- hand-written sm_100a kernels of the differentiable heightfield path.
//
//   hf_pyramid_tile_kernel / hf_pyramid_top_kernel : min/max mip pyramid (the shape's acceleration
//        data; replaces the Embree re-commit / OptiX GAS rebuild of src/render/scene.cpp:343-367)
//   hf_intersect_kernel      : hierarchical DDA + two Moeller-Trumbore triangles per cell
//        (Shape::ray_intersect_preliminary / ray_test / ray_intersect, shape.h:137-197)
//   hf_si_kernel             : Shape::compute_surface_interaction (+ finalize), shape.h:179-183
//   (hf_backward_kernel lives in hf_backward.cu: reverse mode w.r.t. the hit cell's texels, warp-aggregated atomics)
//   hf_adam_kernel           : src/python/python/ad/optimizers.py:263-302
//
// Compiled with -fmad=false: see hf_device.cuh.
#include <stdlib.h>
#include <algorithm>
#include "hf_kernels.h"
#include "hf_walk_common.cuh"

// ------------------------------------------------------------------------------------------
// Pyramid build
// ------------------------------------------------------------------------------------------
#define PT 64                      // cells per tile edge
#define PT_LEVELS 6                // levels 1..6 come out of one tile

__global__ void __launch_bounds__(256) hf_pyramid_tile_kernel(const __grid_constant__ HfParams P, float2 *__restrict__ pyr) {
    __shared__ float s_z[(PT + 1) * (PT + 1)];
    __shared__ float2 s_lv[1024 + 256 + 64 + 16 + 4 + 1];

    const int tx0 = blockIdx.x * PT, ty0 = blockIdx.y * PT;      // first cell == first texel of the tile
    const int tw = min(PT + 1, P.W - tx0), th = min(PT + 1, P.H - ty0);   // valid texels in the tile

    for (int idx = threadIdx.x; idx < (PT + 1) * (PT + 1); idx += blockDim.x) {
        int j = idx / (PT + 1), i = idx - j * (PT + 1);
        float z = 0.f;
        if (i < tw && j < th) z = P.mh * __ldg(P.heights + (size_t) (ty0 + j) * P.W + (tx0 + i));
        s_z[idx] = z;
    }
    __syncthreads();

    // level 1: node covers texels [2lx, 2lx+2] x [2ly, 2ly+2], clamped to the grid
    {
        const int lw = P.lvl_w[1], lh = (P.NY + 1) >> 1;
        const int gx0 = blockIdx.x * 32, gy0 = blockIdx.y * 32;
        for (int idx = threadIdx.x; idx < 1024; idx += blockDim.x) {
            int ly = idx >> 5, lx = idx & 31;
            float mn = HF_INF, mx = -HF_INF;
            if (gx0 + lx < lw && gy0 + ly < lh) {
                int i1 = min(2 * lx + 2, tw - 1), j1 = min(2 * ly + 2, th - 1);
                for (int j = 2 * ly; j <= j1; ++j)
                    for (int i = 2 * lx; i <= i1; ++i) {
                        float z = s_z[j * (PT + 1) + i];
                        mn = fminf(mn, z); mx = fmaxf(mx, z);
                    }
                pyr[(size_t) P.lvl_off[1] + (size_t) (gy0 + ly) * lw + (gx0 + lx)] = make_float2(mn, mx);
            }
            s_lv[idx] = make_float2(mn, mx);     // absent nodes hold (+inf,-inf): neutral for min/max
        }
    }
    __syncthreads();

    // levels 2..6 from shared memory
    int src_off = 0, src_dim = 32;
    for (int l = 2; l <= PT_LEVELS; ++l) {
        const int dim = src_dim >> 1, dst_off = src_off + src_dim * src_dim;
        if (l <= P.Ltop) {
            const int lw = P.lvl_w[l], lh = (P.NY + (1 << l) - 1) >> l;
            const int gx0 = blockIdx.x * dim, gy0 = blockIdx.y * dim;
            for (int idx = threadIdx.x; idx < dim * dim; idx += blockDim.x) {
                int ly = idx / dim, lx = idx - ly * dim;
                float2 a = s_lv[src_off + (2 * ly) * src_dim + 2 * lx], b = s_lv[src_off + (2 * ly) * src_dim + 2 * lx + 1];
                float2 c = s_lv[src_off + (2 * ly + 1) * src_dim + 2 * lx], d = s_lv[src_off + (2 * ly + 1) * src_dim + 2 * lx + 1];
                float mn = fminf(fminf(a.x, b.x), fminf(c.x, d.x)), mx = fmaxf(fmaxf(a.y, b.y), fmaxf(c.y, d.y));
                s_lv[dst_off + idx] = make_float2(mn, mx);
                if (gx0 + lx < lw && gy0 + ly < lh)
                    pyr[(size_t) P.lvl_off[l] + (size_t) (gy0 + ly) * lw + (gx0 + lx)] = make_float2(mn, mx);
            }
        }
        __syncthreads();
        src_off = dst_off; src_dim = dim;
    }
}

// levels PT_LEVELS+1 .. Ltop: one CTA walks up through global memory
__global__ void __launch_bounds__(1024) hf_pyramid_top_kernel(const __grid_constant__ HfParams P, float2 *__restrict__ pyr) {
    for (int l = PT_LEVELS + 1; l <= P.Ltop; ++l) {
        const int sw = P.lvl_w[l - 1], sh = (P.NY + (1 << (l - 1)) - 1) >> (l - 1);
        const int lw = P.lvl_w[l], lh = (P.NY + (1 << l) - 1) >> l;
        const float2 *src = pyr + P.lvl_off[l - 1];
        float2 *dst = pyr + P.lvl_off[l];
        for (int idx = threadIdx.x; idx < lw * lh; idx += blockDim.x) {
            int ny = idx / lw, nx = idx - ny * lw;
            float mn = HF_INF, mx = -HF_INF;
            #pragma unroll
            for (int dj = 0; dj < 2; ++dj)
                #pragma unroll
                for (int di = 0; di < 2; ++di) {
                    int i = 2 * nx + di, j = 2 * ny + dj;
                    if (i < sw && j < sh) { float2 c = src[(size_t) j * sw + i]; mn = fminf(mn, c.x); mx = fmaxf(mx, c.y); }
                }
            dst[idx] = make_float2(mn, mx);
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// Quantised pyramid (hf_qpyr.h): what the production walk reads. Built right after the fp32 pyramid, whose top node
// gives the global z range the 16-bit codes are spread over.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int hf_qpitch(const HfParams &P, int l) { int p = (P.xmask + 1) >> l; return p > 0 ? p : 1; }
__device__ __forceinline__ int hf_qrows(const HfParams &P, int l) { int p = (P.ymask + 1) >> l; return p > 0 ? p : 1; }

#define HF_QBASE_TOP 6             // the base kernel writes levels 0..6
// levels 0, 1, 2: one thread per 4 x 4 block of cells; levels 3..6: reduced inside the CTA
__global__ void __launch_bounds__(256) hf_qpyramid_base_kernel(const __grid_constant__ HfParams P, uint32_t *__restrict__ qpyr) {
    const int Mx = P.xmask + 1, My = P.ymask + 1;
    const int bx = blockIdx.x * 16 + (threadIdx.x & 15), by = blockIdx.y * 16 + (threadIdx.x >> 4);
    const int x0 = bx * 4, y0 = by * 4;
    __shared__ uint32_t s_q[256 + 64 + 16 + 4 + 1];
    const float2 top = P.pyr[P.lvl_off[P.Ltop]];
    const HfQuant Q = hf_quant_make(top.x, top.y);
    uint32_t l1[4] = { HF_Q_EMPTY, HF_Q_EMPTY, HF_Q_EMPTY, HF_Q_EMPTY };
    #pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int y = y0 + j;
        uint32_t row[4];
        #pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int x = x0 + i;
            uint32_t q = HF_Q_EMPTY;
            if (x < P.NX && y < P.NY) {
                const float *r0 = P.heights + (size_t) y * P.W + x;
                const float z00 = P.mh * __ldg(r0), z10 = P.mh * __ldg(r0 + 1), z01 = P.mh * __ldg(r0 + P.W), z11 = P.mh * __ldg(r0 + P.W + 1);
                q = hf_quant_node(fminf(fminf(z00, z10), fminf(z01, z11)), fmaxf(fmaxf(z00, z10), fmaxf(z01, z11)), Q);
            }
            row[i] = q;
            l1[(j >> 1) * 2 + (i >> 1)] = hf_quant_merge(l1[(j >> 1) * 2 + (i >> 1)], q);
        }
        if (y < My && x0 < Mx) {
            uint32_t *dst = qpyr + P.qoff[0] + (size_t) y * Mx + x0;
            if (x0 + 3 < Mx) *reinterpret_cast<uint4 *>(dst) = make_uint4(row[0], row[1], row[2], row[3]);   // Mx % 4 == 0 here
            else { for (int i = 0; i < 4; ++i) if (x0 + i < Mx) dst[i] = row[i]; }
        }
    }
    if (P.QLtop >= 1) {
        const int p1 = hf_qpitch(P, 1), r1 = hf_qrows(P, 1);
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int nx = 2 * bx + (k & 1), ny = 2 * by + (k >> 1);
            if (nx < p1 && ny < r1) qpyr[P.qoff[1] + (size_t) ny * p1 + nx] = l1[k];
        }
    }
    const uint32_t l2 = hf_quant_merge(hf_quant_merge(l1[0], l1[1]), hf_quant_merge(l1[2], l1[3]));
    if (P.QLtop >= 2) {
        const int p2 = hf_qpitch(P, 2), r2 = hf_qrows(P, 2);
        if (bx < p2 && by < r2) qpyr[P.qoff[2] + (size_t) by * p2 + bx] = l2;
    }
    // levels 3..6 of this CTA's 64 x 64 cells through shared memory (16 x 16 level-2 nodes -> 8 x 8 -> 4 x 4 -> 2 x 2 -> 1)
    s_q[threadIdx.x] = l2;
    __syncthreads();
    int src = 0, sdim = 16;
    #pragma unroll
    for (int l = 3; l <= HF_QBASE_TOP; ++l) {
        const int dim = sdim >> 1, dst = src + sdim * sdim;
        if ((int) threadIdx.x < dim * dim) {
            const int ly = threadIdx.x / dim, lx = threadIdx.x - ly * dim;
            const uint32_t q = hf_quant_merge(hf_quant_merge(s_q[src + (2 * ly) * sdim + 2 * lx], s_q[src + (2 * ly) * sdim + 2 * lx + 1]),
                                              hf_quant_merge(s_q[src + (2 * ly + 1) * sdim + 2 * lx], s_q[src + (2 * ly + 1) * sdim + 2 * lx + 1]));
            s_q[dst + threadIdx.x] = q;
            const int nx = blockIdx.x * dim + lx, ny = blockIdx.y * dim + ly;
            if (l <= P.QLtop && nx < hf_qpitch(P, l) && ny < hf_qrows(P, l)) qpyr[P.qoff[l] + (size_t) ny * hf_qpitch(P, l) + nx] = q;
        }
        __syncthreads();
        src = dst; sdim = dim;
    }
}

// levels HF_QBASE_TOP + 1 .. QLtop: one CTA walks up through global memory (level 7 of a 16384^2 field has 128 x 128 nodes)
__global__ void __launch_bounds__(1024) hf_qpyramid_top_kernel(const __grid_constant__ HfParams P, uint32_t *__restrict__ qpyr) {
    for (int l = HF_QBASE_TOP + 1; l <= P.QLtop; ++l) {
        const int pw = hf_qpitch(P, l), ph = hf_qrows(P, l), cw = hf_qpitch(P, l - 1), ch = hf_qrows(P, l - 1);
        const int fx = cw / pw, fy = ch / ph;                    // 2, or 1 once the axis is down to one node
        const uint32_t *src = qpyr + P.qoff[l - 1];
        for (int idx = threadIdx.x; idx < pw * ph; idx += blockDim.x) {
            const int ny = idx / pw, nx = idx - ny * pw;
            uint32_t q = HF_Q_EMPTY;
            for (int dj = 0; dj < fy; ++dj)
                for (int di = 0; di < fx; ++di) q = hf_quant_merge(q, src[(size_t) (ny * fy + dj) * cw + (nx * fx + di)]);
            qpyr[P.qoff[l] + idx] = q;
        }
        __syncthreads();
    }
}

cudaError_t hf_launch_pyramid(const HfParams &P, float2 *pyr, uint32_t *qpyr, cudaStream_t s) {
    dim3 grid((P.NX + PT - 1) / PT, (P.NY + PT - 1) / PT);
    hf_pyramid_tile_kernel<<<grid, 256, 0, s>>>(P, pyr);
    if (P.Ltop > PT_LEVELS) hf_pyramid_top_kernel<<<1, 1024, 0, s>>>(P, pyr);
    const int Mx = P.xmask + 1, My = P.ymask + 1;
    dim3 qgrid((Mx + 63) / 64, (My + 63) / 64);
    hf_qpyramid_base_kernel<<<qgrid, 256, 0, s>>>(P, qpyr);
    if (P.QLtop > HF_QBASE_TOP) hf_qpyramid_top_kernel<<<1, 1024, 0, s>>>(P, qpyr);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Traversal
// ------------------------------------------------------------------------------------------
struct TopCache {               // shared-memory view of pyramid levels >= Ls, plus per-level tables
    const int *off;             // [HF_MAX_LEVELS] node offset of each level
    const int *w;               // [HF_MAX_LEVELS] nodes per row
    const float2 *nodes;        // levels Ls..Ltop
    int Ls, base;               // base = off[Ls]
};

struct HitRec {
    float t, u, v;
    uint32_t prim;
    int cx, cy;
    float z00, z10, z01, z11;
};

__device__ __forceinline__ int floor_clamp(float x, int lo, int hi) {
    if (!(x > (float) lo)) return lo;
    if (!(x < (float) hi)) return hi;
    return __float2int_rd(x);
}

// CANONICAL traversal: mirrors oracle/hf_oracle.c::traverse step for step. It defines N_node / N_leaf of the byte
// model and runs only in the counting entry point; the production walk is hf_traverse.cuh.
// Object-space ray (o, d, maxt). Front-to-back walk: the first cell with a hit is the closest hit.
template <bool COUNT>
__device__ __forceinline__ bool traverse(const HfParams &P, const TopCache &C, f3 o, f3 d, float maxt, HitRec &h,
                                         uint32_t &cn, uint32_t &cl) {
    const int NX = P.NX, NY = P.NY, Ltop = P.Ltop;
    if (d.x == 0.f && d.y == 0.f && d.z == 0.f) return false;
    if (!(maxt >= 0.f)) return false;

    const float ogx = (o.x - P.x0) * P.sgx, ogy = (o.y - P.y0) * P.sgy;
    const float dgx = d.x * P.sgx, dgy = d.y * P.sgy;
    const float oz = o.z, dz = d.z;
    const float idx = __fdiv_rn(1.0f, dgx), idy = __fdiv_rn(1.0f, dgy), idz = __fdiv_rn(1.0f, dz);
    const bool zx = dgx == 0.f, zy = dgy == 0.f;

    float t0 = 0.f, t1 = maxt;
    if (!zx) {
        float a = (0.f - ogx) * idx, b = ((float) NX - ogx) * idx;
        t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    } else if (ogx < 0.f || ogx > (float) NX) return false;
    if (!zy) {
        float a = (0.f - ogy) * idy, b = ((float) NY - ogy) * idy;
        t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    } else if (ogy < 0.f || ogy > (float) NY) return false;
    {
        const float2 top = C.nodes[C.off[Ltop] - C.base];
        const float pad = HF_Z_SLACK * fmaxf(fabsf(top.x), fabsf(top.y)) + HF_Z_SLACK * fabsf(oz);
        if (dz != 0.f) {
            float a = (top.x - pad - oz) * idz, b = (top.y + pad - oz) * idz;
            float tn = fminf(a, b), tf = fmaxf(a, b);
            tn -= fabsf(tn) * HF_Z_SLACK; tf += fabsf(tf) * HF_Z_SLACK;
            t0 = fmaxf(t0, tn); t1 = fminf(t1, tf);
        } else if (oz < top.x - pad || oz > top.y + pad) return false;
    }
    if (!(t0 <= t1)) return false;

    const float epsz = HF_Z_SLACK * (fabsf(oz) + fabsf(t1 * dz));
    const bool sx = dgx > 0.f, sy = dgy > 0.f;

    int ix = floor_clamp(fmaf(t0, dgx, ogx), 0, NX - 1);
    int iy = floor_clamp(fmaf(t0, dgy, ogy), 0, NY - 1);
    int L = Ltop;
    float t = t0;
    bool alive = true;
    const float *__restrict__ H = P.heights;

    for (;;) {
        bool cand = false;
        // phase 1: walk nodes until a leaf cell overlaps the ray's z interval, or the ray ends
        while (alive && !cand) {
            const int nx = ix >> L, ny = iy >> L;
            const int bx = sx ? min((nx + 1) << L, NX) : (nx << L);
            const int by = sy ? min((ny + 1) << L, NY) : (ny << L);
            const float tx = zx ? HF_INF : ((float) bx - ogx) * idx;
            const float ty = zy ? HF_INF : ((float) by - ogy) * idy;
            float te = fminf(tx, ty);
            const bool last = !(te < t1);
            if (last) te = t1;
            const float za = fmaf(t, dz, oz), zb = fmaf(te, dz, oz);
            const float zlo = fminf(za, zb) - epsz, zhi = fmaxf(za, zb) + epsz;

            if (L > 0) {
                const int ni = C.off[L] + ny * C.w[L] + nx;
                const float2 node = (L >= C.Ls) ? C.nodes[ni - C.base] : __ldg(P.pyr + ni);
                if (COUNT) cn++;
                if (zlo <= node.y && zhi >= node.x) { L--; continue; }
            } else {
                const float *row0 = H + (size_t) iy * P.W + ix;
                const float z00 = P.mh * __ldg(row0), z10 = P.mh * __ldg(row0 + 1);
                const float z01 = P.mh * __ldg(row0 + P.W), z11 = P.mh * __ldg(row0 + P.W + 1);
                const float mn = fminf(fminf(z00, z10), fminf(z01, z11));
                const float mx = fmaxf(fmaxf(z00, z10), fmaxf(z01, z11));
                if (COUNT) cl++;
                if (zlo <= mx && zhi >= mn) {
                    cand = true;
                    h.cx = ix; h.cy = iy; h.z00 = z00; h.z10 = z10; h.z01 = z01; h.z11 = z11;
                }
            }
            // advance to the next node (also after a candidate: phase 2 resumes from there on a miss)
            if (last) { alive = false; break; }
            const int ox_ = ix, oy_ = iy;
            const bool stepx = tx <= ty, stepy = ty <= tx;
            if (stepx) ix = sx ? bx : bx - 1;
            else {
                const int lo = nx << L, hi = min(((nx + 1) << L) - 1, NX - 1);
                const int q = floor_clamp(fmaf(te, dgx, ogx), lo, hi);
                ix = sx ? max(q, ox_) : (dgx < 0.f ? min(q, ox_) : ox_);
            }
            if (stepy) iy = sy ? by : by - 1;
            else {
                const int lo = ny << L, hi = min(((ny + 1) << L) - 1, NY - 1);
                const int q = floor_clamp(fmaf(te, dgy, ogy), lo, hi);
                iy = sy ? max(q, oy_) : (dgy < 0.f ? min(q, oy_) : oy_);
            }
            if (ix < 0 || ix >= NX || iy < 0 || iy >= NY) { alive = false; break; }
            t = te;
            if (L < Ltop && ((((ix ^ ox_) | (iy ^ oy_)) >> (L + 1)) != 0)) L++;
        }
        if (!cand) return false;
        // phase 2: the per-cell primitive
        if (cell_intersect(P, h.cx, h.cy, h.z00, h.z10, h.z01, h.z11, o, d, maxt, h.t, h.u, h.v, h.prim)) return true;
    }
}

// ------------------------------------------------------------------------------------------
// Surface interaction (src/render/mesh.cpp:672-903 on the hit triangle in object space, mapped to
// world like src/shapes/rectangle.cpp:250-323; finalize per interaction.h:257-266, 476-499)
// ------------------------------------------------------------------------------------------
#define STF(ptr, val) do { if (ptr) __stcs((ptr) + i, (val)); } while (0)

__device__ __forceinline__ void write_si_invalid(const hf_si &S, size_t i, f3 dw, uint32_t flags) {
    STF(S.t, HF_INF);
    STF(S.px, 0.f); STF(S.py, 0.f); STF(S.pz, 0.f);
    STF(S.nx, 0.f); STF(S.ny, 0.f); STF(S.nz, 0.f);
    STF(S.uvx, 0.f); STF(S.uvy, 0.f);
    STF(S.dpdu_x, 0.f); STF(S.dpdu_y, 0.f); STF(S.dpdu_z, 0.f);
    STF(S.dpdv_x, 0.f); STF(S.dpdv_y, 0.f); STF(S.dpdv_z, 0.f);
    STF(S.shn_x, 0.f); STF(S.shn_y, 0.f); STF(S.shn_z, 0.f);
    STF(S.shs_x, 0.f); STF(S.shs_y, 0.f); STF(S.shs_z, 0.f);
    STF(S.sht_x, 0.f); STF(S.sht_y, 0.f); STF(S.sht_z, 0.f);
    STF(S.wi_x, -dw.x); STF(S.wi_y, -dw.y); STF(S.wi_z, -dw.z);
    STF(S.boundary_test, (flags & HF_RAY_BOUNDARYTEST) ? 1e8f : 0.f);      // interaction.h:497-498
}

__device__ __forceinline__ void write_si(const HfParams &P, const hf_si &S, size_t i, uint32_t flags,
                                         f3 ow, f3 dw, float t_in, float b1, float b2, f3 p0, f3 p1, f3 p2) {
    const float b0 = 1.f - b1 - b2;
    const f3 dp0 = sub3(p1, p0), dp1 = sub3(p2, p0);
    // mesh.cpp:748
    const f3 po = mk3(fmaf(p0.x, b0, fmaf(p1.x, b1, p2.x * b2)),
                      fmaf(p0.y, b0, fmaf(p1.y, b1, p2.y * b2)),
                      fmaf(p0.z, b0, fmaf(p1.z, b1, p2.z * b2)));
    // mesh.cpp:757
    const f3 no = normalize3(cross3(dp0, dp1));
    const f3 pw = xform_point(P.to_world, po);
    f3 nw = normalize3(xform_normal(P.to_object, no));

    const float u0x = tex(p0.x), u0y = tex(p0.y), u1x = tex(p1.x), u1y = tex(p1.y), u2x = tex(p2.x), u2y = tex(p2.y);
    const float uvx = fmaf(u2x, b2, fmaf(u1x, b1, u0x * b0));
    const float uvy = fmaf(u2y, b2, fmaf(u1y, b1, u0y * b0));

    f3 dpdu, dpdv;
    bool have = false;
    if (flags & HF_RAY_DPDUV) {      // mesh.cpp:778-789
        const float d0x = u1x - u0x, d0y = u1y - u0y, d1x = u2x - u0x, d1y = u2y - u0y;
        const float det = fmaf(d0x, d1y, -(d0y * d1x));
        const float inv_det = rcp_llvm(det);
        if (det != 0.f) {
            f3 du = mk3(fmaf(d1y, dp0.x, -(d0y * dp1.x)) * inv_det, fmaf(d1y, dp0.y, -(d0y * dp1.y)) * inv_det,
                        fmaf(d1y, dp0.z, -(d0y * dp1.z)) * inv_det);
            f3 dv = mk3(fmaf(-d1x, dp0.x, d0x * dp1.x) * inv_det, fmaf(-d1x, dp0.y, d0x * dp1.y) * inv_det,
                        fmaf(-d1x, dp0.z, d0x * dp1.z) * inv_det);
            dpdu = xform_vector(P.to_world, du);
            dpdv = xform_vector(P.to_world, dv);
            have = true;
        }
    }
    if (!have) coordinate_system(nw, dpdu, dpdv);       // mesh.cpp:762
    if (P.flip_normals) nw = neg3(nw);                  // mesh.cpp:837-840

    float t = t_in;
    if (flags & HF_RAY_FOLLOWSHAPE)                     // mesh.cpp:751-752
        t = __fsqrt_rn(__fdiv_rn(sqnorm3(sub3(pw, ow)), sqnorm3(dw)));

    STF(S.t, t);
    STF(S.px, pw.x); STF(S.py, pw.y); STF(S.pz, pw.z);
    STF(S.nx, nw.x); STF(S.ny, nw.y); STF(S.nz, nw.z);
    STF(S.uvx, uvx); STF(S.uvy, uvy);
    STF(S.dpdu_x, dpdu.x); STF(S.dpdu_y, dpdu.y); STF(S.dpdu_z, dpdu.z);
    STF(S.dpdv_x, dpdv.x); STF(S.dpdv_y, dpdv.y); STF(S.dpdv_z, dpdv.z);
    STF(S.shn_x, nw.x); STF(S.shn_y, nw.y); STF(S.shn_z, nw.z);

    const bool want_frame = S.shs_x || S.shs_y || S.shs_z || S.sht_x || S.sht_y || S.sht_z || S.wi_x || S.wi_y || S.wi_z;
    if (want_frame) {
        f3 ss = mk3(0.f, 0.f, 0.f), st = mk3(0.f, 0.f, 0.f);
        if (flags & HF_RAY_SHADINGFRAME) {              // interaction.h:257-266
            const float nd = -dot3(nw, dpdu);
            ss = normalize3(mk3(fmaf(nw.x, nd, dpdu.x), fmaf(nw.y, nd, dpdu.y), fmaf(nw.z, nd, dpdu.z)));
            if (dpdu.x == 0.f && dpdu.y == 0.f && dpdu.z == 0.f) { f3 tmp; coordinate_system(nw, ss, tmp); }
            st = cross3(nw, ss);
        }
        const f3 md = neg3(dw);
        STF(S.shs_x, ss.x); STF(S.shs_y, ss.y); STF(S.shs_z, ss.z);
        STF(S.sht_x, st.x); STF(S.sht_y, st.y); STF(S.sht_z, st.z);
        STF(S.wi_x, dot3(md, ss)); STF(S.wi_y, dot3(md, st)); STF(S.wi_z, dot3(md, nw));
    }
    if (S.boundary_test) {
        // mesh.cpp:845-888, the branch of a mesh without vertex normals (the heightfield is flat-shaded): least-squares (u, v) of the
        // hit point inside its triangle, mapped onto an equilateral triangle; distance to the nearest edge, scaled so that the
        // barycentre is at 1. Evaluated on the object-space triangle (the solve is affine-invariant).
        float bt = 0.f;
        if (flags & HF_RAY_BOUNDARYTEST) {
            const f3 rel = sub3(po, p0);
            const float bb1 = dot3(dp0, rel), bb2 = dot3(dp1, rel), a11 = dot3(dp0, dp0), a12 = dot3(dp0, dp1), a22 = dot3(dp1, dp1);
            const float inv_det = rcp_llvm(a11 * a22 - a12 * a12);
            const float u = fmaf(a22, bb1, -(a12 * bb2)) * inv_det;      // fmsub
            const float v = fmaf(-a12, bb1, a11 * bb2) * inv_det;        // fnmadd
            const float hq = 0.5f * 1.7320508075688772f;                 // 0.5f * sqrt(3.f), both rounded to fp32
            // p = tp0 * w + tp1 * u + tp2 * v, tp0 = (0, 0), tp1 = (1, 0), tp2 = (1/2, sqrt(3)/2): the products with the literals 0 and 1
            // fold away in Dr.Jit's constant propagation (w drops out)
            const float qx = u + 0.5f * v, qy = hq * v;
            const float ex[3] = { 1.f, -0.5f, -0.5f }, ey[3] = { 0.f, hq, -hq };            // e0 = tp1 - tp0, e1 = tp2 - tp1, e2 = tp0 - tp2
            const float vx[3] = { qx, qx - 1.f, qx - 0.5f }, vy[3] = { qy, qy, qy - hq };     // p - tp0, p - tp1, p - tp2
            float d2 = HF_INF;
#pragma unroll
            for (int e = 0; e < 3; ++e) {
                const float ve = fmaf(vy[e], ey[e], vx[e] * ex[e]), ee = fmaf(ey[e], ey[e], ex[e] * ex[e]);
                const float c = fminf(fmaxf(__fdiv_rn(ve, ee), 0.f), 1.f);
                const float px = vx[e] - ex[e] * c, py = vy[e] - ey[e] * c;
                d2 = fminf(d2, fmaf(py, py, px * px));
            }
            bt = __fdiv_rn(__fsqrt_rn(d2), __fdiv_rn(1.7320508075688772f, 6.f));
        }
        STF(S.boundary_test, bt);
    }
}

// The common request -- t, p, n, uv only (no dp_du / dp_dv, shading frame, wi, no FollowShape): same arithmetic as write_si for these
// fields, without the 17 null tests and the partial-derivative code of the general writer (3-4 % of the fused kernel's instructions).
__device__ __forceinline__ void write_si_lean(const HfParams &P, const hf_si &S, size_t i, float t, float b1, float b2, f3 p0, f3 p1, f3 p2) {
    const float b0 = 1.f - b1 - b2;
    const f3 dp0 = sub3(p1, p0), dp1 = sub3(p2, p0);
    const f3 po = mk3(fmaf(p0.x, b0, fmaf(p1.x, b1, p2.x * b2)),
                      fmaf(p0.y, b0, fmaf(p1.y, b1, p2.y * b2)),
                      fmaf(p0.z, b0, fmaf(p1.z, b1, p2.z * b2)));
    const f3 no = normalize3(cross3(dp0, dp1));
    const f3 pw = xform_point(P.to_world, po);
    f3 nw = normalize3(xform_normal(P.to_object, no));
    const float u0x = tex(p0.x), u0y = tex(p0.y), u1x = tex(p1.x), u1y = tex(p1.y), u2x = tex(p2.x), u2y = tex(p2.y);
    const float uvx = fmaf(u2x, b2, fmaf(u1x, b1, u0x * b0));
    const float uvy = fmaf(u2y, b2, fmaf(u1y, b1, u0y * b0));
    if (P.flip_normals) nw = neg3(nw);
    STF(S.t, t);
    STF(S.px, pw.x); STF(S.py, pw.y); STF(S.pz, pw.z);
    STF(S.nx, nw.x); STF(S.ny, nw.y); STF(S.nz, nw.z);
    STF(S.uvx, uvx); STF(S.uvy, uvy);
}
__device__ __forceinline__ void write_si_invalid_lean(const hf_si &S, size_t i) {
    STF(S.t, HF_INF);
    STF(S.px, 0.f); STF(S.py, 0.f); STF(S.pz, 0.f);
    STF(S.nx, 0.f); STF(S.ny, 0.f); STF(S.nz, 0.f);
    STF(S.uvx, 0.f); STF(S.uvy, 0.f);
}
static inline bool hf_si_is_lean(const hf_si &S, uint32_t flags) {
    return !(flags & HF_RAY_FOLLOWSHAPE) && !S.dpdu_x && !S.dpdu_y && !S.dpdu_z && !S.dpdv_x && !S.dpdv_y && !S.dpdv_z &&
           !S.shn_x && !S.shn_y && !S.shn_z && !S.shs_x && !S.shs_y && !S.shs_z && !S.sht_x && !S.sht_y && !S.sht_z &&
           !S.wi_x && !S.wi_y && !S.wi_z && !S.boundary_test;
}

// ------------------------------------------------------------------------------------------
// Intersection kernel (persistent CTAs, one ray per thread per tile of HF_BLOCK rays)
// ------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(HF_BLOCK, 4)
hf_intersect_kernel(const __grid_constant__ HfParams P, const HfLaunchInfo li, const hf_rays R, const hf_prelim O,
                    const hf_si S, uint8_t *__restrict__ hit_out, const uint32_t flags, const size_t n,
                    unsigned long long *__restrict__ counters) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2 *s_nodes = reinterpret_cast<float2 *>(smem_raw);                       // 16-byte aligned (bulk copy)
    int *s_off = reinterpret_cast<int *>(smem_raw + (size_t) li.top_nodes * sizeof(float2));
    int *s_w = s_off + HF_MAX_LEVELS;
    __shared__ __align__(8) uint64_t s_bar;

    // stage the top of the pyramid with one bulk async copy (TMA engine), tables by plain stores
    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < HF_MAX_LEVELS) { s_off[threadIdx.x] = P.lvl_off[threadIdx.x]; s_w[threadIdx.x] = P.lvl_w[threadIdx.x]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t bytes = (uint32_t) li.top_nodes * (uint32_t) sizeof(float2);
        mbar_expect_tx(&s_bar, bytes);
        bulk_g2s(s_nodes, P.pyr + li.top_off, bytes, &s_bar);
    }
    {
        uint32_t spins = 0;
        while (!mbar_try_wait(&s_bar, 0)) { if (++spins > (1u << 24)) __trap(); }
    }

    TopCache C;
    C.off = s_off; C.w = s_w; C.nodes = s_nodes; C.Ls = li.top_level_start; C.base = li.top_off;

    unsigned long long acc_nodes = 0, acc_leaves = 0, acc_hits = 0;

    const size_t ntiles = (n + HF_BLOCK - 1) / HF_BLOCK;
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const size_t i = tile * HF_BLOCK + threadIdx.x;
        if (i >= n) continue;
        const f3 ow = mk3(__ldcs(R.ox + i), __ldcs(R.oy + i), __ldcs(R.oz + i));
        const f3 dw = mk3(__ldcs(R.dx + i), __ldcs(R.dy + i), __ldcs(R.dz + i));
        const float maxt = __ldcs(R.maxt + i);
        const bool active = R.active ? (R.active[i] != 0) : true;

        HitRec h;
        uint32_t cn = 0, cl = 0;
        bool hit = false;
        f3 o, d;
        if (active) {
            // Transform::transform_affine(ray), src/shapes/rectangle.cpp:206-214: t is preserved
            o = xform_point(P.to_object, ow);
            d = xform_vector(P.to_object, dw);
            hit = traverse<MODE == HF_MODE_COUNTED>(P, C, o, d, maxt, h, cn, cl);
        }

        if (MODE == HF_MODE_TEST) {
            hit_out[i] = hit ? 1 : 0;
            continue;
        }
        if (O.t) {
            __stcs(O.t + i, hit ? h.t : HF_INF);
            __stcs(O.u + i, hit ? h.u : 0.f);
            __stcs(O.v + i, hit ? h.v : 0.f);
            __stcs(O.prim_index + i, hit ? h.prim : 0u);
        }
        if (MODE == HF_MODE_FUSED_SI) {
            if (hit) {
                // vertices of the hit triangle straight from the cell registers
                const float x0 = vtx(h.cx, P.W, P.rw), x1 = vtx(h.cx + 1, P.W, P.rw);
                const float y0 = vtx(h.cy, P.H, P.rh), y1 = vtx(h.cy + 1, P.H, P.rh);
                const f3 p0 = mk3(x1, y0, h.z10);
                f3 p1, p2;
                if ((h.prim & 1u) == 0u) { p1 = mk3(x0, y1, h.z01); p2 = mk3(x0, y0, h.z00); }
                else { p1 = mk3(x1, y1, h.z11); p2 = mk3(x0, y1, h.z01); }
                write_si(P, S, i, flags, ow, dw, h.t, h.u, h.v, p0, p1, p2);
            } else {
                write_si_invalid(S, i, dw, flags);
            }
        }
        if (MODE == HF_MODE_COUNTED) { acc_nodes += cn; acc_leaves += cl; acc_hits += hit ? 1 : 0; }
    }

    if (MODE == HF_MODE_COUNTED) {
        #pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            acc_nodes += __shfl_xor_sync(FULL_MASK, acc_nodes, s);
            acc_leaves += __shfl_xor_sync(FULL_MASK, acc_leaves, s);
            acc_hits += __shfl_xor_sync(FULL_MASK, acc_hits, s);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(counters + 1, acc_nodes);
            atomicAdd(counters + 2, acc_leaves);
            atomicAdd(counters + 3, acc_hits);
        }
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(counters + 0, (unsigned long long) n);
    }
}

// ------------------------------------------------------------------------------------------
// Production intersection kernel: warp-uniform walk (hf_traverse.cuh) + per-cell primitive + fused surface interaction.
// Persistent CTAs; one ray per thread per tile of HF_BLOCK rays; SoA ray loads / result stores are fully coalesced
// streaming accesses (ld.cs / st.cs: the ray stream must not evict the pyramid and the texels from L1/L2).
// The coarse levels of the pyramid are staged in shared memory by one bulk async copy (TMA engine); the level table
// holds generic pointers, so the walk reads a node with one load whether the level lives in shared or global memory.
// ------------------------------------------------------------------------------------------
// One ray through the production walk: load, transform, walk + per-cell primitive, store (+ fused surface interaction).
// t_start / L_start come from the bundle march (hf_beam_enter) or are 0 / HF_START_LEVEL.
// The result of one ray: preliminary record and (fused modes) the surface interaction of the hit triangle, whose vertices come from
// the candidate cell `c` the walk left in registers.
template <int MODE>
__device__ __forceinline__ void hf_store_result(const HfParams &P, const hf_rays &R, const hf_prelim &O, const hf_si &S, uint8_t *__restrict__ hit_out,
                                                const uint32_t flags, const size_t i, const bool hit, const float ht, const float hu, const float hv,
                                                const uint32_t hprim, const HfCand &c) {
    if (MODE == HF_MODE_TEST) {
        hit_out[i] = hit ? 1 : 0;
        return;
    }
    if (O.t) {
        __stcs(O.t + i, hit ? ht : HF_INF);
        __stcs(O.u + i, hit ? hu : 0.f);
        __stcs(O.v + i, hit ? hv : 0.f);
        __stcs(O.prim_index + i, hit ? hprim : 0u);
    }
    if (MODE == HF_MODE_FUSED_SI || MODE == HF_MODE_FUSED_SI_LEAN) {
        if (hit) {
            // vertices of the hit triangle straight from the candidate-cell registers
            const float x0 = vtx(c.cx, P.W, P.rw), x1 = vtx(c.cx + 1, P.W, P.rw);
            const float y0 = vtx(c.cy, P.H, P.rh), y1 = vtx(c.cy + 1, P.H, P.rh);
            const f3 p0 = mk3(x1, y0, c.z10);
            f3 p1, p2;
            if ((hprim & 1u) == 0u) { p1 = mk3(x0, y1, c.z01); p2 = mk3(x0, y0, c.z00); }
            else { p1 = mk3(x1, y1, c.z11); p2 = mk3(x0, y1, c.z01); }
            if (MODE == HF_MODE_FUSED_SI_LEAN) {
                write_si_lean(P, S, i, ht, hu, hv, p0, p1, p2);
            } else {
                // the world-space ray is only needed by FollowShape and wi: re-read it (L2) instead of keeping six
                // registers alive across the walk
                f3 ow2 = mk3(0.f, 0.f, 0.f), dw2 = ow2;
                if ((flags & HF_RAY_FOLLOWSHAPE) || S.wi_x || S.wi_y || S.wi_z) {
                    ow2 = mk3(__ldcs(R.ox + i), __ldcs(R.oy + i), __ldcs(R.oz + i));
                    dw2 = mk3(__ldcs(R.dx + i), __ldcs(R.dy + i), __ldcs(R.dz + i));
                }
                write_si(P, S, i, flags, ow2, dw2, ht, hu, hv, p0, p1, p2);
            }
        } else if (MODE == HF_MODE_FUSED_SI_LEAN) {
            write_si_invalid_lean(S, i);
        } else {
            f3 dw2 = mk3(0.f, 0.f, 0.f);
            if (S.wi_x || S.wi_y || S.wi_z) dw2 = mk3(__ldcs(R.dx + i), __ldcs(R.dy + i), __ldcs(R.dz + i));
            write_si_invalid(S, i, dw2, flags);
        }
    }
}

template <int MODE>
__device__ __forceinline__ void hf_process_loaded(const HfParams &P, const HfGrid &G, const hf_rays &R, const hf_prelim &O,
                                                  const hf_si &S, uint8_t *__restrict__ hit_out, const uint32_t flags,
                                                  const size_t i, const f3 ow, const f3 dw, const float maxt, const float t_start, const int L_start) {
    const bool active = R.active ? (R.active[i] != 0) : true;

    bool hit = false;
    HfCand c;
    float ht = HF_INF, hu = 0.f, hv = 0.f;
    uint32_t hprim = 0u;
    if (active) {
        // Transform::transform_affine(ray), src/shapes/rectangle.cpp:206-214: t is preserved
        const f3 o = xform_point(P.to_object, ow), d = xform_vector(P.to_object, dw);
        HfWalk w;
        if (hf_walk_init(G, o.x, o.y, o.z, d.x, d.y, d.z, maxt, w, t_start, L_start)) {
            // lanes leave hf_walk_next at different iterations and reconverge here, so the per-cell primitive
            // runs for all candidate lanes of the warp together
            while (hf_walk_next(G, w, c)) {
                if (cell_intersect(P, c.cx, c.cy, c.z00, c.z10, c.z01, c.z11, o, d, maxt, ht, hu, hv, hprim)) { hit = true; break; }
            }
        }
    }

    hf_store_result<MODE>(P, R, O, S, hit_out, flags, i, hit, ht, hu, hv, hprim, c);
}

template <int MODE>
__device__ __forceinline__ void hf_process_ray(const HfParams &P, const HfGrid &G, const hf_rays &R, const hf_prelim &O,
                                               const hf_si &S, uint8_t *__restrict__ hit_out, const uint32_t flags,
                                               const size_t i, const float t_start, const int L_start) {
    const f3 ow = mk3(__ldcs(R.ox + i), __ldcs(R.oy + i), __ldcs(R.oz + i));
    const f3 dw = mk3(__ldcs(R.dx + i), __ldcs(R.dy + i), __ldcs(R.dz + i));
    const float maxt = __ldcs(R.maxt + i);
    hf_process_loaded<MODE>(P, G, R, O, S, hit_out, flags, i, ow, dw, maxt, t_start, L_start);
}

#ifndef HF_MIN_CTAS
#define HF_MIN_CTAS 3              // 80 registers, no spills: measured 3 % faster than 4 CTAs of 64 registers once the loop got shorter
#endif
#ifdef HF_DEBUG_TIMING            // dev builds only (make VARIANT=_dbg EXTRA=-DHF_DEBUG_TIMING): per-warp timeline of the fast kernel
__device__ unsigned long long hf_dbg[8192 * 6];
extern "C" int hf_debug_read(unsigned long long *dst) { return (int) cudaMemcpyFromSymbol(dst, hf_dbg, sizeof(hf_dbg)); }
__device__ __forceinline__ unsigned long long hf_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#endif
#ifndef HF_FAST_TAIL
#define HF_FAST_TAIL 48
#endif
#ifndef HF_FAST_MIDTAIL
#define HF_FAST_MIDTAIL 48
#endif
#ifndef HF_HEAVY_FACTOR
#define HF_HEAVY_FACTOR 8          // adaptive schedule: a group is "heavy" above this many times the mean cost of a group ...
#endif
#ifndef HF_HEAVY_SPAN
#define HF_HEAVY_SPAN 16           // ... and above 1/16 of the time a resident warp is busy in the launch
#endif
#ifndef HF_FAST_GROUPS
#define HF_FAST_GROUPS 4           // 32-ray groups a warp takes off the global counter at a time
#endif
// (Tried in round 2 and removed: copying the next group's rays into shared memory with cp.async while the current group is walked, and
// taking the next fetch index one group early. 3.93 vs 3.81 ms at C3 with the carve-out adjusted: the other 23 warps of the SM already
// cover those waits, the copies and the extra warp synchronisations only add instructions. profiles/r02_notes.md)
// Persistent warps; each takes HF_FAST_GROUPS consecutive 32-ray groups at a time off one global counter: hit and miss regions of a
// ray batch differ ~8x in cost per ray, and a static deal of tiles left the SMs idle 7 % of the kernel (profiles/r01_notes.md).
template <int MODE, bool ADAPT>
__global__ void __launch_bounds__(HF_BLOCK, HF_MIN_CTAS)
hf_intersect_fast_kernel(const __grid_constant__ HfParams P, const HfLaunchInfo li, const hf_rays R, const hf_prelim O,
                         const hf_si S, uint8_t *__restrict__ hit_out, const uint32_t flags, const size_t n,
                         const unsigned long long nbig, const unsigned long long npair, unsigned long long *__restrict__ sched) {
    HfGrid G;
    // ADAPT (hf_set_adaptive_schedule): the groups the previous launch found far more expensive than the rest go out first, one per
    // fetch, and are skipped when their turn comes in the natural order -- a 100 us group that starts at t = 0 costs nothing, one
    // that starts 10 us before the end of a 0.5 ms kernel costs 18 % (profiles/r02_notes.md). The list and the bitmap are written
    // together by one kernel (hf_pick_heavy_kernel), so a group is processed exactly once whatever they hold; any schedule gives
    // the same results. Its state lives in shared memory: the walk below has no register to spare.
    __shared__ unsigned s_nheavy, s_c0[HF_BLOCK / 32];       // (costs in units of 64 cycles: 32 bits last 140 s)
    // ... and a coarse copy of the bitmap (one bit per 2^heavy_shift groups, 8192 bits): the exact bit is a global load on every
    // group's critical path (~550 cycles per group, 4 % of the kernel when measured); with the coarse bit clear it is not needed
    __shared__ uint32_t s_coarse[256];
    if (ADAPT) {
        const unsigned nheavy = min(*li.heavy_n, li.heavy_cap);
        if (threadIdx.x == 0) s_nheavy = nheavy;
        s_coarse[threadIdx.x & 255u] = 0u;
        __syncthreads();
        for (unsigned k = threadIdx.x; k < nheavy; k += HF_BLOCK) {
            const unsigned b = (li.heavy_list[k] >> li.heavy_shift) & 8191u;
            atomicOr(&s_coarse[b >> 5], 1u << (b & 31u));
        }
    }
    hf_stage_pyramid<false>(P, li, G);       // no bundle march here: only the quantised coarse levels are staged (ends with a CTA barrier)

    const unsigned lane = threadIdx.x & 31u;
    const unsigned long long ngroups = (n + 31) / 32;
#ifdef HF_DEBUG_TIMING
    const unsigned long long dbg_t0 = hf_gtime();
    unsigned long long dbg_max = 0, dbg_arg = 0, dbg_n = 0, dbg_last = dbg_t0, dbg_prev_t = 0;
#endif
    if (ADAPT) {
        // 32-bit group ids (< 2^31: bit 31 of g0 marks a fetch from the list) and the clock in shared memory: the walk below has no
        // register to spare (one more live value and the compiler recomputes loop invariants inside the leaf loop). A group's cost
        // is the time since the warp finished the group before it -- one clock read per group -- and the sum of all costs is
        // the time the warps were busy.
        const unsigned ng32 = (unsigned) ngroups, nbig32 = (unsigned) nbig, npair32 = (unsigned) npair;
        if (lane == 0) s_c0[threadIdx.x >> 5] = (unsigned) clock();
        const unsigned c_begin = s_c0[threadIdx.x >> 5];
        for (;;) {
            unsigned t = 0;
            if (lane == 0) t = (unsigned) atomicAdd(sched, 1ull);
            t = __shfl_sync(FULL_MASK, t, 0);
            const unsigned nheavy = s_nheavy;
            unsigned g0, ng;
            if (t < nheavy) {
                g0 = li.heavy_list[t]; ng = g0 < ng32 ? 1u : 0u;       // (an id beyond the batch: a list made for a larger batch of this context)
                g0 |= 0x80000000u;
            } else {
                t -= nheavy;
                if (t < nbig32) { g0 = t * HF_FAST_GROUPS; ng = (unsigned) HF_FAST_GROUPS; }
                else if (t < nbig32 + npair32) { g0 = nbig32 * HF_FAST_GROUPS + (t - nbig32) * 2u; ng = 2u; }
                else { g0 = nbig32 * HF_FAST_GROUPS + npair32 * 2u + (t - nbig32 - npair32); ng = 1u; }
                if (g0 >= ng32) break;
                ng = min(ng, ng32 - g0);
                if (nheavy == 0u) g0 |= 0x80000000u;                     // no list: nothing to skip
            }
            #pragma unroll 1
            for (unsigned g = 0; g < ng; ++g) {
                const unsigned gid = (g0 & 0x7fffffffu) + g;
                if (!(g0 >> 31)) {
                    const unsigned b = (gid >> li.heavy_shift) & 8191u;
                    if (((s_coarse[b >> 5] >> (b & 31u)) & 1u) && ((li.heavy_bits[gid >> 5] >> (gid & 31u)) & 1u)) continue;      // it went out first
                }
                const size_t i = (size_t) gid * 32u + lane;
                if (i < n) hf_process_ray<MODE>(P, G, R, O, S, hit_out, flags, i, 0.f, HF_START_LEVEL);
                __syncwarp();
#ifndef HF_AS_NO_RECORD          // (dev: what the consuming half of the schedule costs on its own)
                if (lane == 0) {
                    const unsigned now = (unsigned) clock();
                    li.grp_cost[gid] = (uint16_t) min((now - s_c0[threadIdx.x >> 5]) >> 6, 65535u);
                    s_c0[threadIdx.x >> 5] = now;
                }
#endif
            }
        }
        if (lane == 0) atomicAdd(li.cost_sum, (unsigned long long) (((unsigned) clock() - c_begin) >> 6));
    } else
    for (;;) {
        unsigned long long t = 0;
        if (lane == 0) t = atomicAdd(sched, 1ull);
        t = __shfl_sync(FULL_MASK, t, 0);
#ifdef HF_DEBUG_TIMING
        {
            const unsigned long long now = hf_gtime();
            if (dbg_n > 0 && now - dbg_last > dbg_max) { dbg_max = now - dbg_last; dbg_arg = dbg_prev_t; }
            dbg_last = now; ++dbg_n; dbg_prev_t = t;
            if (lane == 0) {
                unsigned long long *d = hf_dbg + ((size_t) blockIdx.x * (HF_BLOCK / 32) + (threadIdx.x >> 5)) * 6;
                d[0] = dbg_t0; d[1] = now; d[2] = dbg_n; d[3] = dbg_max; d[4] = dbg_arg; d[5] = t;
            }
        }
#endif
        // the first nbig fetches are HF_FAST_GROUPS groups wide (a warp that stays on neighbouring rays re-uses its L1 lines), the
        // next npair fetches two groups, the closing ones a single group: every phase outlasts the slowest fetch of the one before
        unsigned long long g0;
        unsigned ng;
        if (t < nbig) { g0 = t * HF_FAST_GROUPS; ng = (unsigned) HF_FAST_GROUPS; }
        else if (t < nbig + npair) { g0 = nbig * HF_FAST_GROUPS + (t - nbig) * 2ull; ng = 2u; }
        else { g0 = nbig * HF_FAST_GROUPS + npair * 2ull + (t - nbig - npair); ng = 1u; }
        if (g0 >= ngroups) break;
        #pragma unroll 1
        for (unsigned g = 0; g < ng; ++g) {
            const size_t i = (size_t) (g0 + g) * 32u + lane;
            if (i < n) hf_process_ray<MODE>(P, G, R, O, S, hit_out, flags, i, 0.f, HF_START_LEVEL);
            __syncwarp();
        }
    }
}

// The list for the next launch: groups that cost more than `factor` times the mean AND more than 1 / HF_HEAVY_SPAN of what a resident
// warp spent in the whole launch (only groups that can hold up the end of the kernel are worth taking out of their neighbourhood: a
// listed group loses the L1 lines it shares with the groups next to it), at most `cap` of them. The bitmap is rewritten as a whole
// (nwords words, zeros beyond the batch) so that it always matches the list.
__global__ void __launch_bounds__(256)
hf_pick_heavy_kernel(const uint16_t *__restrict__ cost, const unsigned ngroups, const unsigned long long *__restrict__ cost_sum,
                     unsigned *__restrict__ heavy_n, unsigned *__restrict__ heavy_list, uint32_t *__restrict__ heavy_bits,
                     const unsigned nwords, const unsigned cap, const unsigned factor, const unsigned span, const unsigned resident_warps) {
    const unsigned long long sum = *cost_sum;
    const unsigned long long by_mean = sum / (ngroups ? ngroups : 1u) * factor, by_span = sum / resident_warps / span;
    const unsigned thresh = (unsigned) min((unsigned long long) 65534u, max(by_mean, by_span));
    const unsigned lane = threadIdx.x & 31u;
    for (unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < nwords; w += (gridDim.x * blockDim.x) >> 5) {
        const unsigned g = w * 32u + lane;
        bool h = g < ngroups && cost[g] > thresh;
        if (h) {
            const unsigned slot = atomicAdd(heavy_n, 1u);
            if (slot < cap) heavy_list[slot] = g; else h = false;
        }
        const unsigned bits = __ballot_sync(FULL_MASK, h);
        if (lane == 0) heavy_bits[w] = bits;
    }
}

// ------------------------------------------------------------------------------------------
// Lane-pool intersection kernel for INCOHERENT ray batches (hf_set_ray_coherence(ctx, 0); the reference's Scene::ray_intersect takes
// the same hint as its `coherent` argument, src/render/scene.cpp:141-150).
//
// In the kernel above a warp takes 32 consecutive rays and walks them in lock step until the last one is finished: fine for the
// samples of one pixel, wasteful when the 32 rays have nothing to do with each other (C5: origins on a hemisphere) -- a warp then
// needs 51 lock-step iterations for rays that need 17 each (tools/sim_pool.py). Here every LANE owns its own ray and takes the next
// ray of the warp's pool as soon as its own is finished ("persistent while-while"):
//   refill   idle lanes take consecutive rays of the pool (coalesced loads), transform them and set up their walks;
//   walk     up to HF_POOL_ROUND iterations of hf_walk_iter for the lanes that are searching (a cap: lanes that hold a candidate must
//            not wait for the slowest searcher);
//   test     the per-cell primitive for the lanes that hold a candidate;
//   retire   finished rays drop a 13-word record (ray id, hit, cell) into a per-warp shared-memory queue and the lane is idle again;
//            when 32 records are queued the WHOLE warp turns them into surface interactions and stores them, one record per lane --
//            so the expensive output code runs at 32 lanes, not at the two or three that happen to finish together.
// Results are bit-identical to the kernel above (same walk, same primitive, same surface-interaction code). Measured (profiles/r02_notes.md):
// the instruction model promised -38 %; on a B200 the kernel is 5-18 % faster than the lock-step kernel for intersect-only hemisphere rays
// and about equal with the fused surface interaction -- large fields make incoherent rays latency-bound, and this kernel's per-lane state
// costs the registers that resident warps would need.
// ------------------------------------------------------------------------------------------
#ifndef HF_POOL_ROUND
#define HF_POOL_ROUND 8
#endif
#ifndef HF_POOL_RAYS
#define HF_POOL_RAYS 256           // rays a warp takes off the global counter at a time
#endif
#ifndef HF_POOL_REFILL
#define HF_POOL_REFILL 8           // refill as soon as this many lanes are idle (or nobody is working)
#endif
#define HF_POOL_REC 13             // words per queued result: ray id (2), hit, t, u, v, prim_index, cell (2), its four corner heights

#ifndef HF_POOL_MIN_CTAS
#define HF_POOL_MIN_CTAS 4         // 64 registers (192 bytes of spills): incoherent batches on large fields are latency-bound, resident warps beat
#endif                             // registers -- C5, 2^24 rays, intersect-only Grays/s at 512^2 / 4096^2 / 16384^2: 2 CTAs 5.8 / 3.1 / 2.2, 3: 7.1 / 4.1 / 2.9,
                                   // 4: 7.5 / 4.6 / 3.3, 5: 6.5 / 3.9 / 2.8, 6: 5.4 / 3.2 / 2.3; the lock-step kernel: 6.3 / 4.4 / 3.1
template <int MODE>
__global__ void __launch_bounds__(HF_BLOCK, HF_POOL_MIN_CTAS)
hf_intersect_pool_kernel(const __grid_constant__ HfParams P, const HfLaunchInfo li, const hf_rays R, const hf_prelim O,
                         const hf_si S, uint8_t *__restrict__ hit_out, const uint32_t flags, const size_t n,
                         unsigned long long *__restrict__ sched) {
    HfGrid G;
    hf_stage_pyramid<false>(P, li, G);
    __shared__ uint32_t s_rec[HF_BLOCK / 32][HF_POOL_REC][32];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    uint32_t (*rec)[32] = s_rec[warp];

    // per-lane state
    bool busy = false, have_cand = false, done = false, hit = false;
    size_t ri = 0;
    f3 o = mk3(0.f, 0.f, 0.f), d = o;
    float maxt = 0.f, ht = HF_INF, hu = 0.f, hv = 0.f;
    uint32_t hprim = 0u;
    HfWalk w;
    w.alive = false;
    HfCand c;
    c.cx = c.cy = 0; c.z00 = c.z10 = c.z01 = c.z11 = 0.f;
    // per-warp state (uniform)
    size_t pnext = 0, pend = 0;
    bool more = true;
    unsigned queued = 0;

    auto flush = [&]() {                                    // the whole warp: one queued record per lane -> results in memory
        __syncwarp();
        if (lane < queued) {
            const size_t i = ((size_t) rec[1][lane] << 32) | rec[0][lane];
            const bool h = rec[2][lane] != 0u;
            HfCand cc;
            cc.cx = (int) rec[7][lane]; cc.cy = (int) rec[8][lane];
            cc.z00 = __uint_as_float(rec[9][lane]); cc.z10 = __uint_as_float(rec[10][lane]);
            cc.z01 = __uint_as_float(rec[11][lane]); cc.z11 = __uint_as_float(rec[12][lane]);
            hf_store_result<MODE>(P, R, O, S, hit_out, flags, i, h, __uint_as_float(rec[3][lane]), __uint_as_float(rec[4][lane]),
                                  __uint_as_float(rec[5][lane]), rec[6][lane], cc);
        }
        __syncwarp();
        queued = 0;
    };

    for (;;) {
        // ---- refill
        const unsigned idle = __ballot_sync(FULL_MASK, !busy);
        const unsigned working = __ballot_sync(FULL_MASK, busy && !done);
        if (idle && more && ((unsigned) __popc(idle) >= HF_POOL_REFILL || working == 0u)) {
            if (pnext >= pend) {
                unsigned long long t = 0;
                if (lane == 0) t = atomicAdd(sched, 1ull);
                t = __shfl_sync(FULL_MASK, t, 0);
                pnext = (size_t) t * HF_POOL_RAYS;
                pend = pnext + HF_POOL_RAYS < n ? pnext + HF_POOL_RAYS : n;
                if (pnext >= n) { more = false; pnext = pend = 0; }
            }
            if (more) {
                const size_t mine = pnext + __popc(idle & lt);
                if (!busy && mine < pend) {
                    ri = mine; busy = true; have_cand = false; done = false; hit = false;
                    ht = HF_INF; hu = 0.f; hv = 0.f; hprim = 0u;
                    const f3 ow = mk3(__ldcs(R.ox + ri), __ldcs(R.oy + ri), __ldcs(R.oz + ri));
                    const f3 dw = mk3(__ldcs(R.dx + ri), __ldcs(R.dy + ri), __ldcs(R.dz + ri));
                    maxt = __ldcs(R.maxt + ri);
                    const bool active = R.active ? (R.active[ri] != 0) : true;
                    o = xform_point(P.to_object, ow); d = xform_vector(P.to_object, dw);     // transform_affine: t is preserved
                    w.alive = false;
                    if (!active || !hf_walk_init(G, o.x, o.y, o.z, d.x, d.y, d.z, maxt, w)) done = true;
                }
                const size_t adv = pnext + __popc(idle);
                pnext = adv < pend ? adv : pend;
            }
        }
        if (!more && __ballot_sync(FULL_MASK, busy) == 0u) break;

        // ---- walk: at most HF_POOL_ROUND lock-step iterations, searching lanes only
        #pragma unroll 1
        for (int it = 0; it < HF_POOL_ROUND; ++it) {
            const bool search = busy && !done && !have_cand;
            if (!__any_sync(FULL_MASK, search)) break;
            if (search) {
                if (!w.alive) done = true;
                else {
                    const int st = hf_walk_iter(G, w, c);
                    if (st == 1) have_cand = true;
                    else if (st == 2) done = true;
                }
            }
        }
        // ---- per-cell primitive for the lanes that hold a candidate
        if (__any_sync(FULL_MASK, have_cand)) {
            if (have_cand) {
                have_cand = false;
                if (cell_intersect(P, c.cx, c.cy, c.z00, c.z10, c.z01, c.z11, o, d, maxt, ht, hu, hv, hprim)) { hit = true; done = true; }
                else if (!w.alive) done = true;
            }
        }
        // ---- retire: finished rays queue their record; a full queue is turned into results by the whole warp
        for (;;) {
            const unsigned fin = __ballot_sync(FULL_MASK, busy && done);
            if (fin == 0u) break;
            const unsigned slot = queued + __popc(fin & lt);
            if (busy && done && slot < 32u) {
                rec[0][slot] = (uint32_t) ri; rec[1][slot] = (uint32_t) ((unsigned long long) ri >> 32);
                rec[2][slot] = hit ? 1u : 0u;
                rec[3][slot] = __float_as_uint(ht); rec[4][slot] = __float_as_uint(hu); rec[5][slot] = __float_as_uint(hv); rec[6][slot] = hprim;
                rec[7][slot] = (uint32_t) c.cx; rec[8][slot] = (uint32_t) c.cy;
                rec[9][slot] = __float_as_uint(c.z00); rec[10][slot] = __float_as_uint(c.z10); rec[11][slot] = __float_as_uint(c.z01);
                rec[12][slot] = __float_as_uint(c.z11);
                busy = false; done = false;
            }
            const unsigned total = queued + __popc(fin);
            queued = total < 32u ? total : 32u;
            if (queued == 32u) flush();
        }
    }
    if (queued) flush();
}

// ------------------------------------------------------------------------------------------
// Bundled intersection kernel. A warp owns a tile of 32 x 32 = 1024 consecutive rays:
//   phase A  for each of its 32 groups of 32 consecutive rays, lanes load one ray each and the warp reduces the
//            component-wise bounds of the object-space directions (shuffles); lane g keeps the bounds of group g.
//            A group qualifies as a bundle when all its rays leave from one bit-identical origin (primary rays).
//   phase B  every lane marches ITS bundle through the pyramid (hf_beam_enter): 32 different bundles per warp
//            instruction instead of 32 copies of the same coarse descent.
//   phase C  group by group, every lane walks its own ray from the bundle's entry parameter (broadcast by shuffle).
// Rays that do not form a bundle (different origins, a single active ray) start at 0 as in the plain kernel.
// Phase A reads the rays with default caching so that phase C finds them in L2; phase C streams (ld.cs / st.cs).
// ------------------------------------------------------------------------------------------
#define HF_WARP_TILE 1024
#define HF_SMALL_GROUPS 4          // groups per tile in the closing phase of the schedule
#ifndef HF_BEAM_EXACT_BOUNDS
#define HF_BEAM_EXACT_BOUNDS 0
#endif
// Schedule: tiles come off one global counter. The first `nbig` tiles are 32 groups (1024 rays) wide, the remaining groups are
// handed out HF_SMALL_GROUPS at a time: a 1024-ray tile is ~0.3 ms of one warp, and with only large tiles the SMs sat idle
// 23 % of the kernel waiting for the last ones (profiles/r01_notes.md); the closing phase absorbs that imbalance.
template <int MODE>
__global__ void __launch_bounds__(HF_BLOCK, HF_MIN_CTAS)
hf_intersect_beam_kernel(const __grid_constant__ HfParams P, const HfLaunchInfo li, const hf_rays R, const hf_prelim O,
                         const hf_si S, uint8_t *__restrict__ hit_out, const uint32_t flags, const size_t n,
                         const unsigned long long nbig, unsigned long long *__restrict__ sched) {
    HfGrid G;
    hf_stage_pyramid<true>(P, li, G);

    const unsigned lane = threadIdx.x & 31u;
    const unsigned long long ngroups = (n + 31) / 32;
    for (;;) {
        unsigned long long wt = 0;
        if (lane == 0) wt = atomicAdd(sched, 1ull);
        wt = __shfl_sync(FULL_MASK, wt, 0);
        unsigned long long g0;
        unsigned ng;
        if (wt < nbig) { g0 = wt * 32ull; ng = 32u; }
        else {
            g0 = nbig * 32ull + (wt - nbig) * HF_SMALL_GROUPS;
            if (g0 >= ngroups) break;
            ng = (unsigned) (ngroups - g0 < HF_SMALL_GROUPS ? ngroups - g0 : HF_SMALL_GROUPS);
        }
        const size_t base = (size_t) g0 * 32u;
        // ---- phase A: bundle bounds, lane g <- group g
        HfBeam B;
        B.ogx = B.ogy = B.oz = 0.f; B.dxl = B.dyl = B.dzl = 0.f; B.dxh = B.dyh = B.dzh = 0.f;
        bool bundle = false;
        #pragma unroll 2
        for (unsigned g = 0; g < ng; ++g) {
            const size_t i = base + (size_t) g * 32u + lane;
            bool valid = i < n;
            if (valid && R.active) valid = R.active[i] != 0;
            f3 o = mk3(0.f, 0.f, 0.f), d = o;
            if (valid) {
                const f3 ow = mk3(__ldg(R.ox + i), __ldg(R.oy + i), __ldg(R.oz + i));
                const f3 dw = mk3(__ldg(R.dx + i), __ldg(R.dy + i), __ldg(R.dz + i));
                o = xform_point(P.to_object, ow); d = xform_vector(P.to_object, dw);
            }
            const unsigned vm = __ballot_sync(FULL_MASK, valid);
            const int first = (__ffs((int) vm) - 1) & 31;
            const float ox0 = __shfl_sync(FULL_MASK, o.x, first), oy0 = __shfl_sync(FULL_MASK, o.y, first),
                        oz0 = __shfl_sync(FULL_MASK, o.z, first);
            const bool fin = (d.x == d.x) && (d.y == d.y) && (d.z == d.z);
            const bool same = __all_sync(FULL_MASK, !valid || (o.x == ox0 && o.y == oy0 && o.z == oz0 && fin));
            const float dgx = d.x * P.sgx, dgy = d.y * P.sgy;
#if HF_BEAM_EXACT_BOUNDS
            float xl = valid ? dgx : HF_INF, xh = valid ? dgx : -HF_INF, yl = valid ? dgy : HF_INF, yh = valid ? dgy : -HF_INF;
            float zl = valid ? d.z : HF_INF, zh = valid ? d.z : -HF_INF;
            #pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) {
                xl = fminf(xl, __shfl_xor_sync(FULL_MASK, xl, sft)); xh = fmaxf(xh, __shfl_xor_sync(FULL_MASK, xh, sft));
                yl = fminf(yl, __shfl_xor_sync(FULL_MASK, yl, sft)); yh = fmaxf(yh, __shfl_xor_sync(FULL_MASK, yh, sft));
                zl = fminf(zl, __shfl_xor_sync(FULL_MASK, zl, sft)); zh = fmaxf(zh, __shfl_xor_sync(FULL_MASK, zh, sft));
            }
#else
            // centre (the first valid ray's direction) + radius: |delta| >= 0 orders like its bit pattern, so each radius is
            // ONE integer warp reduction (REDUX) instead of ten shuffles; the box is at most twice the tight one per axis
            const float cx = __shfl_sync(FULL_MASK, dgx, first), cy = __shfl_sync(FULL_MASK, dgy, first), cz = __shfl_sync(FULL_MASK, d.z, first);
            const float ex_ = valid ? fmaxf(fabsf(dgx - cx), fabsf(dgy - cy)) : 0.f, ez_ = valid ? fabsf(d.z - cz) : 0.f;
            float rxy = __uint_as_float(__reduce_max_sync(FULL_MASK, __float_as_uint(ex_)));
            float rz = __uint_as_float(__reduce_max_sync(FULL_MASK, __float_as_uint(ez_)));
            rxy = rxy * 1.0000005f + 1e-30f; rz = rz * 1.0000005f + 1e-30f;          // outward: the subtraction above rounded once
            const float xl = cx - rxy, xh = cx + rxy, yl = cy - rxy, yh = cy + rxy, zl = cz - rz, zh = cz + rz;
#endif
            if (lane == g) {
                bundle = same && (__popc(vm) > 1);
                B.ogx = (ox0 - P.x0) * P.sgx; B.ogy = (oy0 - P.y0) * P.sgy; B.oz = oz0;
                B.dxl = xl; B.dxh = xh; B.dyl = yl; B.dyh = yh; B.dzl = zl; B.dzh = zh;
            }
        }
        // ---- phase B: one bundle per lane
        float t_enter = 0.f;
        int L_start = HF_START_LEVEL;
        if (bundle) t_enter = hf_beam_enter(G, B, L_start);
        __syncwarp();
        // ---- phase C: per-ray walks from the bundle entry
        for (unsigned g = 0; g < ng; ++g) {
            const float ts = __shfl_sync(FULL_MASK, t_enter, g);
            const int ls = __shfl_sync(FULL_MASK, L_start, g);
            const size_t i = base + (size_t) g * 32u + lane;
            if (i < n) hf_process_ray<MODE>(P, G, R, O, S, hit_out, flags, i, ts, ls);
            __syncwarp();
        }
    }
}

// L1 / shared-memory split: the walk kernels use a few KB of shared memory per CTA and live off L1 hits on pyramid nodes and
// texels, so ask for the smallest shared carve-out that still fits the resident CTAs (HF_B200_CARVEOUT overrides, in percent).
template <typename K>
static void hf_prefer_l1(K kernel, size_t smem_per_cta, int ctas) {
    int pct = (int) ((smem_per_cta + 2048) * (size_t) ctas * 100 / (228 * 1024)) + 1;
    if (pct < 25) pct = 25;            // measured on C3: 25 % is the optimum, 10 % already costs resident CTAs (profiles/r01_notes.md)
    if (const char *e = getenv("HF_B200_CARVEOUT")) pct = atoi(e);
    if (pct < 0) return;
    if (pct > 100) pct = 100;
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}

static inline int hf_grid_for(size_t n, int num_sms, int ctas_per_sm) {
    size_t ntiles = (n + HF_BLOCK - 1) / HF_BLOCK;
    size_t cap = (size_t) num_sms * ctas_per_sm;
    return (int) (ntiles < cap ? (ntiles ? ntiles : 1) : cap);
}

cudaError_t hf_launch_pick_heavy(const HfLaunchInfo &li, unsigned units, unsigned resident_warps, cudaStream_t s) {
    static const int env_factor = getenv("HF_B200_HEAVY_FACTOR") ? atoi(getenv("HF_B200_HEAVY_FACTOR")) : HF_HEAVY_FACTOR;    // dev knobs
    static const int env_span = getenv("HF_B200_HEAVY_SPAN") ? atoi(getenv("HF_B200_HEAVY_SPAN")) : HF_HEAVY_SPAN;
    cudaError_t e = cudaMemsetAsync(li.heavy_n, 0, sizeof(unsigned), s);
    if (e != cudaSuccess) return e;
    const unsigned blocks = (li.heavy_words * 32u + 255u) / 256u, capb = (unsigned) li.num_sms * 8u;
    hf_pick_heavy_kernel<<<blocks < capb ? (blocks ? blocks : 1u) : capb, 256, 0, s>>>(li.grp_cost, units, li.cost_sum, li.heavy_n, li.heavy_list, li.heavy_bits,
                                                                                      li.heavy_words, li.heavy_cap, (unsigned) (env_factor > 0 ? env_factor : HF_HEAVY_FACTOR),
                                                                                      (unsigned) (env_span > 0 ? env_span : HF_HEAVY_SPAN), resident_warps ? resident_warps : 1u);
    return cudaGetLastError();
}

cudaError_t hf_launch_intersect(const HfParams &P, const HfLaunchInfo &li, int mode, const hf_rays &rays,
                                const hf_prelim &prelim, const hf_si &si, uint8_t *d_hit, uint32_t flags,
                                size_t n, unsigned long long *d_counters, unsigned long long *d_sched, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    const int grid = hf_grid_for(n, li.num_sms, 8);
    if (mode == HF_MODE_COUNTED) {
        const size_t smem = (size_t) li.top_nodes * sizeof(float2) + 2 * HF_MAX_LEVELS * sizeof(int);
        hf_intersect_kernel<HF_MODE_COUNTED><<<grid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, d_counters);
        return cudaGetLastError();
    }
    if (mode == HF_MODE_FUSED_SI && hf_si_is_lean(si, flags)) mode = HF_MODE_FUSED_SI_LEAN;
    if (li.use_bundles && mode != HF_MODE_TEST) {
        const size_t smem = hf_walk_smem(li, true);
        const unsigned long long ngroups = (n + 31) / 32;
        const unsigned long long nbig = (ngroups - ngroups / 8) / 32;                 // the last eighth (at least) goes out in small tiles
        const unsigned long long ntile = nbig + (ngroups - nbig * 32 + HF_SMALL_GROUPS - 1) / HF_SMALL_GROUPS;
        const size_t nct = (size_t) ((ntile + HF_BLOCK / 32 - 1) / (HF_BLOCK / 32));
        const size_t cap = (size_t) li.num_sms * HF_MIN_CTAS;
        const int bgrid = (int) (nct < cap ? nct : cap);
        unsigned long long *sched = d_sched;
        cudaError_t e = cudaMemsetAsync(sched, 0, sizeof(unsigned long long), s);
        if (e != cudaSuccess) return e;
        if (mode == HF_MODE_PRELIM)
            hf_intersect_beam_kernel<HF_MODE_PRELIM><<<bgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, nbig, sched);
        else if (mode == HF_MODE_FUSED_SI_LEAN)
            hf_intersect_beam_kernel<HF_MODE_FUSED_SI_LEAN><<<bgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, nbig, sched);
        else
            hf_intersect_beam_kernel<HF_MODE_FUSED_SI><<<bgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, nbig, sched);
        return cudaGetLastError();
    }
    if (li.incoherent) {
        // every lane on its own ray (hf_intersect_pool_kernel): 2 CTAs per SM, pools of HF_POOL_RAYS rays off the global counter
        const size_t smem = hf_walk_smem(li, false);
        const size_t npools = (n + HF_POOL_RAYS - 1) / HF_POOL_RAYS, nct = (npools + HF_BLOCK / 32 - 1) / (HF_BLOCK / 32), cap = (size_t) li.num_sms * HF_POOL_MIN_CTAS;
        const int pgrid = (int) (nct < cap ? nct : cap);
        cudaError_t e = cudaMemsetAsync(d_sched, 0, sizeof(unsigned long long), s);
        if (e != cudaSuccess) return e;
        switch (mode) {
            case HF_MODE_PRELIM:
                hf_intersect_pool_kernel<HF_MODE_PRELIM><<<pgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, d_sched); break;
            case HF_MODE_FUSED_SI_LEAN:
                hf_intersect_pool_kernel<HF_MODE_FUSED_SI_LEAN><<<pgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, d_sched); break;
            case HF_MODE_FUSED_SI:
                hf_intersect_pool_kernel<HF_MODE_FUSED_SI><<<pgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, d_sched); break;
            default:
                hf_intersect_pool_kernel<HF_MODE_TEST><<<pgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, d_sched); break;
        }
        return cudaGetLastError();
    }
    static bool carved_dev[64] = { false };     // function attributes are per device: one process may drive several GPUs
    int dev_ = 0;
    cudaGetDevice(&dev_);
    bool &carved = carved_dev[dev_ & 63];
    const size_t smem = hf_walk_smem(li, false);
    if (!carved) {
        carved = true;
        const size_t stat = 1024;                   // static shared memory of the kernel (level tables, barrier; adaptive schedule: + 1 KB coarse bitmap, 2112 bytes in all)
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_PRELIM, false>, smem + stat, HF_MIN_CTAS);
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_PRELIM, true>, smem + stat, HF_MIN_CTAS);
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_FUSED_SI, false>, smem + stat, HF_MIN_CTAS);
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_FUSED_SI, true>, smem + stat, HF_MIN_CTAS);
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_FUSED_SI_LEAN, false>, smem + stat, HF_MIN_CTAS);
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_FUSED_SI_LEAN, true>, smem + stat, HF_MIN_CTAS);
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_TEST, false>, smem + stat, HF_MIN_CTAS);
        hf_prefer_l1(hf_intersect_fast_kernel<HF_MODE_TEST, true>, smem + stat, HF_MIN_CTAS);
    }
    {
        const size_t ngroups = (n + 31) / 32;
        const size_t nwt = (ngroups + HF_FAST_GROUPS - 1) / HF_FAST_GROUPS, nct = (nwt + HF_BLOCK / 32 - 1) / (HF_BLOCK / 32);
        const size_t cap = (size_t) li.num_sms * HF_MIN_CTAS;
        const int fgrid = (int) (nct < cap ? nct : cap);
        // Closing phases. A group whose rays skim down a slope (tens of candidate cells per ray) keeps its warp for ~100 us, and such
        // groups are neighbours: a fetch of four lasts up to ~0.5 ms, of two ~0.25 ms (profiles/r01_notes.md). So the last
        // HF_FAST_TAIL groups per resident warp (~12 us each: ~0.3 ms) go out one at a time, the HF_FAST_MIDTAIL groups per warp
        // before them in pairs (single groups lose the L1 sharing between the two halves of a 64-spp pixel, so they are kept to
        // what the pairs need), and only the rest in fetches of HF_FAST_GROUPS.
        const size_t warps = (size_t) fgrid * (HF_BLOCK / 32);
        static const int env_tail = getenv("HF_B200_TAIL") ? atoi(getenv("HF_B200_TAIL")) : -1;          // dev knobs: groups per resident warp
        static const int env_mid = getenv("HF_B200_MIDTAIL") ? atoi(getenv("HF_B200_MIDTAIL")) : -1;
        // (with the adaptive schedule the ~100 us groups are no longer waiting in these phases: half as long is enough -- 1/8 of C3:
        // 0.529 -> 0.521 ms; 12 / 12 and 4 / 8 are slower again)
        const bool adapt = li.grp_cost && li.heavy_n && li.heavy_list && li.heavy_bits && li.cost_sum;
        const size_t tail = std::min(ngroups, warps * (size_t) (env_tail >= 0 ? env_tail : adapt ? HF_FAST_TAIL / 2 : HF_FAST_TAIL));
        const size_t mid = std::min(ngroups - tail, warps * (size_t) (env_mid >= 0 ? env_mid : adapt ? HF_FAST_MIDTAIL / 2 : HF_FAST_MIDTAIL));
        const unsigned long long nbig = (ngroups - tail - mid) / HF_FAST_GROUPS;
        const unsigned long long npair = (ngroups - tail - nbig * HF_FAST_GROUPS) / 2;
        unsigned long long *sched = d_sched;
        cudaError_t e = cudaMemsetAsync(sched, 0, sizeof(unsigned long long), s);
        if (e != cudaSuccess) return e;
        if (adapt && (e = cudaMemsetAsync(li.cost_sum, 0, sizeof(unsigned long long), s)) != cudaSuccess) return e;
        static const bool env_plain = getenv("HF_B200_AS_PLAIN_KERNEL") != nullptr;      // dev: the list is made but the plain kernel runs
#define HF_FAST_LAUNCH(M) do { \
            if (adapt && !env_plain) hf_intersect_fast_kernel<M, true><<<fgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, nbig, npair, sched); \
            else hf_intersect_fast_kernel<M, false><<<fgrid, HF_BLOCK, smem, s>>>(P, li, rays, prelim, si, d_hit, flags, n, nbig, npair, sched); } while (0)
        switch (mode) {
            case HF_MODE_PRELIM: HF_FAST_LAUNCH(HF_MODE_PRELIM); break;
            case HF_MODE_FUSED_SI_LEAN: HF_FAST_LAUNCH(HF_MODE_FUSED_SI_LEAN); break;
            case HF_MODE_FUSED_SI: HF_FAST_LAUNCH(HF_MODE_FUSED_SI); break;
            default: HF_FAST_LAUNCH(HF_MODE_TEST); break;
        }
#undef HF_FAST_LAUNCH
        if (adapt) {                                 // the list for the next launch over a batch of this size
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            return hf_launch_pick_heavy(li, (unsigned) ngroups, (unsigned) warps, s);
        }
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Stand-alone compute_surface_interaction
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(HF_BLOCK)
hf_si_kernel(const __grid_constant__ HfParams P, const hf_rays R, const hf_prelim PI, const uint32_t flags,
             const hf_si S, const size_t n) {
    for (size_t i = (size_t) blockIdx.x * HF_BLOCK + threadIdx.x; i < n; i += (size_t) gridDim.x * HF_BLOCK) {
        const f3 ow = mk3(__ldcs(R.ox + i), __ldcs(R.oy + i), __ldcs(R.oz + i));
        const f3 dw = mk3(__ldcs(R.dx + i), __ldcs(R.dy + i), __ldcs(R.dz + i));
        const float t = __ldcs(PI.t + i);
        const bool active = R.active ? (R.active[i] != 0) : true;
        if (!active || t == HF_INF) { write_si_invalid(S, i, dw, flags); continue; }
        const float u = __ldcs(PI.u + i), v = __ldcs(PI.v + i);
        const uint32_t prim = __ldcs(PI.prim_index + i);
        f3 p0, p1, p2; uint32_t i0, i1, i2;
        cell_triangle(P, prim, p0, p1, p2, i0, i1, i2);
        write_si(P, S, i, flags, ow, dw, t, u, v, p0, p1, p2);
    }
}

cudaError_t hf_launch_surface_interaction(const HfParams &P, const HfLaunchInfo &li, const hf_rays &rays,
                                          const hf_prelim &pi, uint32_t flags, const hf_si &si, size_t n, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    hf_si_kernel<<<hf_grid_for(n, li.num_sms, 8), HF_BLOCK, 0, s>>>(P, rays, pi, flags, si, n);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Adam (src/python/python/ad/optimizers.py:263-302), fp32, op order of the LLVM lowering
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void adam_one(float &p, float g, float &m, float &v, float lr_t, float b1, float b2,
                                         float omb1, float omb2, float eps, int mask) {
    float m_t = b1 * m + omb1 * g;
    float v_t = b2 * v + omb2 * (g * g);
    const bool nz = g != 0.f;
    if (mask && !nz) { m_t = m; v_t = v; }
    m = m_t; v = v_t;
    float step = __fdiv_rn(lr_t * m_t, __fsqrt_rn(v_t) + eps);
    if (mask && !nz) step = 0.f;
    p = p - step;
}

// Step counter and bias correction live on the DEVICE, so that a CUDA graph which captured an optimisation step replays a correct Adam
// (t advances with every replay): one thread bumps t and looks lr_scale(t) up in a table the host filled once per (beta1, beta2) with
// the reference's double-precision expression (optimizers.py:266-269), then lr_t = lr * lr_scale in fp32 as there.
// Beyond the table (t >= table_len) the same expression is evaluated here in double: for the default betas it has long reached 1, but for
// betas nearer 1 it has not (beta2 = 0.9999: 0.9993 at t = 65535), and the reference keeps counting (optimizers.py:266-269).
__global__ void hf_adam_prep_kernel(int *__restrict__ t, float *__restrict__ lr_t, const float *__restrict__ scale_table, const int table_len,
                                    const float lr, const double b1, const double b2) {
    const int prev = *t;
    const int step = prev < 0x7fffffff ? prev + 1 : prev;
    *t = step;
    float scale;
    if (step < table_len) scale = scale_table[step];
    else scale = (float) (sqrt(1.0 - pow(b2, (double) step)) / (1.0 - pow(b1, (double) step)));
    *lr_t = lr * scale;
}

__global__ void __launch_bounds__(256)
hf_adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v,
               const size_t n, const float *__restrict__ lr_t_ptr, const float b1, const float b2, const float omb1, const float omb2,
               const float eps, const int mask) {
    const float lr_t = __ldg(lr_t_ptr);
    const size_t n4 = n / 4;
    float4 *p4 = reinterpret_cast<float4 *>(p), *m4 = reinterpret_cast<float4 *>(m), *v4 = reinterpret_cast<float4 *>(v);
    const float4 *g4 = reinterpret_cast<const float4 *>(g);
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t) gridDim.x * blockDim.x) {
        float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
        adam_one(pp.x, gg.x, mm.x, vv.x, lr_t, b1, b2, omb1, omb2, eps, mask);
        adam_one(pp.y, gg.y, mm.y, vv.y, lr_t, b1, b2, omb1, omb2, eps, mask);
        adam_one(pp.z, gg.z, mm.z, vv.z, lr_t, b1, b2, omb1, omb2, eps, mask);
        adam_one(pp.w, gg.w, mm.w, vv.w, lr_t, b1, b2, omb1, omb2, eps, mask);
        p4[i] = pp; m4[i] = mm; v4[i] = vv;
    }
    for (size_t i = n4 * 4 + (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        float pp = p[i], mm = m[i], vv = v[i];
        adam_one(pp, g[i], mm, vv, lr_t, b1, b2, omb1, omb2, eps, mask);
        p[i] = pp; m[i] = mm; v[i] = vv;
    }
}

// ------------------------------------------------------------------------------------------
// Multi-GPU over peer memory (hf_peer_attach): the pieces that are not a scatter
// ------------------------------------------------------------------------------------------
// Cross-GPU barrier of one NVSwitch box, entirely on the device (it lives in the stream and in CUDA graphs like any kernel): every
// rank adds 1 to a counter in EVERY rank's memory (red.add at system scope over NVLink), then spins until its own counter has seen
// all ranks for this epoch. The counters only grow (epoch e ends at e * nranks), so there is nothing to reset. The system fences
// order it against the kernels before (their peer writes / atomics) and after it in the stream.
__global__ void hf_peer_barrier_kernel(unsigned *const *__restrict__ peer_counter, unsigned *__restrict__ my_counter, unsigned *__restrict__ my_epoch,
                                       const int nranks) {
    __threadfence_system();
    __shared__ unsigned s_target;
    if (threadIdx.x == 0) { const unsigned e = *my_epoch + 1u; *my_epoch = e; s_target = e * (unsigned) nranks; }
    __syncthreads();
    if ((int) threadIdx.x < nranks) atomicAdd_system(peer_counter[threadIdx.x], 1u);
    if (threadIdx.x == 0) {
        volatile unsigned *c = my_counter;
        unsigned long long spins = 0;
        while ((int) (*c - s_target) < 0) { if (++spins > (1ull << 34)) __trap(); }      // a missing rank: fail loudly instead of hanging the box
    }
    __syncthreads();
    __threadfence_system();
}

// all-gather by peer stores: this rank's slab [lo, hi) of `local` goes to the same offsets of every other rank's buffer
__global__ void __launch_bounds__(256)
hf_peer_push_kernel(const float *__restrict__ local, float *const *__restrict__ peers, const int nranks, const int rank, const size_t lo, const size_t hi) {
    const size_t n4 = (hi - lo) / 4;                      // slabs are multiples of 1024 texels, buffers 256-byte aligned
    const float4 *src = reinterpret_cast<const float4 *>(local + lo);
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t) gridDim.x * blockDim.x) {
        const float4 v = src[i];
        for (int r = 0; r < nranks; ++r)
            if (r != rank) __stcs(reinterpret_cast<float4 *>(peers[r] + lo) + i, v);
    }
    for (size_t i = lo + n4 * 4 + (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (size_t) gridDim.x * blockDim.x)
        for (int r = 0; r < nranks; ++r)
            if (r != rank) peers[r][i] = local[i];
}

// Sharded Adam fused with the all-gather of the parameter: this rank updates the texels of ITS slab (whose gradient the scatter
// kernels have already summed into this rank's grid) and stores every updated height into all ranks' height grids over NVLink.
__global__ void __launch_bounds__(256)
hf_peer_adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v, float *const *__restrict__ peer_heights,
                    const int nranks, const int rank, const size_t lo, const size_t hi, const float *__restrict__ lr_t_ptr, const float b1, const float b2,
                    const float omb1, const float omb2, const float eps, const int mask) {
    const float lr_t = __ldg(lr_t_ptr);
    const size_t n4 = (hi - lo) / 4;
    float4 *p4 = reinterpret_cast<float4 *>(p + lo), *m4 = reinterpret_cast<float4 *>(m + lo), *v4 = reinterpret_cast<float4 *>(v + lo);
    const float4 *g4 = reinterpret_cast<const float4 *>(g + lo);
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t) gridDim.x * blockDim.x) {
        float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
        adam_one(pp.x, gg.x, mm.x, vv.x, lr_t, b1, b2, omb1, omb2, eps, mask);
        adam_one(pp.y, gg.y, mm.y, vv.y, lr_t, b1, b2, omb1, omb2, eps, mask);
        adam_one(pp.z, gg.z, mm.z, vv.z, lr_t, b1, b2, omb1, omb2, eps, mask);
        adam_one(pp.w, gg.w, mm.w, vv.w, lr_t, b1, b2, omb1, omb2, eps, mask);
        p4[i] = pp; m4[i] = mm; v4[i] = vv;
        for (int r = 0; r < nranks; ++r)
            if (r != rank) __stcs(reinterpret_cast<float4 *>(peer_heights[r] + lo) + i, pp);
    }
    for (size_t i = lo + n4 * 4 + (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (size_t) gridDim.x * blockDim.x) {
        float pp = p[i], mm = m[i], vv = v[i];
        adam_one(pp, g[i], mm, vv, lr_t, b1, b2, omb1, omb2, eps, mask);
        p[i] = pp; m[i] = mm; v[i] = vv;
        for (int r = 0; r < nranks; ++r)
            if (r != rank) peer_heights[r][i] = pp;
    }
}

// ---- reduce-scatter by peer STORES (the default peer mode; remote atomics reach only ~2 G/s over NVLink, stores run at link rate) ----
// A. every rank stores the part of its LOCAL gradient grid that belongs to rank r into rank r's staging area (slot = sender).
// A rank's rays come from its image tiles, so its grid is zero over most of the terrain: 128-byte lines that hold only zeros are
// NOT sent. The staging areas are all-zero whenever no reduction is in flight (cleared at export; kernel B clears what it read).
__global__ void __launch_bounds__(256)
hf_peer_send_partials_kernel(const float *__restrict__ local, float *const *__restrict__ peer_stage, const int nranks, const int rank,
                             const size_t slab, const size_t total) {
    const size_t slab4 = slab / 4;                              // slabs are multiples of 1024 texels: `work` is a multiple of the CTA size,
    const size_t work = slab4 * (size_t) (nranks - 1);          // so the trip count is uniform over a CTA and the ballot below is safe
    const unsigned group = 0xffu << (threadIdx.x & 24u);        // the 8 lanes of one 128-byte line
    for (size_t w = (size_t) blockIdx.x * blockDim.x + threadIdx.x; w < work; w += (size_t) gridDim.x * blockDim.x) {
        const int k = (int) (w / slab4);
        const size_t i = w - (size_t) k * slab4;
        const int r = k < rank ? k : k + 1;                     // the k-th OTHER rank
        const size_t src = (size_t) r * slab + 4 * i;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (src + 4 <= total) v = *reinterpret_cast<const float4 *>(local + src);
        else if (src < total) { v.x = local[src]; v.y = src + 1 < total ? local[src + 1] : 0.f; v.z = src + 2 < total ? local[src + 2] : 0.f; }
        const bool nz = v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f;              // (NaN != 0: a poisoned grid travels)
        if (__ballot_sync(0xffffffffu, nz) & group) __stcs(reinterpret_cast<float4 *>(peer_stage[r] + (size_t) rank * slab) + i, v);
    }
}

// B. the owner sums the partial slabs in rank order (own grid in place of its own slot) and stores the sum into every rank's grid:
// reduction and all-gather in one pass. ADAM: the sum feeds the Adam update of the slab instead, and the HEIGHTS go to every rank.
// Every staged value that was not zero is cleared behind the read, which is what lets kernel A skip zero lines next time.
template <bool ADAM>
__global__ void __launch_bounds__(256)
hf_peer_reduce_gather_kernel(float *__restrict__ local_grad, float *__restrict__ my_stage, float *const *__restrict__ peer_out, const int nranks,
                             const int rank, const size_t lo, const size_t hi, const size_t slab, float *__restrict__ p, float *__restrict__ m,
                             float *__restrict__ v, const float *__restrict__ lr_t_ptr, const float b1, const float b2, const float omb1, const float omb2,
                             const float eps, const int mask) {
    const float lr_t = ADAM ? __ldg(lr_t_ptr) : 0.f;
    const size_t n4 = (hi - lo) / 4;                            // lo and the slots are multiples of 1024 texels, the buffers 256-byte aligned
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t) gridDim.x * blockDim.x) {
        // all partials first (independent loads: one memory latency, not nranks of them), then the clearing stores, then the sum in rank order
        float4 a[HF_MAX_PEERS];
        #pragma unroll
        for (int r = 0; r < HF_MAX_PEERS; ++r) {
            a[r] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < nranks) a[r] = r == rank ? reinterpret_cast<const float4 *>(local_grad + lo)[i] : __ldcs(reinterpret_cast<const float4 *>(my_stage + (size_t) r * slab) + i);
        }
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        #pragma unroll
        for (int r = 0; r < HF_MAX_PEERS; ++r) {
            if (r < nranks) {
                if (r != rank && (a[r].x != 0.f || a[r].y != 0.f || a[r].z != 0.f || a[r].w != 0.f))
                    __stcs(reinterpret_cast<float4 *>(my_stage + (size_t) r * slab) + i, make_float4(0.f, 0.f, 0.f, 0.f));
                g.x += a[r].x; g.y += a[r].y; g.z += a[r].z; g.w += a[r].w;
            }
        }
        reinterpret_cast<float4 *>(local_grad + lo)[i] = g;
        float4 outv = g;
        if (ADAM) {
            float4 pp = reinterpret_cast<float4 *>(p + lo)[i], mm = reinterpret_cast<float4 *>(m + lo)[i], vv = reinterpret_cast<float4 *>(v + lo)[i];
            adam_one(pp.x, g.x, mm.x, vv.x, lr_t, b1, b2, omb1, omb2, eps, mask);
            adam_one(pp.y, g.y, mm.y, vv.y, lr_t, b1, b2, omb1, omb2, eps, mask);
            adam_one(pp.z, g.z, mm.z, vv.z, lr_t, b1, b2, omb1, omb2, eps, mask);
            adam_one(pp.w, g.w, mm.w, vv.w, lr_t, b1, b2, omb1, omb2, eps, mask);
            reinterpret_cast<float4 *>(p + lo)[i] = pp; reinterpret_cast<float4 *>(m + lo)[i] = mm; reinterpret_cast<float4 *>(v + lo)[i] = vv;
            outv = pp;
        }
        for (int r = 0; r < nranks; ++r)
            if (r != rank) __stcs(reinterpret_cast<float4 *>(peer_out[r] + lo) + i, outv);
    }
    for (size_t i = 4 * n4 + (size_t) blockIdx.x * blockDim.x + threadIdx.x; lo + i < hi; i += (size_t) gridDim.x * blockDim.x) {      // W*H % 4 texels
        float g = 0.f;
        for (int r = 0; r < nranks; ++r) {
            if (r == rank) { g += local_grad[lo + i]; continue; }
            float *slot = my_stage + (size_t) r * slab + i;
            const float a = __ldcs(slot);
            if (a != 0.f) *slot = 0.f;
            g += a;
        }
        local_grad[lo + i] = g;
        float outv = g;
        if (ADAM) {
            float pp = p[lo + i], mm = m[lo + i], vv = v[lo + i];
            adam_one(pp, g, mm, vv, lr_t, b1, b2, omb1, omb2, eps, mask);
            p[lo + i] = pp; m[lo + i] = mm; v[lo + i] = vv;
            outv = pp;
        }
        for (int r = 0; r < nranks; ++r)
            if (r != rank) __stcs(peer_out[r] + lo + i, outv);
    }
}

cudaError_t hf_launch_peer_send_partials(const float *local, float *const *d_peer_stage, int nranks, int rank, size_t slab, size_t total, int num_sms, cudaStream_t s) {
    if (nranks < 2) return cudaSuccess;
    const size_t work = slab / 4 * (size_t) (nranks - 1), blocks = (work + 255) / 256, cap = (size_t) num_sms * 8;
    hf_peer_send_partials_kernel<<<(int) (blocks < cap ? (blocks ? blocks : 1) : cap), 256, 0, s>>>(local, d_peer_stage, nranks, rank, slab, total);
    return cudaGetLastError();
}
cudaError_t hf_launch_peer_reduce_gather(float *local_grad, float *my_stage, float *const *d_peer_out, int nranks, int rank, size_t lo, size_t hi, size_t slab,
                                         bool adam, float *p, float *m, float *v, int *d_t, float *d_lr_t, const float *d_scale_table, int table_len, float lr,
                                         float beta1, float beta2, float omb1, float omb2, float eps, int mask_updates, double beta1_d, double beta2_d, int num_sms, cudaStream_t s) {
    if (adam) hf_adam_prep_kernel<<<1, 1, 0, s>>>(d_t, d_lr_t, d_scale_table, table_len, lr, beta1_d, beta2_d);
    if (hi > lo) {
        const size_t blocks = ((hi - lo + 3) / 4 + 255) / 256, cap = (size_t) num_sms * 8;
        const int grid = (int) (blocks < cap ? (blocks ? blocks : 1) : cap);
        if (adam) hf_peer_reduce_gather_kernel<true><<<grid, 256, 0, s>>>(local_grad, my_stage, d_peer_out, nranks, rank, lo, hi, slab, p, m, v, d_lr_t, beta1, beta2, omb1, omb2, eps, mask_updates);
        else hf_peer_reduce_gather_kernel<false><<<grid, 256, 0, s>>>(local_grad, my_stage, d_peer_out, nranks, rank, lo, hi, slab, nullptr, nullptr, nullptr, nullptr, 0.f, 0.f, 0.f, 0.f, 0.f, 0);
    }
    return cudaGetLastError();
}

cudaError_t hf_launch_peer_barrier(unsigned *const *d_peer_counter, unsigned *d_my_counter, unsigned *d_my_epoch, int nranks, cudaStream_t s) {
    hf_peer_barrier_kernel<<<1, 32, 0, s>>>(d_peer_counter, d_my_counter, d_my_epoch, nranks);
    return cudaGetLastError();
}
cudaError_t hf_launch_peer_push(const float *local, float *const *d_peers, int nranks, int rank, size_t lo, size_t hi, int num_sms, cudaStream_t s) {
    if (hi <= lo) return cudaSuccess;
    const size_t blocks = ((hi - lo) / 4 + 255) / 256, cap = (size_t) num_sms * 4;
    hf_peer_push_kernel<<<(int) (blocks < cap ? (blocks ? blocks : 1) : cap), 256, 0, s>>>(local, d_peers, nranks, rank, lo, hi);
    return cudaGetLastError();
}
cudaError_t hf_launch_peer_adam(float *p, const float *g, float *m, float *v, float *const *d_peer_heights, int nranks, int rank, size_t lo, size_t hi,
                                int *d_t, float *d_lr_t, const float *d_scale_table, int table_len, float lr, float beta1, float beta2, float omb1,
                                float omb2, float eps, int mask_updates, double beta1_d, double beta2_d, int num_sms, cudaStream_t s) {
    hf_adam_prep_kernel<<<1, 1, 0, s>>>(d_t, d_lr_t, d_scale_table, table_len, lr, beta1_d, beta2_d);
    if (hi > lo) {
        const size_t blocks = ((hi - lo) / 4 + 255) / 256, cap = (size_t) num_sms * 4;
        hf_peer_adam_kernel<<<(int) (blocks < cap ? (blocks ? blocks : 1) : cap), 256, 0, s>>>(p, g, m, v, d_peer_heights, nranks, rank, lo, hi, d_lr_t, beta1,
                                                                                            beta2, omb1, omb2, eps, mask_updates);
    }
    return cudaGetLastError();
}

cudaError_t hf_launch_adam(float *p, const float *g, float *m, float *v, size_t n, int *d_t, float *d_lr_t, const float *d_scale_table,
                           int table_len, float lr, float beta1, float beta2, float omb1, float omb2, float eps, int mask_updates, double beta1_d, double beta2_d,
                           int num_sms, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    hf_adam_prep_kernel<<<1, 1, 0, s>>>(d_t, d_lr_t, d_scale_table, table_len, lr, beta1_d, beta2_d);
    size_t blocks = (n / 4 + 255) / 256;
    size_t cap = (size_t) num_sms * 8;
    int grid = (int) (blocks < cap ? (blocks ? blocks : 1) : cap);
    hf_adam_kernel<<<grid, 256, 0, s>>>(p, g, m, v, n, d_lr_t, beta1, beta2, omb1, omb2, eps, mask_updates);
    return cudaGetLastError();
}
