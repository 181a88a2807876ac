// hf_qpyr.h -- the QUANTISED min/max pyramid the production walk reads (hf_traverse.cuh, walk v2).
//
// One 32-bit node per (level, cell block): low half = code of a lower bound of min z, high half = code of an upper
// bound of max z, z = max_height * h. Codes are 16-bit fixed point over the field's global z range:
//     decode(c) = fmaf(float(c), qs, zb),   (zb, qs) = hf_quant_make(top_min, top_max)
// Encoding is CONSERVATIVE BY CONSTRUCTION: the encoder evaluates the very same decode expression and moves the code
// until decode(lo) <= min and decode(hi) >= max hold in fp32, so a walk that prunes on decoded bounds can never skip a
// cell the exact bounds would keep. What quantisation costs is a slack of <= 2 codes (range / 32640) per bound.
//
// Unlike the fp32 pyramid (levels 1..Ltop, tight pitch, read by the bundle march and the counting kernel), this one
//   * has a level 0 (one node per CELL), so the walk's loop body is the same load + decode at every level and the
//     four texels of a cell are fetched only for actual candidates;
//   * is padded to a power of two per axis (Mx, My >= NX, NY; level l has pitch max(Mx >> l, 1)), so that a ray
//     travelling towards -x can be walked in MIRRORED cell coordinates ix' = Mx-1-ix with node boundaries still at
//     multiples of 2^l: the walk then only ever moves towards +x', +y' and needs no per-axis direction selects.
//     Padding nodes hold HF_Q_EMPTY (inverted range).
// Compiles as CUDA and as plain C++ (host tests build the same pyramid with the same functions).
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define HF_QHD __host__ __device__ __forceinline__
#else
#define HF_QHD static inline
#endif

#define HF_Q_EMPTY 0x0000FFFFu          // lo = 65535, hi = 0
#define HF_Q_MAX_LEVELS 26

struct HfQuant { float zb, qs, rqs; };   // rqs = 1 / qs: the encoder's first guess (the verify loops below make it exact)

HF_QHD HfQuant hf_quant_make(float top_min, float top_max) {
    HfQuant q;
    float qs = (top_max - top_min) * (1.0f / 65280.0f);       // 255 codes of headroom split below / above the range
    // flat (or degenerate) field: every code decodes to ~top_min. The floor is one fp32 ulp of the field's own magnitude (finer codes
    // resolve nothing) and keeps 1 / qs, which the walk multiplies ray heights by (hf_traverse.cuh, HF_CODE_SPACE), far from overflow
    const float qmin = fmaxf(5.9604644775390625e-08f * fmaxf(fabsf(top_min), fabsf(top_max)), 1e-20f);
    if (!(qs > qmin) || !(qs < 1e30f)) qs = qmin;
    q.qs = qs;
    q.rqs = 1.0f / qs;
    q.zb = top_min - 128.0f * qs;
    return q;
}

HF_QHD float hf_quant_decode(uint32_t code, float qs, float zb) { return fmaf((float) code, qs, zb); }

HF_QHD uint32_t hf_quant_lo(float mn, HfQuant q) {            // largest convenient code with decode(code) <= mn
    float f = floorf((mn - q.zb) * q.rqs) - 1.0f;
    int c = !(f > 0.f) ? 0 : (f > 65535.f ? 65535 : (int) f);
    while (c > 0 && !(hf_quant_decode((uint32_t) c, q.qs, q.zb) <= mn)) --c;
    return (uint32_t) c;
}
HF_QHD uint32_t hf_quant_hi(float mx, HfQuant q) {            // smallest convenient code with decode(code) >= mx
    float f = ceilf((mx - q.zb) * q.rqs) + 1.0f;
    int c = !(f > 0.f) ? 0 : (f > 65535.f ? 65535 : (int) f);
    while (c < 65535 && !(hf_quant_decode((uint32_t) c, q.qs, q.zb) >= mx)) ++c;
    return (uint32_t) c;
}
HF_QHD uint32_t hf_quant_node(float mn, float mx, HfQuant q) { return (hf_quant_hi(mx, q) << 16) | hf_quant_lo(mn, q); }
HF_QHD uint32_t hf_quant_merge(uint32_t a, uint32_t b) {      // parent of two nodes (also absorbs HF_Q_EMPTY)
    const uint32_t lo = (a & 0xffffu) < (b & 0xffffu) ? (a & 0xffffu) : (b & 0xffffu);
    const uint32_t hi = (a >> 16) > (b >> 16) ? (a >> 16) : (b >> 16);
    return (hi << 16) | lo;
}

// Layout shared by the builder (hf_kernels.cu), the host API (hf_api.cu) and the host tests.
struct HfQLayout {
    int Mx, My, Ltop;                       // padded cell counts (powers of two), top level (one node)
    uint32_t off[HF_Q_MAX_LEVELS];          // node offset of level l
    int pitch[HF_Q_MAX_LEVELS];             // nodes per row of level l
    int rows[HF_Q_MAX_LEVELS];
    size_t nodes;                           // total node count
};
static inline HfQLayout hf_qlayout(int NX, int NY) {
    HfQLayout q;
    int lx = 0, ly = 0;
    while ((1 << lx) < NX) ++lx;
    while ((1 << ly) < NY) ++ly;
    q.Mx = 1 << lx; q.My = 1 << ly; q.Ltop = lx > ly ? lx : ly;
    if (q.Ltop < 1) q.Ltop = 1;
    size_t o = 0;
    for (int l = 0; l < HF_Q_MAX_LEVELS; ++l) { q.off[l] = 0; q.pitch[l] = 1; q.rows[l] = 1; }
    for (int l = 0; l <= q.Ltop; ++l) {
        q.pitch[l] = (q.Mx >> l) > 0 ? (q.Mx >> l) : 1;
        q.rows[l] = (q.My >> l) > 0 ? (q.My >> l) : 1;
        q.off[l] = (uint32_t) o;
        o += ((size_t) q.pitch[l] * q.rows[l] + 3) & ~size_t(3);      // 16-byte aligned levels
    }
    q.nodes = o;
    return q;
}
