// hf_traverse.cuh -- the production traversal: warp-uniform hierarchical DDA over the min/max pyramid.
//
// Same geometry as the canonical traversal that defines the byte model (oracle/hf_oracle.c::traverse, mirrored on the
// device by traverse_canonical in hf_kernels.cu); what changes is the control structure and the data it reads, which is
// what the SMs pay for (the kernel is bound by instruction issue, profiles/r01_notes.md):
//   * ONE loop body for "descend", "step" and "leaf": every lane of a warp executes the same instructions each
//     iteration and updates its state through selects (the canonical form serialises three divergent paths);
//   * the loop reads the QUANTISED, power-of-two padded pyramid of hf_qpyr.h, which has a level 0: one 4-byte load and
//     the same decode at every level; the four texels of a cell are fetched only when the cell becomes a candidate;
//   * the ray is walked in MIRRORED cell coordinates (ix' = Mx-1-ix when it travels towards -x, same for y) and in a
//     FLIPPED z (z' = -z when it climbs): it then always moves towards +x', +y', -z', so the loop has no per-axis
//     direction selects and no min/max of the two z end points;
//   * the walk starts at level min(Ltop, HF_START_LEVEL) instead of the root (the root-side levels always overlap
//     at the entry point), and after a step it may pop several levels at once (HF_POP_LEVELS) when the step crossed
//     a coarse node boundary;
//   * no per-iteration clamps of the node bounds to the grid: the ray's [t0, t1] is already clipped to the grid slab;
//   * a descending lane clips its entry point to the node's z slab [min, max] (the ray-vs-box clip the root gets, applied
//     at every level), which skips the children in front of the point where the ray can first touch the node's content.
// A lane leaves the loop when it holds a candidate cell whose z range overlaps the ray (the caller then runs the
// per-cell primitive for all such lanes together) or when the ray is finished.
//
// The file compiles as CUDA (device code) and as plain C++ (host): the host build exists ONLY so that tests/ can
// check the walk against brute force on millions of rays without a GPU (tests/trav_host.cpp). The shipped library
// never calls the host instantiation.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>
#include "hf_qpyr.h"

#ifndef HF_STAGE_Q
#define HF_STAGE_Q 1
#endif
#ifdef __CUDACC__
#define HF_HD __device__ __forceinline__
#define HF_LDG(p) __ldg(p)
#define HF_RCP(x) __frcp_rn(x)
#else
#define HF_RCP(x) (1.0f / (x))
#define HF_HD static inline
#define HF_LDG(p) (*(p))
struct float2 { float x, y; };
#endif

#ifndef HF_START_LEVEL
#define HF_START_LEVEL 8
#endif
#ifndef HF_POP_LEVELS
#define HF_POP_LEVELS 1
#endif
#ifndef HF_DESC_LEVELS
#define HF_DESC_LEVELS 2      // levels dropped per descent (the slab clip places the entry cell, so skipping a level is cheap)
#endif

#ifndef HF_EVEN_LEVELS
#define HF_EVEN_LEVELS 1      // 1: the walk only visits even levels (descends and pops two levels at a time)
#endif

#define HF_TRAV_SLACK 9.5367431640625e-07f   // 2^-20, identical to HF_Z_SLACK / HFO_Z_SLACK

// HF_CODE_SPACE = 1: the walk compares the ray's z interval with the node's 16-bit bounds IN CODE UNITS instead of decoding the two
// bounds to object-space floats at every node. z'(t) is carried as  M + (z - zb) / qs  with M = 2^23 (sign-flipped for a climbing
// ray), and a code c becomes the float M + c by ONE byte permute that puts the code into the mantissa of 0x4B000000 (the classic
// int -> float magic number; 0xCB000000 for the flipped sign): per node 2 PRMT + 2 FSETP instead of 2 PRMT + 2 I2F + 2 FFMA + 2 FSETP,
// and nothing left on the conversion (XU) pipe. Conservative: one extra code of slack pays for the integer rounding at M (below).
#ifndef HF_CODE_SPACE
#define HF_CODE_SPACE 1
#endif
#define HF_CODE_MAGIC 8388608.0f             // 2^23: float(M + c) has the bit pattern 0x4B000000 | c for 0 <= c < 2^23

struct __attribute__((aligned(16))) HfLevel {   // one pyramid level; `nodes` may point to global OR shared memory (generic address)
    const float2 *nodes;       // row-major (min,max) of z = max_height*h
    int w;                     // nodes per row
    int pad;
};

struct __attribute__((aligned(16))) HfQLevel {  // one level of the quantised pyramid; `nodes` is a generic address (staged coarse levels: shared memory)
    const uint32_t *nodes;     // row-major, (hi << 16) | lo codes
    int pitch;                 // nodes per row (a power of two)
    int pad;
};

#ifdef __CUDACC__
#define HF_LVL_ENTRIES 26
__shared__ HfLevel hf_s_lvl[HF_LVL_ENTRIES];      // filled by the kernel prologue
__shared__ HfQLevel hf_s_qlvl[HF_LVL_ENTRIES];    // level 0..Ltop of the quantised pyramid
#endif

struct HfGrid {                // what the walk needs to know about the shape
    int W, NX, NY, Ltop;
    float mh;                  // max_height
    float x0, y0, sgx, sgy;    // object -> grid units
    const float *heights;      // (H, W)
    const HfLevel *lvl;        // [level 1..Ltop]; the table itself sits in shared memory on the device
    float top_min, top_max;    // root node
    // quantised pyramid (hf_qpyr.h): what hf_walk_next reads
    const HfQLevel *qlvl;      // [level 0..QLtop] (host build; the device reads hf_s_qlvl through qlvl_saddr)
    uint32_t qlvl_saddr;       // shared-memory address of hf_s_qlvl (device)
    int QLtop;                 // top level of the quantised pyramid (one node)
    int xmask, ymask;          // Mx - 1, My - 1 (padded cell counts are powers of two)
    float qs, zb;              // decode(c) = fmaf(float(c), qs, zb)
    float rqs;                 // 1 / qs (HfQuant::rqs)
};

struct HfWalk {                // per-ray state, lives in registers. x, y are MIRRORED grid units, z is FLIPPED (see top)
    float ogx, ogy, dgx, dgy;  // dgx, dgy >= 0
    float ogxl, ogyl;          // low parts of the mirrored origin: Mx - og is kept EXACTLY as ogx + ogxl (a rounded mirror image is a
                               // different ray: next to a cell boundary a nearly axis-parallel ray would leave its row far too early)
    float dz;                  // <= 0  (HF_CODE_SPACE: in code units per unit of t)
    float idx, idy;            // 1/dg; after hf_walk_init: 0 when the component is zero ...
    float infx, infy;          // ... and +inf here (0 otherwise), so that t_exit = fma(b - og, id, inf) needs no select
    float t1;
    float t;                   // ray parameter at the entry of the current node
    float idz;                 // 1/dz (0 when dz == 0)
    float ozp, ozm;            // oz + epsz, oz - epsz: z(t) padded up / down by the walk's z slack (HF_CODE_SPACE: code units, offset by +-M)
    float ogxb, ogyb;          // og backed off against the direction of travel: cells derived from t are never ahead of the ray
#if HF_CODE_SPACE
    uint32_t kmagic;           // 0x4B000000 (M + code) or 0xCB000000 (-(M + code), flipped ray): upper bytes of a node bound as a float
#else
    float qsf, zbf;            // decode constants of the quantised nodes, negated for a flipped ray
#endif
    int ix, iy, L;             // mirrored cell, level
    int xm, ym;                // mirror masks: real cell = (ix ^ xm, iy ^ ym)
    int xlim, ylim;            // last real cell in mirrored coordinates: beyond it lies padding (or the end of the padded domain)
    uint32_t sel_lo;           // byte selector (prmt) of the node half that bounds z' from below
    bool alive;
};

struct HfCand {                // a leaf cell whose z range overlaps the ray
    int cx, cy;
    float z00, z10, z01, z11;
};

HF_HD int hf_floor_sat(float x) {
#ifdef __CUDACC__
    return __float2int_rd(x);                      // saturating; NaN -> 0
#else
    if (!(x == x)) return 0;
    if (x >= 2147483520.0f) return 2147483647;
    if (x <= -2147483648.0f) return (-2147483647 - 1);
    return (int) floorf(x);
#endif
}
HF_HD int hf_imin(int a, int b) { return a < b ? a : b; }
HF_HD int hf_imax(int a, int b) { return a > b ? a : b; }
HF_HD int hf_msb(unsigned v) {                     // index of the highest set bit, v != 0
#ifdef __CUDACC__
    return 31 - __clz((int) v);
#else
    return 31 - __builtin_clz(v);
#endif
}

// Clip the object-space ray against the shape's box and set up the walk. Returns false when the ray cannot hit.
// Arithmetic identical to the canonical traversal (slab semantics of include/mitsuba/core/bbox.h:303-327).
// t_start / L_start: the walk may begin at a ray parameter > 0 and at a finer level when the caller knows that the ray
// cannot touch the surface before t_start (hf_beam_enter below); t_start = 0, L_start = HF_START_LEVEL is the plain walk.
HF_HD bool hf_walk_init(const HfGrid &G, float ox, float oy, float oz, float dx, float dy, float dz, float maxt, HfWalk &w,
                        float t_start = 0.f, int L_start = HF_START_LEVEL, bool force_level = false) {
    w.alive = false;
    if (dx == 0.f && dy == 0.f && dz == 0.f) return false;
    if (!(maxt >= 0.f)) return false;
    const int NX = G.NX, NY = G.NY;
    float ogx = (ox - G.x0) * G.sgx, ogy = (oy - G.y0) * G.sgy;
    float dgx = dx * G.sgx, dgy = dy * G.sgy;
    float idx = HF_RCP(dgx), idy = HF_RCP(dgy);             // correctly rounded on both sides (== 1.0f / x)
    float idz = HF_RCP(dz);
    float t0 = 0.f, t1 = maxt;
    if (dgx != 0.f) {
        float a = (0.f - ogx) * idx, b = ((float) NX - ogx) * idx;
        t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    } else if (ogx < 0.f || ogx > (float) NX) return false;
    if (dgy != 0.f) {
        float a = (0.f - ogy) * idy, b = ((float) NY - ogy) * idy;
        t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    } else if (ogy < 0.f || ogy > (float) NY) return false;
    {
        const float pad = HF_TRAV_SLACK * fmaxf(fabsf(G.top_min), fabsf(G.top_max)) + HF_TRAV_SLACK * fabsf(oz);
        if (dz != 0.f) {
            float a = (G.top_min - pad - oz) * idz, b = (G.top_max + pad - oz) * idz;
            float tn = fminf(a, b), tf = fmaxf(a, b);
            tn -= fabsf(tn) * HF_TRAV_SLACK; tf += fabsf(tf) * HF_TRAV_SLACK;
            t0 = fmaxf(t0, tn); t1 = fminf(t1, tf);
        } else if (oz < G.top_min - pad || oz > G.top_max + pad) return false;
    }
    if (!(t0 <= t1)) return false;
    w.t1 = t1;
    const float epsz = HF_TRAV_SLACK * (fabsf(oz) + fabsf(t1 * dz));
    // entry cell in REAL coordinates (same arithmetic as the canonical walk), then mirrored below
    int ix, iy;
    const bool late = t_start > t0;          // nothing can be hit before t_start: enter the grid there, one cell back at most
    if (late) {
        if (!(t_start <= t1)) return false;
        t0 = t_start;
    }
    {   // 6 ulp of the largest grid coordinate the ray reaches (in either orientation): a cell derived from a ray parameter
        // is never ahead of the true one; covers the one extra rounding of the mirrored origin
        const float ex = 3.5762786865234375e-07f * ((float) (G.xmask + 1) + fabsf(ogx) + fabsf(t1 * dgx));
        const float ey = 3.5762786865234375e-07f * ((float) (G.ymask + 1) + fabsf(ogy) + fabsf(t1 * dgy));
        // (no back-off along an axis the ray does not move on: its cell there is exact, and a ray lying in a cell boundary
        // plane keeps the side the walk from t = 0 is on)
        const float bx0 = late && dgx != 0.f ? (dgx > 0.f ? ogx - ex : ogx + ex) : ogx;
        const float by0 = late && dgy != 0.f ? (dgy > 0.f ? ogy - ey : ogy + ey) : ogy;
        ix = hf_imin(hf_imax(hf_floor_sat(fmaf(t0, dgx, bx0)), 0), NX - 1);
        iy = hf_imin(hf_imax(hf_floor_sat(fmaf(t0, dgy, by0)), 0), NY - 1);
        // mirror: the walk only ever moves towards +x', +y'
        const bool mx_ = dgx < 0.f, my_ = dgy < 0.f;
        w.xm = mx_ ? G.xmask : 0; w.ym = my_ ? G.ymask : 0;
        w.xlim = mx_ ? G.xmask : NX - 1; w.ylim = my_ ? G.ymask : NY - 1;   // a mirrored ray has the padding behind it
        float ogxl = 0.f, ogyl = 0.f;        // low parts of the mirrored origin: Mx - og is kept EXACTLY as ogx + ogxl (a rounded mirror image
                                             // is a different ray: next to a cell boundary a nearly axis-parallel ray would leave its row far too early)
        if (mx_) {                           // two-sum: M + (-og) = s + e exactly
            const float a = (float) (G.xmask + 1), b = -ogx, sm = a + b, bb = sm - a;
            ogxl = (a - (sm - bb)) + (b - bb); ogx = sm; dgx = -dgx; idx = -idx;
        }
        if (my_) {
            const float a = (float) (G.ymask + 1), b = -ogy, sm = a + b, bb = sm - a;
            ogyl = (a - (sm - bb)) + (b - bb); ogy = sm; dgy = -dgy; idy = -idy;
        }
        w.dgx = dgx; w.dgy = dgy;
        w.ogx = ogx; w.ogy = ogy; w.ogxl = ogxl; w.ogyl = ogyl;
        w.ogxb = ogx - ex; w.ogyb = ogy - ey;
        w.ix = ix ^ w.xm; w.iy = iy ^ w.ym;
    }
    // force_level: the caller knows the ray starts next to the surface (a visibility ray leaving a hit point): start low and climb
    w.L = hf_imin(G.QLtop, (late || force_level) ? L_start : HF_START_LEVEL);
#if HF_EVEN_LEVELS
    w.L &= ~1;
#endif
    w.t = t0;
    {   // flip z for a climbing ray: the walk only ever moves towards -z'
        const bool fz = dz > 0.f;
#if HF_CODE_SPACE
        // code units: zc = (z - zb) * rqs. What must hold for every node the walk prunes: the padded interval [zlo, zhi] contains the
        // true z range of the ray over [t, te] as measured against the bounds the ENCODER guaranteed in fp32 (decode(lo) <= min,
        // decode(hi) >= max with decode = fmaf(code, qs, zb)). Slack, in codes:
        //   * the object-space slack of the float walk, scaled: 2^-20 (|oz - zb| + |t1 dz|) rqs  (rounding of z(t), of t, of rqs);
        //   * the rounding of decode itself, which the float compare shared with the encoder and this one does not: 2^-23 |z|max rqs;
        //   * 1.25: the offset M makes the two roundings (of oz' +- eps + M, and of the fma per node) integer roundings, 0.5 each.
        const float rqs = G.rqs;
        const float oc = (oz - G.zb) * rqs, dzc = dz * rqs;
        const float epsc = HF_TRAV_SLACK * (fabsf(oc) + fabsf(t1 * dzc)) +
                           1.1920928955078125e-07f * fmaxf(fabsf(G.top_min), fabsf(G.top_max)) * rqs + 1.25f;
        const float ocf = fz ? -oc : oc, sM = fz ? -HF_CODE_MAGIC : HF_CODE_MAGIC;
        w.dz = fz ? -dzc : dzc;
        w.idz = dz != 0.f ? (fz ? -idz : idz) * G.qs : 0.f;
        w.ozp = (ocf + epsc) + sM; w.ozm = (ocf - epsc) + sM;
        w.kmagic = fz ? 0xCB000000u : 0x4B000000u;
        w.sel_lo = fz ? 0x7632u : 0x7610u;       // the node half that bounds z' from below: lo codes, or the hi codes of a flipped ray
        (void) epsz;
#else
        const float ozf = fz ? -oz : oz;
        w.dz = fz ? -dz : dz;
        w.idz = dz != 0.f ? (fz ? -idz : idz) : 0.f;
        w.ozp = ozf + epsz; w.ozm = ozf - epsz;
        w.qsf = fz ? -G.qs : G.qs; w.zbf = fz ? -G.zb : G.zb;
        w.sel_lo = fz ? 0x4432u : 0x4410u;
#endif
    }
    const float inf = INFINITY;
    w.infx = dgx != 0.f ? 0.f : inf; w.infy = dgy != 0.f ? 0.f : inf;
    w.idx = dgx != 0.f ? idx : 0.f;
    w.idy = dgy != 0.f ? idy : 0.f;
    w.alive = true;
    return true;
}

#ifdef HF_DEBUG_BOUNDS
#ifdef __CUDACC__
#define HF_BOUNDS_FAIL() __trap()
#else
#include <stdlib.h>
#define HF_BOUNDS_FAIL() abort()
#endif
#endif

HF_HD uint32_t hf_prmt(uint32_t q, uint32_t k, uint32_t sel) {      // __byte_perm(q, k, sel) for the selectors the walk uses
#ifdef __CUDACC__
    return __byte_perm(q, k, sel);
#else
    const uint32_t half = ((sel & 0xffu) == 0x10u) ? (q & 0xffffu) : (q >> 16);
    return (sel & 0xff00u) == 0x7600u ? ((k & 0xffff0000u) | half) : half;      // 0x76xx: upper two bytes from k
#endif
}
HF_HD float hf_bits_float(uint32_t b) {
#ifdef __CUDACC__
    return __uint_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}

// ONE iteration of the walk (the loop body of hf_walk_next): 0 = keep going, 1 = a leaf cell overlaps the ray's z interval (cell and
// texels in `c`; the state has already moved past that cell, so a failed primitive test simply continues), 2 = the ray is finished.
// Kernels that keep every lane of a warp on its own ray (hf_intersect_pool_kernel) call this directly.
HF_HD int hf_walk_iter(const HfGrid &G, HfWalk &w, HfCand &c) {
    const uint32_t sel_hi = w.sel_lo ^ 0x0022u;
    const int L = w.L, ix = w.ix, iy = w.iy;
    const int mask = (1 << L) - 1;
    const int hx = ix | mask, hy = iy | mask;                 // last cell of this node on each axis
    const int bx = hx + 1, by = hy + 1;                       // exit planes of this node
    // (NOT fma(float(b), id, -og * id): for a nearly axis-parallel ray id is huge and the two products cancel catastrophically;
    // the subtraction below is exact next to the ray -- tests/test_traversal_host.py::test_walk_under_an_affine_transform)
    const float tx = fmaf(((float) bx - w.ogx) - w.ogxl, w.idx, w.infx); // the fma only adds +0 or +inf
    const float ty = fmaf(((float) by - w.ogy) - w.ogyl, w.idy, w.infy);
    const float tmin = fminf(tx, ty);
    const bool last = !(tmin < w.t1);
    const float te = fminf(tmin, w.t1);
    // z' never increases along the ray: its padded range over [t, te] is [z'(te) - eps, z'(t) + eps]
    const float zhi = fmaf(w.t, w.dz, w.ozp), zlo = fmaf(te, w.dz, w.ozm);

#ifdef HF_DEBUG_BOUNDS
    // checked build (make VARIANT=_bounds EXTRA=-DHF_DEBUG_BOUNDS; tests/trav_host.cpp defines it too): every node and texel
    // index the walk forms is tested against the layout before it is used (compute-sanitizer is not available on the GPU pool)
    if (L < 0 || L > G.QLtop || ((ix ^ w.xm) >> L) < 0 || ((ix ^ w.xm) >> L) > (G.xmask >> L) ||
        ((iy ^ w.ym) >> L) < 0 || ((iy ^ w.ym) >> L) > (G.ymask >> L)) HF_BOUNDS_FAIL();
#endif
    uint32_t q;
    {
#ifdef __CUDACC__
        uint32_t p0, p1, pitch, pad_ = 0u;
        asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(p0), "=r"(p1), "=r"(pitch), "=r"(pad_) : "r"(G.qlvl_saddr + (uint32_t) L * 16u));
        (void) pad_;
        const uint32_t *nodes = reinterpret_cast<const uint32_t *>(((unsigned long long) p1 << 32) | (unsigned long long) p0);
#if HF_STAGE_Q
        q = nodes[(unsigned) (((iy ^ w.ym) >> L) * (int) pitch + ((ix ^ w.xm) >> L))];       // generic: shared (staged coarse levels) or global
#else
        q = HF_LDG(nodes + (unsigned) (((iy ^ w.ym) >> L) * (int) pitch + ((ix ^ w.xm) >> L)));
#endif
#else
        const HfQLevel lv = G.qlvl[L];
        q = lv.nodes[(unsigned) (((iy ^ w.ym) >> L) * lv.pitch + ((ix ^ w.xm) >> L))];
#endif
    }
#if HF_CODE_SPACE
    const float mn = hf_bits_float(hf_prmt(q, w.kmagic, w.sel_lo));     // +-(M + code): no conversion, no decode
    const float mx = hf_bits_float(hf_prmt(q, w.kmagic, sel_hi));
#else
    const float mn = fmaf((float) hf_prmt(q, 0u, w.sel_lo), w.qsf, w.zbf);
    const float mx = fmaf((float) hf_prmt(q, 0u, sel_hi), w.qsf, w.zbf);
#endif
    const bool overlap = (zlo <= mx) && (zhi >= mn);
    const bool desc = overlap && (L > 0);
    const bool found = overlap && (L == 0);

    // Descending lanes first clip their entry point to the node's z slab: before the ray crosses mx + eps it
    // cannot touch anything inside this node, so the children in front of that point are skipped.
    float ts = (mx - w.ozm) * w.idz;
    ts = fmaf(-fabsf(ts), HF_TRAV_SLACK, ts);                         // never later than the true crossing
    const float tc = fminf(fmaxf(w.t, ts), te);
    // tq: the parameter the next cell is derived from -- the clipped entry (descend) or the node's exit (step)
    const float tq = desc ? tc : te;
    const int qx = hf_floor_sat(fmaf(tq, w.dgx, w.ogxb)), qy = hf_floor_sat(fmaf(tq, w.dgy, w.ogyb));
    // position at tq, kept inside the node and monotone along the ray
    const int rx = hf_imin(hf_imax(qx, ix), hx), ry = hf_imin(hf_imax(qy, iy), hy);
    // a step leaves the node through the plane(s) reached first; a descent stays inside
    const bool stepx = !desc && (tx <= ty), stepy = !desc && (ty <= tx);
    const int nix = stepx ? bx : rx, niy = stepy ? by : ry;
    const bool outside = (nix > w.xlim) || (niy > w.ylim);            // never enters a padding cell (their nodes hold HF_Q_EMPTY)
    const unsigned m = (unsigned) ((nix ^ ix) | (niy ^ iy));
#if HF_EVEN_LEVELS
    const bool up = m > (unsigned) (4 * mask + 3);                    // left the grandparent node: go up two levels
#else
    const bool up = m > (unsigned) (2 * mask + 1);                    // left the parent node: go up one level
#endif

#if !defined(__CUDACC__) && defined(HF_WALK_HIST)
    { extern unsigned long long hf_walk_hist[3][32];
      _Pragma("omp atomic") hf_walk_hist[desc ? 0 : (found ? 2 : 1)][L]++; }
#endif
#if HF_EVEN_LEVELS
    w.L = desc ? L - 2 : (up ? L + 2 : L);                            // levels stay even: a 4 x 4-ary walk
#else
    w.L = desc ? hf_imax(L - HF_DESC_LEVELS, 0) : (up ? L + 1 : L);
#endif
    w.ix = nix;
    w.iy = niy;
    w.t = tq;
    w.alive = desc ? true : !(last || outside);
    if (found) {
        const int cx = ix ^ w.xm, cy = iy ^ w.ym;                     // always a real cell: see `outside`
#ifdef HF_DEBUG_BOUNDS
        if (cx < 0 || cx >= G.NX || cy < 0 || cy >= G.NY) HF_BOUNDS_FAIL();
#endif
        const float *row0 = G.heights + (unsigned) (cy * G.W + cx);   // W*H < 2^32 (checked by hf_create)
        c.z00 = G.mh * HF_LDG(row0); c.z10 = G.mh * HF_LDG(row0 + 1);
        c.z01 = G.mh * HF_LDG(row0 + G.W); c.z11 = G.mh * HF_LDG(row0 + G.W + 1);
        c.cx = cx; c.cy = cy;
        return 1;
    }
    return w.alive ? 0 : 2;
}

// Advance the walk until a leaf cell overlaps the ray's z interval (returns true, cell in `c`, and the state is
// already moved past that cell so that a failed primitive test simply calls again) or the ray is finished (false).
// `iters` (host statistics only) counts loop iterations.
HF_HD bool hf_walk_next(const HfGrid &G, HfWalk &w, HfCand &c
#ifndef __CUDACC__
                        , unsigned *iters = nullptr
#endif
) {
    if (!w.alive) return false;
    for (;;) {
#ifndef __CUDACC__
        if (iters) ++*iters;
#endif
        const int st = hf_walk_iter(G, w, c);
        if (st == 1) return true;
        if (st == 2) return false;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Bundle entry: a conservative lower bound of the first-hit parameter shared by a bundle of rays with ONE origin
// (primary rays of a pixel / an image tile: the 32 consecutive rays a warp loads differ only by a sub-pixel jitter).
//
// One lane marches the whole bundle with interval arithmetic: over the parameter segment [t, te] the bundle sweeps the
// grid-space box  og + [t, te] * [d_lo, d_hi]  (s >= 0, so the extremes sit at the segment ends); the box is covered by at
// most 2 x 2 nodes of a pyramid level whose node edge is at least the box edge, and if the box's z range clears the
// (min, max) of those nodes no ray of the bundle can touch a triangle there. The segment doubles after two successes and
// halves (one level down) after a failure; the march stops when the level is too fine for the bundle's own width. Every
// ray of the bundle then starts its own walk at the returned parameter, a few cells in front of its hit, instead of at
// the grid boundary -- which removes most iterations of the per-ray walk and most of the divergence between lanes.
// Conservative by construction: padding of 1/32 cell + 2^-18 relative on the footprint, HF_TRAV_SLACK on z.
// ------------------------------------------------------------------------------------------------------------------
#ifndef HF_BEAM_START_LEVEL
#define HF_BEAM_START_LEVEL 8
#endif
#ifndef HF_BEAM_MAX_ITERS
#define HF_BEAM_MAX_ITERS 96
#endif
#ifndef HF_BEAM_ROOM_FRAC
#define HF_BEAM_ROOM_FRAC 0.25f
#endif
#define HF_BEAM_ALL_MISS 3.0e38f

struct HfBeam {
    float ogx, ogy, oz;                     // common origin: grid units in x, y; object space z
    float dxl, dxh, dyl, dyh, dzl, dzh;     // component-wise bounds of (dgx, dgy, dz) over the bundle
};

// slab [lo, hi] against origin o and direction interval [dl, dh]: narrows [ta, tb] to parameters at which SOME ray of the
// bundle can be inside the slab (lower bound of every entry, upper bound of every exit)
HF_HD void hf_beam_slab(float o, float lo, float hi, float dl, float dh, float &ta, float &tb) {
    if (dl > 0.f) {
        if (o < lo) ta = fmaxf(ta, (lo - o) / dh);
        tb = fminf(tb, o < hi ? (hi - o) / dl : -1.f);
    } else if (dh < 0.f) {
        if (o > hi) ta = fmaxf(ta, (o - hi) / -dl);
        tb = fminf(tb, o > lo ? (o - lo) / -dh : -1.f);
    }
    // an interval containing 0 constrains nothing (rays that stay outside the slab are sorted out by their own clip)
}

HF_HD float hf_beam_enter(const HfGrid &G, const HfBeam &b, int &L_start
#ifndef __CUDACC__
                          , unsigned *iters = nullptr
#endif
) {
    const int NX = G.NX, NY = G.NY, Ltop = G.Ltop;
    L_start = hf_imin(Ltop, HF_START_LEVEL);
    float ta = 0.f, tb = INFINITY;
    hf_beam_slab(b.ogx, 0.f, (float) NX, b.dxl, b.dxh, ta, tb);
    hf_beam_slab(b.ogy, 0.f, (float) NY, b.dyl, b.dyh, ta, tb);
    const float zpad = HF_TRAV_SLACK * (fmaxf(fabsf(G.top_min), fabsf(G.top_max)) + fabsf(b.oz));
    hf_beam_slab(b.oz, G.top_min - zpad, G.top_max + zpad, b.dzl, b.dzh, ta, tb);
    ta -= ta * 3.814697265625e-06f; tb += fabsf(tb) * 3.814697265625e-06f;     // 2^-18
    if (!(ta <= tb)) return HF_BEAM_ALL_MISS;
    const float dmax = fmaxf(fmaxf(fabsf(b.dxl), fabsf(b.dxh)), fmaxf(fabsf(b.dyl), fabsf(b.dyh)));
    const float dwid = fmaxf(b.dxh - b.dxl, b.dyh - b.dyl);
    if (!(dmax > 0.f) || !(tb < INFINITY)) return ta;          // vertical bundle or unbounded range: nothing to march
    const float rdmax = 1.0f / dmax;
    const float zmag = fabsf(b.oz) + tb * fmaxf(fabsf(b.dzl), fabsf(b.dzh));
    const float epsz = 2.f * HF_TRAV_SLACK * zmag;
    const float pad = 0.03125f + 3.814697265625e-06f * (fmaxf(fabsf(b.ogx), fabsf(b.ogy)) + tb * dmax);

    int L = hf_imin(Ltop, HF_BEAM_START_LEVEL), grow = 0;
    float t = ta;
    for (int it = 0; it < HF_BEAM_MAX_ITERS; ++it) {
#ifndef __CUDACC__
        if (iters) ++*iters;
#endif
        const float span = (float) (1 << L);
        const float room = span - t * dwid - 1.f;
        if (room < HF_BEAM_ROOM_FRAC * span) break;                         // the level is too fine for the bundle's own width
        const float te = fminf(t + room * rdmax, tb);
        const float xl = b.ogx + fminf(t * b.dxl, te * b.dxl) - pad, xh = b.ogx + fmaxf(t * b.dxh, te * b.dxh) + pad;
        const float yl = b.ogy + fminf(t * b.dyl, te * b.dyl) - pad, yh = b.ogy + fmaxf(t * b.dyh, te * b.dyh) + pad;
        const float zl = b.oz + fminf(t * b.dzl, te * b.dzl) - epsz, zh = b.oz + fmaxf(t * b.dzh, te * b.dzh) + epsz;
        const bool outside = (xh < 0.f) || (xl > (float) NX) || (yh < 0.f) || (yl > (float) NY);
        const int cxl = hf_imin(hf_imax(hf_floor_sat(xl), 0), NX - 1), cxh = hf_imin(hf_imax(hf_floor_sat(xh), 0), NX - 1);
        const int cyl = hf_imin(hf_imax(hf_floor_sat(yl), 0), NY - 1), cyh = hf_imin(hf_imax(hf_floor_sat(yh), 0), NY - 1);
        const unsigned ext = (unsigned) hf_imax(cxh - cxl, cyh - cyl);
        const int Lq = hf_imin(hf_imax(hf_imax(L, ext ? hf_msb(ext) + 1 : 0), 1), Ltop);     // 2^Lq >= ext + 1: at most 2 nodes per axis
#ifdef __CUDACC__
        const HfLevel lv = hf_s_lvl[Lq];
#else
        const HfLevel lv = G.lvl[Lq];
#endif
        const int nxl = cxl >> Lq, nxh = cxh >> Lq, nyl = cyl >> Lq, nyh = cyh >> Lq;
#ifdef HF_DEBUG_BOUNDS
        if (nxl < 0 || nxh >= lv.w || nyl < 0 || nyh > ((NY - 1) >> Lq) || nxh > ((NX - 1) >> Lq)) HF_BOUNDS_FAIL();
#endif
        const float2 a0 = lv.nodes[nyl * lv.w + nxl], a1 = lv.nodes[nyl * lv.w + nxh];
        const float2 a2 = lv.nodes[nyh * lv.w + nxl], a3 = lv.nodes[nyh * lv.w + nxh];
        const float mn = fminf(fminf(a0.x, a1.x), fminf(a2.x, a3.x)), mx = fmaxf(fmaxf(a0.y, a1.y), fmaxf(a2.y, a3.y));
        const bool safe = outside || (zl > mx) || (zh < mn);
        if (safe) {
            t = te;
            if (!(te < tb)) return HF_BEAM_ALL_MISS;
            if (++grow >= 2 && L < Ltop) { ++L; grow = 0; }
        } else {
            grow = 0;
            if (--L < 1) break;
        }
    }
    L_start = hf_imin(hf_imin(hf_imax(L + 1, 1), HF_START_LEVEL), Ltop);
    return fmaxf(t - t * 3.814697265625e-06f, 0.f);
}
