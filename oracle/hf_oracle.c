/*
 * hf_oracle.c -- CPU restatement of the reference arithmetic for the
 * differentiable heightfield path. TEST INFRASTRUCTURE ONLY (see hf_oracle.h).
 *
 * Compile with -ffp-contract=off and without fast-math: the parity target is
 * the reference's llvm_ad_rgb lowering, which never contracts a*b+c, uses a
 * single-rounding fma only where the source says fmadd/fmsub/fnmadd, IEEE
 * division for rcp and sqrt(1/x) for rsqrt
 * (ext/drjit/ext/drjit-core/src/op.cpp:1194-1203, 1231-1236;
 *  ext/drjit/ext/drjit-core/src/llvm_eval.cpp:450-531).
 */
#include "hf_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------------- */
/* Dr.Jit op order, restated                                               */
/* ---------------------------------------------------------------------- */

/* ext/drjit/include/drjit/array_base.h:603-631 : r = a0*b0; r = fma(a1,b1,r); r = fma(a2,b2,r) */
static inline float dot3(const float a[3], const float b[3]) {
    float r = a[0] * b[0];
    r = fmaf(a[1], b[1], r);
    r = fmaf(a[2], b[2], r);
    return r;
}

/* ext/drjit/include/drjit/array_router.h:648-661 : fmsub(a.yzx, b.zxy, a.zxy * b.yzx) */
static inline void cross3(const float a[3], const float b[3], float c[3]) {
    float c0 = fmaf(a[1], b[2], -(a[2] * b[1]));
    float c1 = fmaf(a[2], b[0], -(a[0] * b[2]));
    float c2 = fmaf(a[0], b[1], -(a[1] * b[0]));
    c[0] = c0; c[1] = c1; c[2] = c2;
}

/* array_router.h:629-638 */
static inline float sqnorm3(const float v[3]) {
    float r = v[0] * v[0];
    r = fmaf(v[1], v[1], r);
    r = fmaf(v[2], v[2], r);
    return r;
}

/* op.cpp:1194-1203 */
static inline float rcp_llvm(float x) { return 1.0f / x; }
/* op.cpp:1231-1236: reciprocal, then square root */
static inline float rsqrt_llvm(float x) { return sqrtf(1.0f / x); }

/* array_router.h:644-646 */
static inline void normalize3(const float v[3], float out[3]) {
    float s = rsqrt_llvm(sqnorm3(v));
    out[0] = v[0] * s; out[1] = v[1] * s; out[2] = v[2] * s;
}

/* include/mitsuba/core/transform.h:104-111 (point, affine) */
static inline void xform_point(const float m[12], const float p[3], float out[3]) {
    float r0 = m[3], r1 = m[7], r2 = m[11];
    r0 = fmaf(m[0], p[0], r0); r1 = fmaf(m[4], p[0], r1); r2 = fmaf(m[8],  p[0], r2);
    r0 = fmaf(m[1], p[1], r0); r1 = fmaf(m[5], p[1], r1); r2 = fmaf(m[9],  p[1], r2);
    r0 = fmaf(m[2], p[2], r0); r1 = fmaf(m[6], p[2], r1); r2 = fmaf(m[10], p[2], r2);
    out[0] = r0; out[1] = r1; out[2] = r2;
}

/* transform.h:129-139 (vector): col0*x, fma(col1,y), fma(col2,z) */
static inline void xform_vector(const float m[12], const float v[3], float out[3]) {
    float r0 = m[0] * v[0], r1 = m[4] * v[0], r2 = m[8] * v[0];
    r0 = fmaf(m[1], v[1], r0); r1 = fmaf(m[5], v[1], r1); r2 = fmaf(m[9],  v[1], r2);
    r0 = fmaf(m[2], v[2], r0); r1 = fmaf(m[6], v[2], r1); r2 = fmaf(m[10], v[2], r2);
    out[0] = r0; out[1] = r1; out[2] = r2;
}

/* transform.h:141-153 (normal): columns of the inverse transpose == rows of to_object */
static inline void xform_normal(const float to_object[12], const float n[3], float out[3]) {
    float r0 = to_object[0] * n[0], r1 = to_object[1] * n[0], r2 = to_object[2]  * n[0];
    r0 = fmaf(to_object[4], n[1], r0); r1 = fmaf(to_object[5], n[1], r1); r2 = fmaf(to_object[6],  n[1], r2);
    r0 = fmaf(to_object[8], n[2], r0); r1 = fmaf(to_object[9], n[2], r1); r2 = fmaf(to_object[10], n[2], r2);
    out[0] = r0; out[1] = r1; out[2] = r2;
}

/* include/mitsuba/core/vector.h:116-136 */
static inline float signf_(float x) { return copysignf(1.0f, x); }           /* dr::sign: +-1 by sign bit */
static inline float mulsign_(float a, float b) { return signbit(b) ? -a : a; }
static inline float mulsign_neg_(float a, float b) { return signbit(b) ? a : -a; }
static inline void coordinate_system(const float n[3], float s[3], float t[3]) {
    float sign = signf_(n[2]);
    float a = -rcp_llvm(sign + n[2]);
    float b = n[0] * n[1] * a;
    s[0] = mulsign_((n[0] * n[0]) * a, n[2]) + 1.0f;
    s[1] = mulsign_(b, n[2]);
    s[2] = mulsign_neg_(n[0], n[2]);
    t[0] = b;
    t[1] = fmaf(n[1], n[1] * a, sign);
    t[2] = -n[1];
}

/* ---------------------------------------------------------------------- */
/* The per-cell primitive                                                  */
/* ---------------------------------------------------------------------- */

/* include/mitsuba/render/mesh.h:357-380 */
static inline int moeller_trumbore(const float o[3], const float d[3], float maxt,
                                   const float p0[3], const float p1[3], const float p2[3],
                                   float *t_, float *u_, float *v_) {
    float e1[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] };
    float e2[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
    float pvec[3]; cross3(d, e2, pvec);
    float inv_det = rcp_llvm(dot3(e1, pvec));
    float tvec[3] = { o[0] - p0[0], o[1] - p0[1], o[2] - p0[2] };
    float u = dot3(tvec, pvec) * inv_det;
    int active = (u >= 0.f) && (u <= 1.f);
    float qvec[3]; cross3(tvec, e1, qvec);
    float v = dot3(d, qvec) * inv_det;
    active = active && (v >= 0.f) && (u + v <= 1.f);
    float t = dot3(e2, qvec) * inv_det;
    active = active && (t >= 0.f) && (t <= maxt);
    *t_ = t; *u_ = u; *v_ = v;
    return active;
}

int hfo_moeller_trumbore(const float o[3], const float d[3], float maxt,
                         const float p0[3], const float p1[3], const float p2[3],
                         float *t, float *u, float *v) {
    return moeller_trumbore(o, d, maxt, p0, p1, p2, t, u, v);
}

/* test hooks for the golden vectors (tests/golden/drjit_triangle.npz) */
void hfo_interpolate_and_normal(const float p0[3], const float p1[3], const float p2[3], float b1, float b2,
                                float p_out[3], float n_out[3]) {
    float b0 = 1.f - b1 - b2;
    float dp0[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] };
    float dp1[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
    for (int a = 0; a < 3; ++a) p_out[a] = fmaf(p0[a], b0, fmaf(p1[a], b1, p2[a] * b2));
    float c[3]; cross3(dp0, dp1, c); normalize3(c, n_out);
}
void hfo_xform_point(const float m[12], const float p[3], float out[3]) { xform_point(m, p, out); }
void hfo_xform_vector(const float m[12], const float v[3], float out[3]) { xform_vector(m, v, out); }

/* ---------------------------------------------------------------------- */
/* Heightfield definition (DESIGN.md "Definition")     */
/* ---------------------------------------------------------------------- */

/* X(i) = float(2i-(W-1)) * (1/float(W-1)) : symmetric about 0, monotone */
static inline float vtx(int i, int W, float rw) { return (float) (2 * i - (W - 1)) * rw; }
float hfo_vertex_x(int32_t i, int32_t W) { return vtx(i, W, 1.0f / (float) (W - 1)); }
/* texcoord of a vertex: fmadd(x, 0.5, 0.5) as src/shapes/rectangle.cpp:312-313 */
static inline float tex(float x) { return fmaf(x, 0.5f, 0.5f); }

typedef struct {
    int W, H, NX, NY;
    float rw, rh, mh;
    const float *h;
} grid_t;

static inline grid_t make_grid(const hfo_shape *s) {
    grid_t g;
    g.W = s->W; g.H = s->H; g.NX = s->W - 1; g.NY = s->H - 1;
    g.rw = 1.0f / (float) g.NX; g.rh = 1.0f / (float) g.NY;
    g.mh = s->max_height; g.h = s->heights;
    return g;
}

/* Corner vertices of cell (cx,cy): p00, p10, p01, p11 */
static inline void cell_vertices(const grid_t *g, int cx, int cy,
                                 float p00[3], float p10[3], float p01[3], float p11[3]) {
    float x0 = vtx(cx, g->W, g->rw), x1 = vtx(cx + 1, g->W, g->rw);
    float y0 = vtx(cy, g->H, g->rh), y1 = vtx(cy + 1, g->H, g->rh);
    const float *row0 = g->h + (size_t) cy * g->W + cx, *row1 = row0 + g->W;
    p00[0] = x0; p00[1] = y0; p00[2] = g->mh * row0[0];
    p10[0] = x1; p10[1] = y0; p10[2] = g->mh * row0[1];
    p01[0] = x0; p01[1] = y1; p01[2] = g->mh * row1[0];
    p11[0] = x1; p11[1] = y1; p11[2] = g->mh * row1[1];
}

/* Triangle k of a cell: 0 = (p10, p01, p00), 1 = (p10, p11, p01)
 * (same diagonal and winding as resources/data/common/meshes/rectangle.obj faces "2 3 1", "2 4 3").
 * Also returns the texel (i,j) of each vertex. */
static inline void cell_triangle(const grid_t *g, int cx, int cy, int k,
                                 float p0[3], float p1[3], float p2[3], int vi[3], int vj[3]) {
    float p00[3], p10[3], p01[3], p11[3];
    cell_vertices(g, cx, cy, p00, p10, p01, p11);
    if (k == 0) {
        memcpy(p0, p10, 12); memcpy(p1, p01, 12); memcpy(p2, p00, 12);
        vi[0] = cx + 1; vj[0] = cy; vi[1] = cx; vj[1] = cy + 1; vi[2] = cx; vj[2] = cy;
    } else {
        memcpy(p0, p10, 12); memcpy(p1, p11, 12); memcpy(p2, p01, 12);
        vi[0] = cx + 1; vj[0] = cy; vi[1] = cx + 1; vj[1] = cy + 1; vi[2] = cx; vj[2] = cy + 1;
    }
}

/* Both triangles of a cell; closest wins, tie -> smaller index.
 * mesh.h:220-248 (hit ? t : inf). any_hit: stop at the first. */
static inline int cell_intersect(const grid_t *g, int cx, int cy,
                                 const float o[3], const float d[3], float maxt,
                                 float *t, float *u, float *v, uint32_t *prim) {
    float p00[3], p10[3], p01[3], p11[3];
    cell_vertices(g, cx, cy, p00, p10, p01, p11);
    float t0, u0, v0, t1, u1, v1;
    int h0 = moeller_trumbore(o, d, maxt, p10, p01, p00, &t0, &u0, &v0);
    int h1 = moeller_trumbore(o, d, maxt, p10, p11, p01, &t1, &u1, &v1);
    uint32_t base = 2u * ((uint32_t) cy * (uint32_t) g->NX + (uint32_t) cx);
    if (h1 && (!h0 || t1 < t0)) { *t = t1; *u = u1; *v = v1; *prim = base + 1u; return 1; }
    if (h0)                     { *t = t0; *u = u0; *v = v0; *prim = base;      return 1; }
    return 0;
}

int hfo_cell_intersect(const hfo_shape *s, int32_t cx, int32_t cy, const float o[3], const float d[3], float maxt,
                       float *t, float *u, float *v, uint32_t *prim) {
    grid_t g = make_grid(s);
    return cell_intersect(&g, cx, cy, o, d, maxt, t, u, v, prim);
}

/* ---------------------------------------------------------------------- */
/* Min/max pyramid                                                         */
/* ---------------------------------------------------------------------- */

#define HFO_MAX_LEVELS 32
struct hfo_pyramid {
    int NX, NY, Ltop;               /* level l node covers 2^l x 2^l cells; levels 1..Ltop stored */
    int lw[HFO_MAX_LEVELS], lh[HFO_MAX_LEVELS];
    float *lvl[HFO_MAX_LEVELS];     /* interleaved (min,max) of z = max_height*h */
};

hfo_pyramid *hfo_pyramid_build(const hfo_shape *s) {
    grid_t g = make_grid(s);
    hfo_pyramid *p = (hfo_pyramid *) calloc(1, sizeof(hfo_pyramid));
    p->NX = g.NX; p->NY = g.NY;
    int L = 1;
    while (((g.NX + (1 << L) - 1) >> L) > 1 || ((g.NY + (1 << L) - 1) >> L) > 1) L++;
    p->Ltop = L;
    for (int l = 1; l <= L; ++l) {
        p->lw[l] = (g.NX + (1 << l) - 1) >> l;
        p->lh[l] = (g.NY + (1 << l) - 1) >> l;
        p->lvl[l] = (float *) malloc(sizeof(float) * 2 * (size_t) p->lw[l] * p->lh[l]);
    }
    /* level 1 from texels: node (nx,ny) covers texels [2nx, 2nx+2] x [2ny, 2ny+2], clamped */
    #pragma omp parallel for schedule(static)
    for (int ny = 0; ny < p->lh[1]; ++ny)
        for (int nx = 0; nx < p->lw[1]; ++nx) {
            float mn = INFINITY, mx = -INFINITY;
            int j1 = 2 * ny + 2 < g.H - 1 ? 2 * ny + 2 : g.H - 1;
            int i1 = 2 * nx + 2 < g.W - 1 ? 2 * nx + 2 : g.W - 1;
            for (int j = 2 * ny; j <= j1; ++j)
                for (int i = 2 * nx; i <= i1; ++i) {
                    float z = g.mh * g.h[(size_t) j * g.W + i];
                    mn = fminf(mn, z); mx = fmaxf(mx, z);
                }
            float *o = p->lvl[1] + 2 * ((size_t) ny * p->lw[1] + nx);
            o[0] = mn; o[1] = mx;
        }
    for (int l = 2; l <= L; ++l) {
        const float *src = p->lvl[l - 1];
        int sw = p->lw[l - 1], sh = p->lh[l - 1];
        #pragma omp parallel for schedule(static)
        for (int ny = 0; ny < p->lh[l]; ++ny)
            for (int nx = 0; nx < p->lw[l]; ++nx) {
                float mn = INFINITY, mx = -INFINITY;
                for (int dj = 0; dj < 2; ++dj)
                    for (int di = 0; di < 2; ++di) {
                        int i = 2 * nx + di, j = 2 * ny + dj;
                        if (i < sw && j < sh) {
                            const float *c = src + 2 * ((size_t) j * sw + i);
                            mn = fminf(mn, c[0]); mx = fmaxf(mx, c[1]);
                        }
                    }
                float *o = p->lvl[l] + 2 * ((size_t) ny * p->lw[l] + nx);
                o[0] = mn; o[1] = mx;
            }
    }
    return p;
}

void hfo_pyramid_free(hfo_pyramid *p) {
    if (!p) return;
    for (int l = 0; l < HFO_MAX_LEVELS; ++l) free(p->lvl[l]);
    free(p);
}
int hfo_pyramid_levels(const hfo_pyramid *p) { return p->Ltop; }
void hfo_pyramid_level_dims(const hfo_pyramid *p, int level, int *w, int *h) { *w = p->lw[level]; *h = p->lh[level]; }
const float *hfo_pyramid_level_data(const hfo_pyramid *p, int level) { return p->lvl[level]; }

/* World AABB: union of the 8 corners of the object-space box under to_world
 * (pattern: src/shapes/rectangle.cpp:114-124). out = min xyz, max xyz. */
void hfo_bbox(const hfo_shape *s, const hfo_pyramid *p, float out[6]) {
    grid_t g = make_grid(s);
    float lo[3] = { vtx(0, g.W, g.rw), vtx(0, g.H, g.rh), p->lvl[p->Ltop][0] };
    float hi[3] = { vtx(g.W - 1, g.W, g.rw), vtx(g.H - 1, g.H, g.rh), p->lvl[p->Ltop][1] };
    for (int a = 0; a < 3; ++a) { out[a] = INFINITY; out[3 + a] = -INFINITY; }
    for (int c = 0; c < 8; ++c) {
        float q[3] = { (c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 4) ? hi[2] : lo[2] }, w[3];
        xform_point(s->to_world, q, w);
        for (int a = 0; a < 3; ++a) { out[a] = fminf(out[a], w[a]); out[3 + a] = fmaxf(out[3 + a], w[a]); }
    }
}

/* ---------------------------------------------------------------------- */
/* Ray helpers                                                             */
/* ---------------------------------------------------------------------- */

static inline void load_ray_object(const hfo_shape *s, const hfo_rays *r, size_t i,
                                   float o[3], float d[3], float *maxt) {
    float ow[3] = { r->ox[i], r->oy[i], r->oz[i] };
    float dw[3] = { r->dx[i], r->dy[i], r->dz[i] };
    /* Transform::transform_affine(ray): src/shapes/rectangle.cpp:206-214; t is preserved */
    xform_point(s->to_object, ow, o);
    xform_vector(s->to_object, dw, d);
    *maxt = r->maxt[i];
}

static inline void write_miss(hfo_prelim *out, size_t i) {
    out->t[i] = INFINITY; out->u[i] = 0.f; out->v[i] = 0.f; out->prim_index[i] = 0u;
}

/* ---------------------------------------------------------------------- */
/* Brute force                                                             */
/* ---------------------------------------------------------------------- */

void hfo_intersect_brute(const hfo_shape *s, const hfo_rays *r, size_t n, hfo_prelim *out) {
    grid_t g = make_grid(s);
    #pragma omp parallel for schedule(dynamic, 16)
    for (long long ii = 0; ii < (long long) n; ++ii) {
        size_t i = (size_t) ii;
        if (r->active && !r->active[i]) { write_miss(out, i); continue; }
        float o[3], d[3], maxt;
        load_ray_object(s, r, i, o, d, &maxt);
        float bt = INFINITY, bu = 0.f, bv = 0.f; uint32_t bp = 0u;
        for (int cy = 0; cy < g.NY; ++cy)
            for (int cx = 0; cx < g.NX; ++cx) {
                float t, u, v; uint32_t prim;
                if (cell_intersect(&g, cx, cy, o, d, maxt, &t, &u, &v, &prim) && t < bt) {
                    bt = t; bu = u; bv = v; bp = prim;   /* cells scanned in prim order: strict < keeps the smaller index */
                }
            }
        out->t[i] = bt; out->u[i] = bu; out->v[i] = bv; out->prim_index[i] = bp;
    }
}

/* ---------------------------------------------------------------------- */
/* Canonical hierarchical DDA over the min/max pyramid                     */
/* ---------------------------------------------------------------------- */

/* Relative slack applied to the ray's z interval so that a node which contains a hit
 * is never culled because of rounding in the traversal arithmetic. */
#define HFO_Z_SLACK 9.5367431640625e-07f   /* 2^-20 */

static inline int floor_clamp(float x, int lo, int hi) {
    /* NaN / +-inf safe */
    if (!(x > (float) lo)) return lo;
    if (!(x < (float) hi)) return hi;
    return (int) floorf(x);
}

static int traverse(const grid_t *g, const hfo_pyramid *p,
                    const float o[3], const float d[3], float maxt, int any_hit,
                    float *t_, float *u_, float *v_, uint32_t *prim_,
                    uint32_t *n_node, uint32_t *n_leaf) {
    const int NX = g->NX, NY = g->NY, Ltop = p->Ltop;
    uint32_t cn = 0, cl = 0;
    int result = 0;

    if (d[0] == 0.f && d[1] == 0.f && d[2] == 0.f) goto done;
    if (!(maxt >= 0.f)) goto done;

    /* Ray in grid units: cell boundaries at integers (traversal only; the leaf test uses
     * the object-space vertices) */
    const float x0 = vtx(0, g->W, g->rw), y0 = vtx(0, g->H, g->rh);
    const float sgx = 0.5f * (float) NX, sgy = 0.5f * (float) NY;
    const float ogx = (o[0] - x0) * sgx, ogy = (o[1] - y0) * sgy;
    const float dgx = d[0] * sgx, dgy = d[1] * sgy;
    const float oz = o[2], dz = d[2];
    const float idx = 1.0f / dgx, idy = 1.0f / dgy, idz = 1.0f / dz;

    /* Slab clip against [0,NX]x[0,NY]x[zmin,zmax] (semantics of include/mitsuba/core/bbox.h:303-327,
     * with the zero-slope case handled explicitly) */
    float t0 = 0.f, t1 = maxt;
    if (dgx != 0.f) {
        float a = (0.f - ogx) * idx, b = ((float) NX - ogx) * idx;
        t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    } else if (ogx < 0.f || ogx > (float) NX) goto done;
    if (dgy != 0.f) {
        float a = (0.f - ogy) * idy, b = ((float) NY - ogy) * idy;
        t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    } else if (ogy < 0.f || ogy > (float) NY) goto done;
    {
        const float zmin = p->lvl[Ltop][0], zmax = p->lvl[Ltop][1];
        const float pad = HFO_Z_SLACK * fmaxf(fabsf(zmin), fabsf(zmax)) + HFO_Z_SLACK * fabsf(oz);
        if (dz != 0.f) {
            float a = (zmin - pad - oz) * idz, b = (zmax + pad - oz) * idz;
            float tn = fminf(a, b), tf = fmaxf(a, b);
            /* widen by a few ulp: t0 only seeds the start cell, t1 only ends the walk */
            tn -= fabsf(tn) * HFO_Z_SLACK; tf += fabsf(tf) * HFO_Z_SLACK;
            t0 = fmaxf(t0, tn); t1 = fminf(t1, tf);
        } else if (oz < zmin - pad || oz > zmax + pad) goto done;
    }
    if (!(t0 <= t1)) goto done;

    const float epsz = HFO_Z_SLACK * (fabsf(oz) + fabsf(t1 * dz));
    const int sx = dgx > 0.f, sy = dgy > 0.f;     /* positive-direction flags */

    int ix = floor_clamp(fmaf(t0, dgx, ogx), 0, NX - 1);
    int iy = floor_clamp(fmaf(t0, dgy, ogy), 0, NY - 1);
    int L = Ltop;
    float t = t0;

    for (;;) {
        const int nx = ix >> L, ny = iy >> L;
        /* exit boundaries of this node (clamped to the grid) */
        int bx, by;
        if (sx) { bx = (nx + 1) << L; if (bx > NX) bx = NX; } else bx = nx << L;
        if (sy) { by = (ny + 1) << L; if (by > NY) by = NY; } else by = ny << L;
        const float tx = dgx != 0.f ? ((float) bx - ogx) * idx : INFINITY;
        const float ty = dgy != 0.f ? ((float) by - ogy) * idy : INFINITY;
        float te = fminf(tx, ty);
        const int last = !(te < t1);
        if (last) te = t1;

        const float za = fmaf(t, dz, oz), zb = fmaf(te, dz, oz);
        const float zlo = fminf(za, zb) - epsz, zhi = fmaxf(za, zb) + epsz;

        int overlap;
        if (L > 0) {
            const float *node = p->lvl[L] + 2 * ((size_t) ny * p->lw[L] + nx);
            cn++;
            overlap = (zlo <= node[1]) && (zhi >= node[0]);
            if (overlap) { L--; continue; }
        } else {
            const float *row0 = g->h + (size_t) iy * g->W + ix, *row1 = row0 + g->W;
            const float z00 = g->mh * row0[0], z10 = g->mh * row0[1], z01 = g->mh * row1[0], z11 = g->mh * row1[1];
            const float mn = fminf(fminf(z00, z10), fminf(z01, z11));
            const float mx = fmaxf(fmaxf(z00, z10), fmaxf(z01, z11));
            cl++;
            overlap = (zlo <= mx) && (zhi >= mn);
            if (overlap && cell_intersect(g, ix, iy, o, d, maxt, t_, u_, v_, prim_)) { result = 1; goto done; }
        }
        (void) any_hit;  /* closest == first cell front-to-back; any-hit uses the same walk */

        if (last) goto done;

        /* step to the next node */
        const int ox_ = ix, oy_ = iy;
        const int stepx = tx <= ty, stepy = ty <= tx;
        if (stepx) ix = sx ? bx : bx - 1;
        if (stepy) iy = sy ? by : by - 1;
        if (!stepx) {
            int lo = nx << L, hi = ((nx + 1) << L) - 1; if (hi > NX - 1) hi = NX - 1;
            int q = floor_clamp(fmaf(te, dgx, ogx), lo, hi);
            ix = sx ? (q > ox_ ? q : ox_) : (dgx < 0.f ? (q < ox_ ? q : ox_) : ox_);
        }
        if (!stepy) {
            int lo = ny << L, hi = ((ny + 1) << L) - 1; if (hi > NY - 1) hi = NY - 1;
            int q = floor_clamp(fmaf(te, dgy, ogy), lo, hi);
            iy = sy ? (q > oy_ ? q : oy_) : (dgy < 0.f ? (q < oy_ ? q : oy_) : oy_);
        }
        if (ix < 0 || ix >= NX || iy < 0 || iy >= NY) goto done;
        t = te;
        /* ascend one level when the step left the parent node */
        if (L < Ltop && (((ix >> (L + 1)) != (ox_ >> (L + 1))) || ((iy >> (L + 1)) != (oy_ >> (L + 1))))) L++;
    }

done:
    if (n_node) *n_node = cn;
    if (n_leaf) *n_leaf = cl;
    return result;
}

void hfo_intersect_dda(const hfo_shape *s, const hfo_pyramid *p, const hfo_rays *r, size_t n,
                       hfo_prelim *out, uint32_t *n_node, uint32_t *n_leaf) {
    grid_t g = make_grid(s);
    #pragma omp parallel for schedule(dynamic, 1024)
    for (long long ii = 0; ii < (long long) n; ++ii) {
        size_t i = (size_t) ii;
        uint32_t cn = 0, cl = 0;
        int hit = 0;
        float t = 0, u = 0, v = 0; uint32_t prim = 0;
        if (!r->active || r->active[i]) {
            float o[3], d[3], maxt;
            load_ray_object(s, r, i, o, d, &maxt);
            hit = traverse(&g, p, o, d, maxt, 0, &t, &u, &v, &prim, &cn, &cl);
        }
        if (hit) { out->t[i] = t; out->u[i] = u; out->v[i] = v; out->prim_index[i] = prim; }
        else write_miss(out, i);
        if (n_node) n_node[i] = cn;
        if (n_leaf) n_leaf[i] = cl;
    }
}

void hfo_ray_test(const hfo_shape *s, const hfo_pyramid *p, const hfo_rays *r, size_t n, uint8_t *hit) {
    grid_t g = make_grid(s);
    #pragma omp parallel for schedule(dynamic, 1024)
    for (long long ii = 0; ii < (long long) n; ++ii) {
        size_t i = (size_t) ii;
        int h = 0;
        if (!r->active || r->active[i]) {
            float o[3], d[3], maxt, t, u, v; uint32_t prim;
            load_ray_object(s, r, i, o, d, &maxt);
            h = traverse(&g, p, o, d, maxt, 1, &t, &u, &v, &prim, NULL, NULL);
        }
        hit[i] = (uint8_t) h;
    }
}

/* ---------------------------------------------------------------------- */
/* Surface interaction (src/render/mesh.cpp:672-903 on the hit triangle,  */
/* object space, then mapped to world like src/shapes/rectangle.cpp)      */
/* ---------------------------------------------------------------------- */

static inline void decode_prim(const grid_t *g, uint32_t prim, int *cx, int *cy, int *k) {
    uint32_t cell = prim >> 1;
    *k = (int) (prim & 1u);
    *cy = (int) (cell / (uint32_t) g->NX);
    *cx = (int) (cell % (uint32_t) g->NX);
}

#define ST(ptr, val) do { if (ptr) (ptr)[i] = (val); } while (0)

void hfo_surface_interaction(const hfo_shape *s, const hfo_rays *r, const hfo_prelim *pi,
                             uint32_t flags, size_t n, hfo_si *out) {
    grid_t g = make_grid(s);
    #pragma omp parallel for schedule(static)
    for (long long ii = 0; ii < (long long) n; ++ii) {
        size_t i = (size_t) ii;
        float dw[3] = { r->dx[i], r->dy[i], r->dz[i] };
        int valid = (!r->active || r->active[i]) && pi->t[i] != INFINITY;
        if (!valid) {
            /* dr::zeros<SurfaceInteraction3f>() for masked lanes; t = inf (interaction.h:476-499) */
            ST(out->t, INFINITY);
            ST(out->px, 0.f); ST(out->py, 0.f); ST(out->pz, 0.f);
            ST(out->nx, 0.f); ST(out->ny, 0.f); ST(out->nz, 0.f);
            ST(out->uvx, 0.f); ST(out->uvy, 0.f);
            ST(out->dpdu_x, 0.f); ST(out->dpdu_y, 0.f); ST(out->dpdu_z, 0.f);
            ST(out->dpdv_x, 0.f); ST(out->dpdv_y, 0.f); ST(out->dpdv_z, 0.f);
            ST(out->shn_x, 0.f); ST(out->shn_y, 0.f); ST(out->shn_z, 0.f);
            ST(out->shs_x, 0.f); ST(out->shs_y, 0.f); ST(out->shs_z, 0.f);
            ST(out->sht_x, 0.f); ST(out->sht_y, 0.f); ST(out->sht_z, 0.f);
            ST(out->wi_x, -dw[0]); ST(out->wi_y, -dw[1]); ST(out->wi_z, -dw[2]);
            ST(out->boundary_test, (flags & HFO_RF_BOUNDARYTEST) ? 1e8f : 0.f);   /* interaction.h:497-498 */
            continue;
        }
        int cx, cy, k, vi[3], vj[3];
        decode_prim(&g, pi->prim_index[i], &cx, &cy, &k);
        float p0[3], p1[3], p2[3];
        cell_triangle(&g, cx, cy, k, p0, p1, p2, vi, vj);

        /* mesh.cpp:738-748 */
        float b1 = pi->u[i], b2 = pi->v[i], b0 = 1.f - b1 - b2;
        float dp0[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] };
        float dp1[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
        float po[3];
        for (int a = 0; a < 3; ++a) po[a] = fmaf(p0[a], b0, fmaf(p1[a], b1, p2[a] * b2));
        /* mesh.cpp:757 */
        float c[3], no[3]; cross3(dp0, dp1, c); normalize3(c, no);

        /* world mapping */
        float pw[3], nw[3], nt[3];
        xform_point(s->to_world, po, pw);
        xform_normal(s->to_object, no, nt);
        normalize3(nt, nw);

        /* mesh.cpp:760-789 */
        float uv0[2] = { tex(p0[0]), tex(p0[1]) }, uv1[2] = { tex(p1[0]), tex(p1[1]) }, uv2[2] = { tex(p2[0]), tex(p2[1]) };
        float uv[2];
        for (int a = 0; a < 2; ++a) uv[a] = fmaf(uv2[a], b2, fmaf(uv1[a], b1, uv0[a] * b0));
        float dpdu[3], dpdv[3];
        int have_partials = 0;
        if (flags & HFO_RF_DPDUV) {
            float duv0[2] = { uv1[0] - uv0[0], uv1[1] - uv0[1] }, duv1[2] = { uv2[0] - uv0[0], uv2[1] - uv0[1] };
            float det = fmaf(duv0[0], duv1[1], -(duv0[1] * duv1[0]));
            float inv_det = rcp_llvm(det);
            if (det != 0.f) {
                float du_o[3], dv_o[3];
                for (int a = 0; a < 3; ++a) {
                    du_o[a] = fmaf( duv1[1], dp0[a], -(duv0[1] * dp1[a])) * inv_det;   /* fmsub  */
                    dv_o[a] = fmaf(-duv1[0], dp0[a],   duv0[0] * dp1[a])  * inv_det;   /* fnmadd */
                }
                xform_vector(s->to_world, du_o, dpdu);
                xform_vector(s->to_world, dv_o, dpdv);
                have_partials = 1;
            }
        }
        if (!have_partials) coordinate_system(nw, dpdu, dpdv);   /* mesh.cpp:762 */

        if (s->flip_normals) { nw[0] = -nw[0]; nw[1] = -nw[1]; nw[2] = -nw[2]; }   /* mesh.cpp:837-840 */

        float t = pi->t[i];
        if (flags & HFO_RF_FOLLOWSHAPE) {   /* mesh.cpp:751-752 */
            float ow[3] = { r->ox[i], r->oy[i], r->oz[i] };
            float rel[3] = { pw[0] - ow[0], pw[1] - ow[1], pw[2] - ow[2] };
            t = sqrtf(sqnorm3(rel) / sqnorm3(dw));
        }

        ST(out->t, t);
        ST(out->px, pw[0]); ST(out->py, pw[1]); ST(out->pz, pw[2]);
        ST(out->nx, nw[0]); ST(out->ny, nw[1]); ST(out->nz, nw[2]);
        ST(out->uvx, uv[0]); ST(out->uvy, uv[1]);
        ST(out->dpdu_x, dpdu[0]); ST(out->dpdu_y, dpdu[1]); ST(out->dpdu_z, dpdu[2]);
        ST(out->dpdv_x, dpdv[0]); ST(out->dpdv_y, dpdv[1]); ST(out->dpdv_z, dpdv[2]);
        ST(out->shn_x, nw[0]); ST(out->shn_y, nw[1]); ST(out->shn_z, nw[2]);

        /* interaction.h:257-266, 476-499 */
        float ss[3] = { 0, 0, 0 }, st[3] = { 0, 0, 0 };
        float wi[3];
        if (flags & HFO_RF_SHADINGFRAME) {
            float nd = -dot3(nw, dpdu), q[3];
            for (int a = 0; a < 3; ++a) q[a] = fmaf(nw[a], nd, dpdu[a]);
            normalize3(q, ss);
            if (dpdu[0] == 0.f && dpdu[1] == 0.f && dpdu[2] == 0.f) { float tmp[3]; coordinate_system(nw, ss, tmp); }
            cross3(nw, ss, st);
        }
        {   /* wi = to_local(-ray.d) with whatever frame was requested (s = t = 0 without ShadingFrame) */
            float md[3] = { -dw[0], -dw[1], -dw[2] };
            wi[0] = dot3(md, ss); wi[1] = dot3(md, st); wi[2] = dot3(md, nw);   /* Frame::to_local */
        }
        ST(out->shs_x, ss[0]); ST(out->shs_y, ss[1]); ST(out->shs_z, ss[2]);
        ST(out->sht_x, st[0]); ST(out->sht_y, st[1]); ST(out->sht_z, st[2]);
        ST(out->wi_x, wi[0]); ST(out->wi_y, wi[1]); ST(out->wi_z, wi[2]);

        /* mesh.cpp:845-888, the branch of a mesh WITHOUT vertex normals (the heightfield is flat-shaded): least-squares (u, v) of
           the hit point inside its triangle, mapped onto an equilateral triangle; distance to the nearest edge, scaled so that the
           barycentre is at 1. Evaluated on the object-space triangle (the solve is affine-invariant). */
        float bt = 0.f;
        if (flags & HFO_RF_BOUNDARYTEST) {
            float rel[3] = { po[0] - p0[0], po[1] - p0[1], po[2] - p0[2] };
            float bb1 = dot3(dp0, rel), bb2 = dot3(dp1, rel), a11 = dot3(dp0, dp0), a12 = dot3(dp0, dp1), a22 = dot3(dp1, dp1);
            float inv_det = rcp_llvm(a11 * a22 - a12 * a12);
            float u = fmaf(a22, bb1, -(a12 * bb2)) * inv_det;       /* fmsub  */
            float v = fmaf(-a12, bb1, a11 * bb2) * inv_det;         /* fnmadd */
            const float hq = 0.5f * sqrtf(3.f);
            /* p = tp0 * w + tp1 * u + tp2 * v with tp0 = (0, 0), tp1 = (1, 0), tp2 = (1/2, sqrt(3)/2): the products with the literals 0
               and 1 fold away in Dr.Jit's constant propagation (w drops out) */
            float qx = u + 0.5f * v, qy = hq * v;
            const float ex[3] = { 1.f, -0.5f, -0.5f }, ey[3] = { 0.f, hq, -hq };             /* e0 = tp1 - tp0, e1 = tp2 - tp1, e2 = tp0 - tp2 */
            const float vx[3] = { qx, qx - 1.f, qx - 0.5f }, vy[3] = { qy, qy, qy - hq };      /* p - tp0, p - tp1, p - tp2 */
            float d2 = INFINITY;
            for (int e = 0; e < 3; ++e) {
                float ve = fmaf(vy[e], ey[e], vx[e] * ex[e]), ee = fmaf(ey[e], ey[e], ex[e] * ex[e]);
                float c = fminf(fmaxf(ve / ee, 0.f), 1.f);
                float px = vx[e] - ex[e] * c, py = vy[e] - ey[e] * c;
                d2 = fminf(d2, fmaf(py, py, px * px));
            }
            bt = sqrtf(d2) / (sqrtf(3.f) / 6.f);
        }
        ST(out->boundary_test, bt);
    }
}

/* ---------------------------------------------------------------------- */
/* Reverse mode w.r.t. heights                                             */
/* ---------------------------------------------------------------------- */

typedef struct { double x, y, z; } d3;
static inline d3 D3(double x, double y, double z) { d3 r = { x, y, z }; return r; }
static inline d3 dadd(d3 a, d3 b) { return D3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline d3 dsub(d3 a, d3 b) { return D3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline d3 dscale(d3 a, double s) { return D3(a.x * s, a.y * s, a.z * s); }
static inline double ddot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline d3 dcross(d3 a, d3 b) { return D3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline d3 from_f(const float v[3]) { return D3(v[0], v[1], v[2]); }
static inline double LDA(const float *p, size_t i) { return p ? (double) p[i] : 0.0; }

/* y = M_lin^T x for row-major 3x4 M */
static inline d3 lin_T(const float m[12], d3 x) {
    return D3(m[0] * x.x + m[4] * x.y + m[8] * x.z,
              m[1] * x.x + m[5] * x.y + m[9] * x.z,
              m[2] * x.x + m[6] * x.y + m[10] * x.z);
}
/* y = M_lin x */
static inline d3 lin(const float m[12], d3 x) {
    return D3(m[0] * x.x + m[1] * x.y + m[2] * x.z,
              m[4] * x.x + m[5] * x.y + m[6] * x.z,
              m[8] * x.x + m[9] * x.y + m[10] * x.z);
}

void hfo_backward(const hfo_shape *s, const hfo_rays *r, const hfo_prelim *pi, const hfo_si_adj *adj,
                  uint32_t flags, size_t n, double *grad, int literal) {
    grid_t g = make_grid(s);
    if (flags & HFO_RF_DETACHSHAPE) return;          /* mesh.cpp:713-717: vertices detached -> no height gradient */
    const int follow = (flags & HFO_RF_FOLLOWSHAPE) != 0;

    int nthreads = 1;
    #ifdef _OPENMP
    nthreads = omp_get_max_threads();
    #endif
    (void) nthreads;

    #pragma omp parallel for schedule(static)
    for (long long ii = 0; ii < (long long) n; ++ii) {
        size_t i = (size_t) ii;
        if ((r->active && !r->active[i]) || pi->t[i] == INFINITY) continue;
        int cx, cy, k, vi[3], vj[3];
        decode_prim(&g, pi->prim_index[i], &cx, &cy, &k);
        float p0f[3], p1f[3], p2f[3];
        cell_triangle(&g, cx, cy, k, p0f, p1f, p2f, vi, vj);
        float of[3], df[3], maxt;
        load_ray_object(s, r, i, of, df, &maxt);

        const d3 p0 = from_f(p0f), p1 = from_f(p1f), p2 = from_f(p2f), o = from_f(of), d = from_f(df);
        const d3 e1 = dsub(p1, p0), e2 = dsub(p2, p0);

        /* upstream adjoints, world space */
        double a_t = LDA(adj->t, i);
        d3 a_pw = D3(LDA(adj->px, i), LDA(adj->py, i), LDA(adj->pz, i));
        d3 a_nw = D3(LDA(adj->nx, i), LDA(adj->ny, i), LDA(adj->nz, i));
        double a_uv[2] = { LDA(adj->uvx, i), LDA(adj->uvy, i) };
        d3 a_duw = D3(LDA(adj->dpdu_x, i), LDA(adj->dpdu_y, i), LDA(adj->dpdu_z, i));
        d3 a_dvw = D3(LDA(adj->dpdv_x, i), LDA(adj->dpdv_y, i), LDA(adj->dpdv_z, i));
        if (s->flip_normals) a_nw = dscale(a_nw, -1.0);

        /* world -> object */
        d3 a_p = lin_T(s->to_world, a_pw);
        d3 a_du = lin_T(s->to_world, a_duw), a_dv = lin_T(s->to_world, a_dvw);

        /* primal pieces (double on fp32 inputs) */
        const double u = pi->u[i], v = pi->v[i], b0 = 1.0 - u - v, b1 = u, b2 = v;
        const d3 c = dcross(e1, e2);
        const double clen = sqrt(ddot(c, c));
        const d3 no = dscale(c, 1.0 / clen);
        /* n_w = normalize(TO_lin^T n_o) */
        d3 mvec = lin_T(s->to_object, no);
        const double mlen = sqrt(ddot(mvec, mvec));
        const d3 nw = dscale(mvec, 1.0 / mlen);
        d3 a_m = dscale(dsub(a_nw, dscale(nw, ddot(nw, a_nw))), 1.0 / mlen);
        d3 a_no = lin(s->to_object, a_m);
        d3 a_c = dscale(dsub(a_no, dscale(no, ddot(no, a_no))), 1.0 / clen);
        d3 a_e1 = dcross(e2, a_c), a_e2 = dcross(a_c, e1);

        /* dp_du = (duv1.y*e1 - duv0.y*e2)/det, dp_dv = (-duv1.x*e1 + duv0.x*e2)/det (mesh.cpp:778-789) */
        if (flags & HFO_RF_DPDUV) {
            double uv0[2] = { tex(p0f[0]), tex(p0f[1]) }, uv1[2] = { tex(p1f[0]), tex(p1f[1]) }, uv2[2] = { tex(p2f[0]), tex(p2f[1]) };
            double duv0[2] = { uv1[0] - uv0[0], uv1[1] - uv0[1] }, duv1[2] = { uv2[0] - uv0[0], uv2[1] - uv0[1] };
            double det = duv0[0] * duv1[1] - duv0[1] * duv1[0];
            if (det != 0.0) {
                a_e1 = dadd(a_e1, dadd(dscale(a_du,  duv1[1] / det), dscale(a_dv, -duv1[0] / det)));
                a_e2 = dadd(a_e2, dadd(dscale(a_du, -duv0[1] / det), dscale(a_dv,  duv0[0] / det)));
            }
        }

        d3 a_p0 = D3(0, 0, 0), a_p1 = D3(0, 0, 0), a_p2 = D3(0, 0, 0);

        if (follow) {
            /* mesh.cpp:728-729, 751-752: barycentrics detached, p glued to the triangle,
               t = sqrt(|p-o|^2 / |d|^2) in world space */
            d3 pw_o = dadd(dadd(dscale(p0, b0), dscale(p1, b1)), dscale(p2, b2));
            float pof[3] = { (float) pw_o.x, (float) pw_o.y, (float) pw_o.z }, pwf[3];
            xform_point(s->to_world, pof, pwf);
            d3 relw = D3(pwf[0] - (double) r->ox[i], pwf[1] - (double) r->oy[i], pwf[2] - (double) r->oz[i]);
            d3 dw = D3(r->dx[i], r->dy[i], r->dz[i]);
            double tt = sqrt(ddot(relw, relw) / ddot(dw, dw));
            if (tt > 0.0) a_p = dadd(a_p, lin_T(s->to_world, dscale(relw, a_t / (tt * ddot(dw, dw)))));
            a_p0 = dscale(a_p, b0); a_p1 = dscale(a_p, b1); a_p2 = dscale(a_p, b2);
            /* uv: vertex texcoords are constants, barycentrics detached -> nothing */
        } else if (literal) {
            /* adjoint of the literal op sequence: p = sum b_k p_k, uv = sum b_k uv_k, b from moeller_trumbore */
            a_p0 = dscale(a_p, b0); a_p1 = dscale(a_p, b1); a_p2 = dscale(a_p, b2);
            double uv0[2] = { tex(p0f[0]), tex(p0f[1]) }, uv1[2] = { tex(p1f[0]), tex(p1f[1]) }, uv2[2] = { tex(p2f[0]), tex(p2f[1]) };
            double a_b0 = ddot(p0, a_p) + uv0[0] * a_uv[0] + uv0[1] * a_uv[1];
            double a_b1 = ddot(p1, a_p) + uv1[0] * a_uv[0] + uv1[1] * a_uv[1];
            double a_b2 = ddot(p2, a_p) + uv2[0] * a_uv[0] + uv2[1] * a_uv[1];
            double a_u = a_b1 - a_b0, a_v = a_b2 - a_b0;
            /* reverse of mesh.h:357-380 */
            d3 pvec = dcross(d, e2);
            double det = ddot(e1, pvec), inv = 1.0 / det;
            d3 tvec = dsub(o, p0);
            double aa = ddot(tvec, pvec);
            d3 qvec = dcross(tvec, e1);
            double bb = ddot(d, qvec), cc = ddot(e2, qvec);
            double a_aa = a_u * inv, a_bb = a_v * inv, a_cc = a_t * inv;
            double a_inv = a_u * aa + a_v * bb + a_t * cc;
            double a_det = -a_inv * inv * inv;
            d3 m_e1 = D3(0, 0, 0), m_e2 = D3(0, 0, 0);
            m_e2 = dadd(m_e2, dscale(qvec, a_cc));
            d3 a_q = dadd(dscale(e2, a_cc), dscale(d, a_bb));
            d3 a_tvec = dcross(e1, a_q);
            m_e1 = dadd(m_e1, dcross(a_q, tvec));
            a_tvec = dadd(a_tvec, dscale(pvec, a_aa));
            d3 a_pvec = dadd(dscale(tvec, a_aa), dscale(e1, a_det));
            m_e1 = dadd(m_e1, dscale(pvec, a_det));
            m_e2 = dadd(m_e2, dcross(a_pvec, d));
            a_p0 = dsub(a_p0, a_tvec);
            a_e1 = dadd(a_e1, m_e1); a_e2 = dadd(a_e2, m_e2);
        } else {
            /* closed form: the hit stays on the ray, p = o + t d; uv moves through (u,v) only.
               K [dt,du,dv]^T = w, K = [d,-e1,-e2], w = sum b_k dp_k */
            double uv0[2] = { tex(p0f[0]), tex(p0f[1]) }, uv1[2] = { tex(p1f[0]), tex(p1f[1]) }, uv2[2] = { tex(p2f[0]), tex(p2f[1]) };
            double a_tt = a_t + ddot(d, a_p);
            double a_u = (uv1[0] - uv0[0]) * a_uv[0] + (uv1[1] - uv0[1]) * a_uv[1];
            double a_v = (uv2[0] - uv0[0]) * a_uv[0] + (uv2[1] - uv0[1]) * a_uv[1];
            d3 pvec = dcross(d, e2);
            double det = ddot(e1, pvec);
            d3 a_w = dscale(dadd(dadd(dscale(c, -a_tt), dscale(pvec, -a_u)), dscale(dcross(d, e1), a_v)), 1.0 / det);
            a_p0 = dscale(a_w, b0); a_p1 = dscale(a_w, b1); a_p2 = dscale(a_w, b2);
        }

        a_p1 = dadd(a_p1, a_e1); a_p2 = dadd(a_p2, a_e2); a_p0 = dsub(a_p0, dadd(a_e1, a_e2));

        /* z_k = max_height * h_k */
        double gk[3] = { a_p0.z * (double) g.mh, a_p1.z * (double) g.mh, a_p2.z * (double) g.mh };
        for (int q = 0; q < 3; ++q) {
            double *dst = grad + (size_t) vj[q] * g.W + vi[q];
            #pragma omp atomic
            *dst += gk[q];
        }
    }
}

/* ---------------------------------------------------------------------- */
/* Adam (src/python/python/ad/optimizers.py:263-302), fp32 like llvm_ad_rgb */
/* ---------------------------------------------------------------------- */

void hfo_adam_step(float *p, const float *g, float *m, float *v, size_t n,
                   double lr_d, double b1d, double b2d, double eps_d, int step_t, int mask_updates) {
    /* The hyper-parameters are Python floats (doubles) in optimizers.py: lr_scale (266-269) and (1 - beta) (281-282) are evaluated
       in double and only then become fp32 literals of the traced expression; lr is an opaque fp32 (lr_v). */
    const float lr_scale = (float) (sqrt(1.0 - pow(b2d, step_t)) / (1.0 - pow(b1d, step_t)));
    const float lr_t = (float) lr_d * lr_scale;
    const float beta1 = (float) b1d, beta2 = (float) b2d, eps = (float) eps_d;
    const float omb1 = (float) (1.0 - b1d), omb2 = (float) (1.0 - b2d);
    #pragma omp parallel for schedule(static)
    for (long long ii = 0; ii < (long long) n; ++ii) {
        size_t i = (size_t) ii;
        float gp = g[i];
        float m_t = beta1 * m[i] + omb1 * gp;
        float v_t = beta2 * v[i] + omb2 * (gp * gp);
        int nonzero = gp != 0.f;
        if (mask_updates && !nonzero) { m_t = m[i]; v_t = v[i]; }
        m[i] = m_t; v[i] = v_t;
        float step = lr_t * m_t / (sqrtf(v_t) + eps);
        if (mask_updates && !nonzero) step = 0.f;
        p[i] = p[i] - step;
    }
}

/* ---------------------------------------------------------------------- */
/* Render step on the heightfield alone:        */
/* perspective pinhole (src/sensors/perspective.cpp:198-234), directional */
/* emitter + Lambertian BSDF, direct illumination of the first hit        */
/* (src/integrators/direct.cpp:114-210), optional visibility ray spawned  */
/* as in include/mitsuba/render/interaction.h:136-165. The sensor and     */
/* shading op order is the definition in DESIGN.md section 7.             */
/* ---------------------------------------------------------------------- */
static inline uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
    return x;
}
static inline float hash_uniform(uint32_t seed, uint32_t idx, uint32_t channel) {
    uint32_t h = hash32(idx ^ hash32(seed * 0x9E3779B1u + channel * 0x85EBCA6Bu));
    return (float) (h >> 8) * (1.0f / 16777216.0f);
}

static void camera_dir(const hfo_camera *c, uint32_t pid, uint32_t s, float d[3]) {
    const uint32_t x = pid % c->width, y = pid / c->width;
    const float inv_w = 1.0f / (float) c->width, inv_h = 1.0f / (float) c->height;
    const float tan_x = c->tan_half_fov, tan_y = c->tan_half_fov * ((float) c->height / (float) c->width);
    float jx = 0.5f, jy = 0.5f;
    if (c->spp > 1) {
        const unsigned long long k = (unsigned long long) pid * c->spp + s;
        const uint32_t key = (uint32_t) k ^ ((uint32_t) (k >> 32) * 0x85EBCA6Bu);
        jx = hash_uniform(c->seed, key, 0); jy = hash_uniform(c->seed, key, 1);
    }
    const float fx = ((float) x + jx) * inv_w, fy = ((float) y + jy) * inv_h;
    const float sx = fmaf(2.f, fx, -1.f) * tan_x, sy = fmaf(-2.f, fy, 1.f) * tan_y;
    float v[3];
    for (int a = 0; a < 3; ++a) v[a] = fmaf(sy, c->up[a], fmaf(sx, c->right[a], c->forward[a]));
    normalize3(v, d);
}

void hfo_camera_rays(const hfo_camera *c, const uint32_t *pixels, size_t npix,
                     float *ox, float *oy, float *oz, float *dx, float *dy, float *dz, float *maxt) {
    #pragma omp parallel for schedule(static)
    for (long long kk = 0; kk < (long long) npix; ++kk) {
        const size_t k = (size_t) kk;
        const uint32_t pid = pixels ? pixels[k] : (uint32_t) k;
        for (uint32_t s = 0; s < c->spp; ++s) {
            const size_t i = k * c->spp + s;
            float d[3];
            camera_dir(c, pid, s, d);
            ox[i] = c->origin[0]; oy[i] = c->origin[1]; oz[i] = c->origin[2];
            dx[i] = d[0]; dy[i] = d[1]; dz[i] = d[2]; maxt[i] = INFINITY;
        }
    }
}

/* Per-sample radiance (fp32, the kernel's op order) and the hit triangle of every LIT sample (0xffffffff otherwise).
 * pi (nullable) receives the preliminary intersection of every sample. */
void hfo_render_samples(const hfo_shape *sh, const hfo_pyramid *p, const hfo_camera *c, const hfo_light *l,
                        const uint32_t *pixels, size_t npix, float *value, uint32_t *prim_out, hfo_prelim *pi) {
    grid_t g = make_grid(sh);
    const float scale = (float) ((double) l->albedo / 3.14159265358979323846 * (double) l->irradiance);
    float o_obj[3], l_obj[3];
    xform_point(sh->to_object, c->origin, o_obj);
    xform_vector(sh->to_object, l->direction, l_obj);
    #pragma omp parallel for schedule(dynamic, 16)
    for (long long kk = 0; kk < (long long) npix; ++kk) {
        const size_t k = (size_t) kk;
        const uint32_t pid = pixels ? pixels[k] : (uint32_t) k;
        for (uint32_t s = 0; s < c->spp; ++s) {
            const size_t i = k * c->spp + s;
            float dw[3], d[3], t = 0, u = 0, v = 0; uint32_t prim = 0;
            camera_dir(c, pid, s, dw);
            xform_vector(sh->to_object, dw, d);
            const int hit = traverse(&g, p, o_obj, d, INFINITY, 0, &t, &u, &v, &prim, NULL, NULL);
            if (pi) {
                if (hit) { pi->t[i] = t; pi->u[i] = u; pi->v[i] = v; pi->prim_index[i] = prim; }
                else write_miss(pi, i);
            }
            value[i] = 0.f; prim_out[i] = 0xffffffffu;
            if (!hit) continue;
            int cx, cy, kt, vi[3], vj[3];
            float p0[3], p1[3], p2[3], e1[3], e2[3], cr[3], no[3], nt[3], nw[3];
            decode_prim(&g, prim, &cx, &cy, &kt);
            cell_triangle(&g, cx, cy, kt, p0, p1, p2, vi, vj);
            for (int a = 0; a < 3; ++a) { e1[a] = p1[a] - p0[a]; e2[a] = p2[a] - p0[a]; }
            cross3(e1, e2, cr);
            normalize3(cr, no);
            xform_normal(sh->to_object, no, nt);
            normalize3(nt, nw);
            if (sh->flip_normals) for (int a = 0; a < 3; ++a) nw[a] = -nw[a];
            const float cosl = dot3(nw, l->direction);
            const float mdw[3] = { -dw[0], -dw[1], -dw[2] };
            if (!(dot3(nw, mdw) > 0.f) || !(cosl > 0.f)) continue;      /* one-sided BSDF, src/bsdfs/diffuse.cpp:127-139 */
            if (l->shadow_rays) {
                const float b1 = u, b2 = v, b0 = 1.f - b1 - b2;
                float po[3], pw[3], os[3], oso[3];
                for (int a = 0; a < 3; ++a) po[a] = fmaf(p0[a], b0, fmaf(p1[a], b1, p2[a] * b2));
                xform_point(sh->to_world, po, pw);
                float mag = (1.f + fmaxf(fmaxf(fabsf(pw[0]), fabsf(pw[1])), fabsf(pw[2]))) * 8.940696716308594e-05f;
                mag = mulsign_(mag, cosl);
                for (int a = 0; a < 3; ++a) os[a] = fmaf(mag, nw[a], pw[a]);
                xform_point(sh->to_object, os, oso);
                float t2, u2, v2; uint32_t prim2;
                if (traverse(&g, p, oso, l_obj, 3.4028234663852886e+38f, 1, &t2, &u2, &v2, &prim2, NULL, NULL)) continue;
            }
            value[i] = scale * cosl;
            prim_out[i] = prim;
        }
    }
}

/* Path tracing on the heightfield alone: src/integrators/path.cpp:133-250 for one Lambertian surface
 * under one directional emitter -- emitter sampling at every vertex, cosine-weighted bounce (include/mitsuba/core/warp.h:54-90,
 * 320-328; frame of include/mitsuba/core/frame.h:24-41), throughput *= albedo. The primal arithmetic is fp32 in the kernel's
 * op order. Per sample it returns the radiance, the triangle of every path vertex (0xffffffff past the end) and, in double,
 * the vector g_k with d(radiance)/d(n_k) = g_k (src/python/python/ad/integrators/prb.py:98-254: sampled directions, hit
 * positions and visibility detached): g_k = albedo^k (lit_k scale l + albedo Lo_{k+1} wo_k / dot(n_k, wo_k)). */
#define HFO_MAX_DEPTH 4
static void cosine_hemisphere(float u1, float u2, float out[3]) {
    const float x = fmaf(2.f, u1, -1.f), y = fmaf(2.f, u2, -1.f);
    const int zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    const float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = (0.25f * 3.14159265358979323846f * rp) / r;
    if (q13) phi = 0.5f * 3.14159265358979323846f - phi;
    if (zero) phi = 0.f;
    const float sn = (float) sin((double) phi), cs = (float) cos((double) phi);   /* through double: bit-identical to the GPU */
    const float px = r * cs, py = r * sn;
    const float z2 = 1.f - fmaf(py, py, px * px);
    out[0] = px; out[1] = py; out[2] = sqrtf(fmaxf(z2, 0.f));
}

void hfo_render_path_samples(const hfo_shape *sh, const hfo_pyramid *p, const hfo_camera *c, const hfo_light *l,
                             const uint32_t *pixels, size_t npix, float *value, uint32_t *vprim, double *vadj, double *vinfo) {
    grid_t g = make_grid(sh);
    const float scale = (float) ((double) l->albedo / 3.14159265358979323846 * (double) l->irradiance);
    const int max_depth = l->max_depth < 1 ? 1 : (l->max_depth > HFO_MAX_DEPTH ? HFO_MAX_DEPTH : l->max_depth);
    float o_cam[3], l_obj[3];
    xform_point(sh->to_object, c->origin, o_cam);
    xform_vector(sh->to_object, l->direction, l_obj);
    #pragma omp parallel for schedule(dynamic, 16)
    for (long long kk = 0; kk < (long long) npix; ++kk) {
        const size_t kpx = (size_t) kk;
        const uint32_t pid = pixels ? pixels[kpx] : (uint32_t) kpx;
        for (uint32_t s = 0; s < c->spp; ++s) {
            const size_t i = kpx * c->spp + s;
            const unsigned long long kq = (unsigned long long) pid * c->spp + s;
            const uint32_t key = (uint32_t) kq ^ ((uint32_t) (kq >> 32) * 0x85EBCA6Bu);
            float dw[3], o[3], d[3], maxt = INFINITY, beta = 1.f, L = 0.f;
            camera_dir(c, pid, s, dw);
            memcpy(o, o_cam, sizeof(o));
            xform_vector(sh->to_object, dw, d);
            int nv = 0;
            float ld[HFO_MAX_DEPTH], lit[HFO_MAX_DEPTH], wo_[HFO_MAX_DEPTH][3], cwo_[HFO_MAX_DEPTH];
            for (int k = 0; k < HFO_MAX_DEPTH; ++k) { vprim[i * HFO_MAX_DEPTH + k] = 0xffffffffu; for (int a = 0; a < 3; ++a) vadj[(i * HFO_MAX_DEPTH + k) * 3 + a] = 0.0; }
            for (int k = 0; k < max_depth; ++k) {
                float t = 0, u = 0, v = 0; uint32_t prim = 0;
                if (!traverse(&g, p, o, d, maxt, 0, &t, &u, &v, &prim, NULL, NULL)) break;
                int cx, cy, kt, vi[3], vj[3];
                float p0[3], p1[3], p2[3], e1[3], e2[3], cr[3], no[3], nt[3], nw[3];
                decode_prim(&g, prim, &cx, &cy, &kt);
                cell_triangle(&g, cx, cy, kt, p0, p1, p2, vi, vj);
                for (int a = 0; a < 3; ++a) { e1[a] = p1[a] - p0[a]; e2[a] = p2[a] - p0[a]; }
                cross3(e1, e2, cr); normalize3(cr, no); xform_normal(sh->to_object, no, nt); normalize3(nt, nw);
                if (sh->flip_normals) for (int a = 0; a < 3; ++a) nw[a] = -nw[a];
                const float mdw[3] = { -dw[0], -dw[1], -dw[2] };
                if (!(dot3(nw, mdw) > 0.f)) break;
                const float cosl = dot3(nw, l->direction);
                int is_lit = cosl > 0.f;
                const float b1 = u, b2 = v, b0 = 1.f - b1 - b2;
                float po[3], pw[3], os[3], oso[3];
                for (int a = 0; a < 3; ++a) po[a] = fmaf(p0[a], b0, fmaf(p1[a], b1, p2[a] * b2));
                xform_point(sh->to_world, po, pw);
                const float mag0 = (1.f + fmaxf(fmaxf(fabsf(pw[0]), fabsf(pw[1])), fabsf(pw[2]))) * 8.940696716308594e-05f;
                if (is_lit && l->shadow_rays) {
                    const float mag = mulsign_(mag0, cosl);
                    for (int a = 0; a < 3; ++a) os[a] = fmaf(mag, nw[a], pw[a]);
                    xform_point(sh->to_object, os, oso);
                    float t2, u2, v2; uint32_t prim2;
                    if (traverse(&g, p, oso, l_obj, 3.4028234663852886e+38f, 1, &t2, &u2, &v2, &prim2, NULL, NULL)) is_lit = 0;
                }
                ld[nv] = is_lit ? scale * cosl : 0.f; lit[nv] = is_lit ? 1.f : 0.f;
                wo_[nv][0] = wo_[nv][1] = wo_[nv][2] = 0.f; cwo_[nv] = 0.f;
                vprim[i * HFO_MAX_DEPTH + nv] = prim;
                L = fmaf(beta, ld[nv], L);
                int go_on = k + 1 < max_depth;
                float wo[3] = { 0, 0, 0 };
                if (go_on) {
                    const float u1 = hash_uniform(c->seed, key, 2u + 2u * (uint32_t) k), u2 = hash_uniform(c->seed, key, 3u + 2u * (uint32_t) k);
                    float loc[3], fs[3], ft[3];
                    cosine_hemisphere(u1, u2, loc);
                    coordinate_system(nw, fs, ft);
                    for (int a = 0; a < 3; ++a) wo[a] = fmaf(nw[a], loc[2], fmaf(ft[a], loc[1], fs[a] * loc[0]));
                    const float cwo = dot3(nw, wo);
                    go_on = cwo > 0.f;
                    if (go_on) { memcpy(wo_[nv], wo, sizeof(wo)); cwo_[nv] = cwo; }
                }
                ++nv;
                if (!go_on) break;
                beta = beta * l->albedo;
                const float mag = mulsign_(mag0, cwo_[nv - 1]);
                for (int a = 0; a < 3; ++a) os[a] = fmaf(mag, nw[a], pw[a]);
                xform_point(sh->to_object, os, o);
                xform_vector(sh->to_object, wo, d);
                memcpy(dw, wo, sizeof(wo));
                maxt = 3.4028234663852886e+38f;
            }
            value[i] = L;
            double run = 0.0, lo_next[HFO_MAX_DEPTH];
            for (int k = HFO_MAX_DEPTH - 1; k >= 0; --k) { lo_next[k] = run; if (k < nv) run = (double) ld[k] + (double) l->albedo * run; }
            double bk = 1.0;
            if (vinfo) for (int k = 0; k < HFO_MAX_DEPTH; ++k) {        /* lit, wo (3), dot(n, wo) of every vertex: what a test needs to rebuild L(normals) */
                double *q = vinfo + (i * HFO_MAX_DEPTH + k) * 5;
                q[0] = k < nv ? lit[k] : 0.0; q[1] = k < nv ? wo_[k][0] : 0.0; q[2] = k < nv ? wo_[k][1] : 0.0;
                q[3] = k < nv ? wo_[k][2] : 0.0; q[4] = k < nv ? cwo_[k] : 0.0;
            }
            for (int k = 0; k < nv; ++k) {
                const double kd = bk * (double) lit[k] * (double) scale;
                const double kw = cwo_[k] > 0.f ? bk * (double) l->albedo * lo_next[k] / (double) cwo_[k] : 0.0;
                for (int a = 0; a < 3; ++a) vadj[(i * HFO_MAX_DEPTH + k) * 3 + a] = kd * (double) l->direction[a] + kw * (double) wo_[k][a];
                bk *= (double) l->albedo;
            }
        }
    }
}

void hfo_set_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void) n;
#endif
}
int hfo_get_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
