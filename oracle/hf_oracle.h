/*
 * hf_oracle.h -- CPU oracle for the differentiable heightfield path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing in the product package may include, link
 * or execute this. It is used by tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py, as the checker and as the
 * timed CPU baseline.
 *
 * Parity status: the per-cell triangle primitive is PINNED by the reference's
 * own known-answer tests (src/render/tests/test_mesh.py:257-292, 458-531,
 * 536-671) and by bit-comparison against the reference's Dr.Jit headers
 * (oracle/_ref, see oracle/Makefile). The heightfield level (layout,
 * triangulation, traversal, texel adjoint) is "parity unpinned": the
 * `heightfield` plugin is absent from the reference snapshot,
 * so that level is specified in DESIGN.md and enforced by brute-force-vs-DDA
 * and finite-difference-vs-adjoint self-consistency.
 *
 * All reference citations are relative to /root/reference.
 */
#ifndef HF_ORACLE_H
#define HF_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* RayFlags bit values, include/mitsuba/render/interaction.h:19-69 */
enum {
    HFO_RF_MINIMAL       = 0x1,
    HFO_RF_UV            = 0x2,
    HFO_RF_DPDUV         = 0x4,
    HFO_RF_SHADINGFRAME  = 0x8,
    HFO_RF_BOUNDARYTEST  = 0x40,
    HFO_RF_FOLLOWSHAPE   = 0x80,
    HFO_RF_DETACHSHAPE   = 0x100,
    HFO_RF_ALL           = 0x2 | 0x4 | 0x8
};

typedef struct {
    int32_t W, H;            /* texels; cells are (W-1) x (H-1) */
    float   max_height;
    int32_t flip_normals;
    float   to_world[12];    /* row-major 3x4 affine */
    float   to_object[12];   /* row-major 3x4 affine, inverse of to_world */
    const float *heights;    /* row-major (H, W), row index = y */
} hfo_shape;

/* SoA ray batch (host pointers) */
typedef struct {
    const float *ox, *oy, *oz, *dx, *dy, *dz, *maxt;
    const uint8_t *active;   /* nullable */
} hfo_rays;

typedef struct {
    float *t, *u, *v;
    uint32_t *prim_index;
} hfo_prelim;

/* Any pointer may be NULL (field skipped). World space. */
typedef struct {
    float *t;
    float *px, *py, *pz;
    float *nx, *ny, *nz;
    float *uvx, *uvy;
    float *dpdu_x, *dpdu_y, *dpdu_z;
    float *dpdv_x, *dpdv_y, *dpdv_z;
    float *shn_x, *shn_y, *shn_z;
    float *shs_x, *shs_y, *shs_z;
    float *sht_x, *sht_y, *sht_z;
    float *wi_x, *wi_y, *wi_z;
    float *boundary_test;    /* RayFlags::BoundaryTest (mesh.cpp:845-900, interaction.h:497-498); 0 without the flag */
} hfo_si;

/* Upstream adjoints; any pointer may be NULL (treated as zero). */
typedef struct {
    const float *t;
    const float *px, *py, *pz;
    const float *nx, *ny, *nz;
    const float *uvx, *uvy;
    const float *dpdu_x, *dpdu_y, *dpdu_z;
    const float *dpdv_x, *dpdv_y, *dpdv_z;
} hfo_si_adj;

typedef struct hfo_pyramid hfo_pyramid;

/* -- single-primitive entry points (known-answer tests) ------------------- */

/* include/mitsuba/render/mesh.h:357-380, op order of
 * ext/drjit/include/drjit/array_base.h:603-631, array_router.h:629-661,
 * LLVM lowering of rcp per ext/drjit/ext/drjit-core/src/op.cpp:1194-1203. */
int hfo_moeller_trumbore(const float o[3], const float d[3], float maxt,
                         const float p0[3], const float p1[3], const float p2[3],
                         float *t, float *u, float *v);

/* src/render/mesh.cpp:738-757 and include/mitsuba/core/transform.h:104-139 (golden-vector hooks) */
void hfo_interpolate_and_normal(const float p0[3], const float p1[3], const float p2[3], float b1, float b2,
                                float p_out[3], float n_out[3]);
void hfo_xform_point(const float m[12], const float p[3], float out[3]);
void hfo_xform_vector(const float m[12], const float v[3], float out[3]);

/* The per-cell primitive on cell (cx, cy): both triangles, closest wins, tie -> smaller index.
 * o, d are OBJECT-space. Returns 1 on a hit. (Hook for tests/trav_host.cpp.) */
int hfo_cell_intersect(const hfo_shape *s, int32_t cx, int32_t cy, const float o[3], const float d[3], float maxt,
                       float *t, float *u, float *v, uint32_t *prim);

/* Vertex / texcoord synthesis of the heightfield definition */
float hfo_vertex_x(int32_t i, int32_t W);

/* -- acceleration structure ---------------------------------------------- */
hfo_pyramid *hfo_pyramid_build(const hfo_shape *s);
void  hfo_pyramid_free(hfo_pyramid *p);
int   hfo_pyramid_levels(const hfo_pyramid *p);              /* number of stored levels (1..Ltop) */
void  hfo_pyramid_level_dims(const hfo_pyramid *p, int level, int *w, int *h);
const float *hfo_pyramid_level_data(const hfo_pyramid *p, int level); /* interleaved (min,max) */
void  hfo_bbox(const hfo_shape *s, const hfo_pyramid *p, float out[6]);

/* -- queries -------------------------------------------------------------- */
/* Brute force over every cell: closest hit, tie-break smaller t then smaller prim_index. */
void hfo_intersect_brute(const hfo_shape *s, const hfo_rays *r, size_t n, hfo_prelim *out);

/* Canonical hierarchical DDA. n_node / n_leaf (nullable) receive per-ray counts. */
void hfo_intersect_dda(const hfo_shape *s, const hfo_pyramid *p, const hfo_rays *r, size_t n,
                       hfo_prelim *out, uint32_t *n_node, uint32_t *n_leaf);
void hfo_ray_test(const hfo_shape *s, const hfo_pyramid *p, const hfo_rays *r, size_t n, uint8_t *hit);

void hfo_surface_interaction(const hfo_shape *s, const hfo_rays *r, const hfo_prelim *pi,
                             uint32_t ray_flags, size_t n, hfo_si *out);

/* Reverse mode w.r.t. heights. grad is double[H*W], accumulated (not zeroed).
 * literal != 0: adjoint of the literal op sequence (mesh.cpp:728-789);
 * literal == 0: closed form (p = o + t d). Both in double on fp32 primals. */
void hfo_backward(const hfo_shape *s, const hfo_rays *r, const hfo_prelim *pi, const hfo_si_adj *adj,
                  uint32_t ray_flags, size_t n, double *grad, int literal);

/* src/python/python/ad/optimizers.py:263-302 (fp32, Dr.Jit LLVM lowering) */
void hfo_adam_step(float *p, const float *g, float *m, float *v, size_t n,
                   double lr, double beta1, double beta2, double eps, int step_t, int mask_updates);

/* -- render step on the heightfield alone (sensor + direct illumination) -------- */
typedef struct {
    float origin[3], right[3], up[3], forward[3];
    float tan_half_fov;
    uint32_t width, height, spp, seed;
} hfo_camera;          /* perspective pinhole, src/sensors/perspective.cpp:198-234 */
typedef struct {
    float direction[3];    /* unit, towards the emitter */
    float irradiance, albedo;
    int32_t shadow_rays;
    int32_t max_depth;     /* 0/1 direct illumination; 2..4 path tracing (src/integrators/path.cpp) */
} hfo_light;           /* src/emitters/directional.cpp + src/bsdfs/diffuse.cpp */

void hfo_camera_rays(const hfo_camera *c, const uint32_t *pixels, size_t npix,
                     float *ox, float *oy, float *oz, float *dx, float *dy, float *dz, float *maxt);
void hfo_render_samples(const hfo_shape *s, const hfo_pyramid *p, const hfo_camera *c, const hfo_light *l,
                        const uint32_t *pixels, size_t npix, float *value, uint32_t *prim_out, hfo_prelim *pi);

/* vprim: uint32[n][4], vadj: double[n][4][3] (d radiance / d n_k), n = npix * spp */
void hfo_render_path_samples(const hfo_shape *s, const hfo_pyramid *p, const hfo_camera *c, const hfo_light *l,
                             const uint32_t *pixels, size_t npix, float *value, uint32_t *vprim, double *vadj,
                             double *vinfo /* nullable: double[n][4][5] = lit, wo.xyz, dot(n, wo) */);

void hfo_set_threads(int n);
int  hfo_get_threads(void);

#ifdef __cplusplus
}
#endif
#endif
