/*
 * hf_b200.h -- C ABI of the B200-native differentiable `heightfield` shape.
 *
 * This is the drop-in boundary: every entry point replaces one method (or one accel hook)
 * that a Mitsuba 3 `Shape` plugin provides. Citations are into the reference tree
 * (/root/reference): include/mitsuba/render/shape.h unless noted. The reference reaches the
 * hardware through Dr.Jit tracing + Embree user geometry / OptiX custom primitives
 * (src/render/shape.cpp:109-360); this library replaces that route for this one shape with
 * hand-written sm_100a kernels. There is no CPU fallback: every compute entry point fails with
 * HF_ERR_CUDA when no CUDA device is usable.
 *
 * Conventions
 *   - Plain C types only. Pointers named d_* (and all pointers inside hf_rays / hf_prelim /
 *     hf_si / hf_si_adj) are CUDA DEVICE pointers unless the entry point ends in `_host`.
 *   - Structure-of-arrays, fp32 IEEE, indices uint32. A miss is encoded as t = +inf
 *     (include/mitsuba/render/interaction.h:639-646).
 *   - All launches are asynchronous on the `stream` argument (a cudaStream_t passed as void*;
 *     NULL = the legacy default stream).
 *   - Return value: 0 = HF_OK, otherwise an hf_status; hf_last_error() gives the message of the
 *     last failure on the calling thread. No exception crosses the ABI.
 *   - A context is bound to one device and is not thread-safe (one context per GPU).
 */
#ifndef HF_B200_H
#define HF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HF_B200_ABI_VERSION 4

typedef enum {
    HF_OK = 0,
    HF_ERR_INVALID = 1,      /* bad argument (reference: Throw(...) in the plugin constructor) */
    HF_ERR_CUDA = 2,         /* CUDA runtime failure, or no usable device */
    HF_ERR_NCCL = 3,         /* NCCL missing or failed */
    HF_ERR_STATE = 4         /* call order (e.g. query before heights were set) */
} hf_status;

/* RayFlags, bit-identical to include/mitsuba/render/interaction.h:19-69 */
enum {
    HF_RAY_MINIMAL      = 0x1,
    HF_RAY_UV           = 0x2,
    HF_RAY_DPDUV        = 0x4,
    HF_RAY_SHADINGFRAME = 0x8,
    HF_RAY_BOUNDARYTEST = 0x40,   /* fill hf_si.boundary_test (reparameterised integrators, src/python/python/ad/reparam.py) */
    HF_RAY_FOLLOWSHAPE  = 0x80,
    HF_RAY_DETACHSHAPE  = 0x100,
    HF_RAY_ALL          = 0x2 | 0x4 | 0x8
};

typedef struct hf_ctx hf_ctx;

/* Plugin constructor arguments (reference: Shape(props) src/render/shape.cpp:22-78 consumes
 * `to_world`; pattern for gridded data: src/textures/bitmap.cpp:124-143, 259-263). */
typedef struct {
    uint32_t width, height;        /* texels; cells are (width-1) x (height-1); both >= 2 */
    float    max_height;           /* z = max_height * h */
    float    to_world[12];         /* row-major 3x4 affine */
    float    to_object[12];        /* used iff has_to_object != 0; else fp64 inverse of to_world */
    int32_t  has_to_object;
    int32_t  flip_normals;
    int32_t  device;               /* CUDA ordinal */
} hf_desc;

typedef struct {
    const float *ox, *oy, *oz, *dx, *dy, *dz, *maxt;
    const uint8_t *active;         /* nullable: all active */
} hf_rays;

/* PreliminaryIntersection3f (interaction.h:586-691): shape_index is the constant (uint32)-1,
 * shape = this, instance = null -- not stored. */
typedef struct {
    float *t, *u, *v;
    uint32_t *prim_index;          /* 2*(cy*(width-1)+cx) + k */
} hf_prelim;

/* SurfaceInteraction3f fields in world space; NULL pointers are skipped. */
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
    /* RayFlags::BoundaryTest (src/render/mesh.cpp:845-888, the branch of a mesh without vertex normals): 0 on an edge of the hit
     * triangle, 1 at its barycentre; 1e8 where the ray misses (interaction.h:497-498); 0 when the flag is not set. Detached: the
     * backward entry points have no adjoint for it, as in the reference. (ABI 4: the struct grew by this member.) */
    float *boundary_test;
} hf_si;

/* Upstream adjoints of SurfaceInteraction3f fields; NULL = zero. sh_frame.n shares n. */
typedef struct {
    const float *t;
    const float *px, *py, *pz;
    const float *nx, *ny, *nz;
    const float *uvx, *uvy;
    const float *dpdu_x, *dpdu_y, *dpdu_z;
    const float *dpdv_x, *dpdv_y, *dpdv_z;
} hf_si_adj;

/* src/python/python/ad/optimizers.py:263-302. The hyper-parameters are doubles because they are Python floats there:
 * (1 - beta) and the bias correction are evaluated in double before they become fp32 (0.001f, not 1 - 0.999f). */
typedef struct {
    double lr, beta1, beta2, epsilon;
    int32_t mask_updates;
} hf_adam;

typedef struct {
    uint64_t rays, nodes, leaves, hits;   /* accumulated by the *_counted intersect entry */
    uint64_t segments;                    /* ray segments (camera, bounce, visibility) traced by the path-tracing render entries (max_depth > 1) */
} hf_stats;

/* ---- life cycle ---------------------------------------------------------------------- */
const char *hf_last_error(void);
int hf_abi_version(void);

/* plugin constructor / destructor (class.h:195-224, plugin.cpp:174-200) */
int hf_create(hf_ctx **out, const hf_desc *desc);
int hf_destroy(hf_ctx *ctx);

/* ---- traverse()-exposed parameters (shape.h:533-550; linearcurve.cpp:364-382) -------- */
/* "heights": row-major (height, width). src may be host or device memory. */
int hf_set_heights(hf_ctx *ctx, const float *src, int src_on_device, void *stream);
int hf_heights(hf_ctx *ctx, float **d_heights);          /* device pointer, caller may write in place */
int hf_grad(hf_ctx *ctx, float **d_grad);                /* d(loss)/d(heights), same layout */
int hf_zero_grad(hf_ctx *ctx, void *stream);
int hf_set_to_world(hf_ctx *ctx, const float to_world[12], const float *to_object_or_null);
int hf_set_max_height(hf_ctx *ctx, float max_height);
/* parameters_changed(): rebuild the min/max pyramid, refresh the bbox (shape.h:534; the
 * reference instead re-commits Embree / rebuilds the OptiX GAS, scene.cpp:343-367) */
int hf_parameters_changed(hf_ctx *ctx, void *stream);
int hf_parameters_grad_enabled(hf_ctx *ctx, int *enabled); /* shape.h:550 */
int hf_set_grad_enabled(hf_ctx *ctx, int enabled);

/* ---- queries ------------------------------------------------------------------------- */
int hf_bbox(hf_ctx *ctx, float out_min_max[6]);          /* shape.h:253 (synchronises) */
int hf_primitive_count(hf_ctx *ctx, uint32_t *count);    /* shape.h:432: 2*(W-1)*(H-1) */
int hf_size(hf_ctx *ctx, uint32_t *width, uint32_t *height);
/* acceleration data of the shape (what Embree / OptiX hold for the reference, shape.h:446-474): level l in
 * 1..levels covers 2^l x 2^l cells per node; *d_nodes = device pointer to width*height interleaved (min,max)
 * pairs of z = max_height*h, row-major. levels/width/height/d_nodes are each nullable. */
int hf_pyramid_level(hf_ctx *ctx, uint32_t level, uint32_t *levels, uint32_t *width, uint32_t *height,
                     const float **d_nodes);
/* the quantised copy of that data which the production walk reads (csrc/hf_qpyr.h): level l in 0..levels (level 0 = one node per
 * cell), padded to a power of two per axis: pitch x rows uint32 nodes, row-major, (hi << 16) | lo with
 * decode(code) = fmaf(float(code), qs, zb) a lower (lo) / upper (hi) bound of the node's z range; padding nodes hold 0x0000FFFF.
 * Synchronises (qs, zb derive from the root node on the device). Every output is nullable. */
int hf_qpyramid_level(hf_ctx *ctx, uint32_t level, uint32_t *levels, uint32_t *pitch, uint32_t *rows, const uint32_t **d_nodes,
                      float *qs, float *zb);

/* Hint for the ray-array entry points below, the `coherent` argument of the reference's Scene::ray_intersect / ray_intersect_preliminary /
 * ray_test (src/render/scene.cpp:141-170; Embree picks its traversal by it): coherent != 0 (default) = consecutive rays are neighbours
 * (samples of a pixel, an image tile): a warp walks 32 consecutive rays in lock step; coherent == 0 = unrelated rays: every lane keeps
 * its own ray and refills from a pool (5-18 % on intersect-only hemisphere rays, measured). Results are bit-identical either way. */
int hf_set_ray_coherence(hf_ctx *ctx, int coherent);

/* Adaptive schedule (off by default) for the self-scheduling kernels behind hf_ray_intersect / hf_ray_intersect_preliminary (coherent
 * batches, unit of work = a group of 32 rays) and behind the render steps below (unit = a pixel with >= 32 samples). An optimisation
 * traces (nearly) the same rays every iteration: each launch records what every unit cost, and the next launch over a batch of the same
 * size hands out the units far above the mean FIRST (rays that skim down a slope keep a warp ~50x longer than the mean; at the end of
 * a kernel such a unit leaves the rest of the GPU idle). Only launches of 2^14 units or more and short enough for their end to matter
 * (<= 320 units per resident warp) take part. Results do not depend on the schedule. The launches of a context that use it
 * must be ordered with respect to each other (one stream, or one CUDA graph), like the launches that scatter into its gradient grid.
 * No reference counterpart (Dr.Jit sizes its launches per kernel, ext/drjit/ext/drjit-core/src/cuda_eval.cpp). */
int hf_set_adaptive_schedule(hf_ctx *ctx, int on);
/* what the last such launch left for the next one: `listed` of its `units` go first (0 / 0: nothing learnt yet). kind 0: ray batches,
 * kind 1: the pixel lists of the render steps. Synchronises. */
int hf_adaptive_schedule_info(hf_ctx *ctx, int kind, uint32_t *listed, uint32_t *units);

/* shape.h:137-138 ray_intersect_preliminary */
int hf_ray_intersect_preliminary(hf_ctx *ctx, const hf_rays *rays, const hf_prelim *out, size_t n, void *stream);
/* shape.h:153 ray_test */
int hf_ray_test(hf_ctx *ctx, const hf_rays *rays, uint8_t *d_hit, size_t n, void *stream);
/* shape.h:179-183 compute_surface_interaction + interaction.h:476-499 finalize_surface_interaction */
int hf_compute_surface_interaction(hf_ctx *ctx, const hf_rays *rays, const hf_prelim *pi, uint32_t ray_flags,
                                   const hf_si *out, size_t n, void *stream);
/* shape.h:197 ray_intersect = preliminary + surface interaction in one launch (prelim_out nullable) */
int hf_ray_intersect(hf_ctx *ctx, const hf_rays *rays, uint32_t ray_flags, const hf_prelim *prelim_out,
                     const hf_si *out, size_t n, void *stream);
/* reverse-mode AD of compute_surface_interaction w.r.t. "heights" (mesh.cpp:695-789 on the hit
 * triangle; gather adjoint = scatter_reduce(Add), ext/drjit/src/autodiff/autodiff.cpp:1309-1328) */
int hf_backward(hf_ctx *ctx, const hf_rays *rays, const hf_prelim *pi, const hf_si_adj *adj, uint32_t ray_flags,
                size_t n, void *stream);

/* hf_ray_intersect_preliminary without a context and without any device allocation. No counterpart in the reference (there every
 * array, the acceleration structures included, comes out of Dr.Jit's allocator); added for callers that cannot go to the CUDA allocator:
 * a plugin whose host application has given the device-memory budget to its own pool and forbids a foreign cudaMalloc, a loop with a
 * latency bound that cannot take cudaMalloc's device-wide synchronisation, a process whose memory was partitioned up front. hf_create
 * is out of reach for them (it allocates the grids), so this call takes the description, the caller's own DEVICE copy of the heights
 * (row-major (height, width), read only), the rays and the result arrays, and keeps the acceleration data of the walk in a fixed
 * array inside the library image: HF_STATIC_SCRATCH_BYTES (9 MiB) of device memory on every GPU this entry point has run on, held
 * until the process exits. Limit: the pyramids of the field must fit that array -- every field up to 1025 x 1025 texels does; a
 * larger one is refused with HF_ERR_INVALID and a message that names both sizes (use a context then). Results are bit-identical to
 * hf_ray_intersect_preliminary on a context with the same description and heights.
 *   - Each call clears its part of the array and rebuilds the pyramids from d_heights, in order on `stream`, before the walk: the
 *     heights may change between calls, and nothing carries over from one call to the next.
 *   - The array is one per GPU for the whole process. Calls that can overlap on the GPU (two streams with no event between them,
 *     two host threads) race on it; the caller serialises them.
 *   - The first call that targets a device looks up the array's device address and loads the kernels; make it before any stream
 *     capture begins. desc->device selects the GPU (it is made current).
 *   - It calls no cudaMalloc / cudaFree and no allocating library routine; the only memory it writes is the array and `out`. */
#define HF_STATIC_SCRATCH_BYTES ((size_t) 9 << 20)
int hf_ray_intersect_preliminary_static(const hf_desc *desc, const float *d_heights, const hf_rays *rays, const hf_prelim *out,
                                        size_t n, void *stream);

/* same as hf_ray_intersect_preliminary, also accumulating traversal counters (slower) */
int hf_ray_intersect_preliminary_counted(hf_ctx *ctx, const hf_rays *rays, const hf_prelim *out, size_t n, void *stream);
int hf_stats_get(hf_ctx *ctx, hf_stats *out, int reset);

/* ---- optimiser + multi-GPU (caller side in the reference) ----------------------------- */
int hf_adam_step(hf_ctx *ctx, const hf_adam *cfg, void *stream);   /* updates heights in place, t += 1 */
int hf_adam_reset(hf_ctx *ctx, void *stream);

/* one NCCL communicator per context; the unique id is created on rank 0 and shipped by the caller */
int hf_comm_unique_id(uint8_t out_id[128]);
/* With nranks > 1 hf_comm_init moves the gradient grid into NCCL-allocated, communicator-registered memory (in-switch reduction on
   NVSwitch boxes): the pointer hf_grad() returns changes, fetch it again. HF_B200_NCCL_MEM=0 keeps the original buffer. */
int hf_comm_init(hf_ctx *ctx, int rank, int nranks, const uint8_t id[128]);
int hf_comm_destroy(hf_ctx *ctx);
int hf_allreduce_grad(hf_ctx *ctx, void *stream);                  /* ncclAllReduce(grad, fp32, sum) */

/* ---- multi-GPU over peer memory (one NVSwitch box, one process per GPU) -----------------------------------------------------
 * The fused alternative to "scatter, then ONE ncclAllReduce of the grid": the ranks map each other's grids (CUDA IPC over NVLink),
 * the grid is cut into nranks slabs, rank r OWNS slab r, and the collective is written as kernels of this library:
 *   mode 1 (default)  every kernel scatters into its LOCAL grid as on one GPU; hf_peer_reduce_grad then (A) stores each partial slab
 *                     into its owner's staging area (peer stores at link rate), (barrier), (B) the owner sums the partials in rank
 *                     order and stores the sum into EVERY rank's grid -- reduce-scatter and all-gather in two kernels; with
 *                     hf_peer_adam_step kernel B applies Adam to the slab instead and stores the updated HEIGHTS into every rank's
 *                     height grid (reduce-scatter -> sharded Adam -> all-gather, the gradient never travels back).
 *   mode 0            every scatter of every kernel goes straight to the owner's grid with red.global.add.f32 over NVLink (the
 *                     reduce-scatter happens inside the producing kernel); measured SLOWER than mode 1 and than NCCL: remote atomics
 *                     reach ~2 G/s (profiles/r02_notes.md). Kept for the record and the tests.
 *   set-up   every rank: hf_peer_export -> 256 bytes; ship them (like the NCCL id); every rank: hf_peer_attach(rank, nranks, all)
 *   a step   hf_peer_zero_grad -> forward / backward / render entries -> hf_peer_reduce_grad | hf_peer_adam_step [-> hf_parameters_changed]
 * All of it is stream-ordered and capturable: the cross-GPU barrier is a kernel (counters in peer memory; hf_peer_barrier exports it).
 * While attached use hf_peer_zero_grad, not hf_zero_grad. hf_peer_export moves the gradient grid out of NCCL-registered memory if
 * hf_comm_init put it there (CUDA IPC shares cudaMalloc memory): fetch hf_grad() again. Results: every rank ends with bit-identical
 * grids (one owner sums in a fixed order and everybody receives its stores). */
#define HF_PEER_HANDLE_BYTES 256
int hf_peer_export(hf_ctx *ctx, uint8_t out_handles[HF_PEER_HANDLE_BYTES]);
int hf_peer_attach(hf_ctx *ctx, int rank, int nranks, const uint8_t *all_handles /* nranks x HF_PEER_HANDLE_BYTES, rank-major */);
int hf_peer_set_mode(hf_ctx *ctx, int mode);
int hf_peer_detach(hf_ctx *ctx);
int hf_peer_barrier(hf_ctx *ctx, void *stream);
int hf_peer_zero_grad(hf_ctx *ctx, void *stream);
int hf_peer_reduce_grad(hf_ctx *ctx, void *stream);                 /* afterwards every rank's gradient grid holds the sum over ranks */
int hf_peer_adam_step(hf_ctx *ctx, const hf_adam *cfg, void *stream); /* afterwards every rank's heights hold the Adam update of the summed gradient */

/* ---- render step on the heightfield alone ----------------------------------------------
 * The callers of the shape in the reference, restated for ONE heightfield lit by ONE directional emitter and seen by
 * ONE perspective sensor, so that an inverse-rendering step runs without materialising rays or interaction records:
 *   sensor      perspective pinhole, src/sensors/perspective.cpp:198-234 (origin = camera position, normalised direction,
 *               maxt = +inf); sample index = pixel * spp + sample as in src/render/integrator.cpp:251-265; the sub-pixel
 *               position comes from an integer hash of (seed, index) instead of the PCG32 sampler stream
 *   emitter     directional, src/emitters/directional.cpp (irradiance E arriving along `direction`)
 *   bsdf        Lambertian, src/bsdfs/diffuse.cpp: f * cos = albedo / pi * max(0, n . wo)
 *   integrator  direct illumination of the first hit, src/integrators/direct.cpp:114-210; with shadow_rays != 0 the
 *               emitter is tested for visibility with a ray spawned as in include/mitsuba/render/interaction.h:136-165
 *   film        box reconstruction filter: a sample contributes to its own pixel only (src/render/imageblock.cpp:119-132
 *               with a 0.5-radius box), pixel = mean of its spp samples
 *   adjoint     PRB-style replay (src/python/python/ad/integrators/common.py:784-970): d(image) arrives per pixel, is
 *               spread to the samples, and reaches the heights through the normal of the hit triangle (visibility and
 *               the hit position are detached, as in prb_basic.py without reparameterisation)
 * Images are single-channel fp32, row-major (height, width), indexed by pixel id y * width + x. `d_pixels` lists the
 * pixel ids this context renders (image-tile shard of a rank); NULL = every pixel, npix = width * height. */
typedef struct {
    float origin[3];               /* world space */
    float right[3], up[3], forward[3];   /* orthonormal camera frame, world space */
    float tan_half_fov;            /* of the horizontal field of view */
    uint32_t width, height, spp;
    uint32_t seed;
} hf_camera;

typedef struct {
    float direction[3];            /* unit vector FROM the surface TOWARDS the emitter, world space */
    float irradiance;              /* E */
    float albedo;                  /* Lambertian reflectance */
    int32_t shadow_rays;           /* != 0: one visibility ray per shaded vertex */
    int32_t max_depth;             /* 0 or 1: direct illumination of the first hit (direct.cpp); k in 2..4: path tracing with
                                      k - 1 cosine-sampled diffuse bounces and emitter sampling at every vertex
                                      (src/integrators/path.cpp:133-250); adjoint as in prb.py:98-254 (sampling detached) */
} hf_light;

/* the rays the render entry points generate, written out as an hf_rays-shaped SoA (index = k * spp + sample for the k-th
 * listed pixel); lets a caller run the same batch through hf_ray_intersect / hf_backward */
int hf_render_rays(hf_ctx *ctx, const hf_camera *cam, const uint32_t *d_pixels, size_t npix, const hf_rays *d_out, void *stream);
/* primal image. d_prim (nullable): npix * spp hit triangles (0xffffffff = miss / unlit) for hf_render_backward */
int hf_render_forward(hf_ctx *ctx, const hf_camera *cam, const hf_light *light, const uint32_t *d_pixels, size_t npix,
                      float *d_image, uint32_t *d_prim, void *stream);
/* d(loss)/d(heights) += J^T d_adj_image. d_prim from hf_render_forward of the SAME heights, or NULL to re-trace */
int hf_render_backward(hf_ctx *ctx, const hf_camera *cam, const hf_light *light, const uint32_t *d_pixels, size_t npix,
                       const float *d_adj_image, const uint32_t *d_prim, void *stream);
/* forward + L2 loss against d_target + backward in ONE pass (a warp owns a pixel, so the pixel's adjoint
 * 2 (I - T) / (width * height) is known before its samples leave the registers). Writes the pixels it rendered to
 * d_image (nullable) and ADDS sum (I - T)^2 / (width * height) over them to *d_loss (nullable, device double). */
int hf_render_step_l2(hf_ctx *ctx, const hf_camera *cam, const hf_light *light, const uint32_t *d_pixels, size_t npix,
                      const float *d_target, float *d_image, double *d_loss, void *stream);

/* ---- host-buffer entry points (the llvm_ad_rgb caller keeps its arrays in host memory) ---
 * rays / outputs / adjoints are HOST pointers (pinned memory gives full PCIe bandwidth; use
 * hf_host_alloc). The call stages chunks through device buffers, overlapping H2D, kernels and
 * D2H on internal streams, and returns after everything has landed.
 * Ordering contract: the internal streams are non-blocking. These entry points start behind (1) everything queued on the legacy
 * default stream when they are called and (2) the most recent hf_set_heights / hf_parameters_changed / hf_zero_grad / hf_adam_step of
 * the context, whatever stream that call was given (if those calls went to several streams, or into a stream capture, the entry
 * point synchronises the device instead). Anything else the caller has in flight on a non-blocking stream of its own -- writes
 * through the pointers hf_heights() / hf_grad() return, hf_backward scattering into the grid -- is the caller's to order (an
 * event on the legacy stream, or a synchronisation) before the call. */
int hf_host_alloc(void **out, size_t bytes);
int hf_host_free(void *p);
int hf_ray_intersect_host(hf_ctx *ctx, const hf_rays *h_rays, uint32_t ray_flags, const hf_prelim *h_prelim_out,
                          const hf_si *h_out, size_t n);
/* forward + backward in one pass over host rays and host adjoints; the gradient is accumulated into
 * the context's device grad grid and, when h_grad_out != NULL, copied out (height*width floats). */
int hf_forward_backward_host(hf_ctx *ctx, const hf_rays *h_rays, const hf_si_adj *h_adj, uint32_t ray_flags,
                             const hf_prelim *h_prelim_out, const hf_si *h_out, float *h_grad_out, size_t n);

/* the optimisation step (hf_render_step_l2) for a caller whose arrays live in host memory: h_heights (nullable: keep the device copy;
 * otherwise uploaded and the acceleration data rebuilt), the target image and the pixel list (nullable = every pixel) go in; the image
 * (nullable; pixels not rendered are 0), the gradient grid d(loss)/d(heights) (nullable; height*width floats) and the loss (nullable)
 * come out. Zeroes the gradient grid first. Returns after everything has landed. */
int hf_render_step_l2_host(hf_ctx *ctx, const hf_camera *cam, const hf_light *light, const uint32_t *h_pixels, size_t npix,
                           const float *h_heights, const float *h_target, float *h_image_out, float *h_grad_out, double *h_loss_out);

#ifdef __cplusplus
}
#endif
#endif /* HF_B200_H */
