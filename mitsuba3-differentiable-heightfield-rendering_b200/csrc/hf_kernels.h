// hf_kernels.h -- launchers of the sm_100a kernels (internal; the public boundary is include/hf_b200.h)
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/hf_b200.h"
#include "hf_device.cuh"

struct HfLaunchInfo {
    int num_sms;
    int top_level_start;     // Ls: pyramid levels >= Ls are staged in shared memory
    int top_nodes;           // node count of levels Ls..Ltop (padded to even)
    int top_off;             // node offset of level Ls
    int use_bundles;         // 1: bundled kernel (hf_intersect_beam_kernel), 0: plain per-ray kernel
    int incoherent;          // 1: the batch's rays are unrelated (hf_set_ray_coherence): lane-pool kernel (hf_intersect_pool_kernel)
    // quantised pyramid: the coarse EVEN levels >= q_stage_level (the levels every walk starts from) are staged in shared memory
    int q_stage_level;       // lowest staged level (even), > QLtop: nothing staged
    int q_stage_bytes;       // shared memory they take (each level padded to 16 bytes)
    unsigned long long *seg_counter;   // ray segments traced by the path kernels (camera, bounce and visibility rays): hf_stats.segments
    // Adaptive schedule of the per-ray kernel (hf_set_adaptive_schedule): every launch records what each 32-ray group cost, a small
    // kernel behind it lists the groups far above the mean, and the NEXT launch over a batch of the same size starts with those.
    // All null / 0 when the launch is not part of it.
    uint16_t *grp_cost;                // out: cycles / 64 (saturating) of every group of this launch
    unsigned long long *cost_sum;      // out: their sum
    unsigned *heavy_n;                 // in: groups listed by the previous launch (a device scalar: graph replays follow it)
    unsigned *heavy_list;              // in: their ids; out (pick kernel): the list for the next launch
    uint32_t *heavy_bits;              // in: one bit per group, set for the listed ones (always rewritten as a whole, heavy_words words)
    unsigned heavy_cap, heavy_words, heavy_shift;   // heavy_shift: (groups of the batch) >> heavy_shift <= 8192 (the kernel's coarse bitmap)
};
#ifndef HF_STAGE_Q
#define HF_STAGE_Q 1
#endif
#ifndef HF_QSTAGE_CAP
#define HF_QSTAGE_CAP 17536  // bytes per CTA: levels 6, 8, 10, 12 of a 4096^2 field (16 KB + 1 KB + 64 B + 16 B); 16384^2 stages 8..14
#endif

enum HfIntersectMode { HF_MODE_PRELIM = 0, HF_MODE_FUSED_SI = 1, HF_MODE_TEST = 2, HF_MODE_COUNTED = 3,
                       HF_MODE_FUSED_SI_LEAN = 4 };   // internal: FUSED_SI when only t, p, n, uv are requested (chosen by hf_launch_intersect)

// adaptive schedule: the heavy list for the next launch from the unit costs this launch has recorded (li.grp_cost etc. must be set)
cudaError_t hf_launch_pick_heavy(const HfLaunchInfo &li, unsigned units, unsigned resident_warps, cudaStream_t s);
cudaError_t hf_launch_pyramid(const HfParams &P, float2 *pyr, uint32_t *qpyr, cudaStream_t s);
cudaError_t hf_launch_intersect(const HfParams &P, const HfLaunchInfo &li, int mode, const hf_rays &rays,
                                const hf_prelim &prelim, const hf_si &si, uint8_t *d_hit, uint32_t flags,
                                size_t n, unsigned long long *d_counters, unsigned long long *d_sched, cudaStream_t s);
cudaError_t hf_launch_surface_interaction(const HfParams &P, const HfLaunchInfo &li, const hf_rays &rays,
                                          const hf_prelim &pi, uint32_t flags, const hf_si &si, size_t n,
                                          cudaStream_t s);
cudaError_t hf_launch_backward(const HfParams &P, const HfLaunchInfo &li, const hf_rays &rays, const hf_prelim &pi,
                               const hf_si_adj &adj, uint32_t flags, float *grad, size_t n, cudaStream_t s);
cudaError_t hf_launch_adam(float *p, const float *g, float *m, float *v, size_t n, int *d_t, float *d_lr_t, const float *d_scale_table,
                           int table_len, float lr, float beta1, float beta2, float omb1, float omb2, float eps, int mask_updates, double beta1_d, double beta2_d,
                           int num_sms, cudaStream_t s);

// multi-GPU over peer memory (hf_peer_attach)
cudaError_t hf_launch_peer_barrier(unsigned *const *d_peer_counter, unsigned *d_my_counter, unsigned *d_my_epoch, int nranks, cudaStream_t s);
cudaError_t hf_launch_peer_push(const float *local, float *const *d_peers, int nranks, int rank, size_t lo, size_t hi, int num_sms, cudaStream_t s);
cudaError_t hf_launch_peer_adam(float *p, const float *g, float *m, float *v, float *const *d_peer_heights, int nranks, int rank, size_t lo, size_t hi,
                                int *d_t, float *d_lr_t, const float *d_scale_table, int table_len, float lr, float beta1, float beta2, float omb1,
                                float omb2, float eps, int mask_updates, double beta1_d, double beta2_d, int num_sms, cudaStream_t s);

cudaError_t hf_launch_peer_send_partials(const float *local, float *const *d_peer_stage, int nranks, int rank, size_t slab, size_t total, int num_sms, cudaStream_t s);
cudaError_t hf_launch_peer_reduce_gather(float *local_grad, float *my_stage, float *const *d_peer_out, int nranks, int rank, size_t lo, size_t hi, size_t slab,
                                         bool adam, float *p, float *m, float *v, int *d_t, float *d_lr_t, const float *d_scale_table, int table_len, float lr,
                                         float beta1, float beta2, float omb1, float omb2, float eps, int mask_updates, double beta1_d, double beta2_d, int num_sms, cudaStream_t s);

// hf_render.cu: mode 0 = forward image (+ hit triangles), 1 = fused forward + L2 adjoint + backward, 2 = backward by re-tracing
#define HF_RENDER_MAX_SPP_HOST 256
cudaError_t hf_launch_render(const HfParams &P, const HfLaunchInfo &li, int mode, const hf_camera &cam, const hf_light &light,
                             const uint32_t *d_pixels, size_t npix, float *d_image, uint32_t *d_prim, const float *d_target,
                             const float *d_adj_image, float *d_grad, double *d_loss, unsigned long long *d_sched, cudaStream_t s);
cudaError_t hf_launch_render_backward_prim(const HfParams &P, const HfLaunchInfo &li, const hf_camera &cam, const hf_light &light,
                                           const uint32_t *d_pixels, size_t npix, const float *d_adj_image, const uint32_t *d_prim,
                                           float *d_grad, cudaStream_t s);
cudaError_t hf_launch_render_rays(const HfParams &P, const HfLaunchInfo &li, const hf_camera &cam, const uint32_t *d_pixels, size_t npix,
                                  const hf_rays &out, cudaStream_t s);
cudaError_t hf_launch_l2_adjoint(const HfParams &P, const HfLaunchInfo &li, const hf_camera &cam, const uint32_t *d_pixels, size_t npix,
                                 const float *d_image, const float *d_target, float *d_adj, double *d_loss, cudaStream_t s);
#define HF_MAX_DEPTH_HOST 4
