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
};

enum HfIntersectMode { HF_MODE_PRELIM = 0, HF_MODE_FUSED_SI = 1, HF_MODE_TEST = 2, HF_MODE_COUNTED = 3 };

cudaError_t hf_launch_pyramid(const HfParams &P, float2 *pyr, cudaStream_t s);
cudaError_t hf_launch_intersect(const HfParams &P, const HfLaunchInfo &li, int mode, const hf_rays &rays,
                                const hf_prelim &prelim, const hf_si &si, uint8_t *d_hit, uint32_t flags,
                                size_t n, unsigned long long *d_counters, cudaStream_t s);
cudaError_t hf_launch_surface_interaction(const HfParams &P, const HfLaunchInfo &li, const hf_rays &rays,
                                          const hf_prelim &pi, uint32_t flags, const hf_si &si, size_t n,
                                          cudaStream_t s);
cudaError_t hf_launch_backward(const HfParams &P, const HfLaunchInfo &li, const hf_rays &rays, const hf_prelim &pi,
                               const hf_si_adj &adj, uint32_t flags, float *grad, size_t n, cudaStream_t s);
cudaError_t hf_launch_adam(float *p, const float *g, float *m, float *v, size_t n, float lr_t, float beta1,
                           float beta2, float omb1, float omb2, float eps, int mask_updates, int num_sms,
                           cudaStream_t s);
