// hf_walk_common.cuh -- device helpers shared by the kernels that walk the pyramid (hf_kernels.cu, hf_render.cu):
// mbarrier / bulk-copy PTX, the kernel prologue that stages the coarse pyramid levels in shared memory, and the
// warp-aggregated atomic add of the gradient scatter.
#pragma once
#include "hf_kernels.h"
#include "hf_traverse.cuh"

#ifndef HF_BLOCK
#define HF_BLOCK 256
#endif
#define FULL_MASK 0xffffffffu

// ------------------------------------------------------------------------------------------
// small PTX helpers: mbarrier + 1-D bulk copy (TMA engine, SASS UBLKCP) for pyramid staging
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

// Kernel prologue shared by the walk kernels: level tables in shared memory; the coarse levels are staged next to them by bulk
// async copies (TMA engine, SASS UBLKCP) that complete on one mbarrier:
//   * WANT_FP32: levels >= li.top_level_start of the fp32 pyramid (read by the bundle march, hf_beam_enter);
//   * the coarse EVEN levels >= li.q_stage_level of the quantised pyramid (read by every walk: each ray starts at level
//     min(QLtop, HF_START_LEVEL) and descends two levels at a time, so its first descents are shared-memory loads).
// Both level tables hold GENERIC pointers: the walk reads a node with one load whether its level lives in shared or global memory.
// Dynamic shared memory: [fp32 nodes: li.top_nodes * 8 bytes if WANT_FP32][quantised levels: li.q_stage_bytes].
template <bool WANT_FP32>
__device__ __forceinline__ void hf_stage_pyramid(const HfParams &P, const HfLaunchInfo &li, HfGrid &G) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HfLevel *s_lvl = hf_s_lvl;                                                                // static __shared__ (hf_traverse.cuh)
    const bool fp32_staged = WANT_FP32 && li.top_nodes > 0;
    float2 *s_nodes = reinterpret_cast<float2 *>(smem_raw);                                   // levels >= li.top_level_start
    uint32_t *s_q = reinterpret_cast<uint32_t *>(smem_raw + (fp32_staged ? (size_t) li.top_nodes * sizeof(float2) : 0));
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_qoff[HF_MAX_LEVELS];             // byte offset of a staged level inside s_q

    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        uint32_t o = 0;
        for (int l = li.q_stage_level; l <= P.QLtop; l += 2) {
            s_qoff[l] = o;
            const uint32_t pitch = ((P.xmask + 1) >> l) > 0 ? ((P.xmask + 1) >> l) : 1, rows = ((P.ymask + 1) >> l) > 0 ? ((P.ymask + 1) >> l) : 1;
            o += ((pitch * rows + 3u) & ~3u) * 4u;
        }
    }
    __syncthreads();
    if (threadIdx.x < HF_MAX_LEVELS) {
        const int l = threadIdx.x;
        HfLevel e;
        e.nodes = (fp32_staged && l >= li.top_level_start) ? (s_nodes + (P.lvl_off[l] - li.top_off)) : (P.pyr + P.lvl_off[l]);
        e.w = P.lvl_w[l]; e.pad = 0;
        s_lvl[l] = e;
        HfQLevel qe;
        const bool staged = HF_STAGE_Q && l >= li.q_stage_level && l <= P.QLtop && !(l & 1);
        qe.nodes = staged ? reinterpret_cast<const uint32_t *>(reinterpret_cast<const unsigned char *>(s_q) + s_qoff[l]) : (P.qpyr + P.qoff[l]);
        qe.pitch = ((P.xmask + 1) >> l) > 0 ? ((P.xmask + 1) >> l) : 1; qe.pad = 0;
        hf_s_qlvl[l] = qe;
    }
    __syncthreads();
    const uint32_t fp32_bytes = fp32_staged ? (uint32_t) li.top_nodes * (uint32_t) sizeof(float2) : 0u;
    const uint32_t q_bytes = HF_STAGE_Q ? (uint32_t) li.q_stage_bytes : 0u;
    if (fp32_bytes + q_bytes > 0u) {
        if (threadIdx.x == 0) {
            mbar_expect_tx(&s_bar, fp32_bytes + q_bytes);
            if (fp32_bytes) bulk_g2s(s_nodes, P.pyr + li.top_off, fp32_bytes, &s_bar);
            if (q_bytes)
                for (int l = li.q_stage_level; l <= P.QLtop; l += 2) {
                    const uint32_t pitch = ((P.xmask + 1) >> l) > 0 ? ((P.xmask + 1) >> l) : 1, rows = ((P.ymask + 1) >> l) > 0 ? ((P.ymask + 1) >> l) : 1;
                    bulk_g2s(reinterpret_cast<unsigned char *>(s_q) + s_qoff[l], P.qpyr + P.qoff[l], ((pitch * rows + 3u) & ~3u) * 4u, &s_bar);
                }
        }
        uint32_t spins = 0;
        while (!mbar_try_wait(&s_bar, 0)) { if (++spins > (1u << 24)) __trap(); }
    }

    G.W = P.W; G.NX = P.NX; G.NY = P.NY; G.Ltop = P.Ltop; G.mh = P.mh;
    G.x0 = P.x0; G.y0 = P.y0; G.sgx = P.sgx; G.sgy = P.sgy;
    G.heights = P.heights; G.lvl = s_lvl;
    {
        const float2 top = s_lvl[P.Ltop].nodes[0];
        G.top_min = top.x; G.top_max = top.y;
        const HfQuant Q = hf_quant_make(top.x, top.y);         // the same expression the builder used
        G.qs = Q.qs; G.zb = Q.zb; G.rqs = Q.rqs;
    }
    G.qlvl = hf_s_qlvl; G.qlvl_saddr = smem_u32(hf_s_qlvl);
    G.QLtop = P.QLtop; G.xmask = P.xmask; G.ymask = P.ymask;
}
// dynamic shared memory a walk kernel needs
static inline size_t hf_walk_smem(const HfLaunchInfo &li, bool want_fp32) {
    return (want_fp32 ? (size_t) li.top_nodes * sizeof(float2) : 0) + (HF_STAGE_Q ? (size_t) li.q_stage_bytes : 0);
}

// One RED per distinct address per warp: match.any on the texel index, then a log-step segmented
// reduction inside each peer group (cf. the reference JIT's reduce_add_f32,
// ext/drjit/ext/drjit-core/src/cuda_eval.cpp:805-888, which walks its peers serially).
__device__ __forceinline__ void warp_aggregated_add(const HfParams &P, float *__restrict__ grad, uint32_t idx, float val, bool valid) {
    const unsigned act = __ballot_sync(FULL_MASK, valid);
    if (!valid) return;
    const unsigned lane = threadIdx.x & 31u;
    unsigned peers = __match_any_sync(act, idx);
    const bool leader = (peers & ((1u << lane) - 1u)) == 0u;
    unsigned rel = __popc(peers & ((1u << lane) - 1u));        // my rank inside the group
    peers &= (0xfffffffeu << lane);                            // peers above me
    while (__any_sync(act, peers != 0u)) {
        const int next = __ffs((int) peers);                   // 1-based lane of the next peer above, 0 = none
        const float other = __shfl_sync(act, val, next ? next - 1 : (int) lane);
        if (next) val += other;
        // ranks that are odd retire: drop them from everybody's peer set
        const unsigned retired = __ballot_sync(act, (rel & 1u) != 0u);
        peers &= ~retired;
        rel >>= 1;
    }
    if (leader) hf_grad_add(P, grad, idx, val);
}

