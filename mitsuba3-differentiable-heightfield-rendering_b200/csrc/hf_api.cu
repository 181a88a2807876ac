// hf_api.cu -- context object and the C ABI of include/hf_b200.h (host side, C++).
//
// One hf_ctx == one `heightfield` shape instance on one GPU. It owns the traverse()-exposed
// parameter ("heights"), its gradient, the Adam state and the shape's acceleration data
// (min/max pyramid), i.e. what the reference keeps in Dr.Jit arrays plus Embree/OptiX handles
// (src/render/shape.cpp:109-360). No CPU fallback exists: without a CUDA device every entry
// point that computes returns HF_ERR_CUDA.
#include <dlfcn.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "hf_kernels.h"
#include "hf_qpyr.h"

// NVTX ranges around the phases of a step (pyramid build, intersect, backward, collective, Adam, the host pipelines, the render
// steps), the counterpart of the reference's ScopedPhase profiler (include/mitsuba/core/profiler.h:10-112): they name the host-side
// spans of an nsys / ncu timeline of the 8-GPU step. NVTX 3 is header-only; without a tool attached a range costs one pointer test.
#include <nvtx3/nvToolsExt.h>
struct HfRange {
    explicit HfRange(const char *name) { nvtxRangePushA(name); }
    ~HfRange() { nvtxRangePop(); }
    HfRange(const HfRange &) = delete;
    HfRange &operator=(const HfRange &) = delete;
};
#define HF_RANGE() HfRange hf_range_(__func__)

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(HF_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

#define CHECK_CTX(ctx) do { if (!(ctx)) return fail(HF_ERR_INVALID, "null context"); } while (0)

extern "C" const char *hf_last_error(void) { return g_err; }
extern "C" int hf_abi_version(void) { return HF_B200_ABI_VERSION; }

// ------------------------------------------------------------------------------------------
// NCCL through dlopen: the library has no link-time dependency on NCCL and loads without it
// ------------------------------------------------------------------------------------------
typedef struct { char internal[128]; } hf_nccl_uid;
typedef void *hf_nccl_comm;
struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(hf_nccl_uid *) = nullptr;
    int (*CommInitRank)(hf_nccl_comm *, int, hf_nccl_uid, int) = nullptr;
    int (*CommDestroy)(hf_nccl_comm) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, hf_nccl_comm, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    // optional (NCCL >= 2.19): buffers the switch can reduce in place (NVLS) need ncclMemAlloc + ncclCommRegister
    int (*MemAlloc)(void **, size_t) = nullptr;
    int (*MemFree)(void *) = nullptr;
    int (*CommRegister)(hf_nccl_comm, void *, size_t, void **) = nullptr;
    int (*CommDeregister)(hf_nccl_comm, void *) = nullptr;
};
static NcclApi g_nccl;

static int nccl_load() {
    if (g_nccl.handle) return HF_OK;
    const char *names[] = { "libnccl.so.2", "libnccl.so" };
    void *h = nullptr;
    for (const char *nm : names) { h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
    if (!h) return fail(HF_ERR_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
    NcclApi a;
    a.handle = h;
    a.GetUniqueId = (int (*)(hf_nccl_uid *)) dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (int (*)(hf_nccl_comm *, int, hf_nccl_uid, int)) dlsym(h, "ncclCommInitRank");
    a.CommDestroy = (int (*)(hf_nccl_comm)) dlsym(h, "ncclCommDestroy");
    a.AllReduce = (int (*)(const void *, void *, size_t, int, int, hf_nccl_comm, cudaStream_t)) dlsym(h, "ncclAllReduce");
    a.GetErrorString = (const char *(*) (int) ) dlsym(h, "ncclGetErrorString");
    a.MemAlloc = (int (*)(void **, size_t)) dlsym(h, "ncclMemAlloc");
    a.MemFree = (int (*)(void *)) dlsym(h, "ncclMemFree");
    a.CommRegister = (int (*)(hf_nccl_comm, void *, size_t, void **)) dlsym(h, "ncclCommRegister");
    a.CommDeregister = (int (*)(hf_nccl_comm, void *)) dlsym(h, "ncclCommDeregister");
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.GetErrorString)
        return fail(HF_ERR_NCCL, "libnccl is missing a required symbol");
    g_nccl = a;
    return HF_OK;
}
#define NC(call)                                                                                  \
    do {                                                                                          \
        int r_ = (call);                                                                          \
        if (r_ != 0) return fail(HF_ERR_NCCL, "%s failed: %s", #call, g_nccl.GetErrorString(r_)); \
    } while (0)

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
#define HF_SCHED_SLOTS 256            // every eager launch of a self-scheduling kernel gets its own counter (launches on different streams overlap)
#define HF_SCHED_CAPTURE_SLOTS 1024   // launches recorded into a CUDA graph get a counter of their own for good: a graph bakes the pointer in
                                      // and may replay beside eager launches long after the eager ring has wrapped around
#define HF_ADAM_TABLE 65536            // bias corrections tabulated per (beta1, beta2) with the host's double arithmetic; later steps evaluate it on the device
#define HF_TOP_NODE_CAP 1366          // (32^2 + 16^2 + ... + 1) rounded up to even: 10.9 KB of shared memory
#ifndef HF_HOST_CHUNK_LOG2
#define HF_HOST_CHUNK_LOG2 21
#endif
#define HF_HOST_CHUNK (size_t(1) << HF_HOST_CHUNK_LOG2)
#ifndef HF_HOST_SLOTS
#define HF_HOST_SLOTS 3
#endif
// staging slab of one chunk, in fields of HF_HOST_CHUNK floats: rays | prelim | surface interaction | adjoints | active mask.
// The field counts are DERIVED from the ABI structs, so a field added to hf_si / hf_si_adj cannot be left without a slot.
#define RAY_NFIELDS 7
#define PRE_NFIELDS 4
#define SI_NFIELDS ((int) (sizeof(hf_si) / sizeof(float *)))
#define ADJ_NFIELDS ((int) (sizeof(hf_si_adj) / sizeof(const float *)))
static_assert(sizeof(hf_si) == 28 * sizeof(float *), "hf_si: 28 float* members (t, p3, n3, uv2, dpdu3, dpdv3, shn3, shs3, sht3, wi3, boundary_test)");
static_assert(sizeof(hf_si_adj) == 15 * sizeof(const float *), "hf_si_adj: 15 members");
static_assert(sizeof(hf_prelim) == PRE_NFIELDS * sizeof(void *), "hf_prelim: 4 members");
#define SLAB_RAY 0
#define SLAB_PRE (SLAB_RAY + RAY_NFIELDS)
#define SLAB_SI (SLAB_PRE + PRE_NFIELDS)
#define SLAB_ADJ (SLAB_SI + SI_NFIELDS)
#define SLAB_ACTIVE (SLAB_ADJ + ADJ_NFIELDS)
#define HF_SLAB_FLOATS (SLAB_ACTIVE + 1)

// adaptive schedule of one kind of launch (hf_set_adaptive_schedule): cost per unit of work of the last launch, the heavy list made from it
struct HfAdaptive {
    size_t cap = 0, units = 0;            // units the buffers hold / units of the batch the list was made from
    uint16_t *d_cost = nullptr;
    unsigned *d_list = nullptr, *d_n = nullptr;
    uint32_t *d_bits = nullptr;
    unsigned long long *d_sum = nullptr;
    std::vector<void *> retired;          // outgrown buffers: CUDA graphs recorded earlier may still hold them
    void release() {
        cudaFree(d_cost); cudaFree(d_bits); cudaFree(d_list);
        for (void *p : retired) cudaFree(p);
        retired.clear(); d_cost = nullptr; d_bits = nullptr; d_list = nullptr; d_n = nullptr; d_sum = nullptr; cap = units = 0;
    }
};

struct HostSlot {
    float *slab = nullptr;            // HF_HOST_CHUNK * HF_SLAB_FLOATS floats
    cudaEvent_t in_done = nullptr, k_done = nullptr, out_done = nullptr;   // H2D landed / kernels finished / D2H landed
};

struct hf_ctx {
    int device = 0;
    int num_sms = 0;
    uint32_t W = 0, H = 0;
    HfParams P;
    HfLaunchInfo li;
    size_t pyr_nodes = 0;
    float *d_heights = nullptr, *d_grad = nullptr, *d_m = nullptr, *d_v = nullptr;
    bool grad_nccl = false;           // d_grad comes from ncclMemAlloc (hf_comm_init moved it there)
    void *grad_reg = nullptr;         // ncclCommRegister handle of d_grad
    float2 *d_pyr = nullptr;
    uint32_t *d_qpyr = nullptr;      // quantised pyramid (hf_qpyr.h)
    size_t qpyr_nodes = 0;
    unsigned long long *d_counters = nullptr;     // [0..3] traversal statistics, [8..8+HF_SCHED_SLOTS) tile counters of the persistent kernels
    unsigned sched_next = 0, sched_capture_next = 0;
    float *d_scratch = nullptr;       // two images (path-traced L2 step)
    size_t scratch_bytes = 0;
    bool have_heights = false, accel_valid = false;
    int grad_enabled = 1;
    // Adam: step counter + lr_t on the device (graph replays advance t), bias-correction table per (beta1, beta2)
    int *d_adam_t = nullptr;
    float *d_adam_lr_t = nullptr, *d_adam_table = nullptr;      // d_adam_table: the table of the betas in use (one of adam_tables)
    double tab_b1 = -1.0, tab_b2 = -1.0;
    struct AdamTable { double b1, b2; float *d; };
    std::vector<AdamTable> adam_tables;                         // one per (beta1, beta2) ever used, never rewritten: a CUDA graph keeps the pointer it captured
    hf_nccl_comm comm = nullptr;
    int nranks = 1, rank = 0;
    HostSlot slots[HF_HOST_SLOTS];
    cudaStream_t s_in = nullptr, s_k = nullptr, s_out = nullptr;   // host pipeline: one stream per copy engine + one for the kernels
    cudaEvent_t ev_entry = nullptr;
    cudaEvent_t ev_accel = nullptr;   // recorded behind the last pyramid build (hf_bbox waits for it instead of the whole device)
    bool accel_event_ok = false;
    bool slots_ready = false;
    // what the *_host entry points have to wait for besides the legacy stream (note_mutation / wait_mutations below)
    cudaEvent_t ev_mut = nullptr;
    cudaStream_t mut_stream = nullptr;
    bool mut_pending = false, mut_sync = false;
    // multi-GPU over peer memory (hf_peer_attach): mapped grids of the other ranks, the device-side tables of those pointers, the barrier state
    int peer_n = 1, peer_rank = 0;
    size_t slab_elems = 0;
    float *peer_grad[HF_MAX_PEERS] = { nullptr }, *peer_heights[HF_MAX_PEERS] = { nullptr };
    unsigned *peer_counter[HF_MAX_PEERS] = { nullptr };
    float **d_peer_grad_tab = nullptr, **d_peer_heights_tab = nullptr;
    unsigned **d_peer_counter_tab = nullptr;
    unsigned *d_xcounter = nullptr;       // [0] = this rank's barrier counter (peers add to it), [64] = the epoch (local)
    int peer_mode = 1;                    // 1: scatter locally, reduce-scatter by peer stores (default); 0: scatter to the owner with remote atomics
    float *d_stage = nullptr;             // staging area of this rank: slot r = rank r's partial sums for MY slab
    float *peer_stage[HF_MAX_PEERS] = { nullptr };
    float **d_peer_stage_tab = nullptr;
    // adaptive schedule of the per-ray kernel (hf_set_adaptive_schedule): cost per 32-ray group of the last launch, the heavy list made from it
    int adaptive = 0;
    HfAdaptive as_rays, as_pixels;        // units = 32-ray groups of a ray batch / pixels of a pixel list (render step)
    // host-buffer render step: device copies of the caller's target / image / pixel list, and the loss
    float *d_hs_target = nullptr, *d_hs_image = nullptr;
    uint32_t *d_hs_pixels = nullptr;
    double *d_hs_loss = nullptr;
    size_t hs_img_elems = 0, hs_pix_elems = 0;
};

static void invert_affine(const float m[12], float out[12]) {
    // fp64 inverse of the 3x4 affine, rounded to fp32 (definition of to_object)
    double a[3][3], t[3];
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) a[r][c] = m[r * 4 + c]; t[r] = m[r * 4 + 3]; }
    double det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                 a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    double id = 1.0 / det, inv[3][3];
    inv[0][0] = (a[1][1] * a[2][2] - a[1][2] * a[2][1]) * id;
    inv[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) * id;
    inv[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) * id;
    inv[1][0] = (a[1][2] * a[2][0] - a[1][0] * a[2][2]) * id;
    inv[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) * id;
    inv[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) * id;
    inv[2][0] = (a[1][0] * a[2][1] - a[1][1] * a[2][0]) * id;
    inv[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) * id;
    inv[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) * id;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) out[r * 4 + c] = (float) inv[r][c];
        out[r * 4 + 3] = (float) -(inv[r][0] * t[0] + inv[r][1] * t[1] + inv[r][2] * t[2]);
    }
}

static int set_transform(hf_ctx *c, const float to_world[12], const float *to_object) {
    for (int i = 0; i < 12; ++i) if (!isfinite(to_world[i])) return fail(HF_ERR_INVALID, "to_world has a non-finite entry");
    memcpy(c->P.to_world, to_world, sizeof(float) * 12);
    if (to_object) memcpy(c->P.to_object, to_object, sizeof(float) * 12);
    else invert_affine(to_world, c->P.to_object);
    for (int i = 0; i < 12; ++i) if (!isfinite(c->P.to_object[i])) return fail(HF_ERR_INVALID, "to_world is singular");
    return HF_OK;
}

// Tile counter of one launch of a self-scheduling (persistent) kernel; the launcher zeroes it in stream order. Eager launches take
// the next slot of a ring (two launches that overlap on different streams are never 256 launches apart). A launch that is being
// RECORDED into a CUDA graph gets a slot from a pool that is never handed out again: the graph keeps the pointer for its lifetime
// and may replay at any time beside eager launches of the same context.
static unsigned long long *next_sched(hf_ctx *c, cudaStream_t s) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess) { cudaGetLastError(); st = cudaStreamCaptureStatusNone; }
    if (st == cudaStreamCaptureStatusActive) {
        if (c->sched_capture_next >= HF_SCHED_CAPTURE_SLOTS) return nullptr;
        return c->d_counters + 8 + HF_SCHED_SLOTS + c->sched_capture_next++;
    }
    return c->d_counters + 8 + (c->sched_next++ % HF_SCHED_SLOTS);
}
#define SCHED(c, s) ({ unsigned long long *p_ = next_sched((c), (cudaStream_t) (s)); \
    if (!p_) return fail(HF_ERR_STATE, "this context has recorded more than %d self-scheduling launches into CUDA graphs", HF_SCHED_CAPTURE_SLOTS); p_; })

// Launch info of one self-scheduling launch over `units` units of work (32-ray groups, pixels): c->li plus, when the adaptive schedule
// is on and the launch is of the right size, its buffers. A batch of another size than the last one starts without a list (and
// makes the next one); while a CUDA graph is being recorded nothing may be allocated or reset behind the stream's back, so such a
// launch simply runs the plain schedule.
#define HF_AS_MIN_UNITS 16384u
#define HF_AS_LIST_CAP 4096u
#define HF_AS_MAX_UNITS_PER_WARP 320     // of the 24 resident warps per SM
static HfLaunchInfo adaptive_launch_info(hf_ctx *c, HfAdaptive &A, size_t units, cudaStream_t s) {
    HfLaunchInfo li = c->li;
    if (!c->adaptive || li.incoherent || li.use_bundles || units < HF_AS_MIN_UNITS) return li;
    // The closing phase of a launch weighs ~1 / (units per resident warp); keeping the books costs the kernel ~1 % whatever its
    // length. Measured on C3 (profiles/r02_notes.md): 1/8 of the ray batch (74 groups per warp) -11 %, 1/4 -5 %, 1/2 (295) -1.5 %, the
    // whole batch (590) +1 %. So: only launches short enough for the end to matter.
    static const int env_upw = getenv("HF_B200_AS_MAX_UPW") ? atoi(getenv("HF_B200_AS_MAX_UPW")) : HF_AS_MAX_UNITS_PER_WARP;     // dev knob
    if (units > (size_t) c->num_sms * 24 * (size_t) (env_upw > 0 ? env_upw : HF_AS_MAX_UNITS_PER_WARP) || units >= (size_t(1) << 31)) return li;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess) { cudaGetLastError(); st = cudaStreamCaptureStatusNone; }
    const bool capturing = st == cudaStreamCaptureStatusActive;
    if (units > A.cap) {
        if (capturing) return li;
        const size_t cap = (units + 1023) & ~size_t(1023);
        uint16_t *cost = nullptr; unsigned *list = nullptr; uint32_t *bits = nullptr;
        if (cudaMalloc(&cost, cap * sizeof(uint16_t)) != cudaSuccess || cudaMalloc(&bits, cap / 32 * sizeof(uint32_t)) != cudaSuccess ||
            cudaMalloc(&list, (HF_AS_LIST_CAP + 4) * sizeof(unsigned)) != cudaSuccess) {
            cudaGetLastError(); cudaFree(cost); cudaFree(bits); cudaFree(list);
            return li;                                   // a schedule hint must not fail a launch
        }
        if (A.d_cost) { A.retired.push_back(A.d_cost); A.retired.push_back(A.d_bits); A.retired.push_back(A.d_list); }
        A.d_cost = cost; A.d_bits = bits; A.d_list = list;
        A.d_n = list + HF_AS_LIST_CAP;                   // the count and the cost sum live behind the list
        A.d_sum = reinterpret_cast<unsigned long long *>(list + HF_AS_LIST_CAP + 2);
        A.cap = cap; A.units = 0;
    }
    if (A.units != units) {
        if (capturing) return li;
        if (cudaMemsetAsync(A.d_n, 0, sizeof(unsigned), s) != cudaSuccess) { cudaGetLastError(); return li; }
        A.units = units;
    }
    li.grp_cost = A.d_cost; li.cost_sum = A.d_sum; li.heavy_n = A.d_n; li.heavy_list = A.d_list; li.heavy_bits = A.d_bits;
    li.heavy_cap = HF_AS_LIST_CAP; li.heavy_words = (unsigned) (A.cap / 32);
    li.heavy_shift = 0;
    while (((units - 1) >> li.heavy_shift) >= 8192) li.heavy_shift++;
    return li;
}
static inline HfLaunchInfo launch_info_for(hf_ctx *c, size_t n, cudaStream_t s) { return adaptive_launch_info(c, c->as_rays, (n + 31) / 32, s); }
static inline HfLaunchInfo launch_info_for_pixels(hf_ctx *c, size_t npix, cudaStream_t s) { return adaptive_launch_info(c, c->as_pixels, npix, s); }

// Ordering of the host-buffer entry points (hf_ray_intersect_host, hf_forward_backward_host, hf_render_step_l2_host) behind the calls
// that change what they read. Those entry points run on internal NON-BLOCKING streams, which wait for nothing by themselves; they
// order themselves behind the legacy default stream with an event. A caller that uploads heights, rebuilds the pyramid, clears the
// gradient grid or steps the optimiser on a non-blocking stream of its own would otherwise be overtaken. So each of those calls
// leaves an event behind itself on its stream and the host entries wait for it. One event stands for one stream: mutators on two
// different streams since the last host entry, or inside a stream capture (such an event cannot be waited for from outside), fall
// back to one device-wide synchronisation at the next host entry.
static void note_mutation(hf_ctx *c, cudaStream_t s) {
    if (s == nullptr || s == cudaStreamLegacy) return;            // the host entries wait for the legacy stream anyway
    if (s == c->s_in || s == c->s_k || s == c->s_out) return;     // a host entry's own streams: it returns only after they have drained
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess) { cudaGetLastError(); c->mut_sync = true; return; }
    if (st == cudaStreamCaptureStatusActive) { c->mut_sync = true; return; }
    if (!c->ev_mut && cudaEventCreateWithFlags(&c->ev_mut, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); c->ev_mut = nullptr; c->mut_sync = true; return; }
    if (c->mut_pending && c->mut_stream != s) c->mut_sync = true;
    if (cudaEventRecord(c->ev_mut, s) != cudaSuccess) { cudaGetLastError(); c->mut_sync = true; return; }
    c->mut_pending = true; c->mut_stream = s;
}
static int wait_mutations(hf_ctx *c, cudaStream_t a, cudaStream_t b, cudaStream_t d) {
    if (c->mut_sync) {
        CU(cudaDeviceSynchronize());
    } else if (c->mut_pending) {
        for (cudaStream_t s : { a, b, d }) if (s) CU(cudaStreamWaitEvent(s, c->ev_mut, 0));
    }
    c->mut_sync = false; c->mut_pending = false;
    return HF_OK;
}

static void layout_pyramid(hf_ctx *c) {
    HfParams &P = c->P;
    int L = 1;
    while (((P.NX + (1 << L) - 1) >> L) > 1 || ((P.NY + (1 << L) - 1) >> L) > 1) L++;
    P.Ltop = L;
    size_t off = 0;
    std::vector<size_t> cnt(HF_MAX_LEVELS, 0);
    for (int l = 0; l < HF_MAX_LEVELS; ++l) { P.lvl_off[l] = 0; P.lvl_w[l] = 0; }
    for (int l = 1; l <= L; ++l) {
        int w = (P.NX + (1 << l) - 1) >> l, h = (P.NY + (1 << l) - 1) >> l;
        P.lvl_off[l] = (int) off; P.lvl_w[l] = w;
        cnt[l] = ((size_t) w * h + 1) & ~size_t(1);     // even node count: every level starts 16-byte aligned
        off += cnt[l];
    }
    c->pyr_nodes = off;
    {   // quantised, padded pyramid read by the production walk
        const HfQLayout q = hf_qlayout(P.NX, P.NY);
        P.QLtop = q.Ltop; P.xmask = q.Mx - 1; P.ymask = q.My - 1;
        for (int l = 0; l < HF_MAX_LEVELS; ++l) P.qoff[l] = q.off[l];
        c->qpyr_nodes = q.nodes;
    }
    // levels >= Ls are staged in shared memory by the intersection kernel
    int Ls = L;
    size_t top = cnt[L];
    while (Ls > 1 && top + cnt[Ls - 1] <= HF_TOP_NODE_CAP) { Ls--; top += cnt[Ls]; }
    c->li.top_level_start = Ls;
    c->li.top_nodes = (int) top;
    c->li.top_off = P.lvl_off[Ls];
    c->li.num_sms = c->num_sms;
    {   // coarse even levels of the quantised pyramid that fit the shared-memory budget of a walk kernel, finest first
        const HfQLayout q = hf_qlayout(P.NX, P.NY);
        int lo = (q.Ltop & ~1) + 2;
        size_t bytes = 0;
        for (int l = q.Ltop & ~1; l >= 0; l -= 2) {
            const size_t b = (((size_t) q.pitch[l] * q.rows[l] + 3) & ~size_t(3)) * sizeof(uint32_t);
            if (bytes + b > HF_QSTAGE_CAP) break;
            bytes += b; lo = l;
        }
        const char *e = getenv("HF_B200_QSTAGE");
        if (!HF_STAGE_Q || (e && e[0] == '0')) { lo = q.Ltop + 2; bytes = 0; }
        c->li.q_stage_level = lo; c->li.q_stage_bytes = (int) bytes;
    }
    {   // HF_B200_BUNDLES=1 selects the bundled kernel (hf_intersect_beam_kernel). Measured on C3 (profiles/): it removes 40 % of
        // the walk's instructions but pays them back in the shuffle reductions that find the bundle bounds of a generic ray
        // array, so the plain kernel stays the default for ray-array entry points; the fused render step, which knows its
        // bundles analytically, always marches them.
        const char *e = getenv("HF_B200_BUNDLES");
        c->li.use_bundles = (e && e[0] == '1') ? 1 : 0;
        c->li.incoherent = 0;
    }
}

static void peer_release_fwd(hf_ctx *c);

extern "C" int hf_create(hf_ctx **out, const hf_desc *desc) {
    if (!out || !desc) return fail(HF_ERR_INVALID, "null argument");
    *out = nullptr;
    if (desc->width < 2 || desc->height < 2) return fail(HF_ERR_INVALID, "heightfield needs at least 2x2 texels (got %ux%u)", desc->width, desc->height);
    if (desc->width > (1u << 23) || desc->height > (1u << 23)) return fail(HF_ERR_INVALID, "heightfield side exceeds 2^23 texels");
    if ((uint64_t) (desc->width - 1) * (desc->height - 1) > 0x7fffffffull) return fail(HF_ERR_INVALID, "prim_index would overflow uint32");
    if (!isfinite(desc->max_height)) return fail(HF_ERR_INVALID, "max_height is not finite");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(HF_ERR_CUDA, "no CUDA device available (%s); this library has no CPU path", e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
    if (desc->device < 0 || desc->device >= ndev) return fail(HF_ERR_INVALID, "device ordinal %d out of range", desc->device);
    CU(cudaSetDevice(desc->device));

    hf_ctx *c = new (std::nothrow) hf_ctx();
    if (!c) return fail(HF_ERR_INVALID, "out of host memory");
    c->device = desc->device;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, desc->device);
    if (e != cudaSuccess) { delete c; return fail(HF_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e)); }
    c->num_sms = prop.multiProcessorCount;
    c->W = desc->width; c->H = desc->height;

    HfParams &P = c->P;
    memset(&P, 0, sizeof(P));
    P.W = (int) desc->width; P.H = (int) desc->height; P.NX = P.W - 1; P.NY = P.H - 1;
    P.flip_normals = desc->flip_normals ? 1 : 0;
    P.mh = desc->max_height;
    P.rw = 1.0f / (float) P.NX; P.rh = 1.0f / (float) P.NY;
    P.x0 = (float) (-(P.W - 1)) * P.rw; P.y0 = (float) (-(P.H - 1)) * P.rh;
    P.sgx = 0.5f * (float) P.NX; P.sgy = 0.5f * (float) P.NY;
    { const HfDivMagic dm = hf_divmagic((uint32_t) P.NX); P.nx_magic = dm.m; P.nx_shift = dm.s; }
    int rc = set_transform(c, desc->to_world, desc->has_to_object ? desc->to_object : nullptr);
    if (rc) { delete c; return rc; }
    layout_pyramid(c);

    const size_t texels = (size_t) c->W * c->H;
    cudaError_t e1 = cudaMalloc(&c->d_heights, texels * sizeof(float));
    cudaError_t e2 = cudaMalloc(&c->d_grad, texels * sizeof(float));
    cudaError_t e3 = cudaMalloc(&c->d_pyr, c->pyr_nodes * sizeof(float2));
    cudaError_t e4 = cudaMalloc(&c->d_counters, (8 + HF_SCHED_SLOTS + HF_SCHED_CAPTURE_SLOTS) * sizeof(unsigned long long));
    if (c->qpyr_nodes >= (size_t(1) << 32)) { hf_destroy(c); return fail(HF_ERR_INVALID, "%ux%u heightfield: padded acceleration pyramid exceeds 2^32 nodes", c->W, c->H); }
    cudaError_t e5 = cudaMalloc(&c->d_qpyr, c->qpyr_nodes * sizeof(uint32_t));
    if (e1 || e2 || e3 || e4 || e5) { hf_destroy(c); return fail(HF_ERR_CUDA, "cudaMalloc failed for a %ux%u heightfield", c->W, c->H); }
    cudaMemset(c->d_grad, 0, texels * sizeof(float));
    cudaMemset(c->d_pyr, 0, c->pyr_nodes * sizeof(float2));
    cudaMemset(c->d_counters, 0, (8 + HF_SCHED_SLOTS + HF_SCHED_CAPTURE_SLOTS) * sizeof(unsigned long long));
    P.heights = c->d_heights;
    P.pyr = c->d_pyr;
    P.qpyr = c->d_qpyr;
    c->li.seg_counter = c->d_counters + 4;
    *out = c;
    return HF_OK;
}

extern "C" int hf_destroy(hf_ctx *c) {
    if (!c) return HF_OK;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    peer_release_fwd(c);
    if (c->comm && c->grad_reg && g_nccl.CommDeregister) g_nccl.CommDeregister(c->comm, c->grad_reg);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    if (c->grad_nccl && g_nccl.MemFree) { g_nccl.MemFree(c->d_grad); c->d_grad = nullptr; }
    for (auto &s : c->slots) {
        if (s.slab) cudaFree(s.slab);
        if (s.in_done) cudaEventDestroy(s.in_done);
        if (s.k_done) cudaEventDestroy(s.k_done);
        if (s.out_done) cudaEventDestroy(s.out_done);
    }
    if (c->ev_entry) cudaEventDestroy(c->ev_entry);
    if (c->ev_accel) cudaEventDestroy(c->ev_accel);
    if (c->ev_mut) cudaEventDestroy(c->ev_mut);
    if (c->s_in) cudaStreamDestroy(c->s_in);
    if (c->s_k) cudaStreamDestroy(c->s_k);
    if (c->s_out) cudaStreamDestroy(c->s_out);
    c->as_rays.release(); c->as_pixels.release();
    cudaFree(c->d_heights); cudaFree(c->d_grad); cudaFree(c->d_m); cudaFree(c->d_v); cudaFree(c->d_pyr); cudaFree(c->d_qpyr); cudaFree(c->d_counters); cudaFree(c->d_scratch);
    cudaFree(c->d_adam_t); cudaFree(c->d_adam_lr_t);
    for (auto &e : c->adam_tables) cudaFree(e.d);
    cudaFree(c->d_hs_target); cudaFree(c->d_hs_image); cudaFree(c->d_hs_pixels); cudaFree(c->d_hs_loss);
    delete c;
    return HF_OK;
}

// ------------------------------------------------------------------------------------------
// parameters
// ------------------------------------------------------------------------------------------
extern "C" int hf_set_heights(hf_ctx *c, const float *src, int src_on_device, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (!src) return fail(HF_ERR_INVALID, "null heights");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(c->d_heights, src, (size_t) c->W * c->H * sizeof(float),
                       src_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, (cudaStream_t) stream));
    note_mutation(c, (cudaStream_t) stream);
    c->have_heights = true;
    c->accel_valid = false;
    return HF_OK;
}

extern "C" int hf_heights(hf_ctx *c, float **d) {
    CHECK_CTX(c);
    if (!d) return fail(HF_ERR_INVALID, "null out pointer");
    *d = c->d_heights;
    c->have_heights = true;       // the caller may fill the buffer in place, then call hf_parameters_changed
    return HF_OK;
}

extern "C" int hf_grad(hf_ctx *c, float **d) {
    CHECK_CTX(c);
    if (!d) return fail(HF_ERR_INVALID, "null out pointer");
    *d = c->d_grad;
    return HF_OK;
}

extern "C" int hf_zero_grad(hf_ctx *c, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    CU(cudaSetDevice(c->device));
    CU(cudaMemsetAsync(c->d_grad, 0, (size_t) c->W * c->H * sizeof(float), (cudaStream_t) stream));
    note_mutation(c, (cudaStream_t) stream);
    return HF_OK;
}

extern "C" int hf_set_to_world(hf_ctx *c, const float to_world[12], const float *to_object) {
    CHECK_CTX(c);
    if (!to_world) return fail(HF_ERR_INVALID, "null to_world");
    return set_transform(c, to_world, to_object);
}

extern "C" int hf_set_max_height(hf_ctx *c, float mh) {
    CHECK_CTX(c);
    if (!isfinite(mh)) return fail(HF_ERR_INVALID, "max_height is not finite");
    c->P.mh = mh;
    c->accel_valid = false;
    return HF_OK;
}

extern "C" int hf_parameters_changed(hf_ctx *c, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (!c->have_heights) return fail(HF_ERR_STATE, "heights were never set");
    CU(cudaSetDevice(c->device));
    CU(hf_launch_pyramid(c->P, c->d_pyr, c->d_qpyr, (cudaStream_t) stream));
    if (!c->ev_accel) CU(cudaEventCreateWithFlags(&c->ev_accel, cudaEventDisableTiming));
    {   // an event recorded inside a stream capture cannot be waited for from the host: hf_bbox then synchronises the device instead
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing((cudaStream_t) stream, &st) != cudaSuccess) { cudaGetLastError(); st = cudaStreamCaptureStatusNone; }
        c->accel_event_ok = (st != cudaStreamCaptureStatusActive);
        if (c->accel_event_ok) CU(cudaEventRecord(c->ev_accel, (cudaStream_t) stream));
    }
    note_mutation(c, (cudaStream_t) stream);
    c->accel_valid = true;
    return HF_OK;
}

extern "C" int hf_set_ray_coherence(hf_ctx *c, int coherent) {
    CHECK_CTX(c);
    c->li.incoherent = coherent ? 0 : 1;
    return HF_OK;
}

extern "C" int hf_set_adaptive_schedule(hf_ctx *c, int on) {
    CHECK_CTX(c);
    c->adaptive = on ? 1 : 0;
    c->as_rays.units = c->as_pixels.units = 0;       // whatever was learnt is forgotten: the next launches start new lists
    return HF_OK;
}

extern "C" int hf_adaptive_schedule_info(hf_ctx *c, int kind, uint32_t *listed, uint32_t *groups) {
    CHECK_CTX(c);
    if (kind != 0 && kind != 1) return fail(HF_ERR_INVALID, "hf_adaptive_schedule_info: kind 0 (ray batches) or 1 (pixel lists of the render step)");
    unsigned n = 0;
    const HfAdaptive &A = kind == 1 ? c->as_pixels : c->as_rays;
    if (c->adaptive && A.d_n && A.units) {
        CU(cudaSetDevice(c->device));
        CU(cudaDeviceSynchronize());
        CU(cudaMemcpy(&n, A.d_n, sizeof(n), cudaMemcpyDeviceToHost));
        if (n > HF_AS_LIST_CAP) n = HF_AS_LIST_CAP;
    }
    if (listed) *listed = n;
    if (groups) *groups = c->adaptive ? (uint32_t) A.units : 0u;
    return HF_OK;
}

extern "C" int hf_parameters_grad_enabled(hf_ctx *c, int *enabled) {
    CHECK_CTX(c);
    if (!enabled) return fail(HF_ERR_INVALID, "null out pointer");
    *enabled = c->grad_enabled;
    return HF_OK;
}

extern "C" int hf_set_grad_enabled(hf_ctx *c, int enabled) {
    CHECK_CTX(c);
    c->grad_enabled = enabled ? 1 : 0;
    return HF_OK;
}

// ------------------------------------------------------------------------------------------
// queries
// ------------------------------------------------------------------------------------------
#define NEED_ACCEL(c) do { if (!(c)->accel_valid) return fail(HF_ERR_STATE, "acceleration data is stale: call hf_parameters_changed after setting heights / max_height"); } while (0)

static inline void xform_point_host(const float m[12], const float p[3], float out[3]) {
    for (int r = 0; r < 3; ++r) {
        float v = m[r * 4 + 3];
        v = fmaf(m[r * 4 + 0], p[0], v); v = fmaf(m[r * 4 + 1], p[1], v); v = fmaf(m[r * 4 + 2], p[2], v);
        out[r] = v;
    }
}

extern "C" int hf_bbox(hf_ctx *c, float out[6]) {
    CHECK_CTX(c);
    if (!out) return fail(HF_ERR_INVALID, "null out pointer");
    NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    float2 top;
    if (c->accel_event_ok) CU(cudaEventSynchronize(c->ev_accel));      // the pyramid build this bbox derives from (whatever stream it ran on)
    else CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(&top, c->d_pyr + c->P.lvl_off[c->P.Ltop], sizeof(float2), cudaMemcpyDeviceToHost));
    const HfParams &P = c->P;
    float lo[3] = { P.x0, P.y0, top.x };
    float hi[3] = { (float) (P.W - 1) * P.rw, (float) (P.H - 1) * P.rh, top.y };
    for (int a = 0; a < 3; ++a) { out[a] = INFINITY; out[3 + a] = -INFINITY; }
    for (int k = 0; k < 8; ++k) {
        float q[3] = { (k & 1) ? hi[0] : lo[0], (k & 2) ? hi[1] : lo[1], (k & 4) ? hi[2] : lo[2] }, w[3];
        xform_point_host(P.to_world, q, w);
        for (int a = 0; a < 3; ++a) { out[a] = fminf(out[a], w[a]); out[3 + a] = fmaxf(out[3 + a], w[a]); }
    }
    return HF_OK;
}

extern "C" int hf_primitive_count(hf_ctx *c, uint32_t *count) {
    CHECK_CTX(c);
    if (!count) return fail(HF_ERR_INVALID, "null out pointer");
    *count = 2u * (c->W - 1u) * (c->H - 1u);
    return HF_OK;
}

extern "C" int hf_size(hf_ctx *c, uint32_t *w, uint32_t *h) {
    CHECK_CTX(c);
    if (w) *w = c->W;
    if (h) *h = c->H;
    return HF_OK;
}

extern "C" int hf_qpyramid_level(hf_ctx *c, uint32_t level, uint32_t *levels, uint32_t *pitch, uint32_t *rows, const uint32_t **d_nodes,
                                 float *qs, float *zb) {
    CHECK_CTX(c);
    const HfParams &P = c->P;
    if (levels) *levels = (uint32_t) P.QLtop;
    if (level > (uint32_t) P.QLtop) {
        if (pitch || rows || d_nodes) return fail(HF_ERR_INVALID, "quantised pyramid level %u out of range 0..%d", level, P.QLtop);
    } else {
        if (pitch) *pitch = (uint32_t) std::max((P.xmask + 1) >> level, 1);
        if (rows) *rows = (uint32_t) std::max((P.ymask + 1) >> level, 1);
        if (d_nodes) *d_nodes = c->d_qpyr + P.qoff[level];
    }
    if (qs || zb) {
        NEED_ACCEL(c);
        CU(cudaSetDevice(c->device));
        float2 top;
        if (c->accel_event_ok) CU(cudaEventSynchronize(c->ev_accel));
        else CU(cudaDeviceSynchronize());
        CU(cudaMemcpy(&top, c->d_pyr + P.lvl_off[P.Ltop], sizeof(float2), cudaMemcpyDeviceToHost));
        const HfQuant q = hf_quant_make(top.x, top.y);
        if (qs) *qs = q.qs;
        if (zb) *zb = q.zb;
    }
    return HF_OK;
}

extern "C" int hf_pyramid_level(hf_ctx *c, uint32_t level, uint32_t *levels, uint32_t *w, uint32_t *h, const float **d_nodes) {
    CHECK_CTX(c);
    const HfParams &P = c->P;
    if (levels) *levels = (uint32_t) P.Ltop;
    if (level < 1 || level > (uint32_t) P.Ltop) {
        if (w || h || d_nodes) return fail(HF_ERR_INVALID, "pyramid level %u out of range 1..%d", level, P.Ltop);
        return HF_OK;
    }
    if (w) *w = (uint32_t) P.lvl_w[level];
    if (h) *h = (uint32_t) ((P.NY + (1 << level) - 1) >> level);
    if (d_nodes) *d_nodes = reinterpret_cast<const float *>(c->d_pyr + P.lvl_off[level]);
    return HF_OK;
}

static int check_rays(const hf_rays *r) {
    if (!r) return fail(HF_ERR_INVALID, "null rays");
    if (!r->ox || !r->oy || !r->oz || !r->dx || !r->dy || !r->dz || !r->maxt) return fail(HF_ERR_INVALID, "ray batch has a null component");
    return HF_OK;
}
static int check_prelim(const hf_prelim *p) {
    if (!p || !p->t || !p->u || !p->v || !p->prim_index) return fail(HF_ERR_INVALID, "preliminary intersection has a null component");
    return HF_OK;
}

extern "C" int hf_ray_intersect_preliminary(hf_ctx *c, const hf_rays *rays, const hf_prelim *out, size_t n, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (n == 0) return HF_OK;      // empty batch: nothing to do, pointers may be null
    int rc;
    if ((rc = check_rays(rays)) || (rc = check_prelim(out))) return rc;
    NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    hf_si none; memset(&none, 0, sizeof(none));
    CU(hf_launch_intersect(c->P, launch_info_for(c, n, (cudaStream_t) stream), HF_MODE_PRELIM, *rays, *out, none, nullptr, 0, n, c->d_counters, SCHED(c, stream),
                           (cudaStream_t) stream));
    return HF_OK;
}

extern "C" int hf_ray_intersect_preliminary_counted(hf_ctx *c, const hf_rays *rays, const hf_prelim *out, size_t n, void *stream) {
    CHECK_CTX(c);
    if (n == 0) return HF_OK;      // empty batch: nothing to do, pointers may be null
    int rc;
    if ((rc = check_rays(rays)) || (rc = check_prelim(out))) return rc;
    NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    hf_si none; memset(&none, 0, sizeof(none));
    CU(hf_launch_intersect(c->P, c->li, HF_MODE_COUNTED, *rays, *out, none, nullptr, 0, n, c->d_counters, SCHED(c, stream), (cudaStream_t) stream));
    return HF_OK;
}

extern "C" int hf_stats_get(hf_ctx *c, hf_stats *out, int reset) {
    CHECK_CTX(c);
    if (!out) return fail(HF_ERR_INVALID, "null out pointer");
    CU(cudaSetDevice(c->device));
    unsigned long long h[5];
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(h, c->d_counters, sizeof(h), cudaMemcpyDeviceToHost));
    out->rays = h[0]; out->nodes = h[1]; out->leaves = h[2]; out->hits = h[3]; out->segments = h[4];
    if (reset) CU(cudaMemset(c->d_counters, 0, sizeof(h)));
    return HF_OK;
}

extern "C" int hf_ray_test(hf_ctx *c, const hf_rays *rays, uint8_t *d_hit, size_t n, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (n == 0) return HF_OK;      // empty batch: nothing to do, pointers may be null
    int rc;
    if ((rc = check_rays(rays))) return rc;
    if (!d_hit) return fail(HF_ERR_INVALID, "null hit buffer");
    NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    hf_si none; memset(&none, 0, sizeof(none));
    hf_prelim nop; memset(&nop, 0, sizeof(nop));
    CU(hf_launch_intersect(c->P, c->li, HF_MODE_TEST, *rays, nop, none, d_hit, 0, n, c->d_counters, SCHED(c, stream), (cudaStream_t) stream));
    return HF_OK;
}

static int check_flags(uint32_t flags) {
    // src/render/mesh.cpp:709-711
    if ((flags & HF_RAY_DETACHSHAPE) && (flags & HF_RAY_FOLLOWSHAPE))
        return fail(HF_ERR_INVALID, "Invalid combination of RayFlags: DetachShape | FollowShape");
    return HF_OK;
}

extern "C" int hf_compute_surface_interaction(hf_ctx *c, const hf_rays *rays, const hf_prelim *pi, uint32_t flags,
                                              const hf_si *out, size_t n, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (n == 0) return HF_OK;      // empty batch: nothing to do, pointers may be null
    int rc;
    if ((rc = check_rays(rays)) || (rc = check_prelim(pi)) || (rc = check_flags(flags))) return rc;
    if (!out) return fail(HF_ERR_INVALID, "null surface interaction");
    if (!c->have_heights) return fail(HF_ERR_STATE, "heights were never set");
    CU(cudaSetDevice(c->device));
    CU(hf_launch_surface_interaction(c->P, c->li, *rays, *pi, flags, *out, n, (cudaStream_t) stream));
    return HF_OK;
}

extern "C" int hf_ray_intersect(hf_ctx *c, const hf_rays *rays, uint32_t flags, const hf_prelim *prelim_out,
                                const hf_si *out, size_t n, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (n == 0) return HF_OK;      // empty batch: nothing to do, pointers may be null
    int rc;
    if ((rc = check_rays(rays)) || (rc = check_flags(flags))) return rc;
    if (!out) return fail(HF_ERR_INVALID, "null surface interaction");
    if (prelim_out && (rc = check_prelim(prelim_out))) return rc;
    NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    hf_prelim nop; memset(&nop, 0, sizeof(nop));
    CU(hf_launch_intersect(c->P, launch_info_for(c, n, (cudaStream_t) stream), HF_MODE_FUSED_SI, *rays, prelim_out ? *prelim_out : nop, *out, nullptr, flags, n,
                           c->d_counters, SCHED(c, stream), (cudaStream_t) stream));
    return HF_OK;
}

extern "C" int hf_backward(hf_ctx *c, const hf_rays *rays, const hf_prelim *pi, const hf_si_adj *adj, uint32_t flags,
                           size_t n, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (n == 0) return HF_OK;      // empty batch: nothing to do, pointers may be null
    int rc;
    if ((rc = check_rays(rays)) || (rc = check_prelim(pi)) || (rc = check_flags(flags))) return rc;
    if (!adj) return fail(HF_ERR_INVALID, "null adjoints");
    if (!c->have_heights) return fail(HF_ERR_STATE, "heights were never set");
    if (!c->grad_enabled) return HF_OK;        // parameters_grad_enabled() == false: nothing to accumulate
    CU(cudaSetDevice(c->device));
    CU(hf_launch_backward(c->P, c->li, *rays, *pi, *adj, flags, c->d_grad, n, (cudaStream_t) stream));
    return HF_OK;
}

// ------------------------------------------------------------------------------------------
// optimiser + collective
// ------------------------------------------------------------------------------------------
extern "C" int hf_adam_reset(hf_ctx *c, void *stream) {
    CHECK_CTX(c);
    CU(cudaSetDevice(c->device));
    const size_t bytes = (size_t) c->W * c->H * sizeof(float);
    if (!c->d_m) { CU(cudaMalloc(&c->d_m, bytes)); }
    if (!c->d_v) { CU(cudaMalloc(&c->d_v, bytes)); }
    if (!c->d_adam_t) { CU(cudaMalloc(&c->d_adam_t, sizeof(int))); CU(cudaMalloc(&c->d_adam_lr_t, sizeof(float))); }
    CU(cudaMemsetAsync(c->d_m, 0, bytes, (cudaStream_t) stream));
    CU(cudaMemsetAsync(c->d_v, 0, bytes, (cudaStream_t) stream));
    CU(cudaMemsetAsync(c->d_adam_t, 0, sizeof(int), (cudaStream_t) stream));
    return HF_OK;
}

// lr_scale(t) = sqrt(1 - beta2^t) / (1 - beta1^t) for t = 1 .. HF_ADAM_TABLE - 1, in double, rounded once (entry 0 unused)
static int ensure_adam_table(hf_ctx *c, double b1, double b2, void *stream) {
    if (c->d_adam_table && c->tab_b1 == b1 && c->tab_b2 == b2) return HF_OK;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing((cudaStream_t) stream, &st) == cudaSuccess && st == cudaStreamCaptureStatusActive)
        return fail(HF_ERR_STATE, "Adam: run one eager hf_adam_step with these betas before capturing it into a CUDA graph");
    for (const auto &e : c->adam_tables)
        if (e.b1 == b1 && e.b2 == b2) { c->d_adam_table = e.d; c->tab_b1 = b1; c->tab_b2 = b2; return HF_OK; }
    if (c->adam_tables.size() >= 64) return fail(HF_ERR_STATE, "Adam: more than 64 distinct (beta1, beta2) pairs on one context");
    std::vector<float> tab(HF_ADAM_TABLE);
    tab[0] = 1.f;
    for (int t = 1; t < HF_ADAM_TABLE; ++t) tab[t] = (float) (sqrt(1.0 - pow(b2, t)) / (1.0 - pow(b1, t)));
    float *d = nullptr;
    CU(cudaMalloc(&d, HF_ADAM_TABLE * sizeof(float)));
    cudaError_t ce = cudaMemcpy(d, tab.data(), HF_ADAM_TABLE * sizeof(float), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) { cudaFree(d); return fail(HF_ERR_CUDA, "Adam: table upload failed: %s", cudaGetErrorString(ce)); }
    c->adam_tables.push_back({ b1, b2, d });
    c->d_adam_table = d; c->tab_b1 = b1; c->tab_b2 = b2;
    return HF_OK;
}

extern "C" int hf_adam_step(hf_ctx *c, const hf_adam *cfg, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (!cfg) return fail(HF_ERR_INVALID, "null Adam configuration");
    // optimizers.py:243-244
    if (!(cfg->beta1 >= 0. && cfg->beta1 < 1. && cfg->beta2 >= 0. && cfg->beta2 < 1. && cfg->lr > 0. && cfg->epsilon > 0.))
        return fail(HF_ERR_INVALID, "Adam: need 0 <= beta < 1, lr > 0, epsilon > 0");
    if (!c->have_heights) return fail(HF_ERR_STATE, "heights were never set");
    if (!c->d_m || !c->d_v) { int rc = hf_adam_reset(c, stream); if (rc) return rc; }
    CU(cudaSetDevice(c->device));
    // optimizers.py:266-279. The hyper-parameters are Python floats (doubles) there: lr_scale and (1 - beta) are evaluated in double
    // and only then become fp32 literals of the traced expression; lr is an opaque fp32 (lr_v) multiplied by the fp32 lr_scale.
    const double b1 = cfg->beta1, b2 = cfg->beta2;
    { int rc = ensure_adam_table(c, b1, b2, stream); if (rc) return rc; }
    CU(hf_launch_adam(c->d_heights, c->d_grad, c->d_m, c->d_v, (size_t) c->W * c->H, c->d_adam_t, c->d_adam_lr_t, c->d_adam_table,
                      HF_ADAM_TABLE, (float) cfg->lr, (float) b1, (float) b2, (float) (1.0 - b1), (float) (1.0 - b2), (float) cfg->epsilon,
                      cfg->mask_updates, b1, b2, c->num_sms, (cudaStream_t) stream));
    note_mutation(c, (cudaStream_t) stream);
    c->accel_valid = false;       // heights moved: the caller runs hf_parameters_changed (SceneParameters.update)
    return HF_OK;
}

extern "C" int hf_comm_unique_id(uint8_t out_id[128]) {
    if (!out_id) return fail(HF_ERR_INVALID, "null id buffer");
    int rc = nccl_load();
    if (rc) return rc;
    hf_nccl_uid id;
    NC(g_nccl.GetUniqueId(&id));
    memcpy(out_id, id.internal, 128);
    return HF_OK;
}

extern "C" int hf_comm_init(hf_ctx *c, int rank, int nranks, const uint8_t id_bytes[128]) {
    CHECK_CTX(c);
    if (!id_bytes || nranks < 1 || rank < 0 || rank >= nranks) return fail(HF_ERR_INVALID, "bad communicator arguments");
    int rc = nccl_load();
    if (rc) return rc;
    CU(cudaSetDevice(c->device));
    // With nranks > 1 the gradient grid moves into NCCL's memory below and the cudaMalloc grid is freed. Once hf_peer_export has handed
    // out an IPC handle of that grid, other processes may have it mapped (P.grad_peer, the staging tables): it has to stay where it is.
    const char *env_mem = getenv("HF_B200_NCCL_MEM");
    const bool want_move = nranks > 1 && !(env_mem && env_mem[0] == '0') && !c->grad_nccl;
    if (want_move && (c->peer_n > 1 || c->d_xcounter))
        return fail(HF_ERR_STATE, "hf_comm_init after hf_peer_export: the gradient grid is shared with the peers and cannot move into NCCL memory; "
                                  "call hf_comm_init first (hf_peer_export moves the grid back), or set HF_B200_NCCL_MEM=0");
    if (c->comm) {
        if (c->grad_reg && g_nccl.CommDeregister) g_nccl.CommDeregister(c->comm, c->grad_reg);
        c->grad_reg = nullptr;
        g_nccl.CommDestroy(c->comm); c->comm = nullptr;
    }
    hf_nccl_uid id;
    memcpy(id.internal, id_bytes, 128);
    NC(g_nccl.CommInitRank(&c->comm, nranks, id, rank));
    c->nranks = nranks; c->rank = rank;
    // The gradient grid is the one buffer that crosses NVLink every step. Move it into NCCL-allocated memory and register it with the
    // communicator, so that on an NVSwitch box the allreduce can run in the switch on the user buffer (NVLS) instead of being staged
    // through NCCL's own buffers. The pointer hf_grad() returns CHANGES here. HF_B200_NCCL_MEM=0 keeps the cudaMalloc buffer.
    const char *e = getenv("HF_B200_NCCL_MEM");
    if (nranks > 1 && !(e && e[0] == '0') && g_nccl.MemAlloc && g_nccl.MemFree && g_nccl.CommRegister) {
        const size_t bytes = (size_t) c->W * c->H * sizeof(float);
        if (!c->grad_nccl) {
            void *nb = nullptr;
            if (g_nccl.MemAlloc(&nb, bytes) == 0 && nb) {
                cudaError_t ce = cudaDeviceSynchronize();           // (scatters into the old grid on other streams)
                if (ce == cudaSuccess) ce = cudaMemcpy(nb, c->d_grad, bytes, cudaMemcpyDeviceToDevice);
                if (ce != cudaSuccess) { g_nccl.MemFree(nb); return fail(HF_ERR_CUDA, "hf_comm_init: moving the gradient grid into NCCL memory failed: %s", cudaGetErrorString(ce)); }
                cudaFree(c->d_grad);
                c->d_grad = (float *) nb; c->grad_nccl = true;
            } else cudaGetLastError();
        }
        if (c->grad_nccl && g_nccl.CommRegister(c->comm, c->d_grad, bytes, &c->grad_reg) != 0) c->grad_reg = nullptr;
    }
    return HF_OK;
}

extern "C" int hf_comm_destroy(hf_ctx *c) {
    CHECK_CTX(c);
    if (c->comm) {
        // the registration handle belongs to the communicator: drop it first (the buffer itself stays in ncclMemAlloc memory and is
        // registered again by the next hf_comm_init)
        if (c->grad_reg && g_nccl.CommDeregister) g_nccl.CommDeregister(c->comm, c->grad_reg);
        c->grad_reg = nullptr;
        NC(g_nccl.CommDestroy(c->comm)); c->comm = nullptr;
    }
    c->nranks = 1; c->rank = 0;
    return HF_OK;
}

extern "C" int hf_allreduce_grad(hf_ctx *c, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    if (!c->comm) return fail(HF_ERR_STATE, "hf_comm_init was not called");
    CU(cudaSetDevice(c->device));
    // ncclFloat32 = 7, ncclSum = 0
    NC(g_nccl.AllReduce(c->d_grad, c->d_grad, (size_t) c->W * c->H, 7, 0, c->comm, (cudaStream_t) stream));
    return HF_OK;
}

// ------------------------------------------------------------------------------------------
// Multi-GPU over peer memory: the gradient's reduce-scatter happens INSIDE the scatter kernels (every red.global.add.f32 goes to the
// rank that owns the texel, over NVLink), the all-gather is a store kernel (gradient) or is fused into the sharded Adam (heights).
// One process per GPU: the grids are shared with CUDA IPC handles the caller ships between the ranks (like the NCCL unique id).
// ------------------------------------------------------------------------------------------
// HF_PEER_HANDLE_BYTES (include/hf_b200.h): four cudaIpcMemHandle_t -- gradient grid, heights, barrier counter, staging area
static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");

static void peer_release(hf_ctx *c);
static void peer_release_fwd(hf_ctx *c) {
    peer_release(c);
    cudaFree(c->d_peer_grad_tab); cudaFree(c->d_peer_heights_tab); cudaFree(c->d_peer_counter_tab); cudaFree(c->d_xcounter);
    cudaFree(c->d_peer_stage_tab); cudaFree(c->d_stage);
    c->d_peer_grad_tab = nullptr; c->d_peer_heights_tab = nullptr; c->d_peer_counter_tab = nullptr; c->d_xcounter = nullptr;
    c->d_peer_stage_tab = nullptr; c->d_stage = nullptr;
}
static void peer_release(hf_ctx *c) {
    for (int r = 0; r < HF_MAX_PEERS; ++r) {
        if (r != c->peer_rank || c->peer_n <= 1) {
            if (c->peer_grad[r] && c->peer_grad[r] != c->d_grad) cudaIpcCloseMemHandle(c->peer_grad[r]);
            if (c->peer_heights[r] && c->peer_heights[r] != c->d_heights) cudaIpcCloseMemHandle(c->peer_heights[r]);
            if (c->peer_counter[r] && c->peer_counter[r] != c->d_xcounter) cudaIpcCloseMemHandle(c->peer_counter[r]);
            if (c->peer_stage[r] && c->peer_stage[r] != c->d_stage) cudaIpcCloseMemHandle(c->peer_stage[r]);
        }
        c->peer_grad[r] = nullptr; c->peer_heights[r] = nullptr; c->peer_counter[r] = nullptr; c->peer_stage[r] = nullptr;
    }
    c->peer_n = 1; c->peer_rank = 0; c->slab_elems = 0;
    c->P.peer_n = 1;
    for (int r = 0; r < HF_MAX_PEERS; ++r) c->P.grad_peer[r] = nullptr;
}

extern "C" int hf_peer_export(hf_ctx *c, uint8_t out[HF_PEER_HANDLE_BYTES]) {
    CHECK_CTX(c);
    if (!out) return fail(HF_ERR_INVALID, "null handle buffer");
    CU(cudaSetDevice(c->device));
    if (c->grad_nccl) {
        // CUDA IPC exports cudaMalloc memory only: bring the gradient grid back from ncclMemAlloc memory (hf_grad() changes again)
        // Everything that still uses the old grid (on any stream: non-blocking streams do not wait for the legacy stream) must be done
        // before the copy, and the copy before the free: ncclMemFree unmaps at once, unlike cudaFree (found at C4: a 1 GiB grid).
        const size_t bytes = (size_t) c->W * c->H * sizeof(float);
        float *nb = nullptr;
        CU(cudaDeviceSynchronize());
        CU(cudaMalloc(&nb, bytes));
        CU(cudaMemcpy(nb, c->d_grad, bytes, cudaMemcpyDeviceToDevice));
        CU(cudaDeviceSynchronize());               // (a device-to-device cudaMemcpy returns before the copy has finished)
        if (c->comm && c->grad_reg && g_nccl.CommDeregister) g_nccl.CommDeregister(c->comm, c->grad_reg);
        c->grad_reg = nullptr;
        if (g_nccl.MemFree) g_nccl.MemFree(c->d_grad);
        c->d_grad = nb; c->grad_nccl = false;
    }
    if (!c->d_xcounter) {
        CU(cudaMalloc(&c->d_xcounter, size_t(2) << 20));            // an allocation of its own (IPC maps whole allocations)
        CU(cudaMemset(c->d_xcounter, 0, size_t(2) << 20));
    } else {
        // A new group starts from a clean barrier: counter [0] and epoch [64] count in units of the OLD group's size (the counter
        // stands at epoch * old_nranks), so a group of another size would either walk through its first barrier or never leave it.
        // Every rank exports before any rank attaches, so the exchange of the handles is the host-side barrier between the last
        // kernel of the old group (finished: the device is synchronised first) and the first of the new one.
        CU(cudaDeviceSynchronize());
        CU(cudaMemset(c->d_xcounter, 0, 65 * sizeof(unsigned)));
    }
    {                                                               // one slab per sender, whatever the number of ranks will be
        const size_t bytes = std::max(((size_t) c->W * c->H + (size_t) HF_MAX_PEERS * 2048) * sizeof(float), size_t(2) << 20);
        if (!c->d_stage) CU(cudaMalloc(&c->d_stage, bytes));
        // all-zero before anybody can reach it: the senders skip the zero lines of their partial sums, and the owner clears what it has
        // read, so the area is all-zero again after every reduction
        CU(cudaMemset(c->d_stage, 0, bytes));
        CU(cudaDeviceSynchronize());                                // ... and finished before the handle leaves this process
    }
    cudaIpcMemHandle_t h[4];
    CU(cudaIpcGetMemHandle(&h[0], c->d_grad));
    CU(cudaIpcGetMemHandle(&h[1], c->d_heights));
    CU(cudaIpcGetMemHandle(&h[2], c->d_xcounter));
    CU(cudaIpcGetMemHandle(&h[3], c->d_stage));
    memcpy(out, h, sizeof(h));
    return HF_OK;
}

extern "C" int hf_peer_attach(hf_ctx *c, int rank, int nranks, const uint8_t *all_handles) {
    CHECK_CTX(c);
    if (!all_handles || nranks < 1 || nranks > HF_MAX_PEERS || rank < 0 || rank >= nranks)
        return fail(HF_ERR_INVALID, "hf_peer_attach: need 1 <= nranks <= %d, 0 <= rank < nranks and the handles of every rank", HF_MAX_PEERS);
    if (!c->d_xcounter) return fail(HF_ERR_STATE, "hf_peer_attach: call hf_peer_export on this context first");
    CU(cudaSetDevice(c->device));
    CU(cudaDeviceSynchronize());
    peer_release(c);
    c->peer_rank = rank;
    for (int r = 0; r < nranks; ++r) {
        if (r == rank) { c->peer_grad[r] = c->d_grad; c->peer_heights[r] = c->d_heights; c->peer_counter[r] = c->d_xcounter; c->peer_stage[r] = c->d_stage; continue; }
        cudaIpcMemHandle_t h[4];
        memcpy(h, all_handles + (size_t) r * HF_PEER_HANDLE_BYTES, sizeof(h));
        void *pg = nullptr, *ph = nullptr, *pc = nullptr, *ps = nullptr;
        cudaError_t e1 = cudaIpcOpenMemHandle(&pg, h[0], cudaIpcMemLazyEnablePeerAccess);
        cudaError_t e2 = e1 == cudaSuccess ? cudaIpcOpenMemHandle(&ph, h[1], cudaIpcMemLazyEnablePeerAccess) : e1;
        cudaError_t e3 = e2 == cudaSuccess ? cudaIpcOpenMemHandle(&pc, h[2], cudaIpcMemLazyEnablePeerAccess) : e2;
        if (e3 == cudaSuccess) e3 = cudaIpcOpenMemHandle(&ps, h[3], cudaIpcMemLazyEnablePeerAccess);
        if (e3 != cudaSuccess) {
            cudaGetLastError();
            if (pg) cudaIpcCloseMemHandle(pg);
            if (ph) cudaIpcCloseMemHandle(ph);
            if (pc) cudaIpcCloseMemHandle(pc);
            c->peer_n = r;                         // release what was opened so far
            peer_release(c);
            return fail(HF_ERR_CUDA, "hf_peer_attach: cannot map rank %d's grids over CUDA IPC (%s); is peer access between the GPUs available?", r, cudaGetErrorString(e3));
        }
        c->peer_grad[r] = (float *) pg; c->peer_heights[r] = (float *) ph; c->peer_counter[r] = (unsigned *) pc; c->peer_stage[r] = (float *) ps;
    }
    c->peer_n = nranks;
    const size_t total = (size_t) c->W * c->H;
    c->slab_elems = (((total + nranks - 1) / nranks) + 1023) & ~size_t(1023);
    if (!c->d_peer_grad_tab) {
        CU(cudaMalloc(&c->d_peer_grad_tab, HF_MAX_PEERS * sizeof(float *)));
        CU(cudaMalloc(&c->d_peer_heights_tab, HF_MAX_PEERS * sizeof(float *)));
        CU(cudaMalloc(&c->d_peer_counter_tab, HF_MAX_PEERS * sizeof(unsigned *)));
        CU(cudaMalloc(&c->d_peer_stage_tab, HF_MAX_PEERS * sizeof(float *)));
    }
    CU(cudaMemcpy(c->d_peer_stage_tab, c->peer_stage, HF_MAX_PEERS * sizeof(float *), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_peer_grad_tab, c->peer_grad, HF_MAX_PEERS * sizeof(float *), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_peer_heights_tab, c->peer_heights, HF_MAX_PEERS * sizeof(float *), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_peer_counter_tab, c->peer_counter, HF_MAX_PEERS * sizeof(unsigned *), cudaMemcpyHostToDevice));
    const HfDivMagic dm = hf_divmagic((uint32_t) c->slab_elems);
    c->P.slab_magic = dm.m; c->P.slab_shift = dm.s;
    for (int r = 0; r < HF_MAX_PEERS; ++r) c->P.grad_peer[r] = r < nranks ? c->peer_grad[r] : nullptr;
    c->P.peer_n = c->peer_mode == 0 ? nranks : 1;         // mode 1: the kernels scatter into the local grid
    return HF_OK;
}

extern "C" int hf_peer_set_mode(hf_ctx *c, int mode) {
    CHECK_CTX(c);
    if (mode != 0 && mode != 1) return fail(HF_ERR_INVALID, "hf_peer_set_mode: 0 (remote atomics to the owner) or 1 (local scatter + reduce-scatter by peer stores)");
    c->peer_mode = mode;
    if (c->peer_n > 1) c->P.peer_n = mode == 0 ? c->peer_n : 1;
    return HF_OK;
}

extern "C" int hf_peer_detach(hf_ctx *c) {
    CHECK_CTX(c);
    CU(cudaSetDevice(c->device));
    CU(cudaDeviceSynchronize());
    peer_release(c);
    return HF_OK;
}

#define NEED_PEERS(c) do { if ((c)->peer_n <= 1) return fail(HF_ERR_STATE, "hf_peer_attach was not called (or with a single rank)"); } while (0)
static inline void peer_slab(const hf_ctx *c, size_t &lo, size_t &hi) {
    const size_t total = (size_t) c->W * c->H;
    lo = std::min(total, (size_t) c->peer_rank * c->slab_elems);
    hi = std::min(total, lo + c->slab_elems);
}

extern "C" int hf_peer_barrier(hf_ctx *c, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    NEED_PEERS(c);
    CU(cudaSetDevice(c->device));
    CU(hf_launch_peer_barrier(c->d_peer_counter_tab, c->d_xcounter, c->d_xcounter + 64, c->peer_n, (cudaStream_t) stream));
    return HF_OK;
}
static int peer_barrier(hf_ctx *c, cudaStream_t st);

static int peer_barrier(hf_ctx *c, cudaStream_t st) {
    CU(hf_launch_peer_barrier(c->d_peer_counter_tab, c->d_xcounter, c->d_xcounter + 64, c->peer_n, st));
    return HF_OK;
}

extern "C" int hf_peer_zero_grad(hf_ctx *c, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    NEED_PEERS(c);
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t) stream;
    if (c->peer_mode == 1) {              // every rank scatters into its own (whole) grid
        CU(cudaMemsetAsync(c->d_grad, 0, (size_t) c->W * c->H * sizeof(float), st));
        return HF_OK;
    }
    size_t lo, hi;                        // remote atomics: the owner clears its slab, and nobody may scatter before every slab is clear
    peer_slab(c, lo, hi);
    if (hi > lo) CU(cudaMemsetAsync(c->d_grad + lo, 0, (hi - lo) * sizeof(float), st));
    return peer_barrier(c, st);
}

static int peer_reduce(hf_ctx *c, const hf_adam *cfg, cudaStream_t st) {
    size_t lo, hi;
    peer_slab(c, lo, hi);
    const size_t total = (size_t) c->W * c->H;
    int rc;
    const bool adam = cfg != nullptr;
    const double b1 = adam ? cfg->beta1 : 0.0, b2 = adam ? cfg->beta2 : 0.0;
    if (c->peer_mode == 1) {
        // A: partial slabs to their owners' staging areas (stores at link rate); barrier; B: the owner sums the partials in rank order
        // and stores the sum (or, with Adam, the updated heights) into every rank's grid -- reduction and all-gather in one pass
        CU(hf_launch_peer_send_partials(c->d_grad, c->d_peer_stage_tab, c->peer_n, c->peer_rank, c->slab_elems, total, c->num_sms, st));
        if ((rc = peer_barrier(c, st))) return rc;
        CU(hf_launch_peer_reduce_gather(c->d_grad, c->d_stage, adam ? c->d_peer_heights_tab : c->d_peer_grad_tab, c->peer_n, c->peer_rank, lo, hi, c->slab_elems,
                                        adam, c->d_heights, c->d_m, c->d_v, c->d_adam_t, c->d_adam_lr_t, c->d_adam_table, HF_ADAM_TABLE,
                                        adam ? (float) cfg->lr : 0.f, (float) b1, (float) b2, (float) (1.0 - b1), (float) (1.0 - b2),
                                        adam ? (float) cfg->epsilon : 0.f, adam ? cfg->mask_updates : 0, b1, b2, c->num_sms, st));
    } else {
        // the scatter kernels already summed every contribution into the owner's slab: wait for all of them, then all-gather by stores
        if ((rc = peer_barrier(c, st))) return rc;
        if (adam)
            CU(hf_launch_peer_adam(c->d_heights, c->d_grad, c->d_m, c->d_v, c->d_peer_heights_tab, c->peer_n, c->peer_rank, lo, hi, c->d_adam_t, c->d_adam_lr_t,
                                   c->d_adam_table, HF_ADAM_TABLE, (float) cfg->lr, (float) b1, (float) b2, (float) (1.0 - b1), (float) (1.0 - b2),
                                   (float) cfg->epsilon, cfg->mask_updates, b1, b2, c->num_sms, st));
        else
            CU(hf_launch_peer_push(c->d_grad, c->d_peer_grad_tab, c->peer_n, c->peer_rank, lo, hi, c->num_sms, st));
    }
    return peer_barrier(c, st);           // every rank's stores have landed: the grids (heights) are complete and identical everywhere
}

extern "C" int hf_peer_reduce_grad(hf_ctx *c, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    NEED_PEERS(c);
    CU(cudaSetDevice(c->device));
    return peer_reduce(c, nullptr, (cudaStream_t) stream);
}

extern "C" int hf_peer_adam_step(hf_ctx *c, const hf_adam *cfg, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    NEED_PEERS(c);
    if (!cfg) return fail(HF_ERR_INVALID, "null Adam configuration");
    if (!(cfg->beta1 >= 0. && cfg->beta1 < 1. && cfg->beta2 >= 0. && cfg->beta2 < 1. && cfg->lr > 0. && cfg->epsilon > 0.))
        return fail(HF_ERR_INVALID, "Adam: need 0 <= beta < 1, lr > 0, epsilon > 0");
    if (!c->have_heights) return fail(HF_ERR_STATE, "heights were never set");
    if (!c->d_m || !c->d_v) { int rc = hf_adam_reset(c, stream); if (rc) return rc; }
    CU(cudaSetDevice(c->device));
    { int rc = ensure_adam_table(c, cfg->beta1, cfg->beta2, stream); if (rc) return rc; }
    int rc = peer_reduce(c, cfg, (cudaStream_t) stream);
    c->accel_valid = false;       // heights moved on every rank: hf_parameters_changed next
    return rc;
}

// ------------------------------------------------------------------------------------------
// render step (sensor + direct illumination + film + adjoint on the heightfield alone)
// ------------------------------------------------------------------------------------------
static int check_camera(const hf_camera *cam, bool needs_stash) {
    if (!cam) return fail(HF_ERR_INVALID, "null camera");
    if (cam->width == 0 || cam->height == 0 || cam->spp == 0) return fail(HF_ERR_INVALID, "camera: width, height and spp must be positive");
    if ((uint64_t) cam->width * cam->height > 0xffffffffull) return fail(HF_ERR_INVALID, "camera: pixel ids would overflow uint32");
    if (!(cam->tan_half_fov > 0.f) || !isfinite(cam->tan_half_fov)) return fail(HF_ERR_INVALID, "camera: tan_half_fov must be positive and finite");
    if (needs_stash && cam->spp > HF_RENDER_MAX_SPP_HOST)
        return fail(HF_ERR_INVALID, "camera: spp = %u exceeds %d for the adjoint passes (split the samples over several seeds)", cam->spp, HF_RENDER_MAX_SPP_HOST);
    // the bundle march relies on an orthonormal frame (|fwd + sx right + sy up| >= 1)
    const float *r = cam->right, *up = cam->up, *f = cam->forward;
    auto dot = [](const float *a, const float *b) { return (double) a[0] * b[0] + (double) a[1] * b[1] + (double) a[2] * b[2]; };
    for (int k = 0; k < 3; ++k) if (!isfinite(cam->origin[k]) || !isfinite(r[k]) || !isfinite(up[k]) || !isfinite(f[k])) return fail(HF_ERR_INVALID, "camera: non-finite entry");
    if (fabs(dot(r, r) - 1) > 1e-4 || fabs(dot(up, up) - 1) > 1e-4 || fabs(dot(f, f) - 1) > 1e-4 || fabs(dot(r, up)) > 1e-4 ||
        fabs(dot(r, f)) > 1e-4 || fabs(dot(up, f)) > 1e-4)
        return fail(HF_ERR_INVALID, "camera: right / up / forward must be orthonormal");
    return HF_OK;
}
static int check_light(const hf_light *l) {
    if (!l) return fail(HF_ERR_INVALID, "null light");
    const double n2 = (double) l->direction[0] * l->direction[0] + (double) l->direction[1] * l->direction[1] + (double) l->direction[2] * l->direction[2];
    if (!(fabs(n2 - 1) < 1e-4)) return fail(HF_ERR_INVALID, "light: direction must be a unit vector");
    if (!isfinite(l->irradiance) || !isfinite(l->albedo)) return fail(HF_ERR_INVALID, "light: non-finite irradiance / albedo");
    if (l->max_depth < 0 || l->max_depth > HF_MAX_DEPTH_HOST) return fail(HF_ERR_INVALID, "light: max_depth must be in 0..%d", HF_MAX_DEPTH_HOST);
    return HF_OK;
}
static int check_pixels(const hf_camera *cam, const uint32_t *d_pixels, size_t npix) {
    const size_t all = (size_t) cam->width * cam->height;
    if (!d_pixels && npix != all) return fail(HF_ERR_INVALID, "npix must be width * height when no pixel list is given");
    if (npix > all) return fail(HF_ERR_INVALID, "more listed pixels than the image has");
    return HF_OK;
}

extern "C" int hf_render_rays(hf_ctx *c, const hf_camera *cam, const uint32_t *d_pixels, size_t npix, const hf_rays *d_out, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    int rc;
    if ((rc = check_camera(cam, false)) || (rc = check_pixels(cam, d_pixels, npix))) return rc;
    if (npix == 0) return HF_OK;
    if ((rc = check_rays(d_out))) return rc;
    CU(cudaSetDevice(c->device));
    CU(hf_launch_render_rays(c->P, c->li, *cam, d_pixels, npix, *d_out, (cudaStream_t) stream));
    return HF_OK;
}

extern "C" int hf_render_forward(hf_ctx *c, const hf_camera *cam, const hf_light *light, const uint32_t *d_pixels, size_t npix,
                                 float *d_image, uint32_t *d_prim, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    int rc;
    if ((rc = check_camera(cam, false)) || (rc = check_light(light)) || (rc = check_pixels(cam, d_pixels, npix))) return rc;
    if (npix == 0) return HF_OK;
    if (!d_image) return fail(HF_ERR_INVALID, "null image");
    if (d_prim && light->max_depth > 1) return fail(HF_ERR_INVALID, "hit triangles are only stored for max_depth <= 1 (paths are replayed)");
    NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    CU(hf_launch_render(c->P, launch_info_for_pixels(c, npix, (cudaStream_t) stream), 0, *cam, *light, d_pixels, npix, d_image, d_prim, nullptr, nullptr, nullptr, nullptr,
                        SCHED(c, stream), (cudaStream_t) stream));
    return HF_OK;
}

extern "C" int hf_render_backward(hf_ctx *c, const hf_camera *cam, const hf_light *light, const uint32_t *d_pixels, size_t npix,
                                  const float *d_adj_image, const uint32_t *d_prim, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    int rc;
    if ((rc = check_camera(cam, d_prim == nullptr)) || (rc = check_light(light)) || (rc = check_pixels(cam, d_pixels, npix))) return rc;
    if (npix == 0) return HF_OK;
    if (!d_adj_image) return fail(HF_ERR_INVALID, "null adjoint image");
    if (!c->grad_enabled) return HF_OK;
    if (d_prim && light->max_depth > 1) return fail(HF_ERR_INVALID, "hit triangles are only stored for max_depth <= 1 (paths are replayed)");
    CU(cudaSetDevice(c->device));
    if (d_prim) {
        if (!c->have_heights) return fail(HF_ERR_STATE, "heights were never set");
        CU(hf_launch_render_backward_prim(c->P, c->li, *cam, *light, d_pixels, npix, d_adj_image, d_prim, c->d_grad, (cudaStream_t) stream));
    } else {
        NEED_ACCEL(c);
        CU(hf_launch_render(c->P, launch_info_for_pixels(c, npix, (cudaStream_t) stream), 2, *cam, *light, d_pixels, npix, nullptr, nullptr, nullptr, d_adj_image, c->d_grad, nullptr,
                            SCHED(c, stream), (cudaStream_t) stream));
    }
    return HF_OK;
}

extern "C" int hf_render_step_l2(hf_ctx *c, const hf_camera *cam, const hf_light *light, const uint32_t *d_pixels, size_t npix,
                                 const float *d_target, float *d_image, double *d_loss, void *stream) {
    HF_RANGE();
    CHECK_CTX(c);
    int rc;
    if ((rc = check_camera(cam, true)) || (rc = check_light(light)) || (rc = check_pixels(cam, d_pixels, npix))) return rc;
    if (npix == 0) return HF_OK;
    if (!d_target) return fail(HF_ERR_INVALID, "null target image");
    NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    if (light->max_depth > 1) {
        // paths: forward image -> L2 adjoint image -> replayed adjoint (three launches; the path vertices do not fit the
        // shared-memory stash of the one-kernel step)
        const size_t img_bytes = (size_t) cam->width * cam->height * sizeof(float);
        if (c->scratch_bytes < 2 * img_bytes) {
            if (c->d_scratch) CU(cudaFree(c->d_scratch));
            c->d_scratch = nullptr; c->scratch_bytes = 0;
            CU(cudaMalloc(&c->d_scratch, 2 * img_bytes));
            c->scratch_bytes = 2 * img_bytes;
        }
        float *img = d_image ? d_image : c->d_scratch, *adj = c->d_scratch + (size_t) cam->width * cam->height;
        cudaStream_t st = (cudaStream_t) stream;
        CU(hf_launch_render(c->P, c->li, 0, *cam, *light, d_pixels, npix, img, nullptr, nullptr, nullptr, nullptr, nullptr, SCHED(c, st), st));
        CU(hf_launch_l2_adjoint(c->P, c->li, *cam, d_pixels, npix, img, d_target, adj, d_loss, st));
        if (c->grad_enabled)
            CU(hf_launch_render(c->P, c->li, 2, *cam, *light, d_pixels, npix, nullptr, nullptr, nullptr, adj, c->d_grad, nullptr, SCHED(c, st), st));
        return HF_OK;
    }
    CU(hf_launch_render(c->P, launch_info_for_pixels(c, npix, (cudaStream_t) stream), 1, *cam, *light, d_pixels, npix, d_image, nullptr, d_target, nullptr, c->d_grad, d_loss,
                        SCHED(c, stream), (cudaStream_t) stream));
    return HF_OK;
}

// ------------------------------------------------------------------------------------------
// host-buffer entry points: chunked, double-buffered H2D -> kernels -> D2H
// ------------------------------------------------------------------------------------------
extern "C" int hf_host_alloc(void **out, size_t bytes) {
    if (!out) return fail(HF_ERR_INVALID, "null out pointer");
    CU(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return HF_OK;
}
extern "C" int hf_host_free(void *p) {
    if (p) CU(cudaFreeHost(p));
    return HF_OK;
}

static int ensure_slots(hf_ctx *c) {
    if (c->slots_ready) return HF_OK;
    CU(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->s_k, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&c->ev_entry, cudaEventDisableTiming));
    for (auto &s : c->slots) {
        CU(cudaMalloc(&s.slab, HF_HOST_CHUNK * HF_SLAB_FLOATS * sizeof(float)));
        CU(cudaEventCreateWithFlags(&s.in_done, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s.k_done, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s.out_done, cudaEventDisableTiming));
    }
    c->slots_ready = true;
    return HF_OK;
}

// Chunks of HF_HOST_CHUNK rays flow through HF_HOST_SLOTS device slabs on THREE streams, one per engine: all host->device copies
// on s_in, all kernels on s_k, all device->host copies on s_out, chained per slab by events (in_done -> kernels -> k_done -> D2H ->
// out_done -> the slab's next H2D). Both copy engines and the SMs are busy at the same time, and the step is bound by the larger
// of the two PCIe directions (the reference's host-array caller: src/render/scene_embree.inl:261-326).
static int host_pipeline(hf_ctx *c, const hf_rays *hr, const hf_si_adj *hadj, uint32_t flags, const hf_prelim *hpre,
                         const hf_si *hsi, float *h_grad_out, size_t n, bool backward) {
    int rc;
    if (n == 0 && !h_grad_out) return HF_OK;
    if (n > 0 && ((rc = check_rays(hr)) || (rc = check_flags(flags)))) return rc;
    if (n > 0 && hpre && (rc = check_prelim(hpre))) return rc;
    if (n > 0 && backward && !hadj) return fail(HF_ERR_INVALID, "null adjoints");
    if (n > 0) NEED_ACCEL(c);
    CU(cudaSetDevice(c->device));
    if ((rc = ensure_slots(c))) return rc;
    // order after whatever the caller queued on the legacy default stream (heights upload, pyramid build, zero_grad): an event
    // there, waited for by the pipeline's streams -- no device-wide synchronisation
    CU(cudaEventRecord(c->ev_entry, (cudaStream_t) 0));
    CU(cudaStreamWaitEvent(c->s_in, c->ev_entry, 0));
    CU(cudaStreamWaitEvent(c->s_k, c->ev_entry, 0));
    CU(cudaStreamWaitEvent(c->s_out, c->ev_entry, 0));
    if ((rc = wait_mutations(c, c->s_in, c->s_k, c->s_out))) return rc;

    if (n > 0) {
        float *const *hsi_f = reinterpret_cast<float *const *>(hsi);
        const float *const *hadj_f = reinterpret_cast<const float *const *>(hadj);
        const float *hray_f[RAY_NFIELDS] = { hr->ox, hr->oy, hr->oz, hr->dx, hr->dy, hr->dz, hr->maxt };
        float *const *hpre_f = reinterpret_cast<float *const *>(hpre);

        size_t chunk_idx = 0;
        for (size_t base = 0; base < n; base += HF_HOST_CHUNK, ++chunk_idx) {
            const size_t m = std::min(HF_HOST_CHUNK, n - base);
            HostSlot &s = c->slots[chunk_idx % HF_HOST_SLOTS];
            float *slab = s.slab;
            auto field = [&](int k) { return slab + (size_t) k * HF_HOST_CHUNK; };

            // ---- host -> device (waits until the slab's previous results have left)
            if (chunk_idx >= HF_HOST_SLOTS) CU(cudaStreamWaitEvent(c->s_in, s.out_done, 0));
            hf_rays dr; memset(&dr, 0, sizeof(dr));
            const float **drf = reinterpret_cast<const float **>(&dr);
            for (int k = 0; k < RAY_NFIELDS; ++k) {
                CU(cudaMemcpyAsync(field(SLAB_RAY + k), hray_f[k] + base, m * sizeof(float), cudaMemcpyHostToDevice, c->s_in));
                drf[k] = field(SLAB_RAY + k);
            }
            if (hr->active) {
                uint8_t *da = reinterpret_cast<uint8_t *>(field(SLAB_ACTIVE));
                CU(cudaMemcpyAsync(da, hr->active + base, m, cudaMemcpyHostToDevice, c->s_in));
                dr.active = da;
            }
            hf_si_adj dadj; memset(&dadj, 0, sizeof(dadj));
            const float **dadjf = reinterpret_cast<const float **>(&dadj);
            if (backward) {
                for (int k = 0; k < ADJ_NFIELDS; ++k) if (hadj_f[k]) {
                    CU(cudaMemcpyAsync(field(SLAB_ADJ + k), hadj_f[k] + base, m * sizeof(float), cudaMemcpyHostToDevice, c->s_in));
                    dadjf[k] = field(SLAB_ADJ + k);
                }
            }
            CU(cudaEventRecord(s.in_done, c->s_in));

            // ---- kernels
            hf_prelim dp;
            float **dpf = reinterpret_cast<float **>(&dp);
            for (int k = 0; k < PRE_NFIELDS; ++k) dpf[k] = field(SLAB_PRE + k);
            hf_si dsi; memset(&dsi, 0, sizeof(dsi));
            float **dsif = reinterpret_cast<float **>(&dsi);
            if (hsi) for (int k = 0; k < SI_NFIELDS; ++k) if (hsi_f[k]) dsif[k] = field(SLAB_SI + k);
            CU(cudaStreamWaitEvent(c->s_k, s.in_done, 0));
            CU(hf_launch_intersect(c->P, c->li, HF_MODE_FUSED_SI, dr, dp, dsi, nullptr, flags, m, c->d_counters, SCHED(c, c->s_k), c->s_k));
            if (backward && c->grad_enabled)
                CU(hf_launch_backward(c->P, c->li, dr, dp, dadj, flags, c->d_grad, m, c->s_k));
            CU(cudaEventRecord(s.k_done, c->s_k));

            // ---- device -> host
            CU(cudaStreamWaitEvent(c->s_out, s.k_done, 0));
            if (hpre) for (int k = 0; k < PRE_NFIELDS; ++k)
                CU(cudaMemcpyAsync(hpre_f[k] + base, dpf[k], m * sizeof(float), cudaMemcpyDeviceToHost, c->s_out));
            if (hsi) for (int k = 0; k < SI_NFIELDS; ++k) if (hsi_f[k])
                CU(cudaMemcpyAsync(hsi_f[k] + base, dsif[k], m * sizeof(float), cudaMemcpyDeviceToHost, c->s_out));
            CU(cudaEventRecord(s.out_done, c->s_out));
        }
    }
    if (h_grad_out) {
        // the gradient grid is complete when the last kernel has finished: s_out waits for s_k, then copies
        CU(cudaEventRecord(c->ev_entry, c->s_k));
        CU(cudaStreamWaitEvent(c->s_out, c->ev_entry, 0));
        CU(cudaMemcpyAsync(h_grad_out, c->d_grad, (size_t) c->W * c->H * sizeof(float), cudaMemcpyDeviceToHost, c->s_out));
    }
    CU(cudaStreamSynchronize(c->s_in));
    CU(cudaStreamSynchronize(c->s_k));
    CU(cudaStreamSynchronize(c->s_out));
    return HF_OK;
}

// The optimisation step of BASELINE.json configs[2] for a caller whose arrays live in HOST memory: what crosses PCIe is the parameter
// (heights, optional), the target image and the pixel list on the way in, and the image / gradient grid / loss on the way out --
// no ray or interaction record exists anywhere. Everything runs on the pipeline's kernel stream, ordered after the legacy stream.
extern "C" int hf_render_step_l2_host(hf_ctx *c, const hf_camera *cam, const hf_light *light, const uint32_t *h_pixels, size_t npix,
                                      const float *h_heights, const float *h_target, float *h_image_out, float *h_grad_out,
                                      double *h_loss_out) {
    HF_RANGE();
    CHECK_CTX(c);
    int rc;
    if ((rc = check_camera(cam, true)) || (rc = check_light(light))) return rc;
    const size_t img = (size_t) cam->width * cam->height;
    if (!h_pixels && npix != img) return fail(HF_ERR_INVALID, "npix must be width * height when no pixel list is given");
    if (npix > img) return fail(HF_ERR_INVALID, "more listed pixels than the image has");
    if (!h_target) return fail(HF_ERR_INVALID, "null target image");
    CU(cudaSetDevice(c->device));
    if ((rc = ensure_slots(c))) return rc;
    if (c->hs_img_elems < img) {
        cudaFree(c->d_hs_target); cudaFree(c->d_hs_image); c->d_hs_target = c->d_hs_image = nullptr; c->hs_img_elems = 0;
        CU(cudaMalloc(&c->d_hs_target, img * sizeof(float)));
        CU(cudaMalloc(&c->d_hs_image, img * sizeof(float)));
        c->hs_img_elems = img;
    }
    if (h_pixels && c->hs_pix_elems < npix) {
        cudaFree(c->d_hs_pixels); c->d_hs_pixels = nullptr; c->hs_pix_elems = 0;
        CU(cudaMalloc(&c->d_hs_pixels, npix * sizeof(uint32_t)));
        c->hs_pix_elems = npix;
    }
    if (!c->d_hs_loss) CU(cudaMalloc(&c->d_hs_loss, sizeof(double)));
    cudaStream_t st = c->s_k;
    CU(cudaEventRecord(c->ev_entry, (cudaStream_t) 0));
    CU(cudaStreamWaitEvent(st, c->ev_entry, 0));
    if ((rc = wait_mutations(c, st, nullptr, nullptr))) return rc;
    if (h_heights) {
        CU(cudaMemcpyAsync(c->d_heights, h_heights, (size_t) c->W * c->H * sizeof(float), cudaMemcpyHostToDevice, st));
        c->have_heights = true;
        if ((rc = hf_parameters_changed(c, st))) return rc;
    }
    NEED_ACCEL(c);
    CU(cudaMemcpyAsync(c->d_hs_target, h_target, img * sizeof(float), cudaMemcpyHostToDevice, st));
    if (h_pixels) CU(cudaMemcpyAsync(c->d_hs_pixels, h_pixels, npix * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(c->d_grad, 0, (size_t) c->W * c->H * sizeof(float), st));
    CU(cudaMemsetAsync(c->d_hs_loss, 0, sizeof(double), st));
    if (h_image_out && npix != img) CU(cudaMemsetAsync(c->d_hs_image, 0, img * sizeof(float), st));
    if (npix > 0 && (rc = hf_render_step_l2(c, cam, light, h_pixels ? c->d_hs_pixels : nullptr, npix, c->d_hs_target, c->d_hs_image, c->d_hs_loss, st))) return rc;
    if (h_image_out) CU(cudaMemcpyAsync(h_image_out, c->d_hs_image, img * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (h_loss_out) CU(cudaMemcpyAsync(h_loss_out, c->d_hs_loss, sizeof(double), cudaMemcpyDeviceToHost, st));
    if (h_grad_out) CU(cudaMemcpyAsync(h_grad_out, c->d_grad, (size_t) c->W * c->H * sizeof(float), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return HF_OK;
}

extern "C" int hf_ray_intersect_host(hf_ctx *c, const hf_rays *h_rays, uint32_t flags, const hf_prelim *h_prelim_out,
                                     const hf_si *h_out, size_t n) {
    HF_RANGE();
    CHECK_CTX(c);
    return host_pipeline(c, h_rays, nullptr, flags, h_prelim_out, h_out, nullptr, n, false);
}

extern "C" int hf_forward_backward_host(hf_ctx *c, const hf_rays *h_rays, const hf_si_adj *h_adj, uint32_t flags,
                                        const hf_prelim *h_prelim_out, const hf_si *h_out, float *h_grad_out, size_t n) {
    HF_RANGE();
    CHECK_CTX(c);
    return host_pipeline(c, h_rays, h_adj, flags, h_prelim_out, h_out, h_grad_out, n, true);
}

// ------------------------------------------------------------------------------------------
// hf_ray_intersect_preliminary_static: the preliminary intersection for callers that must not allocate device memory
// ------------------------------------------------------------------------------------------
// Who needs this. A context (hf_create) owns cudaMalloc'ed grids, so it is of no use to code that is not allowed to reach the CUDA
// allocator at all: a plugin inside a host application that has handed the whole device-memory budget to its own pool (Dr.Jit's
// caching allocator in Mitsuba is one) and treats a foreign cudaMalloc as a bug; a frame / control loop with a latency bound, where
// cudaMalloc's device-wide synchronisation is not tolerable; a process whose memory is partitioned ahead of time (MPS limits,
// several tenants on one GPU) and that may not grow. For them the acceleration data of the walk (the fp32 min/max pyramid, the
// quantised pyramid and the work counter of the self-scheduling kernel) lives in the array below, which is part of the library image:
// the driver places it when the module is loaded on a GPU and it stays there until the process ends, on every GPU that runs this
// entry point. The caller supplies the heights, the rays and the result arrays; nothing is allocated or freed here, neither by
// CUDA nor by any library on top of it (the few bytes of host heap the layout helper takes are the only allocation of any kind).
//
// HF_STATIC_SCRATCH_BYTES bounds the field: whatever needs more is rejected with HF_ERR_INVALID before any device work. 9 MiB hold
// every field of up to 1025 x 1025 texels (1024 x 1024 cells: 2.7 MiB fp32 pyramid + 5.4 MiB quantised pyramid) and any other shape
// whose padded pyramids fit (the quantised one pads each axis to a power of two).
//
// `g_static_scratch` is a device symbol. In host code the name is only a handle for the runtime (its host-side shadow), never a
// device address, so the launchers get the pointer cudaGetSymbolAddress returns; it is looked up once per device, with that device
// current, on the first call that targets the device. That first call also loads the module and is best made before any stream
// capture begins.
//
// The array is shared by every call in the process and keeps what the previous call left in it. Each call therefore clears the part
// it uses (in stream order) and rebuilds both pyramids from the caller's heights before it walks them. Two calls that may run on
// the GPU at the same time (different streams without an event between them, different host threads) would build and read the same
// bytes: they race, and the caller has to serialise them.
#include <mutex>
#define HF_STATIC_MAX_DEVICES 64
static __device__ __align__(256) unsigned char g_static_scratch[HF_STATIC_SCRATCH_BYTES];

struct HfStaticDevice { unsigned char *base = nullptr; int num_sms = 0; };
static HfStaticDevice g_static_dev[HF_STATIC_MAX_DEVICES];

// once per device, with that device current
static int static_scratch_for(int device, HfStaticDevice *out) {
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    HfStaticDevice &d = g_static_dev[device];
    int rc = HF_OK;
    if (!d.base) {
        void *p = nullptr;
        int sms = 0;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || !p || sms <= 0) rc = fail(HF_ERR_CUDA, "static scratch of device %d: %s", device, e != cudaSuccess ? cudaGetErrorString(e) : "no address");
        else { d.num_sms = sms; d.base = static_cast<unsigned char *>(p); }
    }
    *out = d;
    return rc;
}

extern "C" int hf_ray_intersect_preliminary_static(const hf_desc *desc, const float *d_heights, const hf_rays *rays, const hf_prelim *out,
                                                   size_t n, void *stream) {
    HF_RANGE();
    if (!desc) return fail(HF_ERR_INVALID, "null heightfield description");
    if (desc->width < 2 || desc->height < 2) return fail(HF_ERR_INVALID, "heightfield needs at least 2x2 texels (got %ux%u)", desc->width, desc->height);
    if (desc->width > (1u << 23) || desc->height > (1u << 23)) return fail(HF_ERR_INVALID, "heightfield side exceeds 2^23 texels");
    if (!isfinite(desc->max_height)) return fail(HF_ERR_INVALID, "max_height is not finite");
    if (!d_heights) return fail(HF_ERR_INVALID, "null heights");

    // the same description -> parameter block -> pyramid layout as hf_create, in a context object that lives on this stack frame and
    // owns nothing (its device pointers are the caller's heights and slices of the static array)
    hf_ctx c{};
    HfParams &P = c.P;
    memset(&P, 0, sizeof(P));
    c.W = desc->width; c.H = desc->height;
    P.W = (int) desc->width; P.H = (int) desc->height; P.NX = P.W - 1; P.NY = P.H - 1;
    P.flip_normals = desc->flip_normals ? 1 : 0;
    P.mh = desc->max_height;
    P.rw = 1.0f / (float) P.NX; P.rh = 1.0f / (float) P.NY;
    P.x0 = (float) (-(P.W - 1)) * P.rw; P.y0 = (float) (-(P.H - 1)) * P.rh;
    P.sgx = 0.5f * (float) P.NX; P.sgy = 0.5f * (float) P.NY;
    { const HfDivMagic dm = hf_divmagic((uint32_t) P.NX); P.nx_magic = dm.m; P.nx_shift = dm.s; }
    int rc = set_transform(&c, desc->to_world, desc->has_to_object ? desc->to_object : nullptr);
    if (rc) return rc;

    // the size check comes before the layout is even formed for fields that are far too large (2^23 x 2^23 texels)
    if ((uint64_t) (desc->width - 1) * (desc->height - 1) > HF_STATIC_SCRATCH_BYTES / sizeof(uint32_t))
        return fail(HF_ERR_INVALID, "%ux%u heightfield is beyond the static scratch of %zu bytes (any field up to 1025x1025 texels fits); use hf_create",
                    desc->width, desc->height, HF_STATIC_SCRATCH_BYTES);
    layout_pyramid(&c);
    const size_t counter_bytes = 256;                                              // [0..7] statistics block, [8] the work counter
    const size_t pyr_bytes = (c.pyr_nodes * sizeof(float2) + 255) & ~size_t(255);
    const size_t qpyr_bytes = (c.qpyr_nodes * sizeof(uint32_t) + 255) & ~size_t(255);
    const size_t need = counter_bytes + pyr_bytes + qpyr_bytes;
    if (need > HF_STATIC_SCRATCH_BYTES)
        return fail(HF_ERR_INVALID, "%ux%u heightfield needs %zu bytes of acceleration data, the static scratch holds %zu (any field up to 1025x1025 texels fits); use hf_create",
                    desc->width, desc->height, need, HF_STATIC_SCRATCH_BYTES);
    if (n == 0) return HF_OK;      // empty batch: nothing to do, pointers may be null
    if ((rc = check_rays(rays)) || (rc = check_prelim(out))) return rc;

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(HF_ERR_CUDA, "no CUDA device available (%s); this library has no CPU path", e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
    if (desc->device < 0 || desc->device >= ndev || desc->device >= HF_STATIC_MAX_DEVICES) return fail(HF_ERR_INVALID, "device ordinal %d out of range", desc->device);
    CU(cudaSetDevice(desc->device));
    HfStaticDevice dev;
    if ((rc = static_scratch_for(desc->device, &dev))) return rc;

    unsigned long long *counters = reinterpret_cast<unsigned long long *>(dev.base);
    float2 *pyr = reinterpret_cast<float2 *>(dev.base + counter_bytes);
    uint32_t *qpyr = reinterpret_cast<uint32_t *>(dev.base + counter_bytes + pyr_bytes);
    c.device = desc->device;
    c.num_sms = dev.num_sms;
    c.li.num_sms = dev.num_sms;
    c.li.seg_counter = counters + 4;
    P.heights = d_heights;
    P.pyr = pyr;
    P.qpyr = qpyr;

    const cudaStream_t st = (cudaStream_t) stream;
    CU(cudaMemsetAsync(dev.base, 0, need, st));                // the previous call's counters and nodes
    CU(hf_launch_pyramid(P, pyr, qpyr, st));
    hf_si none; memset(&none, 0, sizeof(none));
    CU(hf_launch_intersect(P, c.li, HF_MODE_PRELIM, *rays, *out, none, nullptr, 0, n, counters, counters + 8, st));
    return HF_OK;
}
