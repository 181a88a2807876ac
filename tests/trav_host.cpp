// trav_host.cpp -- HOST instantiation of the production walk (csrc/hf_traverse.cuh) for CPU-side testing.
// TEST INFRASTRUCTURE: compiled by tests/test_traversal_host.py with g++ and linked against the oracle, whose
// per-cell primitive and pyramid it borrows; the shipped CUDA library contains none of this.
//
// It answers two questions without a GPU: (1) is the warp-uniform walk conservative -- same closest hit as the
// oracle's brute force / canonical DDA on millions of random + adversarial rays; (2) how many loop iterations and
// primitive rounds does a warp of 32 consecutive rays need under a given HF_START_LEVEL / HF_POP_LEVELS.
#include <stdlib.h>
#include <string.h>
#include <vector>

#ifdef HF_WALK_HIST
unsigned long long hf_walk_hist[3][32];
extern "C" unsigned long long *trav_host_hist(void) { return &hf_walk_hist[0][0]; }
#endif
#define HF_DEBUG_BOUNDS 1     // the host build always checks the node / texel indices the walk forms
#include "../mitsuba3-differentiable-heightfield-rendering_b200/csrc/hf_traverse.cuh"
#include "../oracle/hf_oracle.h"

#ifndef HF_HOST_ROUNDS
#define HF_HOST_ROUNDS 8      // dev statistics: walk iterations recorded for the first HF_HOST_ROUNDS candidate rounds of a ray
#endif


// Quantised pyramid (csrc/hf_qpyr.h) built on the host with the very functions the device builder uses.
struct HostQPyr {
    std::vector<uint32_t> nodes;
    std::vector<HfQLevel> lvl;
    HfQLayout lay;
    void build(const hfo_shape *s, float top_min, float top_max) {
        const int NX = s->W - 1, NY = s->H - 1;
        lay = hf_qlayout(NX, NY);
        nodes.assign(lay.nodes, HF_Q_EMPTY);
        const HfQuant Q = hf_quant_make(top_min, top_max);
        for (int y = 0; y < NY; ++y)
            for (int x = 0; x < NX; ++x) {
                const float *r0 = s->heights + (size_t) y * s->W + x;
                const float z00 = s->max_height * r0[0], z10 = s->max_height * r0[1], z01 = s->max_height * r0[s->W], z11 = s->max_height * r0[s->W + 1];
                nodes[lay.off[0] + (size_t) y * lay.pitch[0] + x] =
                    hf_quant_node(fminf(fminf(z00, z10), fminf(z01, z11)), fmaxf(fmaxf(z00, z10), fmaxf(z01, z11)), Q);
            }
        for (int l = 1; l <= lay.Ltop; ++l) {
            const int fx = lay.pitch[l - 1] / lay.pitch[l], fy = lay.rows[l - 1] / lay.rows[l];
            for (int ny = 0; ny < lay.rows[l]; ++ny)
                for (int nx = 0; nx < lay.pitch[l]; ++nx) {
                    uint32_t q = HF_Q_EMPTY;
                    for (int dj = 0; dj < fy; ++dj)
                        for (int di = 0; di < fx; ++di)
                            q = hf_quant_merge(q, nodes[lay.off[l - 1] + (size_t) (ny * fy + dj) * lay.pitch[l - 1] + (nx * fx + di)]);
                    nodes[lay.off[l] + (size_t) ny * lay.pitch[l] + nx] = q;
                }
        }
        lvl.resize(lay.Ltop + 1);
        for (int l = 0; l <= lay.Ltop; ++l) { lvl[l].nodes = nodes.data() + lay.off[l]; lvl[l].pitch = lay.pitch[l]; lvl[l].pad = 0; }
    }
    void attach(HfGrid &G, float top_min, float top_max) const {
        const HfQuant Q = hf_quant_make(top_min, top_max);
        G.qlvl = lvl.data(); G.qlvl_saddr = 0; G.QLtop = lay.Ltop; G.xmask = lay.Mx - 1; G.ymask = lay.My - 1; G.qs = Q.qs; G.zb = Q.zb; G.rqs = Q.rqs;
    }
};

// rounds (nullable): uint32[n][8], loop iterations spent in each call of hf_walk_next (candidate round)
extern "C" int trav_host_intersect(const hfo_shape *s, const hfo_rays *r, size_t n, hfo_prelim *out,
                                   uint32_t *iters, uint32_t *cands, uint32_t *rounds) {
    hfo_pyramid *p = hfo_pyramid_build(s);
    const int Ltop = hfo_pyramid_levels(p);
    std::vector<HfLevel> lvl(Ltop + 1);
    lvl[0].nodes = nullptr; lvl[0].w = 0;
    for (int l = 1; l <= Ltop; ++l) {
        int w, h;
        hfo_pyramid_level_dims(p, l, &w, &h);
        lvl[l].nodes = reinterpret_cast<const float2 *>(hfo_pyramid_level_data(p, l));   // interleaved (min,max)
        lvl[l].w = w;
    }
    HfGrid G;
    G.W = s->W; G.NX = s->W - 1; G.NY = s->H - 1; G.Ltop = Ltop; G.mh = s->max_height;
    const float rw = 1.0f / (float) G.NX, rh = 1.0f / (float) G.NY;
    G.x0 = (float) (-(s->W - 1)) * rw; G.y0 = (float) (-(s->H - 1)) * rh;
    G.sgx = 0.5f * (float) G.NX; G.sgy = 0.5f * (float) G.NY;
    G.heights = s->heights; G.lvl = lvl.data();
    G.top_min = lvl[Ltop].nodes[0].x; G.top_max = lvl[Ltop].nodes[0].y;
    HostQPyr qp; qp.build(s, G.top_min, G.top_max); qp.attach(G, G.top_min, G.top_max);

    #pragma omp parallel for schedule(dynamic, 1024)
    for (long long ii = 0; ii < (long long) n; ++ii) {
        const size_t i = (size_t) ii;
        unsigned it = 0, nc = 0;
        float bt = INFINITY, bu = 0.f, bv = 0.f; uint32_t bp = 0;
        if (!r->active || r->active[i]) {
            const float *m = s->to_object;
            const float ow[3] = { r->ox[i], r->oy[i], r->oz[i] }, dw[3] = { r->dx[i], r->dy[i], r->dz[i] };
            float o[3], d[3];
            for (int a = 0; a < 3; ++a) {        // transform.h:104-111, 129-139
                float v = m[4 * a + 3];
                v = fmaf(m[4 * a + 0], ow[0], v); v = fmaf(m[4 * a + 1], ow[1], v); v = fmaf(m[4 * a + 2], ow[2], v);
                o[a] = v;
                float q = m[4 * a + 0] * dw[0];
                q = fmaf(m[4 * a + 1], dw[1], q); q = fmaf(m[4 * a + 2], dw[2], q);
                d[a] = q;
            }
            HfWalk w; HfCand c;
            if (hf_walk_init(G, o[0], o[1], o[2], d[0], d[1], d[2], r->maxt[i], w)) {
                for (;;) {
                    const unsigned before = it;
                    const bool more = hf_walk_next(G, w, c, &it);
                    if (rounds && nc < HF_HOST_ROUNDS) rounds[HF_HOST_ROUNDS * i + nc] = it - before;
                    if (!more) break;
                    ++nc;
                    float t, u, v; uint32_t prim;
                    if (hfo_cell_intersect(s, c.cx, c.cy, o, d, r->maxt[i], &t, &u, &v, &prim)) { bt = t; bu = u; bv = v; bp = prim; break; }
                }
            }
        }
        out->t[i] = bt; out->u[i] = bu; out->v[i] = bv; out->prim_index[i] = bp;
        if (iters) iters[i] = it;
        if (cands) cands[i] = nc;
    }
    hfo_pyramid_free(p);
    return 0;
}

static uint32_t *g_beam_rounds = nullptr;      // dev statistics: uint32[n][8] iterations per candidate round (bundled variant)
extern "C" void trav_host_set_beam_rounds(uint32_t *p) { g_beam_rounds = p; }

// Bundled variant: every 32 consecutive rays with a bit-identical origin form a bundle; hf_beam_enter gives the bundle's
// entry parameter and each ray walks from there. beam_iters (nullable): uint32[n/32] iterations of the bundle march.
extern "C" int trav_host_intersect_beam(const hfo_shape *s, const hfo_rays *r, size_t n, hfo_prelim *out,
                                        uint32_t *iters, uint32_t *cands, uint32_t *beam_iters, float *t_enter_out) {
    hfo_pyramid *p = hfo_pyramid_build(s);
    const int Ltop = hfo_pyramid_levels(p);
    std::vector<HfLevel> lvl(Ltop + 1);
    lvl[0].nodes = nullptr; lvl[0].w = 0;
    for (int l = 1; l <= Ltop; ++l) {
        int w, h;
        hfo_pyramid_level_dims(p, l, &w, &h);
        lvl[l].nodes = reinterpret_cast<const float2 *>(hfo_pyramid_level_data(p, l));
        lvl[l].w = w;
    }
    HfGrid G;
    G.W = s->W; G.NX = s->W - 1; G.NY = s->H - 1; G.Ltop = Ltop; G.mh = s->max_height;
    const float rw = 1.0f / (float) G.NX, rh = 1.0f / (float) G.NY;
    G.x0 = (float) (-(s->W - 1)) * rw; G.y0 = (float) (-(s->H - 1)) * rh;
    G.sgx = 0.5f * (float) G.NX; G.sgy = 0.5f * (float) G.NY;
    G.heights = s->heights; G.lvl = lvl.data();
    G.top_min = lvl[Ltop].nodes[0].x; G.top_max = lvl[Ltop].nodes[0].y;
    HostQPyr qp; qp.build(s, G.top_min, G.top_max); qp.attach(G, G.top_min, G.top_max);
    const float *m = s->to_object;
    auto xf = [&](size_t i, float o[3], float d[3]) {
        const float ow[3] = { r->ox[i], r->oy[i], r->oz[i] }, dw[3] = { r->dx[i], r->dy[i], r->dz[i] };
        for (int a = 0; a < 3; ++a) {
            float v = m[4 * a + 3];
            v = fmaf(m[4 * a + 0], ow[0], v); v = fmaf(m[4 * a + 1], ow[1], v); v = fmaf(m[4 * a + 2], ow[2], v);
            o[a] = v;
            float q = m[4 * a + 0] * dw[0];
            q = fmaf(m[4 * a + 1], dw[1], q); q = fmaf(m[4 * a + 2], dw[2], q);
            d[a] = q;
        }
    };
    const long long ngroups = (long long) ((n + 31) / 32);
    #pragma omp parallel for schedule(dynamic, 64)
    for (long long g = 0; g < ngroups; ++g) {
        const size_t i0 = (size_t) g * 32, i1 = i0 + 32 < n ? i0 + 32 : n;
        // bundle bounds (what the kernel gets from warp shuffles)
        bool same = true; int nact = 0;
        HfBeam b; b.dxl = b.dyl = b.dzl = INFINITY; b.dxh = b.dyh = b.dzh = -INFINITY; b.ogx = b.ogy = b.oz = 0.f;
        float o0[3] = { 0, 0, 0 };
        for (size_t i = i0; i < i1; ++i) {
            if (r->active && !r->active[i]) continue;
            float o[3], d[3]; xf(i, o, d);
            if (nact == 0) { o0[0] = o[0]; o0[1] = o[1]; o0[2] = o[2]; }
            else if (memcmp(o, o0, sizeof(o0)) != 0) same = false;
            const float dgx = d[0] * G.sgx, dgy = d[1] * G.sgy;
            b.dxl = fminf(b.dxl, dgx); b.dxh = fmaxf(b.dxh, dgx); b.dyl = fminf(b.dyl, dgy); b.dyh = fmaxf(b.dyh, dgy);
            b.dzl = fminf(b.dzl, d[2]); b.dzh = fmaxf(b.dzh, d[2]);
            if (!(d[0] == d[0] && d[1] == d[1] && d[2] == d[2])) same = false;
            ++nact;
        }
        float t_enter = 0.f; int L_start = HF_START_LEVEL; unsigned bit = 0;
        if (same && nact > 1) {
            b.ogx = (o0[0] - G.x0) * G.sgx; b.ogy = (o0[1] - G.y0) * G.sgy; b.oz = o0[2];
            t_enter = hf_beam_enter(G, b, L_start, &bit);
        }
        if (beam_iters) beam_iters[g] = bit;
        if (t_enter_out) t_enter_out[g] = t_enter;
        for (size_t i = i0; i < i1; ++i) {
            unsigned it = 0, nc = 0;
            float bt = INFINITY, bu = 0.f, bv = 0.f; uint32_t bp = 0;
            if (!r->active || r->active[i]) {
                float o[3], d[3]; xf(i, o, d);
                HfWalk w; HfCand c;
                if (hf_walk_init(G, o[0], o[1], o[2], d[0], d[1], d[2], r->maxt[i], w, t_enter, L_start)) {
                    for (;;) {
                        const unsigned before = it;
                        const bool more = hf_walk_next(G, w, c, &it);
                        if (g_beam_rounds && nc < 8) g_beam_rounds[8 * i + nc] = it - before;
                        if (!more) break;
                        ++nc;
                        float t, u, v; uint32_t prim;
                        if (hfo_cell_intersect(s, c.cx, c.cy, o, d, r->maxt[i], &t, &u, &v, &prim)) { bt = t; bu = u; bv = v; bp = prim; break; }
                    }
                }
            }
            out->t[i] = bt; out->u[i] = bu; out->v[i] = bv; out->prim_index[i] = bp;
            if (iters) iters[i] = it;
            if (cands) cands[i] = nc;
        }
    }
    hfo_pyramid_free(p);
    return 0;
}
