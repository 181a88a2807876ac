// hf_render.cu -- the render step on the heightfield alone (include/hf_b200.h, "render step" section; SURVEY.md 8(f)
// rows 1-2): perspective pinhole rays generated in registers, the production walk (hf_traverse.cuh), Lambert shading
// under one directional emitter (optionally with a visibility ray), box-filter pixel means, and the adjoint of all of it
// with respect to the height texels -- without a ray, an intersection record or a surface interaction ever touching HBM.
//
// Work decomposition: a warp owns a tile of listed pixels, fetched from a global counter: HF_RENDER_TILE rounds of 32 rays, i.e.
// HF_RENDER_TILE pixels at >= 32 spp and up to 32 pixels below (half and quarter tiles in the closing phases of the schedule).
//   phase B  all 32 lanes march bundles through the pyramid: each pixel of the tile is split into 8 / 4 / 2 / 1 sub-rectangles of
//            its jitter square (direction bounds = the corners of the sub-rectangle, known analytically) -> entry parameter
//            and start level for the samples that fall into it;
//   phase C  round by round, the 32 lanes trace 32 samples (of one pixel, or of 32 / seg pixels when spp < 32) from their entry
//            parameters, shade them, and reduce each pixel's mean with shuffles in a fixed order (deterministic image);
//   epilogue the pixel's adjoint is known as soon as its mean is (L2 step) or is read from the adjoint image
//            (backward); the samples' hit triangles wait in shared memory, every distinct triangle of a round gets
//            ONE set of three red.global.add.f32 (match.any on the triangle id; all samples of a pixel carry the
//            same adjoint, so a group's contribution is count x value).
//
// Arithmetic of the shape part is the reference's op order (hf_device.cuh); the sensor / shading part is this
// file's own definition, restated op for op by the oracle (oracle/hf_oracle.c, hfo_render_*).
// Compiled with -fmad=false like the rest: every fma below is explicit.
#include <string.h>
#include <stdlib.h>
#include "hf_walk_common.cuh"

#define HF_RENDER_MAX_SPP 256          // samples of one pixel whose hit triangles wait in shared memory
#define HF_NO_PRIM 0xffffffffu
#ifndef HF_RENDER_TILE
#define HF_RENDER_TILE 4               // 32-ray rounds per warp tile: small tiles keep the tail of the dynamic schedule short
#endif

#ifndef HF_RENDER_TAIL
#define HF_RENDER_TAIL 24              // quarter tiles per resident warp that go out one at a time at the end (a pixel whose rays skim down a slope takes ~0.2 ms)
#endif
#ifndef HF_RENDER_MIDTAIL
#define HF_RENDER_MIDTAIL 24           // quarter tiles per resident warp that go out in pairs before those
#endif
#if HF_RENDER_TILE != 2 && HF_RENDER_TILE != 4 && HF_RENDER_TILE != 8 && HF_RENDER_TILE != 16 && HF_RENDER_TILE != 32
#error "HF_RENDER_TILE must be 2, 4, 8, 16 or 32"
#endif

struct HfRenderArgs {
    hf_camera cam;
    float lx, ly, lz;                  // unit vector towards the emitter, world space
    float scale;                       // albedo / pi * irradiance
    float albedo;
    int max_depth;
    int shadow;
    float inv_w, inv_h, tan_x, tan_y;  // 1/width, 1/height, tan(fov/2), tan(fov/2) * height / width
    float inv_spp, inv_npix;           // 1/spp, 1/(width*height)
    f3 o_obj;                          // camera position in object space
    uint32_t seed_h0, seed_h1;         // lowbias32(seed * 0x9E3779B1 + channel * 0x85EBCA6B) for the two jitter channels (same for every sample)
    const uint32_t *pixels;            // nullable
    size_t npix;
    int tail, midtail;                 // closing phases of the schedule: quarter tiles per resident warp (HF_RENDER_TAIL / _MIDTAIL)
};

// ---- sampling ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t hf_hash32(uint32_t x) {       // lowbias32
    x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
    return x;
}
// U[0,1) with 24 bits from (seed, index, channel) -- same stream as workloads.hash_uniform
__device__ __forceinline__ float hf_hash_uniform(uint32_t seed, uint32_t idx, uint32_t channel) {
    const uint32_t h = hf_hash32(idx ^ hf_hash32(seed * 0x9E3779B1u + channel * 0x85EBCA6Bu));
    return (float) (h >> 8) * (1.0f / 16777216.0f);
}

// Film-plane coordinates of a sub-pixel position -> un-normalised world direction fwd + sx right + sy up
__device__ __forceinline__ f3 hf_film_dir(const HfRenderArgs &A, float fx, float fy) {
    const float sx = fmaf(2.f, fx, -1.f) * A.tan_x;
    const float sy = fmaf(-2.f, fy, 1.f) * A.tan_y;
    return mk3(fmaf(sy, A.cam.up[0], fmaf(sx, A.cam.right[0], A.cam.forward[0])),
               fmaf(sy, A.cam.up[1], fmaf(sx, A.cam.right[1], A.cam.forward[1])),
               fmaf(sy, A.cam.up[2], fmaf(sx, A.cam.right[2], A.cam.forward[2])));
}

// perspective.cpp:212-231 for a pinhole without clipping planes: o = camera position, d = normalised direction
__device__ __forceinline__ void hf_camera_jitter(const HfRenderArgs &A, uint32_t pid, uint32_t s, float &jx, float &jy) {
    jx = 0.5f; jy = 0.5f;
    if (A.cam.spp > 1) {
        const unsigned long long k = (unsigned long long) pid * A.cam.spp + s;
        const uint32_t key = (uint32_t) k ^ ((uint32_t) (k >> 32) * 0x85EBCA6Bu);
        jx = (float) (hf_hash32(key ^ A.seed_h0) >> 8) * (1.0f / 16777216.0f);      // == hf_hash_uniform(seed, key, 0 / 1)
        jy = (float) (hf_hash32(key ^ A.seed_h1) >> 8) * (1.0f / 16777216.0f);
    }
}
__device__ __forceinline__ f3 hf_camera_dir_at(const HfRenderArgs &A, uint32_t pid, float jx, float jy) {
    const uint32_t x = pid % A.cam.width, y = pid / A.cam.width;
    const float fx = ((float) x + jx) * A.inv_w, fy = ((float) y + jy) * A.inv_h;
    return normalize3(hf_film_dir(A, fx, fy));
}
__device__ __forceinline__ f3 hf_camera_dir_xy(const HfRenderArgs &A, uint32_t x, uint32_t y, float jx, float jy) {   // pixel (x, y) given
    const float fx = ((float) x + jx) * A.inv_w, fy = ((float) y + jy) * A.inv_h;
    return normalize3(hf_film_dir(A, fx, fy));
}
__device__ __forceinline__ f3 hf_camera_dir(const HfRenderArgs &A, uint32_t pid, uint32_t s) {
    float jx, jy;
    hf_camera_jitter(A, pid, s, jx, jy);
    return hf_camera_dir_at(A, pid, jx, jy);
}

// ---- one sample: trace, shade -------------------------------------------------------------------------------------
// Closest hit of the object-space ray (o, d) from parameter t_start on; fills the hit triangle's vertices.
__device__ __forceinline__ bool hf_trace(const HfParams &P, const HfGrid &G, f3 o, f3 d, float maxt, float t_start, int L_start,
                                         float &ht, float &hu, float &hv, uint32_t &hprim, f3 &p0, f3 &p1, f3 &p2,
                                         const unsigned conv = 0u) {
    HfWalk w;
    HfCand c;
    bool hit = false;
    if (hf_walk_init(G, o.x, o.y, o.z, d.x, d.y, d.z, maxt, w, t_start, L_start)) {
        while (hf_walk_next(G, w, c)) {
            if (cell_intersect(P, c.cx, c.cy, c.z00, c.z10, c.z01, c.z11, o, d, maxt, ht, hu, hv, hprim)) { hit = true; break; }
        }
    }
    // conv = the lanes that entered together: without this barrier the lanes that hit in different primitive rounds run
    // the shading below in separate groups of ~8 (profiles/r01_notes.md)
    if (conv) __syncwarp(conv);
    if (hit) {
        const float x0 = vtx(c.cx, P.W, P.rw), x1 = vtx(c.cx + 1, P.W, P.rw);
        const float y0 = vtx(c.cy, P.H, P.rh), y1 = vtx(c.cy + 1, P.H, P.rh);
        p0 = mk3(x1, y0, c.z10);
        if ((hprim & 1u) == 0u) { p1 = mk3(x0, y1, c.z01); p2 = mk3(x0, y0, c.z00); }
        else { p1 = mk3(x1, y1, c.z11); p2 = mk3(x0, y1, c.z01); }
    }
    return hit;
}

#ifndef HF_SHADOW_START_LEVEL
#define HF_SHADOW_START_LEVEL 2        // visibility rays start ON the surface: every level above overlaps at the origin anyway
#endif
// any hit (Shape::ray_test): the walk's first hit cell decides
__device__ __forceinline__ bool hf_occluded(const HfParams &P, const HfGrid &G, f3 o, f3 d, float maxt) {
    HfWalk w;
    HfCand c;
    float t, u, v; uint32_t prim;
    if (hf_walk_init(G, o.x, o.y, o.z, d.x, d.y, d.z, maxt, w, 0.f, HF_SHADOW_START_LEVEL, true))
        while (hf_walk_next(G, w, c))
            if (cell_intersect(P, c.cx, c.cy, c.z00, c.z10, c.z01, c.z11, o, d, maxt, t, u, v, prim)) return true;
    return false;
}

// World-space geometric normal of triangle (p0, p1, p2): mesh.cpp:757 then the normal transform of rectangle.cpp:281-286
__device__ __forceinline__ f3 hf_world_normal(const HfParams &P, f3 p0, f3 p1, f3 p2) {
    const f3 no = normalize3(cross3(sub3(p1, p0), sub3(p2, p0)));
    f3 nw = normalize3(xform_normal(P.to_object, no));
    if (P.flip_normals) nw = neg3(nw);
    return nw;
}

// Radiance of one camera sample: direct illumination of the first hit (direct.cpp:114-210 with one directional emitter and a
// Lambertian BSDF). Returns the value; `prim` = hit triangle when the sample is lit (its value depends on the heights
// through the normal), HF_NO_PRIM otherwise.
__device__ __forceinline__ float hf_shade_sample(const HfParams &P, const HfGrid &G, const HfRenderArgs &A, f3 dw, float t_start,
                                                 int L_start, uint32_t &prim, const unsigned conv = 0u) {
    prim = HF_NO_PRIM;
    const f3 d = xform_vector(P.to_object, dw);
    float ht, hu, hv; uint32_t hprim;
    f3 p0, p1, p2;
    if (!hf_trace(P, G, A.o_obj, d, HF_INF, t_start, L_start, ht, hu, hv, hprim, p0, p1, p2, conv)) return 0.f;
    const f3 nw = hf_world_normal(P, p0, p1, p2);
    const f3 l = mk3(A.lx, A.ly, A.lz);
    const float cosl = dot3(nw, l);
    // diffuse.cpp:127-139: the BSDF is one-sided, cos(theta_i) = dot(n, -d) and cos(theta_o) = dot(n, l) must both be positive
    if (!(dot3(nw, neg3(dw)) > 0.f) || !(cosl > 0.f)) return 0.f;
    if (A.shadow) {
        // si.spawn_ray(l): interaction.h:136-165; p by the barycentric fma chain of mesh.cpp:748
        const float b1 = hu, b2 = hv, b0 = 1.f - b1 - b2;
        const f3 po = mk3(fmaf(p0.x, b0, fmaf(p1.x, b1, p2.x * b2)), fmaf(p0.y, b0, fmaf(p1.y, b1, p2.y * b2)),
                          fmaf(p0.z, b0, fmaf(p1.z, b1, p2.z * b2)));
        const f3 pw = xform_point(P.to_world, po);
        float mag = (1.f + fmaxf(fmaxf(fabsf(pw.x), fabsf(pw.y)), fabsf(pw.z))) * 8.940696716308594e-05f;   // RayEpsilon = 1500 * 2^-24
        mag = mulsign_(mag, cosl);
        const f3 os = mk3(fmaf(mag, nw.x, pw.x), fmaf(mag, nw.y, pw.y), fmaf(mag, nw.z, pw.z));
        if (hf_occluded(P, G, xform_point(P.to_object, os), xform_vector(P.to_object, l), 3.4028234663852886e+38f)) return 0.f;
    }
    prim = hprim;
    return A.scale * cosl;
}

// d(value)/d(heights) of a lit sample on triangle `prim` for an upstream adjoint a_v of the sample value:
// value = scale * dot(n_w, l), n_w = normalize(TO^T normalize(cross(e1, e2))). Returns the three texel indices / gradients.
__device__ __forceinline__ void hf_shade_adjoint(const HfParams &P, const HfRenderArgs &A, uint32_t prim, float a_v,
                                                 uint32_t &i0, uint32_t &i1, uint32_t &i2, float &g0, float &g1, float &g2) {
    f3 p0, p1, p2;
    cell_triangle(P, prim, p0, p1, p2, i0, i1, i2);
    const f3 e1 = sub3(p1, p0), e2 = sub3(p2, p0);
    const f3 c = cross3(e1, e2);
    const float k = a_v * A.scale;
    f3 a_nw = mk3(k * A.lx, k * A.ly, k * A.lz);
    if (P.flip_normals) a_nw = neg3(a_nw);
    const float inv_clen = rsqrtf(dot3(c, c));
    const f3 no = mk3(c.x * inv_clen, c.y * inv_clen, c.z * inv_clen);
    const f3 m = lin_T(P.to_object, no);
    const float inv_mlen = rsqrtf(dot3(m, m));
    const f3 nw = mk3(m.x * inv_mlen, m.y * inv_mlen, m.z * inv_mlen);
    const float dn = dot3(nw, a_nw);
    const f3 a_m = mk3((a_nw.x - nw.x * dn) * inv_mlen, (a_nw.y - nw.y * dn) * inv_mlen, (a_nw.z - nw.z * dn) * inv_mlen);
    const f3 a_no = lin(P.to_object, a_m);
    const float dno = dot3(no, a_no);
    const f3 a_c = mk3((a_no.x - no.x * dno) * inv_clen, (a_no.y - no.y * dno) * inv_clen, (a_no.z - no.z * dno) * inv_clen);
    const f3 a_e1 = cross3(e2, a_c), a_e2 = cross3(a_c, e1);
    g0 = P.mh * -(a_e1.z + a_e2.z);
    g1 = P.mh * a_e1.z;
    g2 = P.mh * a_e2.z;
}

// ---- path tracing on the heightfield alone ---------------------------------------------------------
// src/integrators/path.cpp:133-250 for one Lambertian surface under one directional emitter: at every vertex emitter sampling
// (direct term, optional visibility ray), then a cosine-weighted bounce (warp.h:54-90, 320-328 in the frame of
// include/mitsuba/core/frame.h:24-41), throughput *= albedo. No emitter is ever hit by a BSDF sample (delta emitter), a ray
// that leaves the heightfield ends the path. Adjoint as in src/python/python/ad/integrators/prb.py:98-254: the sampled
// directions, hit positions and visibility are detached; the heights act through every vertex normal n_k, via the direct
// term scale * dot(n_k, l) and via the BSDF value albedo / pi * dot(n_k, wo_k) over its detached pdf.
#define HF_MAX_DEPTH 4

struct HfVertex {               // what the adjoint needs of path vertex k
    uint32_t prim;
    float ld;                   // direct radiance leaving the vertex: scale * cos_l when lit, else 0
    float lit;                  // 1 when lit
    float wx, wy, wz, cwo;      // bounce direction (world) and dot(n, wo); cwo = 0 at the last vertex
};

// sin / cos of the concentric-map angle through double precision: the CPU oracle and the GPU then agree to the last bit of
// the fp32 result (libm and CUDA fp32 sincos differ by an ulp, which would send a bounce to another triangle now and then)
__device__ __forceinline__ void hf_sincos_exact(float phi, float &s, float &c) {
    double sd, cd;
    sincos((double) phi, &sd, &cd);
    s = (float) sd; c = (float) cd;
}

// warp::square_to_cosine_hemisphere (warp.h:320-328) over square_to_uniform_disk_concentric (warp.h:54-90)
__device__ __forceinline__ f3 hf_cosine_hemisphere(float u1, float u2) {
    const float x = fmaf(2.f, u1, -1.f), y = fmaf(2.f, u2, -1.f);
    const bool zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    const float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = __fdiv_rn(0.25f * 3.14159265358979323846f * rp, r);
    if (q13) phi = 0.5f * 3.14159265358979323846f - phi;
    if (zero) phi = 0.f;
    float sn, cs;
    hf_sincos_exact(phi, sn, cs);
    const float px = r * cs, py = r * sn;
    const float z2 = 1.f - fmaf(py, py, px * px);
    return mk3(px, py, __fsqrt_rn(fmaxf(z2, 0.f)));
}

// One camera sample through up to A.max_depth vertices. Returns the radiance; fills V[0..nv).
__device__ __forceinline__ float hf_path_sample(const HfParams &P, const HfGrid &G, const HfRenderArgs &A, f3 dw, float t_start, int L_start,
                                                uint32_t pid, uint32_t s, HfVertex *V, int &nv, unsigned &nseg) {
    const f3 l = mk3(A.lx, A.ly, A.lz);
    const unsigned long long kk = (unsigned long long) pid * A.cam.spp + s;
    const uint32_t key = (uint32_t) kk ^ ((uint32_t) (kk >> 32) * 0x85EBCA6Bu);
    f3 o = A.o_obj, d = xform_vector(P.to_object, dw);
    float maxt = HF_INF, beta = 1.f, L = 0.f;
    nv = 0;
    #pragma unroll 1
    for (int k = 0; k < A.max_depth; ++k) {
        float ht, hu, hv; uint32_t hprim;
        f3 p0, p1, p2;
        ++nseg;
        if (!hf_trace(P, G, o, d, maxt, k == 0 ? t_start : 0.f, k == 0 ? L_start : HF_START_LEVEL, ht, hu, hv, hprim, p0, p1, p2)) break;
        const f3 nw = hf_world_normal(P, p0, p1, p2);
        if (!(dot3(nw, neg3(dw)) > 0.f)) break;                      // back side: the one-sided BSDF is zero
        const float cosl = dot3(nw, l);
        bool lit = cosl > 0.f;
        const float b1 = hu, b2 = hv, b0 = 1.f - b1 - b2;            // mesh.cpp:738-748
        const f3 po = mk3(fmaf(p0.x, b0, fmaf(p1.x, b1, p2.x * b2)), fmaf(p0.y, b0, fmaf(p1.y, b1, p2.y * b2)),
                          fmaf(p0.z, b0, fmaf(p1.z, b1, p2.z * b2)));
        const f3 pw = xform_point(P.to_world, po);
        const float mag0 = (1.f + fmaxf(fmaxf(fabsf(pw.x), fabsf(pw.y)), fabsf(pw.z))) * 8.940696716308594e-05f;
        if (lit && A.shadow) {
            const float mag = mulsign_(mag0, cosl);
            const f3 os = mk3(fmaf(mag, nw.x, pw.x), fmaf(mag, nw.y, pw.y), fmaf(mag, nw.z, pw.z));
            ++nseg;
            if (hf_occluded(P, G, xform_point(P.to_object, os), xform_vector(P.to_object, l), 3.4028234663852886e+38f)) lit = false;
        }
        HfVertex v;
        v.prim = hprim; v.lit = lit ? 1.f : 0.f; v.ld = lit ? A.scale * cosl : 0.f;
        v.wx = v.wy = v.wz = 0.f; v.cwo = 0.f;
        L = fmaf(beta, v.ld, L);
        bool go_on = k + 1 < A.max_depth;
        f3 wo = mk3(0.f, 0.f, 0.f);
        if (go_on) {
            const float u1 = hf_hash_uniform(A.cam.seed, key, 2u + 2u * (uint32_t) k), u2 = hf_hash_uniform(A.cam.seed, key, 3u + 2u * (uint32_t) k);
            const f3 loc = hf_cosine_hemisphere(u1, u2);
            f3 fs, ft;
            coordinate_system(nw, fs, ft);                           // Frame3f(n), frame.h:24-41
            wo = mk3(fmaf(nw.x, loc.z, fmaf(ft.x, loc.y, fs.x * loc.x)), fmaf(nw.y, loc.z, fmaf(ft.y, loc.y, fs.y * loc.x)),
                     fmaf(nw.z, loc.z, fmaf(ft.z, loc.y, fs.z * loc.x)));
            const float cwo = dot3(nw, wo);
            go_on = cwo > 0.f;
            if (go_on) { v.wx = wo.x; v.wy = wo.y; v.wz = wo.z; v.cwo = cwo; }
        }
        V[nv++] = v;
        if (!go_on) break;
        beta = beta * A.albedo;
        const float mag = mulsign_(mag0, v.cwo);                     // si.spawn_ray(wo), interaction.h:136-165
        const f3 os = mk3(fmaf(mag, nw.x, pw.x), fmaf(mag, nw.y, pw.y), fmaf(mag, nw.z, pw.z));
        o = xform_point(P.to_object, os); d = xform_vector(P.to_object, wo); dw = wo;
        maxt = 3.4028234663852886e+38f;
    }
    return L;
}

// Bundle of a pixel: every sub-pixel position of pixel (x, y). The un-normalised direction v = fwd + sx right + sy up is affine
// in (sx, sy), so its object-space components are bounded by the four corners, and the march runs on o + s v. The rays
// themselves use d = v / |v|: at parameter t they stand at s = t / |v|, so with |v| >= vmin over the pixel
// (|v|^2 = 1 + sx^2 + sy^2 for an orthonormal frame) every ray is still inside the marched region at t = s_enter * vmin.
// hf_api.cu rejects frames that are not orthonormal to 1e-4; the caller backs the result off by 2e-4 for that.
// (a0, a1) x (b0, b1): the part of the pixel's jitter square the bundle covers (the whole square by default).
__device__ __forceinline__ HfBeam hf_pixel_bundle(const HfParams &P, const HfRenderArgs &A, uint32_t pid, float &vmin,
                                                  float a0 = 0.f, float a1 = 1.f, float b0 = 0.f, float b1 = 1.f) {
    const uint32_t x = pid % A.cam.width, y = pid / A.cam.width;
    const float e = 1e-3f;                                   // 1/1000 pixel of slack around the jitter rectangle
    const float fx0 = ((float) x + (a0 - e)) * A.inv_w, fx1 = ((float) x + (a1 + e)) * A.inv_w;
    const float fy0 = ((float) y + (b0 - e)) * A.inv_h, fy1 = ((float) y + (b1 + e)) * A.inv_h;
    {   // smallest |v| over the pixel
        const float sxa = fmaf(2.f, fx0, -1.f) * A.tan_x, sxb = fmaf(2.f, fx1, -1.f) * A.tan_x;
        const float sya = fmaf(-2.f, fy0, 1.f) * A.tan_y, syb = fmaf(-2.f, fy1, 1.f) * A.tan_y;
        const float mx = (sxa <= 0.f && sxb >= 0.f) ? 0.f : fminf(sxa * sxa, sxb * sxb);
        const float my = (sya >= 0.f && syb <= 0.f) ? 0.f : fminf(sya * sya, syb * syb);
        vmin = sqrtf(1.f + mx + my);
    }
    HfBeam B;
    B.ogx = (A.o_obj.x - P.x0) * P.sgx; B.ogy = (A.o_obj.y - P.y0) * P.sgy; B.oz = A.o_obj.z;
    B.dxl = B.dyl = B.dzl = HF_INF; B.dxh = B.dyh = B.dzh = -HF_INF;
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        const f3 v = xform_vector(P.to_object, hf_film_dir(A, (k & 1) ? fx1 : fx0, (k & 2) ? fy1 : fy0));
        const float gx = v.x * P.sgx, gy = v.y * P.sgy;
        B.dxl = fminf(B.dxl, gx); B.dxh = fmaxf(B.dxh, gx);
        B.dyl = fminf(B.dyl, gy); B.dyh = fmaxf(B.dyh, gy);
        B.dzl = fminf(B.dzl, v.z); B.dzh = fmaxf(B.dzh, v.z);
    }
    return B;
}

// ---- the render kernel ----------------------------------------------------------------------------------------------
enum { HF_RENDER_FWD = 0, HF_RENDER_STEP_L2 = 1, HF_RENDER_BWD = 2 };

// LOW: the image has fewer than 32 spp (several pixels per round, tiles of up to 32 pixels); otherwise the packing is constant-folded
// ADAPT (>= 32 spp only; hf_set_adaptive_schedule, see hf_intersect_fast_kernel): the pixels the previous launch found far more
// expensive than the rest go out first, one per fetch, and are skipped when their turn comes in the natural order.
template <int MODE, bool LOW, bool ADAPT>
__global__ void __launch_bounds__(HF_BLOCK, 3)
hf_render_kernel(const __grid_constant__ HfParams P, const HfLaunchInfo li, const __grid_constant__ HfRenderArgs A,
                 float *__restrict__ image, uint32_t *__restrict__ prim_out, const float *__restrict__ target,
                 const float *__restrict__ adj_image, float *__restrict__ grad, double *__restrict__ loss,
                 unsigned long long *__restrict__ sched) {
    static_assert(!(ADAPT && LOW), "the adaptive schedule's unit is one pixel per round");
    __shared__ uint32_t s_prim[HF_BLOCK / 32][HF_RENDER_MAX_SPP];
    __shared__ unsigned s_nheavy, s_c0[HF_BLOCK / 32], s_begin[HF_BLOCK / 32], s_skip[HF_BLOCK / 32];
    __shared__ uint32_t s_coarse[256];          // one bit per 2^heavy_shift pixel positions: clear = no listed pixel there, no global lookup
    unsigned nheavy = 0u;
    if (ADAPT) {
        nheavy = min(*li.heavy_n, li.heavy_cap);
        s_coarse[threadIdx.x & 255u] = 0u;
        __syncthreads();
        for (unsigned k = threadIdx.x; k < nheavy; k += HF_BLOCK) {
            const unsigned b = (li.heavy_list[k] >> li.heavy_shift) & 8191u;
            atomicOr(&s_coarse[b >> 5], 1u << (b & 31u));
        }
        if (threadIdx.x == 0) s_nheavy = nheavy;
    }
    HfGrid G;
    hf_stage_pyramid<true>(P, li, G);           // (ends with a CTA barrier)

    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t spp = A.cam.spp;
    const uint32_t rounds = (spp + 31u) >> 5;
    uint32_t seg = 32u;                                    // lanes per pixel in a round, pixels per round (phase C)
    if (LOW) while (seg > 1u && (seg >> 1) >= spp) seg >>= 1;
    const uint32_t ppr = 32u / seg;
    // Tile sizes in pixels: a full tile is HF_RENDER_TILE rounds of rays (HF_RENDER_TILE pixels at >= 32 spp, HF_RENDER_TILE * ppr below,
    // 32 at most: one pixel per marching lane), the closing phases hand out half and quarter tiles.
    const unsigned u4 = min((unsigned) HF_RENDER_TILE * ppr, 32u), u2 = max(u4 / 2u, 1u), u1 = max(u4 / 4u, 1u);
    // closing phases of the schedule: the last HF_RENDER_TAIL quarter tiles per resident warp go out one at a time and the
    // HF_RENDER_MIDTAIL quarter tiles per warp before them in pairs, so that every phase outlasts the slowest tile of the one before
    // (a pixel whose rays skim down a slope takes ~0.2 ms, and such pixels are neighbours) whatever the size of the pixel list
    const size_t nwarps = (size_t) gridDim.x * (HF_BLOCK / 32);
    const size_t tailpx = min(A.npix, nwarps * (size_t) A.tail * u1);
    const size_t midpx = min(A.npix - tailpx, nwarps * (size_t) A.midtail * u1);
    const size_t nbig = (A.npix - tailpx - midpx) / u4, rest0 = nbig * u4;
    const size_t npair = (A.npix - tailpx - rest0) / u2, rest1 = rest0 + npair * u2;
    double loss_acc = 0.0;
    if (ADAPT && lane == 0) s_c0[warp] = s_begin[warp] = (unsigned) clock();

    for (;;) {
        unsigned long long tile = 0;
        if (lane == 0) tile = atomicAdd(sched, 1ull);
        tile = __shfl_sync(FULL_MASK, tile, 0);
        size_t k0;
        unsigned cap;
        if (ADAPT && lane == 0) s_skip[warp] = 0u;         // pixels of this tile that went out first (bit j); (the shuffle above has gathered the warp)
        if (ADAPT && tile < s_nheavy) {
            k0 = li.heavy_list[tile]; cap = 1u;
            if (k0 >= A.npix) continue;         // a list made for a longer pixel list of this context
        } else {
            if (ADAPT) tile -= s_nheavy;
            if (tile < nbig) { k0 = (size_t) tile * u4; cap = u4; }
            else if (tile < nbig + npair) { k0 = rest0 + (size_t) (tile - nbig) * u2; cap = u2; }
            else { k0 = rest1 + (size_t) (tile - nbig - npair) * u1; cap = u1; }
            if (k0 >= A.npix) break;
            if (ADAPT && s_nheavy) {
                unsigned skip = 0u;
                for (unsigned j = 0; j < cap; ++j) {
                    const unsigned k = (unsigned) k0 + j, b = (k >> li.heavy_shift) & 8191u;
                    if (k < A.npix && ((s_coarse[b >> 5] >> (b & 31u)) & 1u) && ((li.heavy_bits[k >> 5] >> (k & 31u)) & 1u)) skip |= 1u << j;
                }
                if (skip == (1u << cap) - 1u) continue;
                if (lane == 0) s_skip[warp] = skip;
            }
        }
        if (ADAPT) __syncwarp();
        // ---- phase B: all 32 lanes march: lane = pixel * nsub + sub-rectangle of the pixel's jitter square (sub_x x sub_y; 8 bundles
        // per pixel when the tile has <= 4 pixels, 4 / 2 / 1 for 8 / 16 / 32 pixels). A narrower bundle is marched further down the
        // pyramid, so its samples start their own walks closer to the surface.
        const unsigned nsub = LOW ? min(8u, 32u / cap) : 8u, sub_x = nsub == 8u ? 4u : (nsub >= 2u ? 2u : 1u), sub_y = nsub / sub_x;
        const float rsx = 1.f / (float) sub_x, rsy = 1.f / (float) sub_y;      // 1, 1/2, 1/4: exact
        const size_t kq = k0 + lane / nsub;
        const bool have = lane / nsub < cap && kq < A.npix && !(ADAPT && ((s_skip[warp] >> (lane / nsub)) & 1u));
        const uint32_t my_pid = have ? (A.pixels ? __ldg(A.pixels + kq) : (uint32_t) kq) : 0u;
        float t_enter = 0.f;
        int L_start = HF_START_LEVEL;
        if (have) {
            const unsigned sub = lane % nsub, sbx = sub % sub_x, sby = sub / sub_x;
            float vmin;
            const HfBeam B = hf_pixel_bundle(P, A, my_pid, vmin, (float) sbx * rsx, (float) (sbx + 1) * rsx, (float) sby * rsy, (float) (sby + 1) * rsy);
            t_enter = hf_beam_enter(G, B, L_start);
            if (t_enter < HF_BEAM_ALL_MISS) t_enter = t_enter * vmin * (1.f - 2e-4f);
        }
        __syncwarp();
        // ---- phase C: pixel by pixel; below 32 spp several pixels share a round: `seg` lanes per pixel (spp rounded up to a power
        // of two), 32 / seg pixels at a time, so that a 1 / 2 / 4 / 8 / 16 spp image still fills the warp
        const unsigned npx = (unsigned) min((size_t) cap, A.npix - k0);
        for (unsigned j0 = 0; j0 < npx; j0 += ppr) {
            if (ADAPT && ((s_skip[warp] >> j0) & 1u)) continue;
            const unsigned j = j0 + lane / seg;
            const bool pix_ok = j < npx;
            const unsigned jj = pix_ok ? j : j0;
            const unsigned sl = lane & (seg - 1u);                               // this lane's sample slot within its pixel
            const unsigned segmask = seg == 32u ? FULL_MASK : (((1u << seg) - 1u) << (lane & ~(seg - 1u)));
            const uint32_t pid = __shfl_sync(FULL_MASK, my_pid, jj * nsub);
            const uint32_t px = pid % A.cam.width, py = pid / A.cam.width;       // once per pixel, not per sample
            float acc = 0.f;
            for (uint32_t r = 0; r < rounds; ++r) {
                const uint32_t s = r * 32u + sl;
                const bool on = pix_ok && s < spp;
                uint32_t prim = HF_NO_PRIM;
                float v = 0.f;
                float jx, jy;
                hf_camera_jitter(A, pid, on ? s : 0u, jx, jy);
                // the sub-rectangle this sample's jitter falls into (jx, jy in [0, 1): the products are exact)
                const unsigned src = jj * nsub + min((unsigned) (jy * (float) sub_y), sub_y - 1u) * sub_x
                                     + min((unsigned) (jx * (float) sub_x), sub_x - 1u);
                const float ts = __shfl_sync(FULL_MASK, t_enter, src);
                const int ls = __shfl_sync(FULL_MASK, L_start, src);
                const unsigned conv = __ballot_sync(FULL_MASK, on);
                if (on) v = hf_shade_sample(P, G, A, hf_camera_dir_xy(A, px, py, jx, jy), ts, ls, prim, conv);
                acc += v;
                s_prim[warp][(r * 32u + lane) & (HF_RENDER_MAX_SPP - 1)] = prim;
                if (MODE == HF_RENDER_FWD && prim_out && on)
                    __stcs(prim_out + (k0 + j) * spp + s, prim);
            }
            // pixel mean: lane partials in sample order, then a butterfly over the pixel's lanes -- the same order on every run,
            // and the same sums as a 32-lane butterfly whose other lanes hold +0
            #pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) {
                const float other = __shfl_xor_sync(FULL_MASK, acc, sft);
                if ((unsigned) sft < seg) acc += other;
            }
            const float I = acc * A.inv_spp;
            const bool lead = pix_ok && sl == 0u;                                // one lane per pixel reports it
            if (MODE == HF_RENDER_FWD) {
                if (lead) image[pid] = I;
                if (ADAPT && lane == 0) {
                    const unsigned now = (unsigned) clock();
                    li.grp_cost[k0 + j0] = (uint16_t) min((now - s_c0[warp]) >> 6, 65535u);
                    s_c0[warp] = now;
                }
                continue;
            }
            float a_v;                                     // adjoint of ONE sample value of this lane's pixel
            if (MODE == HF_RENDER_STEP_L2) {
                const float diff = I - __ldg(target + pid);
                if (lead) {
                    if (image) image[pid] = I;
                    loss_acc += (double) diff * (double) diff;
                }
                a_v = (2.f * diff * A.inv_npix) * A.inv_spp;
            } else {
                a_v = __ldg(adj_image + pid) * A.inv_spp;
            }
            if (!pix_ok) a_v = 0.f;
            __syncwarp();
            if (__any_sync(FULL_MASK, a_v != 0.f)) {
                for (uint32_t r = 0; r < rounds; ++r) {
                    const uint32_t prim = s_prim[warp][(r * 32u + lane) & (HF_RENDER_MAX_SPP - 1)];
                    const bool lit = prim != HF_NO_PRIM && a_v != 0.f;
                    const unsigned act = __ballot_sync(FULL_MASK, lit);
                    if (lit) {
                        const unsigned peers = __match_any_sync(act, prim) & segmask;     // same facet AND same pixel (same a_v)
                        uint32_t i0, i1, i2; float g0, g1, g2;
                        hf_shade_adjoint(P, A, prim, a_v, i0, i1, i2, g0, g1, g2);
                        if ((peers & ((1u << lane) - 1u)) == 0u) {            // lowest lane of the group
                            const float cnt = (float) __popc(peers);
                            hf_grad_add(P, grad, i0, cnt * g0); hf_grad_add(P, grad, i1, cnt * g1); hf_grad_add(P, grad, i2, cnt * g2);
                        }
                    }
                }
            }
            __syncwarp();
            if (ADAPT && lane == 0) {            // the pixel's cost: the time since this warp finished the pixel before it
                const unsigned now = (unsigned) clock();
                li.grp_cost[k0 + j0] = (uint16_t) min((now - s_c0[warp]) >> 6, 65535u);
                s_c0[warp] = now;
            }
        }
    }
    if (ADAPT && lane == 0) atomicAdd(li.cost_sum, (unsigned long long) (((unsigned) clock() - s_begin[warp]) >> 6));
    if (MODE == HF_RENDER_STEP_L2 && loss && loss_acc != 0.0) atomicAdd(loss, loss_acc * (double) A.inv_npix);
}

// d/d(heights) of dot(a_nw, n_w(prim)) for an arbitrary world-space adjoint a_nw of the facet normal
__device__ __forceinline__ void hf_normal_adjoint(const HfParams &P, uint32_t prim, f3 a_nw, uint32_t &i0, uint32_t &i1, uint32_t &i2,
                                                  float &g0, float &g1, float &g2) {
    f3 p0, p1, p2;
    cell_triangle(P, prim, p0, p1, p2, i0, i1, i2);
    const f3 e1 = sub3(p1, p0), e2 = sub3(p2, p0);
    const f3 c = cross3(e1, e2);
    if (P.flip_normals) a_nw = neg3(a_nw);
    const float inv_clen = rsqrtf(dot3(c, c));
    const f3 no = mk3(c.x * inv_clen, c.y * inv_clen, c.z * inv_clen);
    const f3 m = lin_T(P.to_object, no);
    const float inv_mlen = rsqrtf(dot3(m, m));
    const f3 nw = mk3(m.x * inv_mlen, m.y * inv_mlen, m.z * inv_mlen);
    const float dn = dot3(nw, a_nw);
    const f3 a_m = mk3((a_nw.x - nw.x * dn) * inv_mlen, (a_nw.y - nw.y * dn) * inv_mlen, (a_nw.z - nw.z * dn) * inv_mlen);
    const f3 a_no = lin(P.to_object, a_m);
    const float dno = dot3(no, a_no);
    const f3 a_c = mk3((a_no.x - no.x * dno) * inv_clen, (a_no.y - no.y * dno) * inv_clen, (a_no.z - no.z * dno) * inv_clen);
    const f3 a_e1 = cross3(e2, a_c), a_e2 = cross3(a_c, e1);
    g0 = P.mh * -(a_e1.z + a_e2.z);
    g1 = P.mh * a_e1.z;
    g2 = P.mh * a_e2.z;
}

// Path-traced image (BWD = false) or its adjoint by replay (BWD = true: the pixel adjoint is read from adj_image, every
// vertex of every path scatters into grad). Same tiling as hf_render_kernel; only the primary segment uses the bundle entry.
template <bool BWD>
#ifndef HF_PATH_MIN_CTAS
#define HF_PATH_MIN_CTAS 3             // 80 registers (68 bytes of spills): the walk waits on L2 at 10 of 32 lanes, resident warps pay -- C4 step 38.9 (2) / 35.7 (3) / 37.2 ms (4)
#endif
__global__ void __launch_bounds__(HF_BLOCK, HF_PATH_MIN_CTAS)
hf_render_path_kernel(const __grid_constant__ HfParams P, const HfLaunchInfo li, const __grid_constant__ HfRenderArgs A,
                      float *__restrict__ image, const float *__restrict__ adj_image, float *__restrict__ grad,
                      unsigned long long *__restrict__ sched) {
    HfGrid G;
    hf_stage_pyramid<true>(P, li, G);
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t spp = A.cam.spp, rounds = (spp + 31u) >> 5;
    uint32_t seg = 32u;                                    // as in hf_render_kernel: lanes per pixel, pixels per round, pixels per tile
    while (seg > 1u && (seg >> 1) >= spp) seg >>= 1;
    const uint32_t ppr = 32u / seg;
    const unsigned T = min((unsigned) HF_RENDER_TILE * ppr, 32u);
    const size_t ntiles = (A.npix + T - 1) / T;
    unsigned nseg = 0;                                     // ray segments this thread traced (statistics: hf_stats.segments)
    for (;;) {
        unsigned long long tile = 0;
        if (lane == 0) tile = atomicAdd(sched, 1ull);
        tile = __shfl_sync(FULL_MASK, tile, 0);
        if (tile >= ntiles) {
            nseg = __reduce_add_sync(FULL_MASK, nseg);
            if (lane == 0 && li.seg_counter && nseg) atomicAdd(li.seg_counter, (unsigned long long) nseg);
            break;
        }
        const size_t kq = (size_t) tile * T + lane;
        const bool have = lane < T && kq < A.npix;
        const uint32_t my_pid = have ? (A.pixels ? __ldg(A.pixels + kq) : (uint32_t) kq) : 0u;
        float t_enter = 0.f;
        int L_start = HF_START_LEVEL;
        if (have) {
            float vmin;
            const HfBeam B = hf_pixel_bundle(P, A, my_pid, vmin);
            t_enter = hf_beam_enter(G, B, L_start);
            if (t_enter < HF_BEAM_ALL_MISS) t_enter = t_enter * vmin * (1.f - 2e-4f);
        }
        __syncwarp();
        const unsigned npx = (unsigned) min((size_t) T, A.npix - (size_t) tile * T);
        for (unsigned j0 = 0; j0 < npx; j0 += ppr) {
            const unsigned j = j0 + lane / seg;
            const bool pix_ok = j < npx;
            const unsigned jj = pix_ok ? j : j0;
            const uint32_t pid = __shfl_sync(FULL_MASK, my_pid, jj);
            const float ts = __shfl_sync(FULL_MASK, t_enter, jj);
            const int ls = __shfl_sync(FULL_MASK, L_start, jj);
            const float a_v = BWD && pix_ok ? __ldg(adj_image + pid) * A.inv_spp : 0.f;
            float acc = 0.f;
            for (uint32_t r = 0; r < rounds; ++r) {
                const uint32_t s = r * 32u + (lane & (seg - 1u));
                HfVertex V[HF_MAX_DEPTH];
                int nv = 0;
                if (pix_ok && s < spp) acc += hf_path_sample(P, G, A, hf_camera_dir(A, pid, s), ts, ls, pid, s, V, nv, nseg);
                if (BWD) {
                    // radiance arriving at vertex k from the later vertices, back to front: Lo_k = ld_k + albedo * Lo_{k+1}
                    float lo_next[HF_MAX_DEPTH];
                    float run = 0.f;
                    for (int k = HF_MAX_DEPTH - 1; k >= 0; --k) {
                        lo_next[k] = run;
                        if (k < nv) run = fmaf(A.albedo, run, V[k].ld);
                    }
                    float beta = 1.f;
                    for (int k = 0; k < HF_MAX_DEPTH; ++k) {            // uniform trip count: the scatter is a warp-wide operation
                        const bool valid = (k < nv) && (a_v != 0.f);
                        uint32_t i0 = 0, i1 = 0, i2 = 0; float g0 = 0.f, g1 = 0.f, g2 = 0.f;
                        if (valid) {
                            const HfVertex v = V[k];
                            const float kd = a_v * beta * v.lit * A.scale;                                   // direct term
                            const float kw = v.cwo > 0.f ? __fdividef(a_v * beta * A.albedo * lo_next[k], v.cwo) : 0.f;   // BSDF weight of the bounce
                            hf_normal_adjoint(P, v.prim, mk3(kd * A.lx + kw * v.wx, kd * A.ly + kw * v.wy, kd * A.lz + kw * v.wz),
                                              i0, i1, i2, g0, g1, g2);
                        }
                        warp_aggregated_add(P, grad, i0, g0, valid);
                        warp_aggregated_add(P, grad, i1, g1, valid);
                        warp_aggregated_add(P, grad, i2, g2, valid);
                        beta *= A.albedo;
                    }
                }
            }
            if (!BWD) {
                #pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1) {
                    const float other = __shfl_xor_sync(FULL_MASK, acc, sft);
                    if ((unsigned) sft < seg) acc += other;
                }
                if (pix_ok && (lane & (seg - 1u)) == 0u) image[pid] = acc * A.inv_spp;
            }
        }
    }
}

// adjoint image and loss of the L2 objective for the listed pixels: adj = 2 (I - T) / (W H), loss += sum (I - T)^2 / (W H)
__global__ void __launch_bounds__(HF_BLOCK)
hf_l2_adjoint_kernel(const __grid_constant__ HfRenderArgs A, const float *__restrict__ image, const float *__restrict__ target,
                     float *__restrict__ adj, double *__restrict__ loss) {
    double acc = 0.0;
    for (size_t k = (size_t) blockIdx.x * HF_BLOCK + threadIdx.x; k < A.npix; k += (size_t) gridDim.x * HF_BLOCK) {
        const uint32_t pid = A.pixels ? __ldg(A.pixels + k) : (uint32_t) k;
        const float diff = image[pid] - target[pid];
        adj[pid] = 2.f * diff * A.inv_npix;
        acc += (double) diff * (double) diff;
    }
    #pragma unroll
    for (int sft = 16; sft > 0; sft >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, sft);
    if (loss && (threadIdx.x & 31) == 0 && acc != 0.0) atomicAdd(loss, acc * (double) A.inv_npix);
}

// Backward from stored hit triangles (no re-trace): sample i = k * spp + s of listed pixel k
__global__ void __launch_bounds__(HF_BLOCK)
hf_render_backward_prim_kernel(const __grid_constant__ HfParams P, const __grid_constant__ HfRenderArgs A,
                               const float *__restrict__ adj_image, const uint32_t *__restrict__ prim_in, float *__restrict__ grad) {
    const size_t n = A.npix * (size_t) A.cam.spp;
    const size_t ntiles = (n + HF_BLOCK - 1) / HF_BLOCK;
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const size_t i = tile * HF_BLOCK + threadIdx.x;
        uint32_t prim = HF_NO_PRIM;
        float a_v = 0.f;
        if (i < n) {
            prim = __ldcs(prim_in + i);
            const size_t k = i / A.cam.spp;
            const uint32_t pid = A.pixels ? __ldg(A.pixels + k) : (uint32_t) k;
            a_v = __ldg(adj_image + pid) * A.inv_spp;
        }
        const bool valid = prim != HF_NO_PRIM && a_v != 0.f;
        uint32_t i0 = 0, i1 = 0, i2 = 0; float g0 = 0.f, g1 = 0.f, g2 = 0.f;
        if (valid) hf_shade_adjoint(P, A, prim, a_v, i0, i1, i2, g0, g1, g2);
        warp_aggregated_add(P, grad, i0, g0, valid);
        warp_aggregated_add(P, grad, i1, g1, valid);
        warp_aggregated_add(P, grad, i2, g2, valid);
    }
}

// The generated rays as an SoA (hf_render_rays)
__global__ void __launch_bounds__(HF_BLOCK)
hf_render_rays_kernel(const __grid_constant__ HfRenderArgs A, float *ox, float *oy, float *oz, float *dx, float *dy, float *dz, float *maxt) {
    const size_t n = A.npix * (size_t) A.cam.spp;
    for (size_t i = (size_t) blockIdx.x * HF_BLOCK + threadIdx.x; i < n; i += (size_t) gridDim.x * HF_BLOCK) {
        const size_t k = i / A.cam.spp;
        const uint32_t s = (uint32_t) (i - k * A.cam.spp);
        const uint32_t pid = A.pixels ? __ldg(A.pixels + k) : (uint32_t) k;
        const f3 d = hf_camera_dir(A, pid, s);
        ox[i] = A.cam.origin[0]; oy[i] = A.cam.origin[1]; oz[i] = A.cam.origin[2];
        dx[i] = d.x; dy[i] = d.y; dz[i] = d.z; maxt[i] = HF_INF;
    }
}

// ---- host side --------------------------------------------------------------------------------------------------------
static HfRenderArgs hf_render_args(const HfParams &P, const hf_camera &cam, const hf_light *light, const uint32_t *d_pixels, size_t npix) {
    HfRenderArgs A;
    memset(&A, 0, sizeof(A));
    A.cam = cam;
    if (light) {
        A.lx = light->direction[0]; A.ly = light->direction[1]; A.lz = light->direction[2];
        A.scale = (float) ((double) light->albedo / 3.14159265358979323846 * (double) light->irradiance);
        A.shadow = light->shadow_rays ? 1 : 0;
        A.albedo = light->albedo;
        A.max_depth = light->max_depth < 1 ? 1 : light->max_depth;
    }
    A.inv_w = 1.0f / (float) cam.width; A.inv_h = 1.0f / (float) cam.height;
    A.tan_x = cam.tan_half_fov;
    A.tan_y = cam.tan_half_fov * ((float) cam.height / (float) cam.width);
    A.inv_spp = 1.0f / (float) cam.spp;
    A.inv_npix = 1.0f / ((float) cam.width * (float) cam.height);
    const float *m = P.to_object;                 // xform_point, same op order as the device
    float r[3];
    for (int a = 0; a < 3; ++a) {
        float v = m[4 * a + 3];
        v = fmaf(m[4 * a + 0], cam.origin[0], v); v = fmaf(m[4 * a + 1], cam.origin[1], v); v = fmaf(m[4 * a + 2], cam.origin[2], v);
        r[a] = v;
    }
    A.o_obj.x = r[0]; A.o_obj.y = r[1]; A.o_obj.z = r[2];
    A.pixels = d_pixels; A.npix = npix;
    static const int env_tail = getenv("HF_B200_RTAIL") ? atoi(getenv("HF_B200_RTAIL")) : -1;          // dev knobs
    static const int env_mid = getenv("HF_B200_RMIDTAIL") ? atoi(getenv("HF_B200_RMIDTAIL")) : -1;
    A.tail = env_tail >= 0 ? env_tail : HF_RENDER_TAIL; A.midtail = env_mid >= 0 ? env_mid : HF_RENDER_MIDTAIL;
    auto lowbias32 = [](uint32_t x) { x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16; return x; };
    A.seed_h0 = lowbias32(cam.seed * 0x9E3779B1u + 0u * 0x85EBCA6Bu);
    A.seed_h1 = lowbias32(cam.seed * 0x9E3779B1u + 1u * 0x85EBCA6Bu);
    return A;
}

static int hf_render_grid(const HfLaunchInfo &li, size_t npix) {
    const size_t ntiles = (npix + HF_RENDER_TILE - 1) / HF_RENDER_TILE, ncta = (ntiles + HF_BLOCK / 32 - 1) / (HF_BLOCK / 32), cap = (size_t) li.num_sms * 3;
    return (int) (ncta < cap ? (ncta ? ncta : 1) : cap);
}

cudaError_t hf_launch_render(const HfParams &P, const HfLaunchInfo &li, int mode, const hf_camera &cam, const hf_light &light,
                             const uint32_t *d_pixels, size_t npix, float *d_image, uint32_t *d_prim, const float *d_target,
                             const float *d_adj_image, float *d_grad, double *d_loss, unsigned long long *d_sched, cudaStream_t s) {
    if (npix == 0) return cudaSuccess;
    HfRenderArgs A = hf_render_args(P, cam, &light, d_pixels, npix);
    cudaError_t e = cudaMemsetAsync(d_sched, 0, sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    const size_t smem = hf_walk_smem(li, true);
    const int grid = hf_render_grid(li, npix);
    if (A.max_depth > 1) {                    // path tracing: forward image or replayed adjoint (the fused L2 step is composed by hf_api.cu)
        const int pgrid = grid < li.num_sms * HF_PATH_MIN_CTAS ? grid : li.num_sms * HF_PATH_MIN_CTAS;
        if (mode == HF_RENDER_FWD)
            hf_render_path_kernel<false><<<pgrid, HF_BLOCK, smem, s>>>(P, li, A, d_image, nullptr, nullptr, d_sched);
        else if (mode == HF_RENDER_BWD)
            hf_render_path_kernel<true><<<pgrid, HF_BLOCK, smem, s>>>(P, li, A, nullptr, d_adj_image, d_grad, d_sched);
        else
            return cudaErrorInvalidValue;
        return cudaGetLastError();
    }
    const bool low = cam.spp < 32u;
    const bool adapt = !low && li.grp_cost && li.heavy_n && li.heavy_list && li.heavy_bits && li.cost_sum;
    if (adapt && (e = cudaMemsetAsync(li.cost_sum, 0, sizeof(unsigned long long), s)) != cudaSuccess) return e;
    if (adapt) {
        // the ~0.2 ms pixels are no longer waiting in the closing phases: a quarter / half of their length is enough, and more pixels
        // go out in tiles of four (1/8 of C3 on one GPU: kernel 0.70 -> 0.64 ms; 2 / 6 and 12 / 12 are slower again)
        if (!getenv("HF_B200_RTAIL")) A.tail = HF_RENDER_TAIL / 4;
        if (!getenv("HF_B200_RMIDTAIL")) A.midtail = HF_RENDER_MIDTAIL / 2;
    }
#define HF_RENDER_LAUNCH(M, ...) do { \
        if (low) hf_render_kernel<M, true, false><<<grid, HF_BLOCK, smem, s>>>(P, li, A, __VA_ARGS__, d_sched); \
        else if (adapt) hf_render_kernel<M, false, true><<<grid, HF_BLOCK, smem, s>>>(P, li, A, __VA_ARGS__, d_sched); \
        else hf_render_kernel<M, false, false><<<grid, HF_BLOCK, smem, s>>>(P, li, A, __VA_ARGS__, d_sched); } while (0)
    if (mode == HF_RENDER_FWD) HF_RENDER_LAUNCH(HF_RENDER_FWD, d_image, d_prim, nullptr, nullptr, nullptr, nullptr);
    else if (mode == HF_RENDER_STEP_L2) HF_RENDER_LAUNCH(HF_RENDER_STEP_L2, d_image, nullptr, d_target, nullptr, d_grad, d_loss);
    else HF_RENDER_LAUNCH(HF_RENDER_BWD, nullptr, nullptr, nullptr, d_adj_image, d_grad, nullptr);
#undef HF_RENDER_LAUNCH
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (adapt) e = hf_launch_pick_heavy(li, (unsigned) npix, (unsigned) grid * (HF_BLOCK / 32), s);       // the list for the next launch over a list of this length
    return e;
}

cudaError_t hf_launch_render_backward_prim(const HfParams &P, const HfLaunchInfo &li, const hf_camera &cam, const hf_light &light,
                                           const uint32_t *d_pixels, size_t npix, const float *d_adj_image, const uint32_t *d_prim,
                                           float *d_grad, cudaStream_t s) {
    if (npix == 0) return cudaSuccess;
    const HfRenderArgs A = hf_render_args(P, cam, &light, d_pixels, npix);
    const size_t n = npix * (size_t) cam.spp, nt = (n + HF_BLOCK - 1) / HF_BLOCK, cap = (size_t) li.num_sms * 8;
    hf_render_backward_prim_kernel<<<(int) (nt < cap ? nt : cap), HF_BLOCK, 0, s>>>(P, A, d_adj_image, d_prim, d_grad);
    return cudaGetLastError();
}

cudaError_t hf_launch_render_rays(const HfParams &P, const HfLaunchInfo &li, const hf_camera &cam, const uint32_t *d_pixels, size_t npix,
                                  const hf_rays &out, cudaStream_t s) {
    if (npix == 0) return cudaSuccess;
    const HfRenderArgs A = hf_render_args(P, cam, nullptr, d_pixels, npix);
    const size_t n = npix * (size_t) cam.spp, nt = (n + HF_BLOCK - 1) / HF_BLOCK, cap = (size_t) li.num_sms * 8;
    hf_render_rays_kernel<<<(int) (nt < cap ? nt : cap), HF_BLOCK, 0, s>>>(A, (float *) out.ox, (float *) out.oy, (float *) out.oz,
                                                                          (float *) out.dx, (float *) out.dy, (float *) out.dz, (float *) out.maxt);
    return cudaGetLastError();
}

cudaError_t hf_launch_l2_adjoint(const HfParams &P, const HfLaunchInfo &li, const hf_camera &cam, const uint32_t *d_pixels, size_t npix,
                                 const float *d_image, const float *d_target, float *d_adj, double *d_loss, cudaStream_t s) {
    if (npix == 0) return cudaSuccess;
    const HfRenderArgs A = hf_render_args(P, cam, nullptr, d_pixels, npix);
    const size_t nt = (npix + HF_BLOCK - 1) / HF_BLOCK, cap = (size_t) li.num_sms * 8;
    hf_l2_adjoint_kernel<<<(int) (nt < cap ? nt : cap), HF_BLOCK, 0, s>>>(A, d_image, d_target, d_adj, d_loss);
    return cudaGetLastError();
}
