// hf_heightfield.hpp -- header-only C++17 host class over the C ABI of hf_b200.h.
//
// `hf::Heightfield` is the host-side object a Mitsuba 3 `Shape` plugin wraps (INTEGRATION.md section 2 shows the plugin): its methods
// carry the names, the order and the argument meaning of the `Shape` virtuals a `heightfield` plugin provides
// (include/mitsuba/render/shape.h in the reference tree):
//
//     ray_intersect_preliminary   shape.h:137-138      bbox                     shape.h:253
//     ray_test                    shape.h:153          primitive_count          shape.h:432
//     compute_surface_interaction shape.h:179-183      traverse                 shape.h:533
//     ray_intersect               shape.h:195-197      parameters_changed       shape.h:534
//                                                      parameters_grad_enabled  shape.h:550
//
// The reference passes Dr.Jit wavefront arrays (Ray3f, Mask, PreliminaryIntersection3f, SurfaceInteraction3f); here the same records
// are structure-of-arrays views of DEVICE memory -- exactly what `dr::eval()`-ed CUDA arrays are (`.data()` of each component) -- so a
// plugin forwards its arrays without a copy. Errors are C++ exceptions, as in the reference (`Throw(...)`, src/render/shape.cpp:36,
// 102): `hf::Error` carries the status and the library's message. No CUDA header is needed: streams are passed as `void *`
// (a cudaStream_t). Nothing here computes: every method is one call into libhf_b200.so; there is no CPU path.
#ifndef HF_HEIGHTFIELD_HPP
#define HF_HEIGHTFIELD_HPP

#include <cstdint>
#include <cstring>
#include <functional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "hf_b200.h"

namespace hf {

/// What `Throw()` is to the reference (include/mitsuba/core/logger.h): the failed call's status + hf_last_error()
class Error : public std::runtime_error {
public:
    Error(int status, const std::string &what) : std::runtime_error(what), m_status(status) {}
    int status() const { return m_status; }
private:
    int m_status;
};

inline void check(int rc) {
    if (rc != HF_OK) throw Error(rc, std::string("heightfield: ") + hf_last_error());
}

/// RayFlags (include/mitsuba/render/interaction.h:19-69), same bit values
enum class RayFlags : uint32_t {
    Empty = 0x0, Minimal = 0x1, UV = 0x2, dPdUV = 0x4, ShadingFrame = 0x8, FollowShape = 0x80, DetachShape = 0x100,
    All = 0x2 | 0x4 | 0x8, AllNonDifferentiable = 0x2 | 0x4 | 0x8 | 0x100
};
constexpr uint32_t operator+(RayFlags f) { return (uint32_t) f; }            // the reference's unary plus (`+RayFlags::All`)

/// ParamFlags (include/mitsuba/core/object.h:275-284)
enum class ParamFlags : uint32_t { Differentiable = 0x0, NonDifferentiable = 0x1, Discontinuous = 0x2 };
constexpr uint32_t operator|(ParamFlags a, ParamFlags b) { return (uint32_t) a | (uint32_t) b; }
constexpr uint32_t operator+(ParamFlags f) { return (uint32_t) f; }

// ---- records: SoA views of device memory (wavefront of `size` lanes) -------------------------------------------------------------
/// Ray3f (include/mitsuba/core/ray.h): o, d, maxt; `time` and `wavelengths` are not read by geometry
struct Ray3f {
    const float *ox = nullptr, *oy = nullptr, *oz = nullptr, *dx = nullptr, *dy = nullptr, *dz = nullptr, *maxt = nullptr;
    size_t size = 0;
    hf_rays c(const uint8_t *active) const { return hf_rays{ ox, oy, oz, dx, dy, dz, maxt, active }; }
};
/// Mask: one byte per lane, nullptr = all lanes active (the reference's default argument `Mask active = true`)
using Mask = const uint8_t *;

/// PreliminaryIntersection3f (interaction.h:586-691): invalid <=> t == +inf (639-646); shape_index is the constant (uint32)-1 and
/// `shape` = the heightfield itself (src/shapes/rectangle.cpp:222-223)
struct PreliminaryIntersection3f {
    float *t = nullptr, *prim_u = nullptr, *prim_v = nullptr;
    uint32_t *prim_index = nullptr;
    static constexpr uint32_t shape_index = 0xFFFFFFFFu;
    const void *shape = nullptr;
    hf_prelim c() const { return hf_prelim{ t, prim_u, prim_v, prim_index }; }
};

/// SurfaceInteraction3f (interaction.h): world space; null members are not materialised (the kernels skip them)
using SurfaceInteraction3f = hf_si;
/// upstream adjoints of a SurfaceInteraction3f (what the AD tape hands to the reverse-mode sweep of src/render/mesh.cpp:695-789)
using SurfaceInteractionAdjoint3f = hf_si_adj;

struct ScalarBoundingBox3f { float min[3], max[3]; };

/// Properties of the plugin constructor (Shape(props), src/render/shape.cpp:22-78; gridded data as src/textures/bitmap.cpp:124-143)
struct Properties {
    uint32_t width = 0, height = 0;         ///< texels (W, H); the tensor is (H, W) row-major, row = y (bitmap.cpp:259-263)
    const float *heights = nullptr;         ///< H * W floats, host or device memory
    bool heights_on_device = false;
    float max_height = 1.f;
    float to_world[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };    ///< row-major 3x4 affine
    const float *to_object = nullptr;       ///< optional 12 floats (m_to_object of the base class); default: fp64 inverse of to_world
    bool flip_normals = false;
    int device = 0;                         ///< CUDA ordinal
    void *stream = nullptr;                 ///< stream the upload and the first acceleration build are queued on
};

/// TraversalCallback (include/mitsuba/core/object.h:293-340), reduced to what this shape exposes: device pointer + extent + flags
struct TraversalCallback {
    virtual ~TraversalCallback() = default;
    /// "heights": float tensor (height, width) in device memory; "max_height": scalar; "to_world": 12 floats (host)
    virtual void put_parameter(const std::string &name, void *ptr, size_t count, uint32_t flags) = 0;
};

class Heightfield {
public:
    explicit Heightfield(const Properties &props) {
        if (!props.heights) throw Error(HF_ERR_INVALID, "heightfield: `heights` is required");
        hf_desc d;
        std::memset(&d, 0, sizeof(d));
        d.width = props.width; d.height = props.height; d.max_height = props.max_height;
        std::memcpy(d.to_world, props.to_world, sizeof(d.to_world));
        if (props.to_object) { std::memcpy(d.to_object, props.to_object, sizeof(d.to_object)); d.has_to_object = 1; }
        d.flip_normals = props.flip_normals ? 1 : 0;
        d.device = props.device;
        check(hf_create(&m_ctx, &d));
        m_max_height = props.max_height;
        std::memcpy(m_to_world, props.to_world, sizeof(m_to_world));
        m_flip_normals = props.flip_normals;
        try {
            check(hf_set_heights(m_ctx, props.heights, props.heights_on_device ? 1 : 0, props.stream));
            check(hf_parameters_changed(m_ctx, props.stream));
        } catch (...) { hf_destroy(m_ctx); m_ctx = nullptr; throw; }
    }
    ~Heightfield() { if (m_ctx) hf_destroy(m_ctx); }
    Heightfield(const Heightfield &) = delete;
    Heightfield &operator=(const Heightfield &) = delete;
    Heightfield(Heightfield &&o) noexcept { *this = std::move(o); }
    Heightfield &operator=(Heightfield &&o) noexcept {
        if (this != &o) {
            if (m_ctx) hf_destroy(m_ctx);
            m_ctx = o.m_ctx; o.m_ctx = nullptr; m_max_height = o.m_max_height; m_flip_normals = o.m_flip_normals;
            std::memcpy(m_to_world, o.m_to_world, sizeof(m_to_world));
        }
        return *this;
    }

    // =============================================================
    //! @{ \name Ray tracing routines (shape.h:137-197)
    // =============================================================

    /// shape.h:137-138. Fills `pi` (t, prim_uv, prim_index) for the `ray.size` lanes; a miss or an inactive lane gets t = +inf
    void ray_intersect_preliminary(const Ray3f &ray, PreliminaryIntersection3f &pi, Mask active = nullptr, void *stream = nullptr) const {
        const hf_rays r = ray.c(active);
        const hf_prelim p = pi.c();
        check(hf_ray_intersect_preliminary(m_ctx, &r, &p, ray.size, stream));
        pi.shape = this;
    }

    /// shape.h:153. `hit`: one byte per lane (device)
    void ray_test(const Ray3f &ray, uint8_t *hit, Mask active = nullptr, void *stream = nullptr) const {
        const hf_rays r = ray.c(active);
        check(hf_ray_test(m_ctx, &r, hit, ray.size, stream));
    }

    /// shape.h:179-183 (+ finalize_surface_interaction, interaction.h:476-499). recursion_depth > 0 returns at once: the shape is
    /// never an instance (src/render/mesh.cpp:681-682)
    void compute_surface_interaction(const Ray3f &ray, const PreliminaryIntersection3f &pi, SurfaceInteraction3f &si,
                                     uint32_t ray_flags = +RayFlags::All, uint32_t recursion_depth = 0, Mask active = nullptr,
                                     void *stream = nullptr) const {
        if (recursion_depth > 0) return;
        const hf_rays r = ray.c(active);
        const hf_prelim p = pi.c();
        check(hf_compute_surface_interaction(m_ctx, &r, &p, ray_flags, &si, ray.size, stream));
    }

    /// shape.h:195-197: preliminary intersection + surface interaction in ONE launch; `pi` (optional) receives the preliminary record
    void ray_intersect(const Ray3f &ray, SurfaceInteraction3f &si, uint32_t ray_flags = +RayFlags::All, Mask active = nullptr,
                       PreliminaryIntersection3f *pi = nullptr, void *stream = nullptr) const {
        const hf_rays r = ray.c(active);
        if (pi) {
            const hf_prelim p = pi->c();
            check(hf_ray_intersect(m_ctx, &r, ray_flags, &p, &si, ray.size, stream));
            pi->shape = this;
        } else {
            check(hf_ray_intersect(m_ctx, &r, ray_flags, nullptr, &si, ray.size, stream));
        }
    }

    //! @}
    // =============================================================
    //! @{ \name Miscellaneous (shape.h:253, 432)
    // =============================================================

    /// shape.h:253: world-space box of every primitive; refreshed by parameters_changed() (it is the root of the min/max pyramid)
    ScalarBoundingBox3f bbox() const {
        float b[6];
        check(hf_bbox(m_ctx, b));
        return ScalarBoundingBox3f{ { b[0], b[1], b[2] }, { b[3], b[4], b[5] } };
    }

    /// shape.h:432: 2 (W - 1) (H - 1) triangles
    uint32_t primitive_count() const { uint32_t n = 0; check(hf_primitive_count(m_ctx, &n)); return n; }

    //! @}
    // =============================================================
    //! @{ \name Parameters (shape.h:533-550; pattern src/shapes/linearcurve.cpp:364-382)
    // =============================================================

    /// shape.h:533: "heights" (Differentiable | Discontinuous), "max_height", "to_world"
    void traverse(TraversalCallback *callback) {
        uint32_t w = 0, h = 0;
        check(hf_size(m_ctx, &w, &h));
        callback->put_parameter("heights", heights(), (size_t) w * h, ParamFlags::Differentiable | ParamFlags::Discontinuous);
        callback->put_parameter("max_height", &m_max_height, 1, +ParamFlags::NonDifferentiable);
        callback->put_parameter("to_world", m_to_world, 12, +ParamFlags::NonDifferentiable);
    }

    /// shape.h:534: after the caller wrote through the pointers traverse() handed out. Rebuilds the acceleration data (min/max pyramid)
    /// and the bbox; the reference re-commits Embree / rebuilds the OptiX GAS here (src/render/scene.cpp:343-367)
    void parameters_changed(const std::vector<std::string> &keys = {}, void *stream = nullptr) {
        auto has = [&](const char *k) { for (auto &s : keys) if (s == k) return true; return keys.empty(); };
        if (has("max_height")) check(hf_set_max_height(m_ctx, m_max_height));
        if (has("to_world")) check(hf_set_to_world(m_ctx, m_to_world, nullptr));
        if (has("heights") || has("max_height")) check(hf_parameters_changed(m_ctx, stream));
    }

    /// shape.h:550
    bool parameters_grad_enabled() const { int e = 0; check(hf_parameters_grad_enabled(m_ctx, &e)); return e != 0; }
    void set_grad_enabled(bool e) { check(hf_set_grad_enabled(m_ctx, e ? 1 : 0)); }

    /// the `coherent` argument of Scene::ray_intersect (src/render/scene.cpp:141-150) for the ray-array entry points
    void set_ray_coherence(bool coherent) { check(hf_set_ray_coherence(m_ctx, coherent ? 1 : 0)); }
    /// launches over the same rays / pixels, iteration after iteration, start with the units of work the previous one found far more
    /// expensive than the rest (no reference counterpart; results do not depend on it)
    void set_adaptive_schedule(bool on) { check(hf_set_adaptive_schedule(m_ctx, on ? 1 : 0)); }

    //! @}
    // =============================================================
    //! @{ \name Reverse mode (what the reference's AD tape does for this shape) and the optimiser
    // =============================================================

    float *heights() const { float *p = nullptr; check(hf_heights(m_ctx, &p)); return p; }       ///< device, (height, width)
    float *grad() const { float *p = nullptr; check(hf_grad(m_ctx, &p)); return p; }              ///< dr::grad(heights), device
    void zero_grad(void *stream = nullptr) { check(hf_zero_grad(m_ctx, stream)); }

    /// reverse-mode sweep of compute_surface_interaction w.r.t. "heights": grad() += J^T adj (src/render/mesh.cpp:695-789 +
    /// scatter_reduce(Add), ext/drjit/src/autodiff/autodiff.cpp:1309-1328). DetachShape: no gradient; FollowShape honoured
    void backward(const Ray3f &ray, const PreliminaryIntersection3f &pi, const SurfaceInteractionAdjoint3f &adj,
                  uint32_t ray_flags = +RayFlags::All, Mask active = nullptr, void *stream = nullptr) {
        const hf_rays r = ray.c(active);
        const hf_prelim p = pi.c();
        check(hf_backward(m_ctx, &r, &p, &adj, ray_flags, ray.size, stream));
    }

    /// mi.ad.Adam.step() for "heights" (src/python/python/ad/optimizers.py:263-302); call parameters_changed({"heights"}) afterwards
    void adam_step(double lr, double beta_1 = 0.9, double beta_2 = 0.999, double epsilon = 1e-8, bool mask_updates = false,
                   void *stream = nullptr) {
        hf_adam cfg{ lr, beta_1, beta_2, epsilon, mask_updates ? 1 : 0 };
        check(hf_adam_step(m_ctx, &cfg, stream));
    }

    //! @}

    float max_height() const { return m_max_height; }
    bool flip_normals() const { return m_flip_normals; }
    hf_ctx *context() const { return m_ctx; }                    ///< for the entry points this class does not wrap (render step, NCCL)

    std::string to_string() const {
        uint32_t w = 0, h = 0;
        hf_size(m_ctx, &w, &h);
        std::ostringstream oss;
        oss << "Heightfield[\n  size = " << w << "x" << h << ",\n  max_height = " << m_max_height << ",\n  primitive_count = "
            << primitive_count() << ",\n  flip_normals = " << (m_flip_normals ? "true" : "false") << "\n]";
        return oss.str();
    }

private:
    hf_ctx *m_ctx = nullptr;
    float m_max_height = 1.f;
    float m_to_world[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
    bool m_flip_normals = false;
};

}  // namespace hf

#endif  // HF_HEIGHTFIELD_HPP
