// abi_smoke.cpp -- the C++ host class of include/hf_heightfield.hpp (hf::Heightfield, the object a Mitsuba `Shape` plugin wraps)
// driven from plain C++ on a GPU: the reference's known answers for the per-cell primitive on the flat 2 x 2 field, which IS
// resources/data/common/meshes/rectangle.obj under the fixed triangulation (src/render/tests/test_mesh.py:257-292 test10,
// :536-671 test16), bbox / primitive_count / traverse / parameters_changed, and the reference's error behaviour (exceptions).
// Built by __graft_entry__.build() (g++ + libcudart), run by tests/test_abi.py::test_cpp_host_class_on_the_gpu (-m gpu).
// TEST CODE: prints "abi_smoke ok" and exits 0, or the failed expectation and exits 1.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "hf_heightfield.hpp"

#define EXPECT(cond) do { if (!(cond)) { std::fprintf(stderr, "abi_smoke: %s:%d: expectation failed: %s\n", __FILE__, __LINE__, #cond); std::exit(1); } } while (0)
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { std::fprintf(stderr, "abi_smoke: %s: %s\n", #call, cudaGetErrorString(e_)); std::exit(1); } } while (0)

static bool close(float a, float b, float tol = 3e-5f) { return std::fabs(a - b) <= tol; }

template <typename T> struct DevBuf {
    T *p = nullptr; size_t n = 0;
    explicit DevBuf(size_t n_) : n(n_) { CU(cudaMalloc(&p, n * sizeof(T))); CU(cudaMemset(p, 0, n * sizeof(T))); }
    DevBuf(const std::vector<T> &h) : DevBuf(h.size()) { CU(cudaMemcpy(p, h.data(), n * sizeof(T), cudaMemcpyHostToDevice)); }
    ~DevBuf() { cudaFree(p); }
    std::vector<T> host() const { std::vector<T> h(n); CU(cudaMemcpy(h.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost)); return h; }
};

struct Recorder : hf::TraversalCallback {
    std::vector<std::string> names; std::vector<uint32_t> flags; void *heights = nullptr; size_t count = 0;
    void put_parameter(const std::string &name, void *ptr, size_t cnt, uint32_t fl) override {
        names.push_back(name); flags.push_back(fl);
        if (name == "heights") { heights = ptr; count = cnt; }
    }
};

int main() {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { std::fprintf(stderr, "abi_smoke: no CUDA device (the library has no CPU path)\n"); return 2; }

    // ---- constructor errors are exceptions carrying the library's message (Shape(props) Throw()s, src/render/shape.cpp:36, 102)
    {
        hf::Properties bad;
        float one = 0.f;
        bad.width = 1; bad.height = 2; bad.heights = &one;
        bool threw = false;
        try { hf::Heightfield h(bad); } catch (const hf::Error &e) { threw = e.status() == HF_ERR_INVALID && std::string(e.what()).find("2x2") != std::string::npos; }
        EXPECT(threw);
    }

    // ---- flat 2 x 2 field == rectangle.obj
    const float zeros[4] = { 0.f, 0.f, 0.f, 0.f };
    hf::Properties props;
    props.width = 2; props.height = 2; props.heights = zeros; props.max_height = 1.f;
    hf::Heightfield shape(props);
    EXPECT(shape.primitive_count() == 2);
    const hf::ScalarBoundingBox3f bb = shape.bbox();
    EXPECT(close(bb.min[0], -1) && close(bb.min[1], -1) && close(bb.min[2], 0) && close(bb.max[0], 1) && close(bb.max[1], 1) && close(bb.max[2], 0));
    EXPECT(shape.parameters_grad_enabled());
    EXPECT(shape.to_string().find("primitive_count = 2") != std::string::npos);

    // test_mesh.py:257-292 (test10): two rays straight up through the two triangles
    const size_t n = 2;
    using VF = std::vector<float>;
    DevBuf<float> ox(VF{ -0.3f, 0.3f }), oy(VF{ -0.3f, 0.3f }), oz(VF{ -10.f, -10.f }), dx(VF{ 0.f, 0.f }), dy(VF{ 0.f, 0.f }), dz(VF{ 1.f, 1.f }),
        maxt(VF{ INFINITY, INFINITY });
    hf::Ray3f ray;
    ray.ox = ox.p; ray.oy = oy.p; ray.oz = oz.p; ray.dx = dx.p; ray.dy = dy.p; ray.dz = dz.p; ray.maxt = maxt.p; ray.size = n;
    DevBuf<float> t(n), u(n), v(n); DevBuf<uint32_t> prim(n);
    hf::PreliminaryIntersection3f pi;
    pi.t = t.p; pi.prim_u = u.p; pi.prim_v = v.p; pi.prim_index = prim.p;
    shape.ray_intersect_preliminary(ray, pi);
    CU(cudaDeviceSynchronize());
    {
        auto ht = t.host(), hu = u.host(), hv = v.host(); auto hp = prim.host();
        EXPECT(close(ht[0], 10) && close(ht[1], 10));
        EXPECT(hp[0] == 0 && hp[1] == 1);
        EXPECT(close(hu[0], 0.35f) && close(hv[0], 0.3f) && close(hu[1], 0.3f) && close(hv[1], 0.35f));
        EXPECT(pi.shape == &shape && hf::PreliminaryIntersection3f::shape_index == 0xFFFFFFFFu);
    }
    DevBuf<uint8_t> hit(n);
    shape.ray_test(ray, hit.p);
    CU(cudaDeviceSynchronize());
    { auto hh = hit.host(); EXPECT(hh[0] == 1 && hh[1] == 1); }

    DevBuf<float> st(n), px(n), py(n), pz(n), nx(n), ny(n), nz(n), uvx(n), uvy(n), dux(n), duy(n), duz(n), dvx(n), dvy(n), dvz(n);
    hf::SurfaceInteraction3f si;
    std::memset(&si, 0, sizeof(si));
    si.t = st.p; si.px = px.p; si.py = py.p; si.pz = pz.p; si.nx = nx.p; si.ny = ny.p; si.nz = nz.p; si.uvx = uvx.p; si.uvy = uvy.p;
    si.dpdu_x = dux.p; si.dpdu_y = duy.p; si.dpdu_z = duz.p; si.dpdv_x = dvx.p; si.dpdv_y = dvy.p; si.dpdv_z = dvz.p;
    shape.compute_surface_interaction(ray, pi, si, +hf::RayFlags::All);
    CU(cudaDeviceSynchronize());
    {
        auto hpx = px.host(), hpy = py.host(), hpz = pz.host(), hux = uvx.host(), huy = uvy.host(), hn = nz.host(), hd = dux.host(), he = dvy.host();
        EXPECT(close(hpx[0], -0.3f) && close(hpy[0], -0.3f) && close(hpz[0], 0) && close(hpx[1], 0.3f) && close(hpy[1], 0.3f));
        EXPECT(close(hux[0], 0.35f) && close(huy[0], 0.35f) && close(hux[1], 0.65f) && close(huy[1], 0.65f));
        EXPECT(close(hn[0], 1) && close(hn[1], 1) && close(hd[0], 2) && close(hd[1], 2) && close(he[0], 2) && close(he[1], 2));
    }
    // ray_intersect == preliminary + surface interaction in one launch: same bits
    DevBuf<float> st2(n), px2(n);
    hf::SurfaceInteraction3f si2;
    std::memset(&si2, 0, sizeof(si2));
    si2.t = st2.p; si2.px = px2.p;
    shape.ray_intersect(ray, si2);
    CU(cudaDeviceSynchronize());
    { auto a = st.host(), b = st2.host(), c = px.host(), d = px2.host(); EXPECT(a[0] == b[0] && a[1] == b[1] && c[0] == d[0] && c[1] == d[1]); }

    // test_mesh.py:536-623 (test16): reverse mode at the corner ray: backward(si.t) puts +1 on the z of vertex (1, 1)
    {
        using V = std::vector<float>;
        DevBuf<float> cox(V{ 0.99999f }), coy(V{ 0.99999f }), coz(V{ -10.f }), cdx(V{ 0.f }), cdy(V{ 0.f }), cdz(V{ 1.f }), cmt(V{ INFINITY });
        hf::Ray3f r1;
        r1.ox = cox.p; r1.oy = coy.p; r1.oz = coz.p; r1.dx = cdx.p; r1.dy = cdy.p; r1.dz = cdz.p; r1.maxt = cmt.p; r1.size = 1;
        DevBuf<float> t1(1), u1(1), v1(1); DevBuf<uint32_t> p1(1);
        hf::PreliminaryIntersection3f pi1;
        pi1.t = t1.p; pi1.prim_u = u1.p; pi1.prim_v = v1.p; pi1.prim_index = p1.p;
        shape.ray_intersect_preliminary(r1, pi1);
        DevBuf<float> one(std::vector<float>{ 1.f });
        hf::SurfaceInteractionAdjoint3f adj;
        std::memset(&adj, 0, sizeof(adj));
        adj.t = one.p;
        shape.zero_grad();
        shape.backward(r1, pi1, adj);
        CU(cudaDeviceSynchronize());
        std::vector<float> g(4);
        CU(cudaMemcpy(g.data(), shape.grad(), 4 * sizeof(float), cudaMemcpyDeviceToHost));
        EXPECT(close(g[0], 0) && close(g[1], 0) && close(g[2], 0) && close(g[3], 1));
        adj.t = nullptr; adj.nx = one.p;                                  // backward(si.n.x): +0.5 / -0.5 (test_mesh.py:600-610)
        shape.zero_grad();
        shape.backward(r1, pi1, adj);
        CU(cudaDeviceSynchronize());
        CU(cudaMemcpy(g.data(), shape.grad(), 4 * sizeof(float), cudaMemcpyDeviceToHost));
        EXPECT(close(g[0], 0) && close(g[1], 0) && close(g[2], 0.5f) && close(g[3], -0.5f));
        // DetachShape | FollowShape is rejected with the reference's message (src/render/mesh.cpp:709-711)
        bool threw = false;
        try { shape.backward(r1, pi1, adj, +hf::RayFlags::DetachShape | +hf::RayFlags::FollowShape); }
        catch (const hf::Error &e) { threw = std::string(e.what()).find("Invalid combination of RayFlags") != std::string::npos; }
        EXPECT(threw);
    }

    // ---- traverse / parameters_changed: raise one texel through the exposed pointer, the bbox follows
    Recorder rec;
    shape.traverse(&rec);
    EXPECT(rec.names.size() == 3 && rec.names[0] == "heights" && rec.count == 4);
    EXPECT(rec.flags[0] == (hf::ParamFlags::Differentiable | hf::ParamFlags::Discontinuous) && rec.flags[1] == +hf::ParamFlags::NonDifferentiable);
    const float raised = 0.5f;
    CU(cudaMemcpy(static_cast<float *>(rec.heights) + 3, &raised, sizeof(float), cudaMemcpyHostToDevice));
    shape.parameters_changed({ "heights" });
    EXPECT(close(shape.bbox().max[2], 0.5f));
    shape.ray_intersect_preliminary(ray, pi);
    CU(cudaDeviceSynchronize());
    { auto ht = t.host(); EXPECT(close(ht[0], 10) && ht[1] > 10.f && ht[1] < 10.5f); }     // triangle 1 now rises towards (1, 1)

    // one Adam step moves the texels against the gradient; the caller then rebuilds the acceleration data
    shape.adam_step(1e-2);
    shape.parameters_changed({ "heights" });
    std::printf("abi_smoke ok\n%s\n", shape.to_string().c_str());
    return 0;
}
