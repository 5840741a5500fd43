// context.cu — the `forma_rt` handle (TraceModel, tracer.rs:12-17, plus its device mirror) and the
// C ABI of include/forma_rt.h.  No CPU fallback: every render entry point launches CUDA kernels.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <numeric>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/forma_rt.h"
#include "json.hpp"
#include "kernels.hpp"
#include "lbvh.hpp"
#include "png.hpp"
#include "postfx.hpp"
#include "../../include/forma_fx.h"
#include "scene.hpp"

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define CU_TRY(expr)                                                                                   \
    do {                                                                                               \
        cudaError_t e__ = (expr);                                                                      \
        if (e__ != cudaSuccess)                                                                        \
            return fail(FORMA_RT_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));          \
    } while (0)

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void**)&p, (n ? n : 1) * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    cudaError_t upload(const std::vector<T>& v, cudaStream_t s) {
        cudaError_t e = ensure(v.size());
        if (e != cudaSuccess || v.empty()) return e;
        return cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s);
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

constexpr int kSmemStageLimit = 96 * 1024;   // stage spheres+tris in shared memory up to this size
// ACCEL_AUTO: LBVH when spheres + 2 x triangles reaches this (a triangle costs about two spheres to scan).  Swept on
// the sphere-field generator at 1080p (tools/accel_threshold_probe.py, profiles/r05z_accel_threshold.txt): 16 spheres scan 2.58 ms
// / tree 3.20 ms, 24: 3.00 / 3.26, 32: 3.42 / 3.40, 64: 5.15 / 3.72.  Round 1's value was 16 primitives, unswept.
constexpr uint32_t kBvhThreshold = 32;
constexpr size_t kRadBudget = 4ull << 30;    // tree mode: bytes of per-path radiance kept between trace and accumulate
                                             // (64 samples of a 1080p frame = 2.1 GB in ONE launch: every launch ends with a drain)

}  // namespace

namespace forma {
int set_last_error(int code, const std::string& msg) { return fail(code, msg); }
}  // namespace forma

extern "C" void forma_rt_destroy(forma_rt* ctx);

struct forma_rt {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_k0 = nullptr, ev_k1 = nullptr;
    uint32_t width = 0, height = 0;
    uint64_t seed = 0;
    uint32_t flags = 0, accel = FORMA_RT_ACCEL_AUTO;
    forma::Scene scene;
    bool scene_dirty = true;
    bool stats_pending = false;
    // render_tree_kernel tuning (FORMA_RT_REFILL / FORMA_RT_LEAF_BATCH override; not part of the ABI)
    uint32_t refill = 20;     // lanes idle before a warp stops walking to refill them
    uint32_t leaf_batch = 8;  // lanes that must stop progressing before the warp looks at its waiting leaves
    uint32_t part_mod = 1, part_rem = 0;  // forma_rt_set_partition: this handle renders the 32x4 tiles t with t % part_mod == part_rem
    int fast_div = -1;   // div_const (trace.cuh): -1 not checked yet, 0 IEEE division, 1 verified multiply form
    uint32_t frame = 0;  // update() call counter -> sample index of the interactive frame
    forma_rt* mt_scene = nullptr;  // render_mt traces a FRESH get_simple_scene() (tracer.rs:105): a child handle holds it
    DevBuf<uint8_t> d_mt_frame;    // one render_mt frame
    DevBuf<float> d_mt_acc;        // save_image_mt's pixels_acc

    // packed scene (host copies kept for re-upload after set_object)
    DevBuf<float4> d_spheres, d_tris, d_planes, d_material;
    DevBuf<int> d_sphere_obj, d_tri_obj, d_plane_obj, d_kindmat, d_ids;
    DevBuf<forma::Segment> d_segments;
    forma::LbvhDevice bvh;
    forma::SceneView view{};

    // frame buffers
    DevBuf<uint8_t> d_rgb;
    DevBuf<float4> d_sum;
    DevBuf<int32_t> d_prim;
    DevBuf<float> d_t;
    DevBuf<float4> d_prog;            // progressive mode: running f32 sum of prog_spp samples per pixel
    uint32_t prog_spp = 0;
    forma_rt_camera prog_cam{};       // camera the running sum belongs to
    uint8_t srgb_lut[256];            // color_utils::srgb_encode as a table (color_utils.rs:49-66)
    DevBuf<float4> d_rad, d_acc;      // tree mode: per-path radiance of a chunk of samples, in-order running sum
    DevBuf<unsigned int> d_job;       // tree mode: job queue head
    DevBuf<unsigned long long> d_walk; // forma_rt_walk_stats_get: node-visit and leaf-test counters
    bool count_walk = false;          // the next render_on_device uses the instrumented tree kernel
    unsigned long long* d_rays = nullptr;  // [0] closest-hit queries, [1] beam-culled paths of the last launch
    unsigned long long* h_rays = nullptr;  // pinned
    unsigned int* h_flag_err = nullptr;    // host-mapped: 1 + index of the first flag a forma_rt_flag_wait_device gave up on
    unsigned int* d_flag_err = nullptr;    // its device alias
    unsigned long long flag_timeout_ns = 30ull * 1000000000ull;  // FORMA_RT_FLAG_TIMEOUT_MS
    std::vector<void*> pinned;             // forma_rt_host_alloc

    std::vector<void*> scratch;
    forma_rt_stats stats{};
    int sm_count = 0, clock_khz = 0;

    ~forma_rt() {
        if (mt_scene) forma_rt_destroy(mt_scene);
        if (device < 0) return;  // host-only handle owns no device resources
        cudaSetDevice(device);
        if (stream) cudaStreamSynchronize(stream);
        for (void* p : scratch) cudaFree(p);
        d_spheres.release(); d_tris.release(); d_planes.release(); d_material.release();
        d_sphere_obj.release(); d_tri_obj.release(); d_plane_obj.release(); d_kindmat.release(); d_ids.release();
        d_segments.release(); d_rgb.release(); d_sum.release(); d_prim.release(); d_t.release();
        d_rad.release(); d_acc.release(); d_job.release(); d_prog.release(); d_walk.release();
        d_mt_frame.release(); d_mt_acc.release();
        bvh.release();
        if (d_rays) cudaFree(d_rays);
        if (h_rays) cudaFreeHost(h_rays);
        if (h_flag_err) cudaFreeHost(h_flag_err);
        for (void* p : pinned) cudaFreeHost(p);
        for (cudaEvent_t e : {ev_begin, ev_end, ev_k0, ev_k1}) if (e) cudaEventDestroy(e);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {

// Scene.objects -> the segment / SoA layout of device_scene.hpp, uploaded on ctx->stream.
int upload_scene(forma_rt* c) {
    using namespace forma;
    const auto& objs = c->scene.objects;
    std::vector<float4> spheres, tris, planes, material(objs.size());
    std::vector<int> sphere_obj, tri_obj, plane_obj, kindmat(objs.size()), ids(objs.size());
    std::vector<Segment> segments;

    size_t i = 0;
    const size_t n = objs.size();
    do {  // one segment per run of non-plane objects, closed by the plane that follows it (if any)
        Segment g;
        g.sph_begin = (int)spheres.size();
        g.tri_begin = (int)(tris.size() / 3);
        size_t j = i;
        for (; j < n && objs[j].kind != FORMA_RT_PLANE; ++j) {
            const forma_rt_object& o = objs[j];
            if (o.kind == FORMA_RT_SPHERE) {
                spheres.push_back(make_float4(o.a[0], o.a[1], o.a[2], o.b[0]));
                sphere_obj.push_back((int)j);
            } else {
                V3 v0{o.a[0], o.a[1], o.a[2]}, v1{o.b[0], o.b[1], o.b[2]}, v2{o.c[0], o.c[1], o.c[2]};
                V3 e1 = v1 - v0, e2 = v2 - v0;
                V3 nrm = unit(cross(e1, e2));
                tris.push_back(make_float4(v0.x, v0.y, v0.z, nrm.x));
                tris.push_back(make_float4(e1.x, e1.y, e1.z, nrm.y));
                tris.push_back(make_float4(e2.x, e2.y, e2.z, nrm.z));
                tri_obj.push_back((int)j);
            }
        }
        g.sph_end = (int)spheres.size();
        g.tri_end = (int)(tris.size() / 3);
        g.plane = -1;
        if (j < n) {  // objs[j] is a plane
            const forma_rt_object& o = objs[j];
            g.plane = (int)plane_obj.size();
            planes.push_back(make_float4(o.a[0], o.a[1], o.a[2], 0.0f));
            planes.push_back(make_float4(o.b[0], o.b[1], o.b[2], 0.0f));
            planes.push_back(make_float4(o.c[0], o.c[1], o.c[2], 0.0f));
            plane_obj.push_back((int)j);
            ++j;
        }
        if (g.sph_end > g.sph_begin || g.tri_end > g.tri_begin || g.plane >= 0) segments.push_back(g);
        i = j;
    } while (i < n);

    uint32_t ns = 0, np = 0, nt = 0;
    for (size_t k = 0; k < n; ++k) {
        const forma_rt_object& o = objs[k];
        material[k] = make_float4(o.color[0], o.color[1], o.color[2], o.fuzz);
        int kind = o.kind == FORMA_RT_SPHERE ? FORMA_RT_SPHERE : (o.kind == FORMA_RT_PLANE ? FORMA_RT_PLANE : FORMA_RT_TRIANGLE);
        kindmat[k] = kind | ((o.material & 0xff) << 8);  // material is a u8 in the reference
        ids[k] = o.id;
        (kind == FORMA_RT_SPHERE ? ns : kind == FORMA_RT_PLANE ? np : nt)++;
    }

    // ---- acceleration structure planning (host): which primitives the LBVH indexes ----------------
    // The tree needs every sphere/triangle in ONE run of the list (planes may precede or follow it).
    int runs = 0;
    for (const Segment& g : segments) runs += (g.sph_end > g.sph_begin || g.tri_end > g.tri_begin) ? 1 : 0;
    const bool want_bvh = c->accel == FORMA_RT_ACCEL_LBVH || (c->accel == FORMA_RT_ACCEL_AUTO && ns + 2 * nt >= kBvhThreshold);
    if (c->accel == FORMA_RT_ACCEL_LBVH && runs > 1)
        return fail(FORMA_RT_EINVAL, "LBVH needs all spheres and triangles in one plane-free run of Scene.objects; use FORMA_RT_ACCEL_BRUTE");
    LbvhInput bin{};
    bool use_bvh = false;
    float bc[3] = {0, 0, 0}, br2 = 0.0f, beps = 0.0f, qstep[3] = {0, 0, 0}, bq = 0.0f, brmax = 0.0f;
    if (want_bvh && runs == 1 && ns + nt >= 2) {
        const int NS = (int)ns, NT = (int)nt, N = NS + NT;
        std::vector<V3> lo(N), hi(N);
        for (int i = 0; i < NS; ++i) {
            // Sphere::hit (sphere.rs:23-51) only uses radius * radius, so a negative radius (the hollow-sphere idiom, a
            // negative scale.x in JSON) is legal and hits like |r|: boxes, pads and the cull use |r|
            const float4 sp = spheres[i];
            const float ar = std::fabs(sp.w);
            lo[i] = {sp.x - ar, sp.y - ar, sp.z - ar};
            hi[i] = {sp.x + ar, sp.y + ar, sp.z + ar};
        }
        for (int t = 0; t < NT; ++t) {
            const float4 a = tris[3 * t], e1 = tris[3 * t + 1], e2 = tris[3 * t + 2];
            const float x[3] = {a.x, a.x + e1.x, a.x + e2.x}, y[3] = {a.y, a.y + e1.y, a.y + e2.y}, z[3] = {a.z, a.z + e1.z, a.z + e2.z};
            lo[NS + t] = {std::min({x[0], x[1], x[2]}), std::min({y[0], y[1], y[2]}), std::min({z[0], z[1], z[2]})};
            hi[NS + t] = {std::max({x[0], x[1], x[2]}), std::max({y[0], y[1], y[2]}), std::max({z[0], z[1], z[2]})};
        }
        auto extent = [&](int i) { return std::max({hi[i].x - lo[i].x, hi[i].y - lo[i].y, hi[i].z - lo[i].z}); };
        // Outsized primitives (a ground sphere of radius 1000 among unit spheres, a 1000-unit floor quad) would
        // blow up the walk's sphere pad (it scales with the squared radius of the tree's bounding sphere) and
        // the Morton grid: they are scanned linearly instead.  "Outsized" = 4x larger than everything else
        // together; merely large primitives stay in the tree (profiles/r01t: with the old 0.25 E rule 162 of
        // scene_01's 350 triangles were scanned by every ray, 85 % of config A's instructions).
        std::vector<char> big(N, 0);
        if (N >= 16) {
            // largest first: a prefix of the order is outsized iff its smallest member exceeds 4x the box around all the others
            std::vector<int> order;
            for (int i = 0; i < N; ++i) if (std::isfinite(extent(i))) order.push_back(i);
            std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return extent(x) > extent(y); });
            const int M = (int)order.size();
            std::vector<float> rest(M + 1, 0.0f);  // rest[k] = extent of the union of order[k..M)
            V3 ulo{FLT_MAX, FLT_MAX, FLT_MAX}, uhi{-FLT_MAX, -FLT_MAX, -FLT_MAX};
            for (int k = M - 1; k >= 0; --k) {
                const int i = order[k];
                ulo = {std::min(ulo.x, lo[i].x), std::min(ulo.y, lo[i].y), std::min(ulo.z, lo[i].z)};
                uhi = {std::max(uhi.x, hi[i].x), std::max(uhi.y, hi[i].y), std::max(uhi.z, hi[i].z)};
                rest[k] = std::max({uhi.x - ulo.x, uhi.y - ulo.y, uhi.z - ulo.z});
            }
            int nbig = 0;  // the longest prefix (<= 64) whose smallest member is still 4x larger than all the rest together
            for (int k = 1; k <= 64 && k + 2 <= M; ++k) if (extent(order[k - 1]) > 4.0f * rest[k]) nbig = k;
            for (int k = 0; k < nbig; ++k) big[order[k]] = 1;
        }
        for (int i = 0; i < N; ++i) if (!std::isfinite(extent(i))) big[i] = 1;  // NaN/inf geometry stays on the exact scan
        // so does a sphere whose 1 / |r| is not finite (r = 0 or denormal): the sphere pad scales with max 1 / |r|
        for (int i = 0; i < NS; ++i) if (!std::isfinite(1.0f / std::fabs(spheres[i].w))) big[i] = 1;
        // stable partition of the packed arrays: big first
        auto partition = [&](int first, int count, int stride, std::vector<float4>& data, std::vector<int>& objs_idx) {
            std::vector<float4> d2; std::vector<int> o2; d2.reserve(data.size()); o2.reserve(objs_idx.size());
            int nb = 0;
            for (int pass = 1; pass >= 0; --pass)
                for (int i = 0; i < count; ++i) if (big[first + i] == pass) {
                    for (int k = 0; k < stride; ++k) d2.push_back(data[stride * i + k]);
                    o2.push_back(objs_idx[i]);
                    nb += pass;
                }
            data.swap(d2); objs_idx.swap(o2);
            return nb;
        };
        std::vector<V3> blo, bhi;  // boxes of what goes into the tree, for the Morton grid / bounding sphere
        for (int i = 0; i < N; ++i) if (!big[i]) { blo.push_back(lo[i]); bhi.push_back(hi[i]); }
        if ((int)blo.size() >= 2) {
            const int big_s = partition(0, NS, 1, spheres, sphere_obj);
            const int big_t = partition(NS, NT, 3, tris, tri_obj);
            V3 clo{FLT_MAX, FLT_MAX, FLT_MAX}, chi{-FLT_MAX, -FLT_MAX, -FLT_MAX}, ulo = clo, uhi = chi;
            for (size_t i = 0; i < blo.size(); ++i) {
                const V3 cc{0.5f * (blo[i].x + bhi[i].x), 0.5f * (blo[i].y + bhi[i].y), 0.5f * (blo[i].z + bhi[i].z)};
                clo = {std::min(clo.x, cc.x), std::min(clo.y, cc.y), std::min(clo.z, cc.z)};
                chi = {std::max(chi.x, cc.x), std::max(chi.y, cc.y), std::max(chi.z, cc.z)};
                ulo = {std::min(ulo.x, blo[i].x), std::min(ulo.y, blo[i].y), std::min(ulo.z, blo[i].z)};
                uhi = {std::max(uhi.x, bhi[i].x), std::max(uhi.y, bhi[i].y), std::max(uhi.z, bhi[i].z)};
            }
            bin.n_spheres = NS; bin.n_tris = NT; bin.big_spheres = big_s; bin.big_tris = big_t;
            bin.n_bvh_spheres = NS - big_s; bin.count = (NS - big_s) + (NT - big_t);
            const float cl[3] = {clo.x, clo.y, clo.z}, ch[3] = {chi.x, chi.y, chi.z};
            // one scale for all three axes: cubic Morton cells, so the tree splits a flat scene along its long axes
            // first instead of spending every third level on the thin one
            const float cext = std::max({ch[0] - cl[0], ch[1] - cl[1], ch[2] - cl[2]});
            const bool iso = !std::getenv("FORMA_RT_MORTON_ANISO");
            for (int k = 0; k < 3; ++k) {
                bin.cmin[k] = cl[k];
                const float e = iso ? cext : ch[k] - cl[k];
                bin.cscale[k] = e > 0.0f ? 1024.0f / e : 0.0f;
            }
            bc[0] = 0.5f * (ulo.x + uhi.x); bc[1] = 0.5f * (ulo.y + uhi.y); bc[2] = 0.5f * (ulo.z + uhi.z);
            const V3 half{0.5f * (uhi.x - ulo.x), 0.5f * (uhi.y - ulo.y), 0.5f * (uhi.z - ulo.z)};
            br2 = 1.0001f * dot(half, half);
            const float scale = std::max({std::fabs(ulo.x), std::fabs(ulo.y), std::fabs(ulo.z), std::fabs(uhi.x), std::fabs(uhi.y), std::fabs(uhi.z)});
            beps = std::ldexp(scale, -16) + 1e-30f;
            // the 16-bit grid the node boxes are stored on: 65533 cells over the tree's box plus one spare cell on
            // either side (lbvh.cu rounds outwards by a whole extra cell); a flat axis gets a token extent
            const float ul[3] = {ulo.x, ulo.y, ulo.z}, uh[3] = {uhi.x, uhi.y, uhi.z};
            for (int k = 0; k < 3; ++k) {
                const float ext = std::max({(uh[k] - ul[k]) * 1.0001f, scale * 1e-6f, 1e-20f});
                qstep[k] = ext / 65533.0f;
                bin.qmin[k] = ul[k] - qstep[k];
                bin.qinv[k] = 1.0f / qstep[k];
            }
            bq = 0.0f;  // max 1/|r| over the spheres of the tree (the outsized ones were moved to the front)
            for (int i = big_s; i < NS; ++i) { bq = std::max(bq, 1.0f / std::fabs(spheres[i].w)); brmax = std::max(brmax, std::fabs(spheres[i].w)); }
            use_bvh = true;
        }
    }

    cudaStream_t s = c->stream;
    CU_TRY(c->d_spheres.upload(spheres, s));
    CU_TRY(c->d_tris.upload(tris, s));
    CU_TRY(c->d_planes.upload(planes, s));
    CU_TRY(c->d_material.upload(material, s));
    CU_TRY(c->d_sphere_obj.upload(sphere_obj, s));
    CU_TRY(c->d_tri_obj.upload(tri_obj, s));
    CU_TRY(c->d_plane_obj.upload(plane_obj, s));
    CU_TRY(c->d_kindmat.upload(kindmat, s));
    CU_TRY(c->d_ids.upload(ids, s));
    CU_TRY(c->d_segments.upload(segments, s));
    CU_TRY(cudaStreamSynchronize(s));  // the host vectors die at return

    SceneView& v = c->view;
    v.spheres = c->d_spheres.p; v.sphere_obj = c->d_sphere_obj.p;
    v.tris = c->d_tris.p; v.tri_obj = c->d_tri_obj.p;
    v.planes = c->d_planes.p; v.plane_obj = c->d_plane_obj.p;
    v.segments = c->d_segments.p; v.material = c->d_material.p;
    v.kindmat = c->d_kindmat.p; v.ids = c->d_ids.p;
    v.n_segments = (int)segments.size();
    v.seg0 = segments.empty() ? Segment{0, 0, 0, 0, -1} : segments[0];
    v.n_spheres = (int)ns; v.n_tris = (int)nt; v.n_planes = (int)np;
    v.bvh_nodes = nullptr;
    v.finite_materials = 1;
    for (const float4& mt : material) if (!(std::isfinite(mt.x) && std::isfinite(mt.y) && std::isfinite(mt.z))) v.finite_materials = 0;

    {   // closest_hit's conservative cull: bounding sphere of the box around every sphere and triangle
        V3 ulo{FLT_MAX, FLT_MAX, FLT_MAX}, uhi{-FLT_MAX, -FLT_MAX, -FLT_MAX};
        bool finite = true;
        auto grow = [&](float x, float y, float z, float r) {
            finite = finite && std::isfinite(x) && std::isfinite(y) && std::isfinite(z) && std::isfinite(r);
            r = std::fabs(r);
            ulo = {std::min(ulo.x, x - r), std::min(ulo.y, y - r), std::min(ulo.z, z - r)};
            uhi = {std::max(uhi.x, x + r), std::max(uhi.y, y + r), std::max(uhi.z, z + r)};
        };
        for (const float4& sp : spheres) grow(sp.x, sp.y, sp.z, sp.w);
        for (size_t t = 0; t + 2 < tris.size(); t += 3) {
            const float4 a = tris[t], e1 = tris[t + 1], e2 = tris[t + 2];
            grow(a.x, a.y, a.z, 0.0f); grow(a.x + e1.x, a.y + e1.y, a.z + e1.z, 0.0f); grow(a.x + e2.x, a.y + e2.y, a.z + e2.z, 0.0f);
        }
        const V3 half{0.5f * (uhi.x - ulo.x), 0.5f * (uhi.y - ulo.y), 0.5f * (uhi.z - ulo.z)};
        const float r2 = 1.0001f * dot(half, half);
        v.cull_center[0] = 0.5f * (ulo.x + uhi.x); v.cull_center[1] = 0.5f * (ulo.y + uhi.y); v.cull_center[2] = 0.5f * (ulo.z + uhi.z);
        v.cull_r2 = (finite && ns + nt > 0 && std::isfinite(r2)) ? r2 : INFINITY;
        if (!(v.cull_r2 < INFINITY)) v.cull_center[0] = v.cull_center[1] = v.cull_center[2] = 0.0f;
    }
    v.bvh_prim = nullptr; v.big_spheres = 0; v.big_tris = 0; v.tree_seg = 0;
    c->stats.bvh_nodes = 0;
    c->stats.accel_in_use = FORMA_RT_ACCEL_BRUTE;
    if (use_bvh) {
        bin.spheres = c->d_spheres.p; bin.tris = c->d_tris.p;
        cudaError_t e = c->bvh.build(bin, s);
        if (e != cudaSuccess) return fail(FORMA_RT_ECUDA, std::string("LBVH build: ") + cudaGetErrorString(e));
        v.bvh_nodes = c->bvh.nodes; v.bvh_prim = c->bvh.prim;
        v.big_spheres = bin.big_spheres; v.big_tris = bin.big_tris;
        for (size_t k = 0; k < segments.size(); ++k)
            if (segments[k].sph_end > segments[k].sph_begin || segments[k].tri_end > segments[k].tri_begin) v.tree_seg = (int)k;
        v.bvh_center[0] = bc[0]; v.bvh_center[1] = bc[1]; v.bvh_center[2] = bc[2];
        v.bvh_r2 = br2; v.bvh_eps = beps; v.bvh_q = bq; v.bvh_rmax = 1.0001f * brmax;
        for (int k = 0; k < 3; ++k) { v.bvh_qmin[k] = bin.qmin[k]; v.bvh_qstep[k] = qstep[k]; }
        c->stats.bvh_nodes = (uint32_t)(bin.count - 1);
        c->stats.accel_in_use = FORMA_RT_ACCEL_LBVH;
    }
    const size_t stage = spheres.size() * sizeof(float4) + tris.size() * sizeof(float4);
    v.smem_bytes = (!v.bvh_nodes && stage > 0 && stage <= (size_t)kSmemStageLimit) ? (int)stage : 0;

    c->stats.n_objects = (uint32_t)n;
    c->stats.n_spheres = ns; c->stats.n_planes = np; c->stats.n_triangles = nt;
    c->scene_dirty = false;
    return FORMA_RT_OK;
}

int ensure_scene(forma_rt* c) {
    if (!c->scene_dirty) return FORMA_RT_OK;
    return upload_scene(c);
}

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
#define NEED_CTX(c) do { if (!(c)) return fail(FORMA_RT_EINVAL, "null forma_rt handle"); if ((c)->device >= 0) cudaSetDevice((c)->device); } while (0)
// every entry point that touches the GPU: a host-only handle fails loudly, there is no CPU path
#define NEED_DEV(c) do { NEED_CTX(c); if ((c)->device < 0) return fail(FORMA_RT_ECUDA, "handle was created with FORMA_RT_HOST_ONLY: no device, and there is no CPU fallback"); } while (0)

extern "C" {

const char* forma_rt_last_error(void) { return g_last_error.c_str(); }
uint32_t forma_rt_abi_version(void) { return FORMA_RT_ABI_VERSION; }

int forma_rt_create(const forma_rt_options* opts, uint32_t width, uint32_t height, forma_rt** out) {
    if (!out) return fail(FORMA_RT_EINVAL, "forma_rt_create: out is null");
    *out = nullptr;
    if (width == 0 || height == 0 || (uint64_t)width * height > (1ull << 31))
        return fail(FORMA_RT_EINVAL, "forma_rt_create: width/height out of range");
    forma_rt_options o{};
    o.struct_size = sizeof o;
    if (opts) {
        if (opts->struct_size != sizeof(forma_rt_options))
            return fail(FORMA_RT_EINVAL, "forma_rt_create: options.struct_size mismatch (ABI skew)");
        o = *opts;
    }
    if (o.accel > FORMA_RT_ACCEL_LBVH) return fail(FORMA_RT_EINVAL, "forma_rt_create: unknown accel");
    if (o.flags & FORMA_RT_HOST_ONLY) {  // scene / camera bookkeeping only (tooling, CPU-side tests)
        forma_rt* c = new (std::nothrow) forma_rt();
        if (!c) return fail(FORMA_RT_ENOMEM, "forma_rt_create: out of host memory");
        c->device = -1; c->width = width; c->height = height;
        c->seed = o.seed; c->flags = o.flags; c->accel = o.accel;
        c->scene.camera = forma::Camera::create(width, height);
        c->scene.objects = forma::builtin_scene(0);
        *out = c;
        return FORMA_RT_OK;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(FORMA_RT_ECUDA, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (o.device < 0 || o.device >= ndev) return fail(FORMA_RT_EINVAL, "forma_rt_create: device ordinal out of range");
    CU_TRY(cudaSetDevice(o.device));

    forma_rt* c = new (std::nothrow) forma_rt();
    if (!c) return fail(FORMA_RT_ENOMEM, "forma_rt_create: out of host memory");
    c->device = o.device; c->width = width; c->height = height;
    c->seed = o.seed; c->flags = o.flags; c->accel = o.accel;
    auto bail = [&](cudaError_t err, const char* what) {
        delete c;
        return fail(FORMA_RT_ECUDA, std::string(what) + ": " + cudaGetErrorString(err));
    };
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    for (cudaEvent_t* ev : {&c->ev_begin, &c->ev_end, &c->ev_k0, &c->ev_k1})
        if ((e = cudaEventCreate(ev)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaMalloc((void**)&c->d_rays, 2 * sizeof(unsigned long long))) != cudaSuccess) return bail(e, "cudaMalloc");
    if ((e = cudaMallocHost((void**)&c->h_rays, 2 * sizeof(unsigned long long))) != cudaSuccess) return bail(e, "cudaMallocHost");
    if ((e = cudaHostAlloc((void**)&c->h_flag_err, sizeof(unsigned int), cudaHostAllocMapped)) != cudaSuccess) return bail(e, "cudaHostAlloc");
    *c->h_flag_err = 0u;
    if ((e = cudaHostGetDevicePointer((void**)&c->d_flag_err, c->h_flag_err, 0)) != cudaSuccess) return bail(e, "cudaHostGetDevicePointer");
    if (const char* r = std::getenv("FORMA_RT_FLAG_TIMEOUT_MS")) {
        const long long v = std::atoll(r);
        if (v > 0) c->flag_timeout_ns = (unsigned long long)v * 1000000ull;
    }
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, o.device);
    cudaDeviceGetAttribute(&c->clock_khz, cudaDevAttrClockRate, o.device);
    c->stats.sm_count = (uint32_t)c->sm_count;
    c->stats.sm_clock_khz = (uint32_t)c->clock_khz;

    if (const char* r = std::getenv("FORMA_RT_REFILL")) {  // tuning knob, not part of the ABI
        const int v = std::atoi(r);
        if (v >= 1 && v <= 32) c->refill = (uint32_t)v;
    }
    if (const char* r = std::getenv("FORMA_RT_LEAF_BATCH")) {
        const int v = std::atoi(r);
        if (v >= 1 && v <= 32) c->leaf_batch = (uint32_t)v;
    }
    forma::srgb_encode_table(c->srgb_lut);
    // Scene::new (scene.rs:10-15): Camera::new + get_simple_scene
    c->scene.camera = forma::Camera::create(width, height);
    c->scene.objects = forma::builtin_scene(0);
    c->scene_dirty = true;
    *out = c;
    return FORMA_RT_OK;
}

void forma_rt_destroy(forma_rt* ctx) { delete ctx; }

int forma_rt_load_builtin(forma_rt* ctx, int which) {
    NEED_CTX(ctx);
    try {
        ctx->scene.objects = forma::builtin_scene(which);
    } catch (const std::exception& ex) {
        return fail(FORMA_RT_EINVAL, ex.what());
    }
    ctx->scene_dirty = true;
    return FORMA_RT_OK;
}

int forma_rt_load_json(forma_rt* ctx, const char* json, size_t len) {
    NEED_CTX(ctx);
    if (!json) return fail(FORMA_RT_EINVAL, "forma_rt_load_json: json is null");
    try {
        std::vector<forma_rt_object> objs;
        forma::JsonCamera cam;
        forma::load_json_scene(json, len, objs, cam);
        ctx->scene.objects.swap(objs);
        ctx->scene.json_camera = cam;
    } catch (const std::exception& ex) {  // the reference panics here (scene_loader.rs:4 unwrap)
        return fail(FORMA_RT_EJSON, ex.what());
    }
    ctx->scene_dirty = true;
    return FORMA_RT_OK;
}

int forma_rt_load_objects(forma_rt* ctx, const forma_rt_object* objs, size_t n) {
    NEED_CTX(ctx);
    if (n && !objs) return fail(FORMA_RT_EINVAL, "forma_rt_load_objects: objs is null");
    for (size_t i = 0; i < n; ++i)
        if (objs[i].kind < FORMA_RT_SPHERE || objs[i].kind > FORMA_RT_TRIANGLE)
            return fail(FORMA_RT_EINVAL, "forma_rt_load_objects: unknown object kind");
    ctx->scene.objects.assign(objs, objs + n);
    ctx->scene_dirty = true;
    return FORMA_RT_OK;
}

int forma_rt_get_objects(forma_rt* ctx, forma_rt_object* objs, size_t cap, size_t* n) {
    NEED_CTX(ctx);
    const auto& v = ctx->scene.objects;
    if (n) *n = v.size();
    if (objs) std::memcpy(objs, v.data(), sizeof(forma_rt_object) * (cap < v.size() ? cap : v.size()));
    return FORMA_RT_OK;
}

int forma_rt_set_object(forma_rt* ctx, uint32_t idx, const float* position, const float* orientation) {
    NEED_CTX(ctx);
    if (idx >= ctx->scene.objects.size()) return fail(FORMA_RT_EINVAL, "forma_rt_set_object: index out of range");
    forma_rt_object& o = ctx->scene.objects[idx];
    if (o.kind != FORMA_RT_PLANE) return FORMA_RT_OK;  // hitable.rs:12-13: default translate/rotate are no-ops
    if (position) std::memcpy(o.a, position, 3 * sizeof(float));       // plane.rs:62-64: assigns
    if (orientation) std::memcpy(o.b, orientation, 3 * sizeof(float)); // plane.rs:66-68: assigns
    ctx->scene_dirty = true;
    return FORMA_RT_OK;
}

// ---- camera ----------------------------------------------------------------------------------------
int forma_rt_camera_new(forma_rt* ctx) {
    NEED_CTX(ctx);
    ctx->scene.camera = forma::Camera::create(ctx->width, ctx->height);
    return FORMA_RT_OK;
}

int forma_rt_camera_look(forma_rt* ctx, const float from[3], const float at[3], const float up[3], float v_fov) {
    NEED_CTX(ctx);
    if (!from || !at || !up) return fail(FORMA_RT_EINVAL, "forma_rt_camera_look: null vector");
    ctx->scene.camera = forma::Camera::look({from[0], from[1], from[2]}, {at[0], at[1], at[2]}, {up[0], up[1], up[2]},
                                            v_fov, (float)ctx->width / (float)ctx->height);
    return FORMA_RT_OK;
}

int forma_rt_camera_json(forma_rt* ctx) {
    NEED_CTX(ctx);
    const forma::JsonCamera& jc = ctx->scene.json_camera;
    if (!jc.present) return fail(FORMA_RT_EINVAL, "forma_rt_camera_json: no JSON scene has been loaded");
    // DESIGN.md "Scene mapping": forward = R*(0,0,1), up = R*(0,1,0), focus distance = |position|
    forma::V3 ax, ay, az;
    forma::quat_axes(jc.rotation, ax, ay, az);
    float focus = forma::length(jc.position);
    if (!(focus > 0.0f)) focus = 1.0f;
    forma::V3 at{jc.position.x + focus * az.x, jc.position.y + focus * az.y, jc.position.z + focus * az.z};
    ctx->scene.camera = forma::Camera::look(jc.position, at, ay, jc.fov, (float)ctx->width / (float)ctx->height);
    return FORMA_RT_OK;
}

int forma_rt_camera_orbit(forma_rt* ctx, const float delta[3]) {
    NEED_CTX(ctx);
    if (!delta) return fail(FORMA_RT_EINVAL, "forma_rt_camera_orbit: null delta");
    ctx->scene.camera.orbit({delta[0], delta[1], delta[2]});
    return FORMA_RT_OK;
}

int forma_rt_camera_translate(forma_rt* ctx, const float delta[3]) {
    NEED_CTX(ctx);
    if (!delta) return fail(FORMA_RT_EINVAL, "forma_rt_camera_translate: null delta");
    ctx->scene.camera.translate({delta[0], delta[1], delta[2]});
    return FORMA_RT_OK;
}

int forma_rt_camera_get(forma_rt* ctx, forma_rt_camera* out) {
    NEED_CTX(ctx);
    if (!out) return fail(FORMA_RT_EINVAL, "forma_rt_camera_get: out is null");
    *out = ctx->scene.camera.c;
    return FORMA_RT_OK;
}

int forma_rt_camera_set(forma_rt* ctx, const forma_rt_camera* cam) {
    NEED_CTX(ctx);
    if (!cam) return fail(FORMA_RT_EINVAL, "forma_rt_camera_set: cam is null");
    ctx->scene.camera.c = *cam;
    return FORMA_RT_OK;
}

}  // extern "C"

// ---- rendering ---------------------------------------------------------------------------------------
namespace {

int flag_timed_out(forma_rt* c) {
    return fail(FORMA_RT_ECUDA, "forma_rt_flag_wait_device: flag " + std::to_string(*c->h_flag_err - 1) +
                " of a wait was not raised within " + std::to_string(c->flag_timeout_ns / 1000000ull) +
                " ms (a peer rank died, or two ranks share one GPU); the frame is incomplete");
}

// Scan mode has two kernels (kernels.cu): a pixel per thread (render_kernel) and (pixel, sample) jobs from a queue
// (render_queue_kernel).  The pixel loop has the least overhead per path - no radiance buffer, no job logic, the
// beam-culled sky loop - and wins wherever paths are short and alike: config C, a 1-spp frame (a queue has nothing to
// refill a lane with), and an object seen from OUTSIDE at any coverage (tools/coverage_probe.py: scene_09's tetrahedron
// filling 0.1 % .. 63 % of a 1080p frame, 1.08 .. 5.7 ms against the queue's 2.5 .. 8.3 ms).  The queue wins when paths
// get trapped and per-pixel cost varies wildly - the eye INSIDE the geometry, a ground sphere or a plane under
// everything, as in every scene of scenes.rs: 600x600x50 spp 5.45 -> 1.08 ms, 1080p x 64 spp 23.5 -> 18.4 ms.  So: the
// queue for >= 4 spp when the scene has planes or non-finite geometry (never culled, order-sensitive), or when the
// eye lies inside the bounding sphere of the spheres and triangles.  FORMA_RT_SCAN_QUEUE=0 / 1 overrides (A/B runs).
bool frame_bounces(const forma_rt* c, uint32_t spp) {
    if (const char* e = std::getenv("FORMA_RT_SCAN_QUEUE")) return std::atoi(e) != 0;
    if (spp < 4) return false;
    const forma::SceneView& v = c->view;
    if (v.n_spheres + v.n_tris + v.n_planes == 0) return false;
    if (v.n_planes != 0 || !(v.cull_r2 < FLT_MAX)) return true;
    const forma_rt_camera& k = c->scene.camera.c;
    double mm = 0.0;
    for (int i = 0; i < 3; ++i) {
        const double m = (double)k.position[i] - (double)v.cull_center[i];
        mm += m * m;
    }
    return !(mm > (double)v.cull_r2);  // NaN counts as inside
}

// One launch of the trace kernel on `stream`; outputs are device pointers.  Records kernel events.
int render_on_device(forma_rt* c, uint32_t spp, uint32_t sample_offset, uint32_t max_depth, uint32_t flags,
                     uint8_t* d_rgb, float4* d_sum, cudaStream_t stream, const float4* d_prior = nullptr,
                     uint32_t spp_prior = 0) {
    if (spp == 0) return fail(FORMA_RT_EINVAL, "render: spp must be >= 1");
    if (max_depth > (uint32_t)forma::kMaxDepthCap)
        return fail(FORMA_RT_EINVAL, "render: max_depth exceeds the attenuation stack (64; the reference uses 50)");
    int rc = ensure_scene(c);
    if (rc) return rc;
    forma::RenderParams P{};
    P.cam = c->scene.camera.c;
    P.scene = c->view;
    P.width = c->width; P.height = c->height;
    P.spp = spp; P.sample_offset = sample_offset; P.max_depth = max_depth; P.flags = flags;
    P.rk.expand(c->seed);
    P.rgb = d_rgb; P.sum = d_sum; P.rays = c->d_rays;
    P.prior = d_prior; P.spp_prior = d_prior ? spp_prior : 0;
    P.part_mod = c->part_mod; P.part_rem = c->part_rem;
    P.part_pitch = (c->width + 31) / 32;  // smallest pitch >= tiles per row that is coprime with part_mod
    while (std::gcd(P.part_pitch, P.part_mod) != 1u) ++P.part_pitch;
    P.queue = (!P.scene.bvh_nodes && frame_bounces(c, spp)) ? 1u : 0u;
    if (P.scene.bvh_nodes || P.queue) {  // job-queue modes: rad holds a chunk of samples of the frame, <= kRadBudget bytes
        const size_t npix = (size_t)c->width * c->height;
        size_t chunk = kRadBudget / (npix * sizeof(float4));
        chunk = std::max<size_t>(1, std::min<size_t>({chunk, spp, 64}));
        CU_TRY(c->d_rad.ensure(npix * chunk));
        CU_TRY(c->d_acc.ensure(npix));
        CU_TRY(c->d_job.ensure(1));
        P.rad = c->d_rad.p; P.acc = c->d_acc.p; P.job_counter = c->d_job.p;
        P.rad_chunk = (uint32_t)chunk;
        P.refill = c->refill; P.leaf_batch = c->leaf_batch;
        P.walk_counters = c->count_walk ? c->d_walk.p : nullptr;
    }
    if (c->fast_div < 0) {  // once per handle: width and height never change
        c->fast_div = 0;
        const char* env = std::getenv("FORMA_RT_FAST_DIV");
        if (!env || std::atoi(env) != 0) {
            CU_TRY(c->d_job.ensure(1));
            CU_TRY(cudaMemsetAsync(c->d_job.p, 0, sizeof(unsigned int), stream));
            CU_TRY(forma::launch_div_verify((float)c->width, (float)c->width + 1.0f, c->d_job.p, c->sm_count, stream));
            CU_TRY(forma::launch_div_verify((float)c->height, (float)c->height + 1.0f, c->d_job.p, c->sm_count, stream));
            unsigned int bad = 1;
            CU_TRY(cudaMemcpyAsync(&bad, c->d_job.p, sizeof bad, cudaMemcpyDeviceToHost, stream));
            CU_TRY(cudaStreamSynchronize(stream));
            c->fast_div = bad == 0 ? 1 : 0;
        }
    }
    P.inv_w = 1.0f / (float)c->width; P.inv_h = 1.0f / (float)c->height;
    P.fast_div = (uint32_t)c->fast_div;
    uint32_t launches = 0;
    CU_TRY(cudaMemsetAsync(c->d_rays, 0, 2 * sizeof(unsigned long long), stream));
    CU_TRY(cudaEventRecord(c->ev_k0, stream));
    CU_TRY(forma::launch_render(P, c->sm_count, stream, &launches));
    CU_TRY(cudaEventRecord(c->ev_k1, stream));
    if (d_rgb && (flags & FORMA_RT_SRGB_OUT)) {
        CU_TRY(forma::launch_srgb_encode(d_rgb, (size_t)c->width * c->height * 3, c->srgb_lut, c->sm_count, stream));
        ++launches;
    }
    c->stats.gpu_launches = launches;
    uint64_t own_pixels = (uint64_t)c->width * c->height;
    if (c->part_mod > 1) {  // pixels of the 32x4 tiles this handle owns
        own_pixels = 0;
        const uint32_t tx = (c->width + 31) / 32, ty = (c->height + 3) / 4;
        for (uint32_t y = 0; y < ty; ++y)
            for (uint32_t x = 0; x < tx; ++x)
                if ((y * P.part_pitch + x) % c->part_mod == c->part_rem)
                    own_pixels += (uint64_t)std::min(32u, c->width - x * 32) * std::min(4u, c->height - y * 4);
    }
    c->stats.paths = own_pixels * spp;
    c->stats_pending = true;  // ray counter + kernel time are read back by whoever synchronises
    return FORMA_RT_OK;
}

// host-facing render: device buffers owned by the handle, D2H of what the caller asked for
int render_to_host(forma_rt* c, uint32_t spp, uint32_t sample_offset, uint32_t max_depth, uint32_t flags,
                   uint8_t* rgb, float* linear_rgba, bool progressive = false) {
    const size_t npix = (size_t)c->width * c->height;
    if (rgb) CU_TRY(c->d_rgb.ensure(npix * 3));
    if (linear_rgba && !progressive) CU_TRY(c->d_sum.ensure(npix));
    if (progressive) CU_TRY(c->d_prog.ensure(npix));
    cudaStream_t s = c->stream;
    CU_TRY(cudaEventRecord(c->ev_begin, s));
    // progressive: the handle's running sum is both the prior and the output (each pixel is read, then written, by one thread)
    float4* d_sum = progressive ? c->d_prog.p : (linear_rgba ? c->d_sum.p : nullptr);
    int rc = render_on_device(c, spp, sample_offset, max_depth, flags, rgb ? c->d_rgb.p : nullptr, d_sum, s,
                              progressive && c->prog_spp > 0 ? c->d_prog.p : nullptr, c->prog_spp);
    if (rc) return rc;
    if (rgb) CU_TRY(cudaMemcpyAsync(rgb, c->d_rgb.p, npix * 3, cudaMemcpyDeviceToHost, s));
    if (linear_rgba) CU_TRY(cudaMemcpyAsync(linear_rgba, d_sum, npix * sizeof(float4), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaMemcpyAsync(c->h_rays, c->d_rays, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaEventRecord(c->ev_end, s));
    CU_TRY(cudaStreamSynchronize(s));
    c->stats.rays = c->h_rays[0]; c->stats.culled_paths = c->h_rays[1];
    c->stats_pending = false;
    CU_TRY(cudaEventElapsedTime(&c->stats.kernel_ms, c->ev_k0, c->ev_k1));
    CU_TRY(cudaEventElapsedTime(&c->stats.total_ms, c->ev_begin, c->ev_end));
    return FORMA_RT_OK;
}

}  // namespace

extern "C" {

int forma_rt_render(forma_rt* ctx, uint32_t spp, uint32_t sample_offset, uint32_t max_depth, uint8_t* rgb,
                    float* linear_rgba) {
    NEED_DEV(ctx);
    return render_to_host(ctx, spp, sample_offset, max_depth, ctx->flags, rgb, linear_rgba);
}

int forma_rt_update(forma_rt* ctx, uint8_t keys, float delta_time, uint8_t* rgb) {
    NEED_DEV(ctx);
    if (!rgb) return fail(FORMA_RT_EINVAL, "forma_rt_update: rgb is null");
    ctx->scene.camera.orbit(forma::keys_to_delta(keys, delta_time));  // tracer.rs:31-52
    int rc = render_to_host(ctx, 1, ctx->frame, 50, ctx->flags, rgb, nullptr);  // tracer.rs:54, MAX_DEPTH :10
    if (rc == FORMA_RT_OK) ctx->frame++;
    return rc;
}

// include/forma_fx.h: trace, widen RGB8 -> RGBA8 and run the filter chain on the tracer's stream; one D2H at the end
int forma_fx_run_traced(forma_fx* fx, forma_rt* ctx, uint32_t spp, uint32_t sample_offset, uint32_t max_depth,
                        float time, uint8_t* rgb, uint8_t* dst_rgba) {
    NEED_DEV(ctx);
    if (!fx || !dst_rgba) return fail(FORMA_RT_EINVAL, "forma_fx_run_traced: null handle or frame");
    if (forma::fx_device(fx) != ctx->device || forma::fx_width(fx) != ctx->width || forma::fx_height(fx) != ctx->height)
        return fail(FORMA_RT_EINVAL, "forma_fx_run_traced: tracer and filter handles differ in device or frame size");
    const size_t npix = (size_t)ctx->width * ctx->height;
    CU_TRY(ctx->d_rgb.ensure(npix * 3));
    cudaStream_t s = ctx->stream;
    CU_TRY(cudaEventRecord(ctx->ev_begin, s));
    int rc = render_on_device(ctx, spp, sample_offset, max_depth, ctx->flags, ctx->d_rgb.p, nullptr, s);
    if (rc) return rc;
    const uint8_t* d_out = nullptr;
    if ((rc = forma::fx_run_from_rgb8(fx, ctx->d_rgb.p, time, s, &d_out))) return rc;
    if (rgb) CU_TRY(cudaMemcpyAsync(rgb, ctx->d_rgb.p, npix * 3, cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaMemcpyAsync(dst_rgba, d_out, npix * 4, cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaMemcpyAsync(ctx->h_rays, ctx->d_rays, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaEventRecord(ctx->ev_end, s));
    CU_TRY(cudaStreamSynchronize(s));
    ctx->stats.rays = ctx->h_rays[0]; ctx->stats.culled_paths = ctx->h_rays[1];
    ctx->stats_pending = false;
    CU_TRY(cudaEventElapsedTime(&ctx->stats.kernel_ms, ctx->ev_k0, ctx->ev_k1));
    CU_TRY(cudaEventElapsedTime(&ctx->stats.total_ms, ctx->ev_begin, ctx->ev_end));
    return FORMA_RT_OK;
}

int forma_rt_set_partition(forma_rt* ctx, uint32_t rank, uint32_t world) {
    NEED_CTX(ctx);
    if (world == 0 || rank >= world) return fail(FORMA_RT_EINVAL, "forma_rt_set_partition: need rank < world");
    ctx->part_mod = world; ctx->part_rem = rank;
    ctx->prog_spp = 0;  // a running sum belongs to one partition
    return FORMA_RT_OK;
}

int forma_rt_accumulate_reset(forma_rt* ctx) {
    NEED_CTX(ctx);
    ctx->prog_spp = 0;
    return FORMA_RT_OK;
}

int forma_rt_accumulate(forma_rt* ctx, uint32_t spp_more, uint32_t max_depth, uint8_t* rgb, float* linear_rgba,
                        uint32_t* spp_total) {
    NEED_DEV(ctx);
    if ((uint64_t)ctx->prog_spp + spp_more > 0x7fffffffu) return fail(FORMA_RT_EINVAL, "forma_rt_accumulate: sample count overflow");
    // samples [prog_spp, prog_spp + spp_more) continue the running sum: bit-identical to one render of the total
    int rc = render_to_host(ctx, spp_more, ctx->prog_spp, max_depth, ctx->flags, rgb, linear_rgba, true);
    if (rc) return rc;
    ctx->prog_spp += spp_more;
    ctx->prog_cam = ctx->scene.camera.c;
    if (spp_total) *spp_total = ctx->prog_spp;
    return FORMA_RT_OK;
}

int forma_rt_update_progressive(forma_rt* ctx, uint8_t keys, float delta_time, uint8_t* rgb, uint32_t* spp_total) {
    NEED_DEV(ctx);
    if (!rgb) return fail(FORMA_RT_EINVAL, "forma_rt_update_progressive: rgb is null");
    ctx->scene.camera.orbit(forma::keys_to_delta(keys, delta_time));  // tracer.rs:31-52, every frame like the reference
    // a still camera and an untouched scene keep refining the frame; anything else starts over at sample 0
    if (ctx->scene_dirty || std::memcmp(&ctx->prog_cam, &ctx->scene.camera.c, sizeof(forma_rt_camera)) != 0) ctx->prog_spp = 0;
    return forma_rt_accumulate(ctx, 1, 50, rgb, nullptr, spp_total);
}

int forma_rt_srgb_encode_device(forma_rt* ctx, void* d_rgb, size_t nbytes, void* stream) {
    NEED_DEV(ctx);
    if (!d_rgb) return fail(FORMA_RT_EINVAL, "forma_rt_srgb_encode_device: d_rgb is null");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    CU_TRY(forma::launch_srgb_encode((uint8_t*)d_rgb, nbytes, ctx->srgb_lut, ctx->sm_count, s));
    return FORMA_RT_OK;
}

int forma_rt_srgb_table(uint8_t table[256]) {
    if (!table) return fail(FORMA_RT_EINVAL, "forma_rt_srgb_table: table is null");
    forma::srgb_encode_table(table);
    return FORMA_RT_OK;
}

int forma_rt_save_image(forma_rt* ctx, uint32_t sample, const char* path) {
    NEED_DEV(ctx);
    if (!path) return fail(FORMA_RT_EINVAL, "forma_rt_save_image: path is null");
    std::vector<uint8_t> rgb((size_t)ctx->width * ctx->height * 3);
    int rc = render_to_host(ctx, sample, 0, 50, ctx->flags, rgb.data(), nullptr);
    if (rc) return rc;
    return forma_rt_write_png(path, rgb.data(), ctx->width, ctx->height);
}

int forma_rt_primary_hits(forma_rt* ctx, int32_t* prim_id, float* t) {
    NEED_DEV(ctx);
    if (!prim_id || !t) return fail(FORMA_RT_EINVAL, "forma_rt_primary_hits: null output");
    int rc = ensure_scene(ctx);
    if (rc) return rc;
    const size_t npix = (size_t)ctx->width * ctx->height;
    CU_TRY(ctx->d_prim.ensure(npix));
    CU_TRY(ctx->d_t.ensure(npix));
    forma::PrimaryParams P{};
    P.cam = ctx->scene.camera.c; P.scene = ctx->view;
    P.width = ctx->width; P.height = ctx->height;
    P.prim_id = ctx->d_prim.p; P.t = ctx->d_t.p;
    cudaStream_t s = ctx->stream;
    CU_TRY(cudaEventRecord(ctx->ev_begin, s));
    CU_TRY(cudaEventRecord(ctx->ev_k0, s));
    CU_TRY(forma::launch_primary_hits(P, s));
    CU_TRY(cudaEventRecord(ctx->ev_k1, s));
    CU_TRY(cudaMemcpyAsync(prim_id, ctx->d_prim.p, npix * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaMemcpyAsync(t, ctx->d_t.p, npix * sizeof(float), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaEventRecord(ctx->ev_end, s));
    CU_TRY(cudaStreamSynchronize(s));
    ctx->stats.rays = npix; ctx->stats.paths = npix; ctx->stats.culled_paths = 0; ctx->stats.gpu_launches = 1;
    CU_TRY(cudaEventElapsedTime(&ctx->stats.kernel_ms, ctx->ev_k0, ctx->ev_k1));
    CU_TRY(cudaEventElapsedTime(&ctx->stats.total_ms, ctx->ev_begin, ctx->ev_end));
    return FORMA_RT_OK;
}

// ---- device-resident entry points ----------------------------------------------------------------------
int forma_rt_render_device(forma_rt* ctx, uint32_t spp, uint32_t sample_offset, uint32_t max_depth, void* d_rgb,
                           void* d_sum_rgba, void* stream) {
    NEED_DEV(ctx);
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    return render_on_device(ctx, spp, sample_offset, max_depth, ctx->flags, (uint8_t*)d_rgb, (float4*)d_sum_rgba, s);
}

int forma_rt_tonemap_device(forma_rt* ctx, const void* d_sum_rgba, uint32_t spp_total, uint64_t pixel_begin,
                            uint64_t pixel_end, void* d_rgb, void* stream) {
    NEED_DEV(ctx);
    if (!d_sum_rgba || !d_rgb || spp_total == 0) return fail(FORMA_RT_EINVAL, "forma_rt_tonemap_device: bad argument");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    CU_TRY(forma::launch_tonemap((const float4*)d_sum_rgba, spp_total, pixel_begin, pixel_end, (uint8_t*)d_rgb,
                                 ctx->sm_count, s));
    return FORMA_RT_OK;
}

int forma_rt_reduce_tonemap_device(forma_rt* ctx, const void* const* d_peer_sums, uint32_t n_peers, uint32_t spp_total,
                                   uint64_t pixel_begin, uint64_t pixel_end, void* d_rgb, void* d_sum_out,
                                   void* stream) {
    NEED_DEV(ctx);
    if (!d_peer_sums || n_peers == 0 || spp_total == 0 || (!d_rgb && !d_sum_out))
        return fail(FORMA_RT_EINVAL, "forma_rt_reduce_tonemap_device: bad argument");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    CU_TRY(forma::launch_reduce_tonemap((const float4* const*)d_peer_sums, (int)n_peers, spp_total, pixel_begin,
                                        pixel_end, (uint8_t*)d_rgb, (float4*)d_sum_out, ctx->sm_count, s));
    return FORMA_RT_OK;
}

int forma_rt_flag_set_device(forma_rt* ctx, void* d_flag, uint32_t value, void* stream) {
    NEED_DEV(ctx);
    if (!d_flag || ((uintptr_t)d_flag & 3u)) return fail(FORMA_RT_EINVAL, "forma_rt_flag_set_device: bad flag pointer");
    CU_TRY(forma::launch_flag_set((uint32_t*)d_flag, value, stream ? (cudaStream_t)stream : ctx->stream));
    return FORMA_RT_OK;
}

int forma_rt_flag_wait_device(forma_rt* ctx, const void* const* d_flags, uint32_t n, uint32_t value, void* stream) {
    NEED_DEV(ctx);
    if (!d_flags || n == 0 || n > 16) return fail(FORMA_RT_EINVAL, "forma_rt_flag_wait_device: 1..16 flags");
    for (uint32_t i = 0; i < n; ++i)
        if (!d_flags[i] || ((uintptr_t)d_flags[i] & 3u)) return fail(FORMA_RT_EINVAL, "forma_rt_flag_wait_device: bad flag pointer");
    if (*ctx->h_flag_err) return flag_timed_out(ctx);  // an earlier wait of this handle gave up: fail instead of queueing more
    CU_TRY(forma::launch_flag_wait((const uint32_t* const*)d_flags, (int)n, value, ctx->flag_timeout_ns, ctx->d_flag_err,
                                   stream ? (cudaStream_t)stream : ctx->stream));
    return FORMA_RT_OK;
}

int forma_rt_ipc_export(forma_rt* ctx, void* d_ptr, uint8_t handle[64]) {
    NEED_DEV(ctx);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaIpcMemHandle_t h;
    CU_TRY(cudaIpcGetMemHandle(&h, d_ptr));
    std::memcpy(handle, &h, 64);
    return FORMA_RT_OK;
}

int forma_rt_ipc_open(forma_rt* ctx, const uint8_t handle[64], void** d_ptr) {
    NEED_DEV(ctx);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    CU_TRY(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return FORMA_RT_OK;
}

int forma_rt_ipc_close(forma_rt* ctx, void* d_ptr) {
    NEED_DEV(ctx);
    CU_TRY(cudaIpcCloseMemHandle(d_ptr));
    return FORMA_RT_OK;
}

int forma_rt_device_alloc(forma_rt* ctx, size_t bytes, void** d_ptr) {
    NEED_DEV(ctx);
    if (!d_ptr) return fail(FORMA_RT_EINVAL, "forma_rt_device_alloc: d_ptr is null");
    CU_TRY(cudaMalloc(d_ptr, bytes ? bytes : 1));
    ctx->scratch.push_back(*d_ptr);
    CU_TRY(cudaMemset(*d_ptr, 0, bytes ? bytes : 1));  // flags start at 0; frames start black
    return FORMA_RT_OK;
}

int forma_rt_device_free(forma_rt* ctx, void* d_ptr) {
    NEED_DEV(ctx);
    for (size_t i = 0; i < ctx->scratch.size(); ++i)
        if (ctx->scratch[i] == d_ptr) {
            ctx->scratch.erase(ctx->scratch.begin() + i);
            CU_TRY(cudaFree(d_ptr));
            return FORMA_RT_OK;
        }
    return fail(FORMA_RT_EINVAL, "forma_rt_device_free: pointer not owned by this handle");
}

int forma_rt_synchronize(forma_rt* ctx) {
    NEED_DEV(ctx);
    CU_TRY(cudaStreamSynchronize(ctx->stream));
    if (*ctx->h_flag_err) return flag_timed_out(ctx);
    return FORMA_RT_OK;
}

int forma_rt_flag_error(forma_rt* ctx, uint32_t* flag_index_plus_one) {
    NEED_DEV(ctx);
    if (flag_index_plus_one) *flag_index_plus_one = *ctx->h_flag_err;
    return *ctx->h_flag_err ? flag_timed_out(ctx) : FORMA_RT_OK;
}

int forma_rt_host_alloc(forma_rt* ctx, size_t bytes, void** h_ptr) {
    NEED_DEV(ctx);
    if (!h_ptr) return fail(FORMA_RT_EINVAL, "forma_rt_host_alloc: h_ptr is null");
    CU_TRY(cudaMallocHost(h_ptr, bytes ? bytes : 1));
    ctx->pinned.push_back(*h_ptr);
    return FORMA_RT_OK;
}

int forma_rt_host_free(forma_rt* ctx, void* h_ptr) {
    NEED_DEV(ctx);
    for (size_t i = 0; i < ctx->pinned.size(); ++i)
        if (ctx->pinned[i] == h_ptr) {
            ctx->pinned.erase(ctx->pinned.begin() + i);
            CU_TRY(cudaFreeHost(h_ptr));
            return FORMA_RT_OK;
        }
    return fail(FORMA_RT_EINVAL, "forma_rt_host_free: pointer not owned by this handle");
}

int forma_rt_stats_get(forma_rt* ctx, forma_rt_stats* out) {
    NEED_CTX(ctx);
    if (!out) return fail(FORMA_RT_EINVAL, "forma_rt_stats_get: out is null");
    if (ctx->stats_pending && ctx->device >= 0) {  // a forma_rt_render_device launch has not been read back yet
        CU_TRY(cudaEventSynchronize(ctx->ev_k1));
        CU_TRY(cudaMemcpyAsync(ctx->h_rays, ctx->d_rays, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
        CU_TRY(cudaStreamSynchronize(ctx->stream));
        ctx->stats.rays = ctx->h_rays[0]; ctx->stats.culled_paths = ctx->h_rays[1];
        CU_TRY(cudaEventElapsedTime(&ctx->stats.kernel_ms, ctx->ev_k0, ctx->ev_k1));
        ctx->stats.total_ms = ctx->stats.kernel_ms;
        ctx->stats_pending = false;
    }
    *out = ctx->stats;
    return FORMA_RT_OK;
}

int forma_rt_walk_stats_get(forma_rt* ctx, uint32_t spp, uint32_t max_depth, forma_rt_walk_stats* out) {
    NEED_DEV(ctx);
    if (!out) return fail(FORMA_RT_EINVAL, "forma_rt_walk_stats_get: out is null");
    forma_rt_walk_stats ws{};
    ws.node_bytes = 32;
    int rc = ensure_scene(ctx);
    if (rc) return rc;
    cudaStream_t s = ctx->stream;
    const size_t npix = (size_t)ctx->width * ctx->height;
    CU_TRY(ctx->d_walk.ensure(2));
    CU_TRY(cudaMemsetAsync(ctx->d_walk.p, 0, 2 * sizeof(unsigned long long), s));
    CU_TRY(ctx->d_sum.ensure(npix));
    ctx->count_walk = true;
    rc = render_on_device(ctx, spp, 0, max_depth, ctx->flags, nullptr, ctx->d_sum.p, s);
    ctx->count_walk = false;
    if (rc) return rc;
    unsigned long long counts[2] = {0, 0}, rays = 0;
    CU_TRY(cudaMemcpyAsync(counts, ctx->d_walk.p, sizeof counts, cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaMemcpyAsync(&rays, ctx->d_rays, sizeof rays, cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaStreamSynchronize(s));
    ctx->stats_pending = false;  // an instrumented frame is not a timed one
    ws.rays = rays; ws.node_visits = counts[0]; ws.leaf_tests = counts[1];
    ws.bvh_nodes = ctx->view.bvh_nodes ? ctx->stats.bvh_nodes : 0;
    ws.leaf_bytes = ctx->stats.n_triangles && !ctx->stats.n_spheres ? 48 : 16;
    {   // streaming read out of L2: 32 MiB (a quarter of the 126 MB L2), one warm-up pass, then 8 timed passes
        const size_t bytes = 32u << 20;
        const uint32_t passes = 8;
        void* buf = nullptr;
        uint32_t* sink = nullptr;
        CU_TRY(cudaMalloc(&buf, bytes + 256));
        sink = (uint32_t*)((char*)buf + bytes);
        cudaError_t e = cudaMemsetAsync(buf, 0x5a, bytes + 256, s);
        size_t per_pass = 0;
        if (e == cudaSuccess) e = forma::launch_l2_read(buf, bytes, 1, sink, ctx->sm_count, s, &per_pass);
        float ms = 0.0f;
        for (int round = 0; round < 5 && e == cudaSuccess; ++round) {  // best of 5 warmed rounds: the peak is a pinned figure
            float r_ms = 0.0f;
            e = cudaEventRecord(ctx->ev_begin, s);
            if (e == cudaSuccess) e = forma::launch_l2_read(buf, bytes, passes, sink, ctx->sm_count, s, nullptr);
            if (e == cudaSuccess) e = cudaEventRecord(ctx->ev_end, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            if (e == cudaSuccess) e = cudaEventElapsedTime(&r_ms, ctx->ev_begin, ctx->ev_end);
            if (e == cudaSuccess && r_ms > 0.0f && (ms == 0.0f || r_ms < ms)) ms = r_ms;
        }
        cudaFree(buf);
        if (e != cudaSuccess) return fail(FORMA_RT_ECUDA, std::string("l2 read microbenchmark: ") + cudaGetErrorString(e));
        ws.l2_read_gbs = ms > 0.0f ? (float)((double)per_pass * passes / (ms * 1e-3) / 1e9) : 0.0f;
    }
    *out = ws;
    return FORMA_RT_OK;
}

int forma_rt_write_png(const char* path, const uint8_t* rgb, uint32_t width, uint32_t height) {
    std::string err;
    if (!forma::write_png_rgb8(path, rgb, width, height, err)) return fail(FORMA_RT_EIO, err);
    return FORMA_RT_OK;
}

}  // extern "C"

// ---- tracer::render_mt / save_image_mt (tracer.rs:83-158) --------------------------------------------------
namespace {

// one render_mt frame into c->d_mt_frame (width * 4 * (height / 4) * 3 bytes) on the handle's stream
int render_mt_on_device(forma_rt* c, uint32_t frame) {
    if (!c->mt_scene) {  // Scene::new's objects = get_simple_scene(), scanned (4 objects)
        forma_rt_options o{};
        o.struct_size = sizeof o;
        o.device = c->device; o.seed = c->seed; o.accel = FORMA_RT_ACCEL_BRUTE;
        int rc = forma_rt_create(&o, c->width, c->height, &c->mt_scene);
        if (rc) return rc;
    }
    int rc = ensure_scene(c->mt_scene);
    if (rc) return rc;
    cudaSetDevice(c->device);
    const size_t rows = 4 * (size_t)(c->height / 4);
    CU_TRY(c->d_mt_frame.ensure(std::max<size_t>(1, rows * c->width * 3)));
    forma::MtParams P{};
    P.cam = c->scene.camera.c;
    P.scene = c->mt_scene->view;
    P.width = c->width; P.height = c->height; P.frame = frame; P.max_depth = 50;  // tracer.rs:10
    P.rk.expand(c->seed);
    P.rgb = c->d_mt_frame.p; P.rays = c->d_rays;
    CU_TRY(forma::launch_render_mt(P, c->stream));
    return FORMA_RT_OK;
}

int average_mt_to_host(forma_rt* c, uint32_t sample, uint8_t* rgb) {
    if (sample == 0) return fail(FORMA_RT_EINVAL, "save_image_mt: sample must be >= 1");
    const size_t n = (size_t)c->width * c->height * 3, n_mt = (size_t)c->width * 4 * (c->height / 4) * 3;
    cudaStream_t s = c->stream;
    CU_TRY(c->d_mt_acc.ensure(n));
    CU_TRY(c->d_rgb.ensure(n));
    CU_TRY(cudaMemsetAsync(c->d_mt_acc.p, 0, n * sizeof(float), s));  // tracer.rs:139: vec![0.0; w * h * 3]
    CU_TRY(cudaMemsetAsync(c->d_rays, 0, 2 * sizeof(unsigned long long), s));
    CU_TRY(cudaEventRecord(c->ev_begin, s));
    CU_TRY(cudaEventRecord(c->ev_k0, s));
    for (uint32_t k = 0; k < sample; ++k) {  // tracer.rs:140-145
        if (int rc = render_mt_on_device(c, k)) return rc;
        CU_TRY(forma::launch_mt_accumulate(c->d_mt_frame.p, c->d_mt_acc.p, sample, n_mt, c->sm_count, s));
    }
    CU_TRY(forma::launch_mt_finish(c->d_mt_acc.p, c->d_rgb.p, n, c->sm_count, s));  // tracer.rs:148-154
    CU_TRY(cudaEventRecord(c->ev_k1, s));
    CU_TRY(cudaMemcpyAsync(rgb, c->d_rgb.p, n, cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaMemcpyAsync(c->h_rays, c->d_rays, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaEventRecord(c->ev_end, s));
    CU_TRY(cudaStreamSynchronize(s));
    c->stats.rays = c->h_rays[0]; c->stats.culled_paths = c->h_rays[1];
    c->stats.paths = (uint64_t)c->width * 4 * (c->height / 4) * sample;
    c->stats.gpu_launches = 2 * sample + 1;
    c->stats_pending = false;
    CU_TRY(cudaEventElapsedTime(&c->stats.kernel_ms, c->ev_k0, c->ev_k1));
    CU_TRY(cudaEventElapsedTime(&c->stats.total_ms, c->ev_begin, c->ev_end));
    return FORMA_RT_OK;
}

}  // namespace

extern "C" {

int forma_rt_render_mt(forma_rt* ctx, uint32_t frame, uint8_t* rgb) {
    NEED_DEV(ctx);
    if (!rgb) return fail(FORMA_RT_EINVAL, "forma_rt_render_mt: rgb is null");
    cudaStream_t s = ctx->stream;
    const size_t n_mt = (size_t)ctx->width * 4 * (ctx->height / 4) * 3;
    CU_TRY(cudaMemsetAsync(ctx->d_rays, 0, 2 * sizeof(unsigned long long), s));
    CU_TRY(cudaEventRecord(ctx->ev_begin, s));
    CU_TRY(cudaEventRecord(ctx->ev_k0, s));
    if (int rc = render_mt_on_device(ctx, frame)) return rc;
    CU_TRY(cudaEventRecord(ctx->ev_k1, s));
    if (n_mt) CU_TRY(cudaMemcpyAsync(rgb, ctx->d_mt_frame.p, n_mt, cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaMemcpyAsync(ctx->h_rays, ctx->d_rays, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaEventRecord(ctx->ev_end, s));
    CU_TRY(cudaStreamSynchronize(s));
    ctx->stats.rays = ctx->h_rays[0]; ctx->stats.culled_paths = ctx->h_rays[1];
    ctx->stats.paths = (uint64_t)ctx->width * 4 * (ctx->height / 4);
    ctx->stats.gpu_launches = n_mt ? 1 : 0;
    ctx->stats_pending = false;
    CU_TRY(cudaEventElapsedTime(&ctx->stats.kernel_ms, ctx->ev_k0, ctx->ev_k1));
    CU_TRY(cudaEventElapsedTime(&ctx->stats.total_ms, ctx->ev_begin, ctx->ev_end));
    return FORMA_RT_OK;
}

int forma_rt_render_mt_average(forma_rt* ctx, uint32_t sample, uint8_t* rgb) {
    NEED_DEV(ctx);
    if (!rgb) return fail(FORMA_RT_EINVAL, "forma_rt_render_mt_average: rgb is null");
    return average_mt_to_host(ctx, sample, rgb);
}

int forma_rt_save_image_mt(forma_rt* ctx, uint32_t sample, const char* path) {
    NEED_DEV(ctx);
    if (!path) return fail(FORMA_RT_EINVAL, "forma_rt_save_image_mt: path is null");
    std::vector<uint8_t> rgb((size_t)ctx->width * ctx->height * 3);
    if (int rc = average_mt_to_host(ctx, sample, rgb.data())) return rc;
    return forma_rt_write_png(path, rgb.data(), ctx->width, ctx->height);
}

}  // extern "C"

// include/forma_rt.h: forma_rt_render_device without a single allocation (no counterpart in the reference).
// Everything that allocates must have happened before: the scene upload, the
// LBVH build and the one-off div_const check are done by any earlier render call of the handle.
extern "C" int forma_rt_render_device_noalloc(forma_rt* ctx, uint32_t spp, uint32_t sample_offset, uint32_t max_depth,
                                              void* d_rgb, void* d_sum_rgba, void* stream) {
    NEED_DEV(ctx);
    if (spp == 0) return fail(FORMA_RT_EINVAL, "forma_rt_render_device_noalloc: spp must be >= 1");
    if (max_depth > (uint32_t)forma::kMaxDepthCap)
        return fail(FORMA_RT_EINVAL, "forma_rt_render_device_noalloc: max_depth exceeds the attenuation stack (64; the reference uses 50)");
    if (!d_rgb && !d_sum_rgba) return fail(FORMA_RT_EINVAL, "forma_rt_render_device_noalloc: no output buffer");
    const uint64_t npix = (uint64_t)ctx->width * ctx->height;
    if (npix > (uint64_t)FORMA_RT_NOALLOC_MAX_PIXELS)
        return fail(FORMA_RT_EINVAL, "forma_rt_render_device_noalloc: " + std::to_string(ctx->width) + " x " + std::to_string(ctx->height) +
                    " = " + std::to_string(npix) + " pixels exceed the static scratch (FORMA_RT_NOALLOC_MAX_PIXELS = " +
                    std::to_string(FORMA_RT_NOALLOC_MAX_PIXELS) + "); use forma_rt_render_device");
    if (ctx->scene_dirty || ctx->fast_div < 0)
        return fail(FORMA_RT_EINVAL, "forma_rt_render_device_noalloc: the scene is not resident on the device yet (uploading it "
                    "allocates): call forma_rt_render / forma_rt_render_device once after every scene change first");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    forma::RenderParams P{};
    P.cam = ctx->scene.camera.c;
    P.scene = ctx->view;
    P.width = ctx->width; P.height = ctx->height;
    P.spp = spp; P.sample_offset = sample_offset; P.max_depth = max_depth; P.flags = ctx->flags;
    P.rk.expand(ctx->seed);
    P.rgb = (uint8_t*)d_rgb; P.sum = (float4*)d_sum_rgba; P.rays = ctx->d_rays;
    P.part_mod = ctx->part_mod; P.part_rem = ctx->part_rem;
    P.part_pitch = (ctx->width + 31) / 32;
    while (std::gcd(P.part_pitch, P.part_mod) != 1u) ++P.part_pitch;
    P.refill = ctx->refill; P.leaf_batch = ctx->leaf_batch;
    P.inv_w = 1.0f / (float)ctx->width; P.inv_h = 1.0f / (float)ctx->height;
    P.fast_div = (uint32_t)ctx->fast_div;
    uint32_t launches = 0;
    CU_TRY(cudaMemsetAsync(ctx->d_rays, 0, 2 * sizeof(unsigned long long), s));
    CU_TRY(cudaEventRecord(ctx->ev_k0, s));
    CU_TRY(forma::launch_render_noalloc(P, ctx->device, ctx->sm_count, s, &launches));
    CU_TRY(cudaEventRecord(ctx->ev_k1, s));
    if (d_rgb && (ctx->flags & FORMA_RT_SRGB_OUT)) {  // as forma_rt_render_device: the table travels as a kernel parameter
        CU_TRY(forma::launch_srgb_encode((uint8_t*)d_rgb, (size_t)npix * 3, ctx->srgb_lut, ctx->sm_count, s));
        ++launches;
    }
    ctx->stats.gpu_launches = launches;
    ctx->stats.paths = npix * spp;  // under forma_rt_set_partition: an upper bound (the handle's own tiles only are traced)
    ctx->stats_pending = true;
    return FORMA_RT_OK;
}
