// trace.cuh — device side of the path: RNG, camera ray, closest hit, scatter, sky.
// Compiled for sm_100a with -fmad=false: rustc never contracts a*b+c, IEEE division
// and square root are nvcc's defaults, so every value below is the f32 the reference would compute
// on the same inputs.  Reference lines are cited per function (paths relative to the fo-rma tree).
#pragma once
#include <cfloat>
#include <cstdint>

#include "device_scene.hpp"

namespace forma {

// ---- Vec3 (cpu_ray_tracer/primitives.rs:50-149) on float3 ----------------------------------------
__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 f3(const float* p) { return make_float3(p[0], p[1], p[2]); }
__device__ __forceinline__ float3 xyz(float4 v) { return make_float3(v.x, v.y, v.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator/(float3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 cross(float3 a, float3 b) {  // primitives.rs:62-68
    return f3(a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float length(float3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ float3 unit(float3 a) { return a / length(a); }

// ---- counter-based RNG (stands in for rand::thread_rng, utility.rs:5,16,33) -----------------------
// Philox4x32-10 (Salmon et al., SC'11).  Key = seed; counter = (pixel, sample, slot, block).
// slot 0 = camera draws of the path, slot 1+d = scatter at depth d.  f32 = (u32 >> 8) * 2^-24.
struct Rand4 {
    float x, y, z, w;
};
// Round keys k + r*W are the same for every call of a launch, so the host expands them once
// (RoundKeys, device_scene.hpp) and the kernel reads them as constant-bank operands: a round is
// two IMAD.WIDE (hi and lo of each product in one instruction) and two 3-input LOP3.
__device__ __forceinline__ Rand4 philox_unit4(const RoundKeys& rk, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ rk.k[2 * r];
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ rk.k[2 * r + 1];
        c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    }
    const float s = 1.0f / 16777216.0f;
    Rand4 o;
    o.x = (float)(c0 >> 8) * s;
    o.y = (float)(c1 >> 8) * s;
    o.z = (float)(c2 >> 8) * s;
    o.w = (float)(c3 >> 8) * s;
    return o;
}

struct PathRng {
    const RoundKeys& rk;
    uint32_t pixel, sample;
    __device__ __forceinline__ Rand4 block(uint32_t slot, uint32_t blk) const {
        return philox_unit4(rk, pixel, sample, slot, blk);
    }
};

// utility.rs:15-25 random_in_unit_sphere: try k = lanes x,y,z of block k of `slot`
__device__ __forceinline__ float3 random_in_unit_sphere(const PathRng& rng, uint32_t slot, const Rand4& first) {
    float3 p = 2.0f * f3(first.x, first.y, first.z) - f3(1.0f, 1.0f, 1.0f);
    uint32_t blk = 0;
    while (dot(p, p) >= 1.0f) {
        ++blk;
        Rand4 r = rng.block(slot, blk);
        p = 2.0f * f3(r.x, r.y, r.z) - f3(1.0f, 1.0f, 1.0f);
    }
    return p;
}

// utility.rs:27-30
__device__ __forceinline__ float3 reflect(float3 v, float3 n) { return v - 2.0f * dot(v, n) * n; }

// utility.rs:37-47
__device__ __forceinline__ bool refract(float3 v, float3 n, float ni_over_nt, float3& refracted) {
    float3 uv = unit(v);
    float dt = dot(uv, n);
    float discriminant = 1.0f - ni_over_nt * ni_over_nt * (1.0f - dt * dt);
    if (discriminant > 0.0f) {
        refracted = ni_over_nt * (uv - dt * n) - sqrtf(discriminant) * n;
        return true;
    }
    return false;
}

// utility.rs:49-54; powi(5) = x * ((x*x)*(x*x)) (compiler-rt __powisf2 order)
__device__ __forceinline__ float schlick(float cosine, float ref_idx) {
    float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
    r0 = r0 * r0;
    float x = 1.0f - cosine;
    float x2 = x * x;
    float x4 = x2 * x2;
    return r0 + (1.0f - r0) * (x * x4);
}

// ---- camera ray (camera.rs:62-72) ------------------------------------------------------------------
// `r0` is block 0 of slot 0: (xi_u, xi_v, lens try 0).  Further lens tries: block j >= 1 holds tries
// 2j-1 (x,y) and 2j (z,w).
__device__ __forceinline__ void camera_ray(const forma_rt_camera& c, const PathRng& rng, const Rand4& r0, float s,
                                           float t, bool pinhole, float3& origin, float3& dir) {
    float3 offset = f3(0.0f, 0.0f, 0.0f);
    if (!pinhole) {
        float px = 2.0f * r0.z - 1.0f, py = 2.0f * r0.w - 1.0f;
        // dot(p,p) with p.z = 2*0 - 0 = 0: (x*x + y*y) + 0*0
        uint32_t blk = 0;
        bool second_half = true;
        Rand4 r = r0;
        while ((px * px + py * py) + 0.0f >= 1.0f) {
            if (second_half) { ++blk; r = rng.block(0, blk); second_half = false; px = 2.0f * r.x - 1.0f; py = 2.0f * r.y - 1.0f; }
            else { second_half = true; px = 2.0f * r.z - 1.0f; py = 2.0f * r.w - 1.0f; }
        }
        float rdx = c.lens_radius * px, rdy = c.lens_radius * py;  // camera.rs:63
        offset = rdx * f3(c.u) + rdy * f3(c.v);                      // camera.rs:64
    }
    float3 pos = f3(c.position);
    origin = pos + offset;                                           // camera.rs:67
    dir = f3(c.lower_left_corner) + s * f3(c.horizontal) + t * f3(c.vertical) - pos - offset;  // :68-70
}

// ---- closest hit (tracer.rs:190-200 over sphere.rs:23-51, plane.rs:24-44, triangle extension) ------
struct HitState {
    float t;        // HitRecord.t as scatter will see it (plane.rs:28 may clobber it, Q5)
    float3 p;       // HitRecord.p
    float3 n;       // HitRecord.normal
    float closest;  // closest_so_far
    int obj;        // list index of temp_obj, -1 = none
};

constexpr float kTMin = 0.001f;  // tracer.rs:191

template <bool SMEM>
__device__ __forceinline__ void closest_hit(const SceneView& S, const float4* __restrict__ sph,
                                            const float4* __restrict__ tri, float3 o, float3 d, HitState& h) {
    h.t = 0.0f; h.p = f3(0.0f, 0.0f, 0.0f); h.n = f3(0.0f, 0.0f, 0.0f);
    h.closest = FLT_MAX; h.obj = -1;
    const float a = dot(d, d);  // sphere.rs:25
    for (int sg = 0; sg < S.n_segments; ++sg) {
        const Segment g = S.segments[sg];
        int win_kind = -1, win_slot = 0;
        // spheres of the segment, list order (sphere.rs:23-51)
        for (int i = g.sph_begin; i < g.sph_end; ++i) {
            const float4 s = sph[i];
            float3 oc = o - xyz(s);
            float b = dot(oc, d);
            float c = dot(oc, oc) - s.w * s.w;
            float disc = b * b - a * c;
            if (disc > 0.0f) {
                float sq = sqrtf(disc);
                float root = (-b - sq) / a;
                if (root > kTMin && root < h.closest) {
                    h.closest = root; win_kind = 0; win_slot = i;
                } else {
                    root = (-b + sq) / a;
                    if (root > kTMin && root < h.closest) { h.closest = root; win_kind = 0; win_slot = i; }
                }
            }
        }
        // triangles of the segment (extension; DESIGN.md "Triangle extension")
        for (int i = g.tri_begin; i < g.tri_end; ++i) {
            const float4 q0 = tri[3 * i], q1 = tri[3 * i + 1], q2 = tri[3 * i + 2];
            float3 e1 = xyz(q1), e2 = xyz(q2);
            float3 pvec = cross(d, e2);
            float det = dot(e1, pvec);
            float3 tvec = o - xyz(q0);
            float uu = dot(tvec, pvec);
            bool pos = det > 0.0f;
            // first half of the inside test before paying for qvec
            if (!(det != 0.0f) || !(pos ? uu >= 0.0f : uu <= 0.0f)) continue;
            float3 qvec = cross(tvec, e1);
            float vv = dot(d, qvec);
            bool inside = pos ? (vv >= 0.0f && uu + vv <= det) : (vv <= 0.0f && uu + vv >= det);
            if (!inside) continue;
            float t = dot(e2, qvec) / det;
            if (t > kTMin) {
                bool take = t < h.closest;
                if (!take && t == h.closest && win_kind == 0) take = S.tri_obj[i] < S.sphere_obj[win_slot];
                if (take) { h.closest = t; win_kind = 2; win_slot = i; }
            }
        }
        if (win_kind >= 0) {  // what the last accepting hit() left in the shared HitRecord
            h.t = h.closest;
            h.p = o + h.t * d;  // ray.rs:30-32
            if (win_kind == 0) {
                const float4 s = sph[win_slot];
                h.n = (h.p - xyz(s)) / s.w;  // sphere.rs:35,44
                h.obj = S.sphere_obj[win_slot];
            } else {
                float3 n = f3(tri[3 * win_slot].w, tri[3 * win_slot + 1].w, tri[3 * win_slot + 2].w);
                if (dot(d, n) > 0.0f) n = n * -1.0f;
                h.n = n;
                h.obj = S.tri_obj[win_slot];
            }
        }
        if (g.plane >= 0) {  // plane.rs:24-44 (Q4: range test on denom; Q5: t, p written before the extent test)
            const float3 position = xyz(S.planes[3 * g.plane]);
            const float3 orientation = xyz(S.planes[3 * g.plane + 1]);
            const float3 size = xyz(S.planes[3 * g.plane + 2]);
            float denom = dot(orientation, d);
            if (denom > kTMin && denom < h.closest) {
                float3 plane_to_ray = position - o;
                h.t = dot(plane_to_ray, orientation) / denom;
                h.p = o + h.t * d;
                if (h.p.x > position.x - size.x && h.p.x < position.x + size.x &&
                    h.p.y > position.y - size.y && h.p.y < position.y + size.y &&
                    h.p.z > position.z - size.z && h.p.z < position.z + size.z) {
                    h.n = orientation * -1.0f;
                    h.closest = h.t;  // tracer.rs:197
                    h.obj = S.plane_obj[g.plane];
                }
            }
        }
    }
}

// ---- scatter (sphere.rs:53-152, plane.rs:46-122) ----------------------------------------------------
// Returns false when the path is absorbed (metal, sphere.rs:104).  `slot` = 1 + depth.
__device__ __forceinline__ bool scatter(const SceneView& S, const PathRng& rng, uint32_t slot, float3 d,
                                        const HitState& h, float3& new_o, float3& new_d) {
    const int km = S.kindmat[h.obj];
    const int kind = km & 0xff, material = km >> 8;
    const Rand4 r = rng.block(slot, 0);
    new_o = h.p;
    if (material == 1) {  // metal: sphere.rs:91-105 == plane.rs:108-122
        float fuzz = S.material[h.obj].w;
        float3 reflected = reflect(unit(d), h.n);
        new_d = reflected + fuzz * random_in_unit_sphere(rng, slot, r);
        return dot(new_d, h.n) > 0.0f;
    }
    if (material == 2 && kind == FORMA_RT_SPHERE) {  // dielectric: sphere.rs:107-145 (spheres only, plane.rs:52-59)
        const float ref_idx = 1.3f;
        float3 outward_normal = f3(0.0f, 0.0f, 0.0f);
        float3 reflected = reflect(d, h.n);
        float ni_over_nt, cosine, reflect_prob;
        float3 refracted = f3(0.0f, 0.0f, 0.0f);
        float dn = dot(d, h.n);
        if (dn > 0.0f) {
            outward_normal = outward_normal - h.n;
            ni_over_nt = ref_idx;
            cosine = ref_idx * dn / length(d);
        } else {
            outward_normal = h.n;
            ni_over_nt = 1.0f / ref_idx;
            cosine = -dn / length(d);
        }
        if (refract(d, outward_normal, ni_over_nt, refracted)) reflect_prob = schlick(cosine, ref_idx);
        else reflect_prob = 1.0f;
        new_d = (r.w < reflect_prob) ? reflected : refracted;  // get_random(): lane w of block 0
        return true;
    }
    // lambertian (sphere.rs:84-89, plane.rs:101-106) and "light" (sphere.rs:147-152): same ray
    float3 target = h.p + h.n + random_in_unit_sphere(rng, slot, r);
    new_d = target - h.p;
    return true;
}

// attenuation of list entry `obj` (sphere.rs:87,101,116,150): colour, or one for a sphere "light"
__device__ __forceinline__ float3 attenuation(const SceneView& S, int obj) {
    const int km = S.kindmat[obj];
    if ((km >> 8) == 3 && (km & 0xff) == FORMA_RT_SPHERE) return f3(1.0f, 1.0f, 1.0f);
    const float4 m = S.material[obj];
    return f3(m.x, m.y, m.z);
}

// tracer.rs:213-217
__device__ __forceinline__ float3 sky(float3 d) {
    float y = d.y / length(d);  // unit_vector().y
    float t = 0.5f * (y + 1.0f);
    return (1.0f - t) * f3(1.0f, 1.0f, 1.0f) + t * f3(0.5f, 0.7f, 1.0f);
}

// get_color (tracer.rs:189-219), recursion unrolled: the attenuations are multiplied innermost
// first, exactly as the recursion returns them, so the list indices of the hit objects are kept on
// a small per-path stack.  `rays` counts closest-hit queries.
template <bool SMEM>
__device__ __forceinline__ float3 get_color(const SceneView& S, const float4* sph, const float4* tri,
                                            const PathRng& rng, float3 o, float3 d, uint32_t max_depth,
                                            uint32_t& rays) {
    int stack[kMaxDepthCap];
    uint32_t depth = 0;
    float3 col;
    for (;;) {
        HitState h;
        ++rays;
        closest_hit<SMEM>(S, sph, tri, o, d, h);
        if (h.obj < 0) { col = sky(d); break; }
        if (depth >= max_depth) { col = f3(0.0f, 0.0f, 0.0f); break; }  // tracer.rs:205,208-209
        float3 no, nd;
        if (!scatter(S, rng, 1u + depth, d, h, no, nd)) { col = f3(0.0f, 0.0f, 0.0f); break; }
        stack[depth] = h.obj;
        ++depth;
        o = no; d = nd;
    }
    while (depth > 0) {  // tracer.rs:206-207 on the way back up
        --depth;
        col = attenuation(S, stack[depth]) * col;
    }
    return col;
}

// Rust `as u8` (tracer.rs:74-76,181-183): truncate, saturate, NaN -> 0
__device__ __forceinline__ uint32_t as_u8(float f) { return min(__float2uint_rz(f), 255u); }

}  // namespace forma
