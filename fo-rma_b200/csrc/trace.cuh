// trace.cuh — device side of the path: RNG, camera ray, closest hit, scatter, sky.
// Compiled for sm_100a with -fmad=false: rustc never contracts a*b+c, IEEE division
// and square root are nvcc's defaults, so every value below is the f32 the reference would compute
// on the same inputs.  Reference lines are cited per function (paths relative to the fo-rma tree).
#pragma once
#include <cfloat>
#include <cstdint>

#include "device_scene.hpp"
#include "divconst.cuh"

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
// Philox4x32-7 (Salmon et al., SC'11: the smallest Crush-resistant round count; the CPU checker draws from the
// same generator with the same round count).  Key = seed; counter = (pixel, sample, slot, block).
// slot 0 = camera draws of the path, slot 1+d = scatter at depth d.  f32 = (u32 >> 8) * 2^-24.
struct Rand4 {
    float x, y, z, w;
};
// Round keys k + r*W are the same for every call of a launch, so the host expands them once
// (RoundKeys, device_scene.hpp) and the kernel reads them as constant-bank operands: a round is
// two IMAD.WIDE (hi and lo of each product in one instruction) and two 3-input LOP3.
#ifndef FORMA_PHILOX_ROUNDS
#define FORMA_PHILOX_ROUNDS 7
#endif
constexpr int kPhiloxRounds = FORMA_PHILOX_ROUNDS;  // RoundKeys holds 10 rounds' keys; the first kPhiloxRounds are used
// the rounds: four u32 words
__device__ __forceinline__ uint4 philox_raw4(const RoundKeys& rk, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
#pragma unroll
    for (int r = 0; r < kPhiloxRounds; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ rk.k[2 * r];
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ rk.k[2 * r + 1];
        c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ Rand4 philox_unit4(const RoundKeys& rk, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
    const uint4 c = philox_raw4(rk, c0, c1, c2, c3);
    const float s = 1.0f / 16777216.0f;
    Rand4 o;
    o.x = (float)(c.x >> 8) * s;
    o.y = (float)(c.y >> 8) * s;
    o.z = (float)(c.z >> 8) * s;
    o.w = (float)(c.w >> 8) * s;
    return o;
}

struct PathRng {
    const RoundKeys& rk;
    uint32_t pixel, sample;
    __device__ __forceinline__ Rand4 block(uint32_t slot, uint32_t blk) const {
        return philox_unit4(rk, pixel, sample, slot, blk);
    }
    // the same block as four 24-bit integers k (as floats): the unit float is k * 2^-24, exactly
    __device__ __forceinline__ float4 block_k(uint32_t slot, uint32_t blk) const {
        const uint4 c = philox_raw4(rk, pixel, sample, slot, blk);
        return make_float4((float)(c.x >> 8), (float)(c.y >> 8), (float)(c.z >> 8), (float)(c.w >> 8));
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
// camera.rs:63-71 for an accepted lens sample (px, py)
__device__ __forceinline__ void camera_ray_at(const forma_rt_camera& c, float px, float py, float s, float t,
                                              bool pinhole, float3& origin, float3& dir) {
    float3 offset = f3(0.0f, 0.0f, 0.0f);
    if (!pinhole) {
        float rdx = c.lens_radius * px, rdy = c.lens_radius * py;  // camera.rs:63
        offset = rdx * f3(c.u) + rdy * f3(c.v);                      // camera.rs:64
    }
    float3 pos = f3(c.position);
    origin = pos + offset;                                           // camera.rs:67
    dir = f3(c.lower_left_corner) + s * f3(c.horizontal) + t * f3(c.vertical) - pos - offset;  // :68-70
}

// utility.rs:4-13 on one RETRY block (tries 2j-1 and 2j, evaluated branch-free): true when one is
// accepted.  dot(p,p) with p.z = 0: (x*x + y*y) + 0*0 == x*x + y*y, a sum of squares is never -0.
__device__ __forceinline__ bool lens_retry(const Rand4& r, float& px, float& py) {
    px = 2.0f * r.x - 1.0f; py = 2.0f * r.y - 1.0f;
    const float qx = 2.0f * r.z - 1.0f, qy = 2.0f * r.w - 1.0f;
    const bool ok1 = !(px * px + py * py >= 1.0f), ok2 = !(qx * qx + qy * qy >= 1.0f);
    if (!ok1) { px = qx; py = qy; }
    return ok1 || ok2;
}

__device__ __forceinline__ void camera_ray(const forma_rt_camera& c, const PathRng& rng, const Rand4& r0, float s,
                                           float t, bool pinhole, float3& origin, float3& dir) {
    float px = 2.0f * r0.z - 1.0f, py = 2.0f * r0.w - 1.0f;
    if (!pinhole && px * px + py * py >= 1.0f) {
        uint32_t blk = 0;
        while (!lens_retry(rng.block(0, ++blk), px, py)) {}
    }
    camera_ray_at(c, px, py, s, t, pinhole, origin, dir);
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

// Winner of the segment being scanned.  The reference scans in list order with strict `<`
// (tracer.rs:196, sphere.rs:32), i.e. it keeps the lexicographic minimum of (t, list index).  Taking
// a candidate iff `t < closest || (t == closest && its list index < the winner's)` gives the same
// result in ANY visiting order, which is what lets spheres, triangles and the LBVH walk be scanned
// separately.  A winner of an earlier segment always has a smaller list index, so ties never beat it.
struct Winner {
    int kind = -1;  // -1 none in this segment, 0 sphere, 2 triangle
    int slot = 0;   // packed index
    int obj = 0;    // list index
};

// Sphere::hit (sphere.rs:23-51).  The candidate root does not depend on t_max: root_one if it
// exceeds t_min, else root_two; if root_one >= t_max then root_two >= root_one fails too.
// Returns the candidate (the caller checks t_min < root <= closest), or -1 when disc <= 0.
__device__ __forceinline__ float sphere_root(const float4 s, float3 o, float3 d, float a) {
    const float3 oc = o - xyz(s);
    const float b = dot(oc, d);
    const float c = dot(oc, oc) - s.w * s.w;
    const float disc = b * b - a * c;
    float root = -1.0f;
    if (disc > 0.0f) {
        const float sq = sqrtf(disc);
        root = (-b - sq) / a;
        if (!(root > kTMin)) root = (-b + sq) / a;
    }
    return root;
}
__device__ __forceinline__ void try_sphere(const SceneView& S, const float4* __restrict__ sph, int i, float3 o,
                                           float3 d, float a, float& closest, Winner& w) {
    const float root = sphere_root(sph[i], o, d, a);
    if (root > kTMin && root <= closest) {
        bool take = root < closest;
        int obj = -1;
        if (!take && w.kind >= 0) { obj = S.sphere_obj[i]; take = obj < w.obj; }
        if (take) { closest = root; w.kind = 0; w.slot = i; w.obj = obj >= 0 ? obj : S.sphere_obj[i]; }
    }
}

// Triangle extension (DESIGN.md "Scene mapping"): division-free inside test, one division for t.
// Returns t for a ray through the triangle (the caller checks t_min < t <= closest), else -1.
__device__ __forceinline__ float triangle_t(const float4* __restrict__ tri, int i, float3 o, float3 d) {
    const float4 q0 = tri[3 * i], q1 = tri[3 * i + 1], q2 = tri[3 * i + 2];
    const float3 e1 = xyz(q1), e2 = xyz(q2);
    const float3 pvec = cross(d, e2);
    const float det = dot(e1, pvec);
    const float3 tvec = o - xyz(q0);
    const float uu = dot(tvec, pvec);
    const bool pos = det > 0.0f;
    float t = -1.0f;
    // first half of the inside test before paying for qvec
    if (det != 0.0f && (pos ? uu >= 0.0f : uu <= 0.0f)) {
        const float3 qvec = cross(tvec, e1);
        const float vv = dot(d, qvec);
        const bool inside = pos ? (vv >= 0.0f && uu + vv <= det) : (vv <= 0.0f && uu + vv >= det);
        if (inside) t = dot(e2, qvec) / det;
    }
    return t;
}
__device__ __forceinline__ void try_triangle(const SceneView& S, const float4* __restrict__ tri, int i, float3 o,
                                             float3 d, float& closest, Winner& w) {
    const float t = triangle_t(tri, i, o, d);
    if (t > kTMin && t <= closest) {
        bool take = t < closest;
        int obj = -1;
        if (!take && w.kind >= 0) { obj = S.tri_obj[i]; take = obj < w.obj; }
        if (take) { closest = t; w.kind = 2; w.slot = i; w.obj = obj >= 0 ? obj : S.tri_obj[i]; }
    }
}

// The same acceptance rule on a UNIFIED primitive index p (< n_spheres: sphere p, else triangle p - n_spheres),
// with the winner kept as one such index (-1 = none): what render_tree_kernel carries through its walk.  The
// list index (Scene.objects position) is only looked up on a tie in t.
__device__ __forceinline__ int prim_obj(const SceneView& S, int p) {
    return p < S.n_spheres ? S.sphere_obj[p] : S.tri_obj[p - S.n_spheres];
}
__device__ __forceinline__ void try_prim(const SceneView& S, const float4* __restrict__ sph, const float4* __restrict__ tri,
                                         int p, float3 o, float3 d, float a, float& closest, int& winner) {
    const float t = p < S.n_spheres ? sphere_root(sph[p], o, d, a) : triangle_t(tri, p - S.n_spheres, o, d);
    if (t > kTMin && t <= closest) {
        bool take = t < closest;
        if (!take && winner >= 0) take = prim_obj(S, p) < prim_obj(S, winner);
        if (take) { closest = t; winner = p; }
    }
}
__device__ __forceinline__ Winner winner_of(const SceneView& S, int p) {
    Winner w;
    if (p >= 0) {
        const bool is_sphere = p < S.n_spheres;
        w.kind = is_sphere ? 0 : 2; w.slot = is_sphere ? p : p - S.n_spheres; w.obj = prim_obj(S, p);
    }
    return w;
}

// What the last accepting hit() of a segment left in the shared HitRecord (sphere.rs:33-36,42-45).
__device__ __forceinline__ void finalize_winner(const float4* __restrict__ sph, const float4* __restrict__ tri,
                                                float3 o, float3 d, const Winner& w, HitState& h) {
    if (w.kind >= 0) {
        h.t = h.closest;
        h.p = o + h.t * d;  // ray.rs:30-32
        if (w.kind == 0) {
            const float4 s = sph[w.slot];
            h.n = (h.p - xyz(s)) / s.w;  // sphere.rs:35,44
        } else {
            float3 n = f3(tri[3 * w.slot].w, tri[3 * w.slot + 1].w, tri[3 * w.slot + 2].w);
            if (dot(d, n) > 0.0f) n = n * -1.0f;
            h.n = n;
        }
        h.obj = w.obj;
    }
}

// Plane::hit (plane.rs:24-44).  Q4: the range test is on denom, never on t; Q5: t and p are written
// before the extent test and stay clobbered on a reject.
__device__ __forceinline__ void plane_hit(const SceneView& S, int k, float3 o, float3 d, HitState& h) {
    const float3 position = xyz(S.planes[3 * k]);
    const float3 orientation = xyz(S.planes[3 * k + 1]);
    const float3 size = xyz(S.planes[3 * k + 2]);
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
            h.obj = S.plane_obj[k];
        }
    }
}

// Conservative whole-scene cull: the line o + t d stays farther than R from the centre C of the
// bounding sphere of every finite sphere and triangle iff a (|oc|^2 - R^2) - (oc.d)^2 > 0.  The margin
// 2^-16 a (|oc|^2 + R^2) exceeds the rounding error of this expression (<= 2^-20 of the same magnitude)
// plus the largest miss distance at which sphere.rs:25-29's disc or the triangle test can still accept
// (same order), so nothing the scan would accept is skipped.  NaN / inf inputs make the comparison
// false -> full scan.
__device__ __forceinline__ bool scene_culled(const SceneView& S, float3 o, float3 d, float a) {
    const float3 oc = o - f3(S.cull_center);
    const float cc = dot(oc, oc), b = dot(oc, d);
    return a * (cc - S.cull_r2) - b * b > 1.52587890625e-05f * (a * (cc + S.cull_r2));
}

// One 32-byte node record as ONE 256-bit load (LDG.E.256, sm_100+): a warp's lanes sit at different nodes,
// so every load instruction costs one L1 look-up per lane, and the walk is bound by exactly those requests
// (profiles/r01o_render_tree_kernel_D_ncu.md).  Layout (lbvh.cu): six words of u16 box planes on the tree's
// 16-bit grid - (Llo.x | Llo.y << 16), (Llo.z | Lhi.x << 16), (Lhi.y | Lhi.z << 16), the same for R - and the
// two child codes.
struct QNode {
    uint32_t w0, w1, w2, w3, w4, w5;
    int cL, cR;
};
__device__ __forceinline__ QNode load_node(const uint4* __restrict__ nodes, int node) {
    QNode n;
    const uint4* p = nodes + 2 * (size_t)node;
    asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(n.w0), "=r"(n.w1), "=r"(n.w2), "=r"(n.w3), "=r"(n.w4), "=r"(n.w5), "=r"(n.cL), "=r"(n.cR) : "l"(p));
    return n;
}

// A ray's view of the tree's grid.  A box plane q (grid units) on axis k lies at qmin_k + q step_k; padded by
// `pad` towards the outside its ray parameter is
//     t = (qmin_k -+ pad + q step_k - o_k) / d_k = q * A_k + B(lo|hi)_k,   A_k = step_k / d_k,
//     Blo_k = (qmin_k - pad - o_k) / d_k,  Bhi_k = (qmin_k + pad - o_k) / d_k
// - one FFMA per plane straight from the 16-bit value.  Exactness: a child box is visited iff the ray's
// interval through the PADDED box reaches [0, closest]; pad = coef(ray) * q_tree + eps with
// coef = 2^-20 L^2, q_tree = max 1/|r| over the spheres of the tree, L = a bound on |o - c| for every sphere the
// walk could still accept: the smallest of (i) |o - C| + R for the tree's bounding sphere (as 2 (|o - C|^2 + R^2)),
// (ii) the distance to the farthest corner of the tree's box, (iii) |d| closest_so_far + r_max at the START of the
// walk - a root <= closest_so_far puts the sphere's centre that close, and on scenes with a floor every downward
// ray arrives with the floor's hit already found (the outsized primitives are scanned first).  It exceeds the largest miss
// distance at which the f32 test of sphere.rs:25-29 can still report disc > 0: term by term (u = 2^-24; b = oc.d
// +-3u |oc||d|, b b +-7u a L^2, c = oc.oc - r r +-5u L^2, a c +-9u a L^2, the final subtraction +-u a L^2) the
// computed disc is within 17 u a L^2 of a (r^2 - rho^2), which is below -2 a r delta for a ray that misses by delta,
// so delta > 8.5 u L^2 / r is safe and the pad is 1.9x that.  Round 1 used (i) alone; (ii) alone is no tighter
// (33.5 against 33.2 visits per ray on config D); (iii) is what shrinks the pad (profiles/r06d_*, r06e_*).
// eps = 2^-16 x scene scale covers the triangle test's rounding and that of the two
// roundings in q * A + B (each <= 1 ulp of |o_k / d_k| in t, i.e. <= 2^-23 |o_k| in space).  The grid itself is
// conservative by construction (lbvh.cu rounds lo down and hi up by a whole extra cell).  So no primitive the
// scan would accept is culled, and the tie rule makes the visiting order irrelevant.
#ifndef FORMA_PAD_NEAR
#define FORMA_PAD_NEAR 1   // 0 = without the closest_so_far bound on |o - c| (A/B switch)
#endif
struct RayGrid {
    float ax, ay, az;        // A_k
    float blx, bly, blz;     // Blo_k
    float bhx, bhy, bhz;     // Bhi_k
};
__device__ __forceinline__ RayGrid ray_grid(const SceneView& S, float3 o, float3 d, float a, float closest) {
    const float huge = 1e30f;
    const float ix = fabsf(d.x) < 1e-30f ? copysignf(huge, d.x) : 1.0f / d.x;
    const float iy = fabsf(d.y) < 1e-30f ? copysignf(huge, d.y) : 1.0f / d.y;
    const float iz = fabsf(d.z) < 1e-30f ? copysignf(huge, d.z) : 1.0f / d.z;
    // |o - c|^2 of every sphere that can still be accepted is bounded three ways; the smallest wins
    const float3 oc = o - f3(S.bvh_center);
    float L2 = 2.0f * (dot(oc, oc) + S.bvh_r2);            // centre inside the tree's bounding sphere
    {
        const float lx = S.bvh_qmin[0] - o.x, hx = __fmaf_rn(65535.0f, S.bvh_qstep[0], S.bvh_qmin[0]) - o.x;
        const float ly = S.bvh_qmin[1] - o.y, hy = __fmaf_rn(65535.0f, S.bvh_qstep[1], S.bvh_qmin[1]) - o.y;
        const float lz = S.bvh_qmin[2] - o.z, hz = __fmaf_rn(65535.0f, S.bvh_qstep[2], S.bvh_qmin[2]) - o.z;
        L2 = fminf(L2, fmaxf(lx * lx, hx * hx) + fmaxf(ly * ly, hy * hy) + fmaxf(lz * lz, hz * hz));  // farthest corner of its box
    }
#if FORMA_PAD_NEAR
    {   // a root <= closest_so_far puts the centre within |d| closest + r_max of the origin (NaN / inf: fminf keeps L2)
        const float ln = __fmaf_rn(sqrtf(a), closest, S.bvh_rmax);
        L2 = fminf(L2, 1.0001f * (ln * ln));
    }
#endif
    const float pad = __fmaf_rn(9.5367432e-7f * L2, S.bvh_q, S.bvh_eps);  // 2^-20 L^2 q + eps
    RayGrid g;
    g.ax = S.bvh_qstep[0] * ix; g.ay = S.bvh_qstep[1] * iy; g.az = S.bvh_qstep[2] * iz;
    g.blx = (S.bvh_qmin[0] - pad - o.x) * ix; g.bhx = (S.bvh_qmin[0] + pad - o.x) * ix;
    g.bly = (S.bvh_qmin[1] - pad - o.y) * iy; g.bhy = (S.bvh_qmin[1] + pad - o.y) * iy;
    g.blz = (S.bvh_qmin[2] - pad - o.z) * iz; g.bhz = (S.bvh_qmin[2] + pad - o.z) * iz;
    return g;
}
// [near, far] of the ray through both padded child boxes of a node, clipped to [0, closest]
__device__ __forceinline__ void slab_pair(const QNode& n, const RayGrid& g, float closest, float& nearL, float& farL,
                                          float& nearR, float& farR) {
    float t0 = __fmaf_rn((float)(n.w0 & 0xffffu), g.ax, g.blx), t1 = __fmaf_rn((float)(n.w1 >> 16), g.ax, g.bhx);
    nearL = fminf(t0, t1); farL = fmaxf(t0, t1);
    t0 = __fmaf_rn((float)(n.w0 >> 16), g.ay, g.bly); t1 = __fmaf_rn((float)(n.w2 & 0xffffu), g.ay, g.bhy);
    nearL = fmaxf(nearL, fminf(t0, t1)); farL = fminf(farL, fmaxf(t0, t1));
    t0 = __fmaf_rn((float)(n.w1 & 0xffffu), g.az, g.blz); t1 = __fmaf_rn((float)(n.w2 >> 16), g.az, g.bhz);
    nearL = fmaxf(fmaxf(nearL, fminf(t0, t1)), 0.0f); farL = fminf(fminf(farL, fmaxf(t0, t1)), closest);
    t0 = __fmaf_rn((float)(n.w3 & 0xffffu), g.ax, g.blx); t1 = __fmaf_rn((float)(n.w4 >> 16), g.ax, g.bhx);
    nearR = fminf(t0, t1); farR = fmaxf(t0, t1);
    t0 = __fmaf_rn((float)(n.w3 >> 16), g.ay, g.bly); t1 = __fmaf_rn((float)(n.w5 & 0xffffu), g.ay, g.bhy);
    nearR = fmaxf(nearR, fminf(t0, t1)); farR = fminf(farR, fmaxf(t0, t1));
    t0 = __fmaf_rn((float)(n.w4 & 0xffffu), g.az, g.blz); t1 = __fmaf_rn((float)(n.w5 >> 16), g.az, g.bhz);
    nearR = fmaxf(fmaxf(nearR, fminf(t0, t1)), 0.0f); farR = fminf(fminf(farR, fmaxf(t0, t1)), closest);
}

// Primary-beam cull, once per pixel.  Every primary ray of pixel (x, y) is (1 - tau) o + tau F, tau >= 0,
// with o on the lens (|o - pos| <= r_o) and F on the pixel's patch of the focus plane (|F - F_c| <= r_F), so
// it stays within |1 - tau| r_o + tau r_F <= r_o + tau (r_o + r_F) of the central ray pos + tau (F_c - pos).
// If that growing ball never reaches the bounding sphere (C, R) of all spheres and triangles -
//   A tau^2 + 2 B tau + Cq > 0 for all tau >= 0,  A = D.D - k^2, B = m.D - R' k, Cq = m.m - R'^2,
//   D = F_c - pos, m = pos - C, k = r_o + r_F, R' = R + r_o -
// no sample of the pixel can hit anything and get_color (tracer.rs:189-219) is its sky branch.  Evaluated
// in f64 with R, r_o, r_F inflated by 2^-10 relative (+ 2^-12 of the distances involved): orders of magnitude
// above the f32 rounding of camera.rs:62-72 and above the miss distance at which sphere.rs:25-29 / the
// triangle test can still accept (see scene_culled).  NaN / inf -> false -> the full scan runs.
__device__ __forceinline__ bool beam_culled(const forma_rt_camera& c, const SceneView& S, uint32_t x, uint32_t y,
                                            uint32_t width, uint32_t height) {
    if (S.n_planes != 0 || !(S.cull_r2 < FLT_MAX)) return false;
    const double w = (double)width, h = (double)height;
    const double sc = ((double)x + 0.5) / w, tc = ((double)(height - y) + 0.5) / h;
    double D[3], m[3], lu = 0.0, lv = 0.0, lh = 0.0, lvv = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double Fc = (double)c.lower_left_corner[i] + sc * (double)c.horizontal[i] + tc * (double)c.vertical[i];
        D[i] = Fc - (double)c.position[i];
        m[i] = (double)c.position[i] - (double)S.cull_center[i];
        lu += (double)c.u[i] * (double)c.u[i]; lv += (double)c.v[i] * (double)c.v[i];
        lh += (double)c.horizontal[i] * (double)c.horizontal[i]; lvv += (double)c.vertical[i] * (double)c.vertical[i];
    }
    const double slack = 1.0 + 9.765625e-4;  // 2^-10
    const double mm = m[0] * m[0] + m[1] * m[1] + m[2] * m[2];
    const double R = sqrt((double)S.cull_r2);
    const double dist = sqrt(mm) + R;
    const double r_o = fabs((double)c.lens_radius) * (sqrt(lu) + sqrt(lv)) * slack + 2.44140625e-4 * dist;
    const double r_F = 0.5 * (sqrt(lh) / w + sqrt(lvv) / h) * slack * slack + 2.44140625e-4 * sqrt(D[0] * D[0] + D[1] * D[1] + D[2] * D[2]);
    const double k = r_o + r_F, Rp = R * slack + r_o + 2.44140625e-4 * dist;
    const double A = D[0] * D[0] + D[1] * D[1] + D[2] * D[2] - k * k;
    const double B = m[0] * D[0] + m[1] * D[1] + m[2] * D[2] - Rp * k;
    const double Cq = mm - Rp * Rp;
    return Cq > 0.0 && A > 0.0 && (B >= 0.0 || B * B < A * Cq);
}

// LBVH walk over the primitives the tree indexes (lbvh.cu), one ray to completion (primary_hits_kernel).
__device__ __forceinline__ void bvh_closest(const SceneView& S, const float4* __restrict__ sph,
                                            const float4* __restrict__ tri, float3 o, float3 d, float a,
                                            HitState& h, Winner& w) {
    const RayGrid g = ray_grid(S, o, d, a, h.closest);  // a = d.d; closest: what the scan has found before the tree
    int stack[64];
    int sp = 0;
    int node = 0;
    for (;;) {
        const QNode n = load_node(S.bvh_nodes, node);
        float nearL, farL, nearR, farR;
        slab_pair(n, g, h.closest, nearL, farL, nearR, farR);
        const bool hitL = nearL <= farL, hitR = nearR <= farR;
        int cL = n.cL, cR = n.cR;
        int next = -1;   // internal node to descend into
        bool have = false;
        if (hitL && hitR) {
            if (nearR < nearL) { const int t = cL; cL = cR; cR = t; }  // cL = nearer child
            // leaves are tested on the spot, internal nodes are descended / pushed
            if (cL < 0) { const int p = ~cL; if (p < S.n_spheres) try_sphere(S, sph, p, o, d, a, h.closest, w); else try_triangle(S, tri, p - S.n_spheres, o, d, h.closest, w); }
            else { next = cL; have = true; }
            if (cR < 0) { const int p = ~cR; if (p < S.n_spheres) try_sphere(S, sph, p, o, d, a, h.closest, w); else try_triangle(S, tri, p - S.n_spheres, o, d, h.closest, w); }
            else if (have) { stack[sp++] = cR; }
            else { next = cR; have = true; }
        } else if (hitL || hitR) {
            const int c = hitL ? cL : cR;
            if (c < 0) { const int p = ~c; if (p < S.n_spheres) try_sphere(S, sph, p, o, d, a, h.closest, w); else try_triangle(S, tri, p - S.n_spheres, o, d, h.closest, w); }
            else { next = c; have = true; }
        }
        if (have) { node = next; continue; }
        if (sp == 0) break;
        node = stack[--sp];
    }
}

// ---- resumable LBVH walk (render_tree_kernel) ------------------------------------------------------
// The same padded walk as bvh_closest, cut into single-node steps so that a warp can advance all its
// lanes together, stop when enough of them have finished, refill those with their next ray, and resume.
// Stack entries carry the entry distance of the pushed child so that taking an entry skips subtrees
// that the hits found meanwhile have put out of range.
constexpr int kWalkDone = (int)0x80000000;  // never a leaf code: leaves are ~l with l < 2^31 - 2
constexpr int kWalkPop = (int)0x80000001;   // convergent walk: this lane takes its next node from the stack

// Walk stack: (child, entry distance) pairs.  The first kSmemStack levels of every thread live in
// shared memory (level-major, one 8-byte word per thread: conflict-free), deeper levels in local
// memory - so the hot part of the stacks does not compete with the tree for L1.  The shared-window
// address of the thread's column is computed ONCE and kept in a register (`sbase`, handed through an
// opaque asm so the compiler cannot rematerialise it from S2R + shifts at every push and pop: that
// cost 8 instructions per stack access in round 1's kernel, profiles/r05a_*).
// 8 levels in shared memory (8 KB per CTA): 16 cost L1 capacity for no gain, 24 lost 8 % (profiles/r04i_occ.txt, r04w)
#ifndef FORMA_SMEM_STACK
#define FORMA_SMEM_STACK 8
#endif
#ifndef FORMA_WALK_PREFETCH
#define FORMA_WALK_PREFETCH 0
#endif
constexpr int kSmemStack = FORMA_SMEM_STACK, kStackCap = 64, kWalkBlock = 128;
struct WalkStack {
    uint32_t sbase;  // shared-window byte address of level 0 of this thread; level l at sbase + l * 8 * kWalkBlock
    int2* local;     // levels >= kSmemStack, indexed by the level itself (the first kSmemStack entries are never touched)
    __device__ __forceinline__ void init(int2* smem_col, int2* local_levels) {
        const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_col);
        asm volatile("mov.u32 %0, %1;" : "=r"(sbase) : "r"(a));
        local = local_levels;
    }
    // entry `sp` <- (child, dist) when `on` (one predicated store; the local levels are a rarely taken branch)
    __device__ __forceinline__ void push_if(bool on, int sp, int child, float dist) const {
        const uint32_t addr = sbase + (uint32_t)sp * (8u * kWalkBlock);
        asm volatile("{ .reg .pred p; setp.ne.b32 p, %3, 0; @p st.shared.v2.b32 [%0], {%1, %2}; }"
                     ::"r"(addr), "r"(child), "r"(__float_as_int(dist)), "r"((int)(on && sp < kSmemStack)) : "memory");
        if (on && sp >= kSmemStack) local[sp] = make_int2(child, __float_as_int(dist));
    }
    // (child, dist) <- entry `sp` when `on`, else both are left as they are
    __device__ __forceinline__ void load_if(bool on, int sp, int& child, float& dist) const {
        const uint32_t addr = sbase + (uint32_t)sp * (8u * kWalkBlock);
        int c = child, f = __float_as_int(dist);
        asm volatile("{ .reg .pred p; setp.ne.b32 p, %3, 0; @p ld.shared.v2.b32 {%0, %1}, [%2]; }"
                     : "+r"(c), "+r"(f) : "r"(addr), "r"((int)(on && sp < kSmemStack)) : "memory");
        if (on && sp >= kSmemStack) { const int2 e = local[sp]; c = e.x; f = e.y; }
        child = c; dist = __int_as_float(f);
    }
};

// A lane that needs a node takes ONE stack entry: the entry's child if its entry distance is still in range,
// else it stays kWalkPop (and takes the next entry in the next iteration); an empty stack ends the walk.
__device__ __forceinline__ void walk_pop(int& node, int& sp, const WalkStack& stack, float closest) {
    const bool popping = node == kWalkPop;
    const bool take = popping && sp > 0;
    sp -= take ? 1 : 0;
    int child = 0;
    float dist = 0.0f;
    stack.load_if(take, sp, child, dist);
    const int next = take ? (dist <= closest ? child : kWalkPop) : kWalkDone;
    node = popping ? next : node;
}

// One internal node: test both child boxes, descend into the nearer hit child (pushing the other), or ask
// for the next stack entry (kWalkPop, taken by the caller's next iteration).
__device__ __forceinline__ void walk_inner(const SceneView& S, const RayGrid& g, int& node, int& sp, const WalkStack& stack,
                                           float closest) {
    const QNode n = load_node(S.bvh_nodes, node);
    float nearL, farL, nearR, farR;
    slab_pair(n, g, closest, nearL, farL, nearR, farR);
    const bool hitL = nearL <= farL, hitR = nearR <= farR;
    const bool both = hitL && hitR;
    const bool r_first = hitR && (!hitL || nearR < nearL);  // which child is entered when at least one is hit
    const int enter = r_first ? n.cR : n.cL, other = r_first ? n.cL : n.cR;
    // both hit: the other child waits on the stack with ITS entry distance, the larger of the two
    stack.push_if(both, sp, other, fmaxf(nearL, nearR));
    sp += both ? 1 : 0;
    node = (hitL || hitR) ? enter : kWalkPop;
#if FORMA_WALK_PREFETCH
    // the child this lane enters is known ~30 instructions (vote, loop, pop logic) before its record is loaded:
    // start the L2 -> L1 transfer now
    if (node >= 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(S.bvh_nodes + 2 * (size_t)node));
#endif
}

// MODE selects what a kernel instantiation carries: the brute-force scan with the primitives staged
// in shared memory, the same scan reading them through L1/L2, or the LBVH walk (plus the linear scan
// of the few "big" primitives the tree leaves out).  The scan kernels contain no tree code at all.
enum : int { kScanSmem = 0, kScanGlobal = 1, kTree = 2 };

template <int MODE>
__device__ __forceinline__ void closest_hit(const SceneView& S, const float4* __restrict__ sph,
                                            const float4* __restrict__ tri, float3 o, float3 d, HitState& h) {
    h.t = 0.0f; h.p = f3(0.0f, 0.0f, 0.0f); h.n = f3(0.0f, 0.0f, 0.0f);
    h.closest = FLT_MAX; h.obj = -1;
    const float a = dot(d, d);  // sphere.rs:25
    const bool culled = scene_culled(S, o, d, a);
    for (int sg = 0; sg < S.n_segments; ++sg) {
        Segment g = S.seg0;  // the first (usually only) segment travels in the constant bank
        if (sg > 0) g = S.segments[sg];
        Winner w;
        int sph_end = g.sph_end, tri_end = g.tri_end;
        bool tree = false;
        if constexpr (MODE == kTree) {
            // the tree holds all but the first big_* primitives of the (single) run
            tree = g.sph_end + g.tri_end > g.sph_begin + g.tri_begin;
            if (tree) { sph_end = S.big_spheres; tri_end = S.big_tris; }
        }
        if (!culled) {
            for (int i = g.sph_begin; i < sph_end; ++i) try_sphere(S, sph, i, o, d, a, h.closest, w);
            for (int i = g.tri_begin; i < tri_end; ++i) try_triangle(S, tri, i, o, d, h.closest, w);
            if constexpr (MODE == kTree) {
                if (tree) bvh_closest(S, sph, tri, o, d, a, h, w);
            }
        }
        finalize_winner(sph, tri, o, d, w, h);
        if (g.plane >= 0) plane_hit(S, g.plane, o, d, h);
    }
}

// ---- scatter (sphere.rs:53-152, plane.rs:46-122) ----------------------------------------------------
// The part of scatter that follows the random draws: `rs` = the accepted random_in_unit_sphere sample (unused by the
// dielectric), `rw` = get_random() (lane w of block 0; used by the dielectric only).  Returns false when the path is
// absorbed (metal, sphere.rs:104).
__device__ __forceinline__ bool scatter_finish(const SceneView& S, float3 d, float3 hp, float3 hn, int hobj, float3 rs,
                                               float rw, float3& new_o, float3& new_d) {
    const int km = S.kindmat[hobj];
    const int kind = km & 0xff, material = km >> 8;
    new_o = hp;
    const bool dielectric = material == 2 && kind == FORMA_RT_SPHERE;
    if (material == 1) {  // metal: sphere.rs:91-105 == plane.rs:108-122
        float fuzz = S.material[hobj].w;
        float3 reflected = reflect(unit(d), hn);
        new_d = reflected + fuzz * rs;
        return dot(new_d, hn) > 0.0f;
    }
    if (dielectric) {  // sphere.rs:107-145 (spheres only, plane.rs:52-59)
        const float ref_idx = 1.3f;
        float3 outward_normal = f3(0.0f, 0.0f, 0.0f);
        float3 reflected = reflect(d, hn);
        float ni_over_nt, cosine, reflect_prob;
        float3 refracted = f3(0.0f, 0.0f, 0.0f);
        float dn = dot(d, hn);
        if (dn > 0.0f) {
            outward_normal = outward_normal - hn;
            ni_over_nt = ref_idx;
            cosine = ref_idx * dn / length(d);
        } else {
            outward_normal = hn;
            ni_over_nt = 1.0f / ref_idx;
            cosine = -dn / length(d);
        }
        if (refract(d, outward_normal, ni_over_nt, refracted)) reflect_prob = schlick(cosine, ref_idx);
        else reflect_prob = 1.0f;
        new_d = (rw < reflect_prob) ? reflected : refracted;  // get_random(): lane w of block 0
        return true;
    }
    // lambertian (sphere.rs:84-89, plane.rs:101-106) and "light" (sphere.rs:147-152): same ray
    float3 target = hp + hn + rs;
    new_d = target - hp;
    return true;
}
// true: list entry `obj` scatters as a dielectric (one get_random() draw, no unit-sphere sample)
__device__ __forceinline__ bool is_dielectric(const SceneView& S, int obj) {
    const int km = S.kindmat[obj];
    return (km >> 8) == 2 && (km & 0xff) == FORMA_RT_SPHERE;
}

// Returns false when the path is absorbed (metal, sphere.rs:104).  `slot` = 1 + depth.
__device__ __forceinline__ bool scatter(const SceneView& S, const PathRng& rng, uint32_t slot, float3 d,
                                        const HitState& h, float3& new_o, float3& new_d) {
    const Rand4 r = rng.block(slot, 0);
    // metal (sphere.rs:97, also at fuzz 0: Q7), lambertian and "light" all draw random_in_unit_sphere: ONE rejection
    // loop for the lanes of all three, instead of one per material branch
    float3 rs = f3(0.0f, 0.0f, 0.0f);
    if (!is_dielectric(S, h.obj)) rs = random_in_unit_sphere(rng, slot, r);
    return scatter_finish(S, d, h.p, h.n, h.obj, rs, r.w, new_o, new_d);
}

// attenuation of list entry `obj` (sphere.rs:87,101,116,150): colour, or one for a sphere "light"
__device__ __forceinline__ float3 attenuation(const SceneView& S, int obj) {
    const int km = S.kindmat[obj];
    if ((km >> 8) == 3 && (km & 0xff) == FORMA_RT_SPHERE) return f3(1.0f, 1.0f, 1.0f);
    const float4 m = S.material[obj];
    return f3(m.x, m.y, m.z);
}

// A path that ended black (absorbed by a metal, sphere.rs:104, or cut at MAX_DEPTH, tracer.rs:205-209)
// stays black on the way back up: finite attenuation x (+-0) = +-0, and the per-pixel sum starts at +0, so
// adding -0 instead of +0 never changes a bit of it.  Only NaN / inf colours could turn the zero into NaN
// (Q10), hence the scene flag.  Saves the 50-level walk back up for every path of a closed scene.
__device__ __forceinline__ bool black_path(const SceneView& S, float3 pc) {
    return S.finite_materials && pc.x == 0.0f && pc.y == 0.0f && pc.z == 0.0f;
}

// tracer.rs:213-217
__device__ __forceinline__ float3 sky(float3 d) {
    float y = d.y / length(d);  // unit_vector().y
    float t = 0.5f * (y + 1.0f);
    return (1.0f - t) * f3(1.0f, 1.0f, 1.0f) + t * f3(0.5f, 0.7f, 1.0f);
}

// Rust `as u8` (tracer.rs:74-76,181-183): truncate, saturate, NaN -> 0
__device__ __forceinline__ uint32_t as_u8(float f) { return min(__float2uint_rz(f), 255u); }

}  // namespace forma
