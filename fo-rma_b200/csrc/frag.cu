// frag.cu — the reference's two fragment-shader ray tracers (src/shaders/rt.wgsl, src/shaders/rt2.wgsl) as one CUDA
// pass per frame, and the forma_frag C ABI (include/forma_frag.h).  SURVEY.md 8(f) row f-4; DESIGN.md 8d.
//
// One thread per pixel.  A warp covers an 8 x 4 pixel tile (neighbouring pixels follow neighbouring paths: 92-95 % of
// the lanes are live per loop trip on the shaders' own 1600 x 1200 frame), a block 4 warps = a 16 x 8 tile.  Both
// scenes are compile-time constants: every sphere centre, radius and material is an immediate operand, nothing is
// read from memory, and the only traffic is the frame going out (16 B f32 + 4 B unorm per pixel, one full 128-byte
// line per tile row for the f32 frame).  The library is built with -fmad=false / IEEE div and sqrt, so each operation
// rounds as the CPU restatement's does and the two agree bit for bit.
//
// What bounds it: rt.wgsl's loop runs up to 1000 trips (rt.wgsl:8,134) and a few pixels use them all (a mirror ray
// offset by EPSILON into a sphere keeps bouncing inside it), one dependent intersect -> reflect chain per trip; the
// frame cannot finish before that chain does.  DESIGN.md 8d has the measured split.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <new>
#include <string>

#include "../../include/forma_frag.h"
#include "postfx.hpp"

namespace {

// ---- fixed f32 log2 / exp2: the recipe of include/forma_frag.h's pow ------------------------------------------------
// log2: reduce to [sqrt(1/2), sqrt(2)), s = f / (2 + f), polynomial in s^2 (fdlibm's log2f scheme and constants),
// product with 1 / ln 2 split hi / lo.
__device__ __forceinline__ float log2_fixed(float x) {
    uint32_t ix = __float_as_uint(x);
    int k = 0;
    if (ix < 0x00800000u) {
        if (ix == 0) return __uint_as_float(0xff800000u);
        k = -25; x *= 33554432.0f; ix = __float_as_uint(x);
    } else if (ix >= 0x7f800000u) {
        if (ix == 0x7f800000u) return x;
        if (ix == 0x80000000u) return __uint_as_float(0xff800000u);
        return __uint_as_float(0x7fc00000u);
    }
    ix += 0x3f800000u - 0x3f3504f3u;
    k += (int)(ix >> 23) - 0x7f;
    ix = (ix & 0x007fffffu) + 0x3f3504f3u;
    x = __uint_as_float(ix);
    const float f = x - 1.0f;
    const float s = f / (2.0f + f);
    const float z = s * s;
    const float w = z * z;
    const float t1 = w * (0.40000972152f + w * 0.24279078841f);
    const float t2 = z * (0.66666662693f + w * 0.28498786688f);
    const float R = t2 + t1;
    const float hfsq = 0.5f * f * f;
    float hi = f - hfsq;
    hi = __uint_as_float(__float_as_uint(hi) & 0xfffff000u);
    const float lo = (f - hi) - hfsq + s * (hfsq + R);
    const float ivln2hi = 1.4428710938f, ivln2lo = -1.7605285393e-04f;
    return (lo + hi) * ivln2lo + lo * ivln2hi + hi * ivln2hi + (float)k;
}

// exp2: k = floor(x + 1/2), r = x - k, e^(r ln 2) by a degree-7 Taylor polynomial (Horner), scaled by 2^k in two steps.
__device__ __forceinline__ float exp2_fixed(float x) {
    if (!(x == x)) return x;
    if (x >= 128.0f) return __uint_as_float(0x7f800000u);
    if (x < -150.0f) return 0.0f;
    const float k = floorf(x + 0.5f);
    const float r = x - k;
    const float t = r * 0.69314718056f;
    float p = 1.0f / 5040.0f;
    p = p * t + 1.0f / 720.0f;
    p = p * t + 1.0f / 120.0f;
    p = p * t + 1.0f / 24.0f;
    p = p * t + 1.0f / 6.0f;
    p = p * t + 0.5f;
    p = p * t + 1.0f;
    p = p * t + 1.0f;
    const int ki = (int)k;
    const int k1 = ki >> 1, k2 = ki - k1;
    return p * __uint_as_float((uint32_t)(k1 + 127) << 23) * __uint_as_float((uint32_t)(k2 + 127) << 23);
}

// not inlined: eight copies of ~75 instructions made `no_instruction` rt2's top stall; one called copy is 4 % faster at 4K
__device__ __noinline__ float pow_wgsl(float x, float y) { return exp2_fixed(y * log2_fixed(x)); }

// ---- WGSL value types and builtins ---------------------------------------------------------------------------------
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 splat(float s) { return V3{s, s, s}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ V3 operator-(float s, V3 a) { return v3(s - a.x, s - a.y, s - a.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 normalize(V3 v) { return v / sqrtf(dot(v, v)); }
__device__ __forceinline__ V3 reflect(V3 e1, V3 e2) { return e1 - (2.0f * dot(e2, e1)) * e2; }
// fmaxf / fminf are PTX max / min: the number when one side is NaN, +0 > -0
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__device__ __forceinline__ float smoothstep(float lo, float hi, float x) {
    const float t = clampf((x - lo) / (hi - lo), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
__device__ __forceinline__ V3 mix(V3 a, V3 b, float t) { return a * (1.0f - t) + b * t; }

struct Ray { V3 origin, direction; };

// per-frame values the host computes once (forma_frag.h: constants are evaluated in f32)
struct FrameConst {
    float res_x, res_y, aspect;
    V3 sun_sky;    // rt.wgsl:105  normalize(vec3(1.0, 0.5, 0.5))
    V3 sun_light;  // rt.wgsl:131  normalize(vec3(1.0, 0.5, 0.0))
    V3 light2;     // rt2.wgsl:144 normalize(vec3(cos(time / 10f), 0.75, sin(time / 10f)))
    V3 ambient2;   // rt2.wgsl:141 vec3(0.6, 0.8, 1.0) * INTENSITY / GAMMA
    float inv_gamma2;  // rt2.wgsl:186 1.0 / GAMMA
};

// ==== rt.wgsl =======================================================================================================
namespace rt1 {
constexpr float EPSILON = 4e-4f;  // rt.wgsl:7
constexpr int SAMPLES = 1000;     // rt.wgsl:8
constexpr int kGroupTail = 4;     // a warp with at most this many live paths finishes them on 8-lane groups

// Hit (rt.wgsl:25-29); material = (color, f0), f0 < 0 = no hit
struct Hit { float t; V3 normal; V3 color; float f0; };

// The four primitive tests of a query are written without branches (selects on the accept condition): they are
// independent of one another, and in one basic block the scheduler interleaves their dependent chains.  That is what
// a lone long-running lane needs - its loop trip is a latency chain, not an issue-slot budget.

// rt.wgsl:77-85 `compare` after rt.wgsl:43-52 `intersect_plane` (Plane(0., vec3(0., 1., 0.), ...), rt.wgsl:94):
// true when `compare` would take the candidate, whose t is returned
__device__ __forceinline__ bool plane_t(float at, const Ray& r, float& t) {
    const V3 n = v3(0.0f, 1.0f, 0.0f);
    const float dotnd = dot(n, r.direction);
    t = -(dot(r.origin, n) + 0.0f) / dotnd;
    return !(dotnd > 0.0f) && t < at;
}

// rt.wgsl:77-85 `compare` after rt.wgsl:54-73 `intersect_sphere`; every sphere has radius 1 (rt.wgsl:91-93)
__device__ __forceinline__ bool sphere_t(float at, const Ray& r, V3 center, float& t) {
    const float radius = 1.0f;
    const V3 op = center - r.origin;
    const float b = dot(op, r.direction);
    const float det = b * b - dot(op, op) + radius * radius;
    // det < 0 is rejected below whatever sq is; 1 keeps sqrtf on its fast path there (a negative, zero or
    // subnormal argument takes the slow one)
    const float sq = sqrtf(det < 0.0f ? 1.0f : det);
    const float t0 = b - sq;
    t = t0 < 0.0f ? b + sq : t0;
    return !(det < 0.0f) && !(t < 0.0f) && t < at;
}

// The same test with the shader's early returns, for the full-warp phase: there the issue slots are the budget, and a
// warp whose lanes all miss a sphere (the sky, the floor) skips its square root and the rest
__device__ __forceinline__ bool sphere_tb(float at, const Ray& r, V3 center, float& t) {
    const V3 op = center - r.origin;
    const float b = dot(op, r.direction);
    const float det = b * b - dot(op, op) + 1.0f * 1.0f;
    if (det < 0.0f) return false;
    const float sq = sqrtf(det);
    t = b - sq;
    if (t < 0.0f) t = b + sq;
    return !(t < 0.0f) && t < at;
}

// the winner's normal and material (rt.wgsl:91-94): centres differ only in x (-2, 0, 2 = 2 * which - 4), radius 1;
// which: 0 plane, 1..3 spheres, < 0 no hit
__device__ __forceinline__ Hit surface(int which, float t, const Ray& ray) {
    Hit hit;
    hit.t = t;
    const V3 center = v3((float)(2 * which - 4), 1.0f, 0.0f);
    const V3 ns = (ray.origin + t * ray.direction - center) / 1.0f;
    hit.normal = which == 0 ? v3(0.0f, 1.0f, 0.0f) : ns;
    hit.color = which == 0 ? v3(0.5f, 0.4f, 0.3f) : v3(which == 1 ? 1.0f : 0.0f, which == 2 ? 1.0f : 0.0f, which == 3 ? 1.0f : 0.0f);
    hit.f0 = which == 0 ? 0.04f : which == 1 ? 0.04f : which == 2 ? 0.1f : 0.5f;
    if (which < 0) { hit.normal = splat(0.0f); hit.color = splat(-1.0f); hit.f0 = -1.0f; }
    return hit;
}

// rt.wgsl:88-102
__device__ __forceinline__ Hit intersect_scene(const Ray& ray) {
    float best = 1e10f, t;
    int which = -1;
    if (plane_t(best, ray, t)) { best = t; which = 0; }
    if (sphere_tb(best, ray, v3(-2.0f, 1.0f, 0.0f), t)) { best = t; which = 1; }
    if (sphere_tb(best, ray, v3(0.0f, 1.0f, 0.0f), t)) { best = t; which = 2; }
    if (sphere_tb(best, ray, v3(2.0f, 1.0f, 0.0f), t)) { best = t; which = 3; }
    return surface(which, best, ray);
}

// the shadow query of rt.wgsl:121-126 only asks whether anything is hit: `compare` accepts a candidate whenever its
// t is below the initial 1e10, and the first acceptance already makes f0 >= 0
__device__ __forceinline__ bool occluded(const Ray& r) {
    float t;
    bool any = plane_t(1e10f, r, t);
    any |= sphere_tb(1e10f, r, v3(-2.0f, 1.0f, 0.0f), t);
    any |= sphere_tb(1e10f, r, v3(0.0f, 1.0f, 0.0f), t);
    any |= sphere_tb(1e10f, r, v3(2.0f, 1.0f, 0.0f), t);
    return any;
}

// rt.wgsl:104-111
__device__ __forceinline__ V3 get_sky_color(V3 d, const FrameConst& fc) {
    const float transition = pow_wgsl(smoothstep(0.02f, 0.5f, d.y), 0.4f);
    const V3 sky = 2e0f * mix(v3(0.52f, 0.77f, 1.0f), v3(0.12f, 0.43f, 1.0f), transition);
    const V3 sun = splat(1.0f) * pow_wgsl(fabsf(dot(d, fc.sun_sky)), 5000.0f);
    return sky + sun;
}

__device__ __forceinline__ float pow5(float x) { return x * x * x * x * x; }  // rt.wgsl:113-115

// rt.wgsl:117-119
__device__ __forceinline__ float fresnel(V3 h, V3 v, float f0) {
    return pow5(1.0f - clampf(dot(h, v), 0.0f, 1.0f)) * (1.0f - f0) + f0;
}

// a pixel's loop state (rt.wgsl:129-131 + the loop counter)
struct Path { Ray ray; V3 accum, attenuation; int i; };

// rt.wgsl:137-149, the light of a hit: `dir` = the direction the ray arrived with, `shadowed` = the shadow query of
// rt.wgsl:122 hit something
__device__ __forceinline__ void shade(Path& p, const Hit& hit, V3 dir, bool shadowed, const FrameConst& fc) {
    const float f = fresnel(hit.normal, -dir, hit.f0);
    V3 incoming = splat(0.0f);
    if (!shadowed) incoming = incoming + clampf(dot(hit.normal, fc.sun_light), 0.0f, 1.0f) * splat(1e3f);
    p.accum = p.accum + hit.color;
    p.accum = p.accum + (1.0f - f) * p.attenuation * hit.color * incoming;
    p.attenuation = p.attenuation * f;
}

// rt.wgsl:150-151, the mirror ray
__device__ __forceinline__ Ray bounce(const Ray& ray, const Hit& hit) {
    const V3 d = reflect(ray.direction, hit.normal);
    return Ray{ray.origin + hit.t * ray.direction + EPSILON * d, d};
}

// rt.wgsl:122 with rt.wgsl:139
__device__ __forceinline__ Ray shadow_ray(const Ray& ray, const Hit& hit, const FrameConst& fc) {
    const V3 hit_pos = ray.origin + hit.t * ray.direction;
    return Ray{hit_pos + EPSILON * fc.sun_light, fc.sun_light};
}

// one trip of rt.wgsl:134-156 on one lane; false when the path has ended
__device__ __forceinline__ bool trip(Path& p, const FrameConst& fc, uint32_t& rays) {
    const Hit hit = intersect_scene(p.ray);
    ++rays;
    if (hit.f0 >= 0.0f) {
        ++rays;
        shade(p, hit, p.ray.direction, occluded(shadow_ray(p.ray, hit, fc)), fc);
        p.ray = bounce(p.ray, hit);
        return ++p.i < SAMPLES;
    }
    p.accum = p.accum + get_sky_color(p.ray.direction, fc);
    return false;
}

// The tail.  rt.wgsl's loop runs up to 1000 trips and the frame holds pixels that use them all: one dependent chain
// per trip (query -> normal -> mirror ray) on a lane whose 31 neighbours have long finished.  Once a warp is down to
// <= kGroupTail live paths, each path moves to a group of 8 lanes, every one of which keeps the whole path state.
// The loop is software-pipelined - the next ray needs only the hit, not its light - so trip k's primary query and trip
// k-1's shadow query are independent, and they run SIDE BY SIDE: lane k of the group's first quad tests primitive k
// (plane, s1, s2, s3) against the ray, lane k of its second quad tests primitive k against the previous hit's shadow
// ray - one pass through the test code for both.  The first quad takes the minimum (ties to the earlier primitive, as
// the sequential `compare` of rt.wgsl:77-85 does) and hands it to the second; a ballot ORs the shadow tests.  The
// light of trip k-1 is added before anything trip k adds, so every sum sees its terms in the shader's order.  Same
// operations on the same values - bit-identical - on a much shorter critical path.
__device__ __forceinline__ void finish_in_groups(unsigned live, bool& alive, Path& p, const FrameConst& fc, uint32_t& rays) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, g = lane >> 3, sub = lane & 3;
    const bool shadow_quad = (lane & 4) != 0;
    const bool has = g < __popc(live);
    const int src = has ? (int)__fns(live, 0, g + 1) : lane;  // the g-th live lane
    Path q;
    q.ray.origin = v3(__shfl_sync(full, p.ray.origin.x, src), __shfl_sync(full, p.ray.origin.y, src), __shfl_sync(full, p.ray.origin.z, src));
    q.ray.direction = v3(__shfl_sync(full, p.ray.direction.x, src), __shfl_sync(full, p.ray.direction.y, src), __shfl_sync(full, p.ray.direction.z, src));
    q.accum = v3(__shfl_sync(full, p.accum.x, src), __shfl_sync(full, p.accum.y, src), __shfl_sync(full, p.accum.z, src));
    q.attenuation = v3(__shfl_sync(full, p.attenuation.x, src), __shfl_sync(full, p.attenuation.y, src), __shfl_sync(full, p.attenuation.z, src));
    q.i = __shfl_sync(full, p.i, src);
    const float inf = __uint_as_float(0x7f800000u);
    const V3 center = v3((float)(2 * sub - 4), 1.0f, 0.0f);  // of sphere `sub` (1..3); lane 0 of a quad has the plane
    bool galive = has, pending = false;  // pending: the previous trip's hit still waits for its light
    Hit ph = surface(-1, 1e10f, q.ray);
    V3 pdir = q.ray.direction;
    Ray psr = Ray{q.ray.origin, fc.sun_light};
    while (__any_sync(full, galive || pending)) {
        const Ray tr = shadow_quad ? psr : q.ray;
        float tp, ts;
        const bool okp = plane_t(1e10f, tr, tp), oks = sphere_t(1e10f, tr, center, ts);
        const bool ok = sub == 0 ? okp : oks;
        float t = ok ? (sub == 0 ? tp : ts) : inf;
        int which = sub;
#pragma unroll
        for (int m = 1; m <= 2; m <<= 1) {  // within each quad
            const float ot = __shfl_xor_sync(full, t, m);
            const int ow = __shfl_xor_sync(full, which, m);
            if (ot < t || (ot == t && ow < which)) { t = ot; which = ow; }
        }
        t = __shfl_sync(full, t, lane & ~4);          // the primary quad's result, for all eight lanes
        which = __shfl_sync(full, which, lane & ~4);
        const unsigned shadow_bits = __ballot_sync(full, shadow_quad && pending && ok);
        if (pending) {
            shade(q, ph, pdir, ((shadow_bits >> (g * 8 + 4)) & 15u) != 0u, fc);
            pending = false;
        }
        if (galive) {
            const bool is_hit = t < inf;
            if ((lane & 7) == 0) rays += is_hit ? 2u : 1u;
            if (is_hit) {
                ph = surface(which, t, q.ray);
                pdir = q.ray.direction;
                psr = shadow_ray(q.ray, ph, fc);
                q.ray = bounce(q.ray, ph);
                pending = true;
                galive = ++q.i < SAMPLES;
            } else {
                q.accum = q.accum + get_sky_color(q.ray.direction, fc);
                galive = false;
            }
        }
    }
    // the r-th live lane's path ran on group r
    const int back = alive ? 8 * __popc(live & ((1u << lane) - 1u)) : lane;
    const V3 done = v3(__shfl_sync(full, q.accum.x, back), __shfl_sync(full, q.accum.y, back), __shfl_sync(full, q.accum.z, back));
    if (alive) p.accum = done;
    alive = false;
}

// rt.wgsl:128-158 `radiance` inside rt.wgsl:161-176 `fs_main`; called by all 32 lanes of a warp (inside = the lane has
// a pixel)
__device__ __forceinline__ V3 fs_main(float px, float py, bool inside, const FrameConst& fc, uint32_t& rays) {
    float ux = 2.0f * px / fc.res_x - 1.0f, uy = 2.0f * py / fc.res_y - 1.0f;
    uy = -uy;
    Path p = {Ray{v3(0.0f, 1.0f, 4.0f), normalize(v3(fc.aspect * ux, uy, -1.5f))}, splat(0.0f), splat(1.0f), 0};
    bool alive = inside;
    for (;;) {
        const unsigned live = __ballot_sync(0xffffffffu, alive);
        if (live == 0u) break;
        if (__popc(live) <= kGroupTail) { finish_in_groups(live, alive, p, fc, rays); break; }
        if (alive) alive = trip(p, fc, rays);
    }
    const V3 accum = p.accum / (float)SAMPLES;
    return accum / 1.0f;
}
}  // namespace rt1

// ==== rt2.wgsl ======================================================================================================
namespace rt2 {
constexpr float EPSILON = 1e-3f;     // rt2.wgsl:49
constexpr float EXPOSURE = 1e-2f;    // rt2.wgsl:52
constexpr float INTENSITY = 100.0f;  // rt2.wgsl:54

// Intersect (rt2.wgsl:31-35); lit = diffuse > 0 || specular > 0 (rt2.wgsl:97-117)
struct Intersect { float len; V3 normal; V3 color; float diffuse, specular; bool lit; };

// rt2.wgsl:56-74, pow(e, 2.0) = e * e; on a hit the sphere replaces whatever `trace` held (no distance test)
__device__ __forceinline__ void sphere(Intersect& a, const Ray& ray, float radius, V3 position, V3 color, float diffuse,
                                       float specular) {
    const V3 oc = position - ray.origin;
    const float l = dot(ray.direction, oc);
    const float det = l * l - dot(oc, oc) + radius * radius;
    if (det < 0.0f) return;
    const float sq = sqrtf(det);
    float len = l - sq;
    if (len < 0.0f) len = l + sq;
    if (len < 0.0f) return;
    a.len = len; a.normal = (ray.origin + len * ray.direction - position) / radius;
    a.color = color; a.diffuse = diffuse; a.specular = specular; a.lit = true;
}

// rt2.wgsl:87-119: plane, s1, s2, s4, s3 — the last one hit wins
__device__ __forceinline__ Intersect trace(const Ray& ray) {
    Intersect it = {0.0f, splat(0.0f), splat(0.0f), 0.0f, 0.0f, false};
    {   // rt2.wgsl:76-85 `intersect_plane` with p1 (rt2.wgsl:94)
        const V3 n = v3(0.0f, 1.0f, 0.0f);
        const float len = -dot(ray.origin, n) / dot(ray.direction, n);
        if (!(len < 0.0f)) { it.len = len; it.normal = n; it.color = splat(1.0f); it.diffuse = 1.0f; it.specular = 0.0f; it.lit = true; }
    }
    sphere(it, ray, 2.0f, v3(-4.0f, 3.0f, 0.0f), v3(1.0f, 0.0f, 0.2f), 1.0f, 0.001f);
    sphere(it, ray, 3.0f, v3(4.0f, 3.0f, 0.0f), v3(0.0f, 0.2f, 1.0f), 1.0f, 0.0f);
    sphere(it, ray, 1.0f, v3(6.0f, 1.0f, 4.0f), v3(0.0f, 1.0f, 0.2f), 0.5f, 0.1f);
    sphere(it, ray, 1.0f, v3(0.5f, 1.0f, 6.0f), v3(1.0f, 1.0f, 1.0f), 0.5f, 0.25f);
    return it;
}

// rt2.wgsl:128-138 `is_miss` of a `trace` result: every primitive's material has diffuse > 0, so the result is the
// all-zero `miss` exactly when nothing was hit
__device__ __forceinline__ bool sphere_any(const Ray& ray, float radius, V3 position) {
    const V3 oc = position - ray.origin;
    const float l = dot(ray.direction, oc);
    const float det = l * l - dot(oc, oc) + radius * radius;
    if (det < 0.0f) return false;
    const float sq = sqrtf(det);
    float len = l - sq;
    if (len < 0.0f) len = l + sq;
    return !(len < 0.0f);
}
__device__ __forceinline__ bool trace_any(const Ray& ray) {
    const V3 n = v3(0.0f, 1.0f, 0.0f);
    bool any = !(-dot(ray.origin, n) / dot(ray.direction, n) < 0.0f);
    any |= sphere_any(ray, 2.0f, v3(-4.0f, 3.0f, 0.0f));
    any |= sphere_any(ray, 3.0f, v3(4.0f, 3.0f, 0.0f));
    any |= sphere_any(ray, 1.0f, v3(6.0f, 1.0f, 4.0f));
    any |= sphere_any(ray, 1.0f, v3(0.5f, 1.0f, 6.0f));
    return any;
}

// rt2.wgsl:140-173 `radiance` inside rt2.wgsl:175-188 `fs_main`
__device__ __forceinline__ V3 fs_main(float px, float py, const FrameConst& fc, uint32_t& rays) {
    float ux = 2.0f * px / fc.res_x - 1.0f, uy = 2.0f * py / fc.res_y - 1.0f;
    ux *= fc.aspect;
    uy = -uy;
    Ray ray = {v3(0.0f, 2.5f, 12.0f), normalize(v3(ux, uy, -1.0f))};
    const V3 light_color = splat(1.0f) * INTENSITY;
    const V3 light_dir = fc.light2;
    V3 color = splat(0.0f), fresnel = splat(0.0f), mask = splat(1.0f);
    for (int i = 0; i <= 16; i += 1) {
        const Intersect hit = trace(ray);
        ++rays;
        if (hit.lit) {
            const V3 r0 = hit.color * hit.specular;
            const float hv = clampf(dot(hit.normal, -ray.direction), 0.0f, 1.0f);
            fresnel = r0 + (1.0f - r0) * pow_wgsl(1.0f - hv, 5.0f);
            mask = mask * fresnel;
            ++rays;
            if (!trace_any(Ray{ray.origin + hit.len * ray.direction + EPSILON * light_dir, light_dir})) {
                color = color + clampf(dot(hit.normal, light_dir), 0.0f, 1.0f) * light_color * hit.color * hit.diffuse *
                                    (1.0f - fresnel) * mask / fresnel;
            }
            const V3 reflection = reflect(ray.direction, hit.normal);
            ray = Ray{ray.origin + hit.len * ray.direction + EPSILON * reflection, reflection};
        } else {
            const V3 spotlight = splat(1e6f) * pow_wgsl(fabsf(dot(ray.direction, light_dir)), 250.0f);
            color = color + mask * (fc.ambient2 + spotlight);
            break;
        }
    }
    const V3 c = color * EXPOSURE;
    return v3(pow_wgsl(c.x, fc.inv_gamma2), pow_wgsl(c.y, fc.inv_gamma2), pow_wgsl(c.z, fc.inv_gamma2));
}
}  // namespace rt2

// float -> unorm8 of a render target: NaN -> 0, clamp, x * 255 + 0.5 truncated
__device__ __forceinline__ uint32_t to_unorm8(float x) {
    if (!(x == x)) return 0u;
    x = x < 0.0f ? 0.0f : x;
    x = x > 1.0f ? 1.0f : x;
    return (uint32_t)(x * 255.0f + 0.5f);
}

__device__ __forceinline__ float srgb_curve(float c) {
    if (c <= 0.0031308f) return 12.92f * c;
    return 1.055f * pow_wgsl(c, 1.0f / 2.4f) - 0.055f;
}

constexpr int kTileW = 16, kTileH = 8, kThreads = kTileW * kTileH;

template <int SHADER, bool SRGB>
__global__ void __launch_bounds__(kThreads) frag_kernel(FrameConst fc, uint32_t width, uint32_t height,
                                                        uint32_t tiles_x, float4* __restrict__ linear,
                                                        uint32_t* __restrict__ rgba8, unsigned long long* __restrict__ n_rays) {
    // lane -> 8 x 4 pixels of the warp's tile; warp -> 2 x 2 within the block's 16 x 8
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t bx = blockIdx.x % tiles_x, by = blockIdx.x / tiles_x;
    const uint32_t x = bx * kTileW + (warp & 1u) * 8u + (lane & 7u);
    const uint32_t y = by * kTileH + (warp >> 1) * 4u + (lane >> 3);
    uint32_t rays = 0;
    const bool inside = x < width && y < height;
    const float px = (float)x + 0.5f, py = (float)y + 0.5f;
    V3 c = splat(0.0f);
    if (SHADER == 1) c = rt1::fs_main(px, py, inside, fc, rays);  // warp-collective
    else if (inside) c = rt2::fs_main(px, py, fc, rays);
    if (inside) {
        const size_t i = (size_t)y * width + x;
        if (linear) linear[i] = make_float4(c.x, c.y, c.z, 1.0f);
        if (rgba8) {
            const uint32_t r = to_unorm8(SRGB ? srgb_curve(c.x) : c.x), g = to_unorm8(SRGB ? srgb_curve(c.y) : c.y),
                           b = to_unorm8(SRGB ? srgb_curve(c.z) : c.z);
            rgba8[i] = r | (g << 8) | (b << 16) | 0xff000000u;
        }
    }
    rays = __reduce_add_sync(0xffffffffu, rays);
    if (lane == 0 && rays) atomicAdd(n_rays, (unsigned long long)rays);
}

}  // namespace

struct forma_frag {
    int device = 0;
    uint32_t width = 0, height = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_k0 = nullptr, ev_k1 = nullptr, ev_end = nullptr;
    float4* d_linear = nullptr;
    uint32_t* d_rgba8 = nullptr;
    unsigned long long* d_rays = nullptr;
    forma_frag_stats stats{};
    bool stats_pending = false;

    size_t pixels() const { return (size_t)width * height; }
    void release_frames() {
        if (d_linear) cudaFree(d_linear);
        if (d_rgba8) cudaFree(d_rgba8);
        d_linear = nullptr; d_rgba8 = nullptr;
    }
    ~forma_frag() {
        cudaSetDevice(device);
        if (stream) cudaStreamSynchronize(stream);
        release_frames();
        if (d_rays) cudaFree(d_rays);
        for (cudaEvent_t e : {ev_begin, ev_k0, ev_k1, ev_end}) if (e) cudaEventDestroy(e);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {

using forma::set_last_error;

#define FRAG_TRY(expr)                                                                                  \
    do {                                                                                                \
        cudaError_t e__ = (expr);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return set_last_error(FORMA_RT_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e__)); \
    } while (0)
#define NEED_FRAG(f) do { if (!(f)) return set_last_error(FORMA_RT_EINVAL, "null forma_frag handle"); cudaSetDevice((f)->device); } while (0)

inline bool size_ok(uint32_t w, uint32_t h) { return w != 0 && h != 0 && (uint64_t)w * h <= (1ull << 28); }

int alloc_frames(forma_frag* f) {
    f->release_frames();
    FRAG_TRY(cudaMalloc((void**)&f->d_linear, f->pixels() * sizeof(float4)));
    FRAG_TRY(cudaMalloc((void**)&f->d_rgba8, f->pixels() * sizeof(uint32_t)));
    return FORMA_RT_OK;
}

// host-side f32 normalize, the order of the device's (v / sqrt(dot)); this file is compiled with contraction off
void normalize3(float x, float y, float z, V3* out) {
    const float len = std::sqrt(x * x + y * y + z * z);
    out->x = x / len; out->y = y / len; out->z = z / len;
}

void light_direction(float time, V3* out) {
    const float a = time / 10.0f;
    normalize3((float)std::cos((double)a), 0.75f, (float)std::sin((double)a), out);
}

int launch(forma_frag* f, int shader, const forma_frag_params* params, uint32_t* d_rgba8, float4* d_linear, cudaStream_t s) {
    if (shader != FORMA_FRAG_RT && shader != FORMA_FRAG_RT2)
        return set_last_error(FORMA_RT_EINVAL, "forma_frag: unknown shader " + std::to_string(shader));
    forma_frag_params p{};
    if (params) p = *params;
    if (p.flags & ~FORMA_FRAG_SRGB_SURFACE) return set_last_error(FORMA_RT_EINVAL, "forma_frag: unknown flag");
    if (!(p.time == p.time) || !(p.res_x == p.res_x) || !(p.res_y == p.res_y) || !(p.aspect == p.aspect))
        return set_last_error(FORMA_RT_EINVAL, "forma_frag: NaN parameter");
    if (((uintptr_t)d_rgba8 & 3u) || ((uintptr_t)d_linear & 15u))
        return set_last_error(FORMA_RT_EINVAL, "forma_frag: misaligned device frame");
    FrameConst fc;
    fc.res_x = p.res_x > 0.0f ? p.res_x : 1600.0f;
    fc.res_y = p.res_y > 0.0f ? p.res_y : 1200.0f;
    fc.aspect = p.aspect > 0.0f ? p.aspect : (shader == FORMA_FRAG_RT ? 4.0f / 3.0f : fc.res_x / fc.res_y);
    normalize3(1.0f, 0.5f, 0.5f, &fc.sun_sky);
    normalize3(1.0f, 0.5f, 0.0f, &fc.sun_light);
    light_direction(p.time, &fc.light2);
    fc.ambient2 = V3{0.6f * 100.0f / 2.2f, 0.8f * 100.0f / 2.2f, 1.0f * 100.0f / 2.2f};
    fc.inv_gamma2 = 1.0f / 2.2f;
    const uint32_t tiles_x = (f->width + kTileW - 1) / kTileW, tiles_y = (f->height + kTileH - 1) / kTileH;
    const bool srgb = (p.flags & FORMA_FRAG_SRGB_SURFACE) != 0;
    FRAG_TRY(cudaMemsetAsync(f->d_rays, 0, sizeof(unsigned long long), s));
    FRAG_TRY(cudaEventRecord(f->ev_k0, s));
    const dim3 grid(tiles_x * tiles_y), block(kThreads);
    if (shader == FORMA_FRAG_RT) {
        if (srgb) frag_kernel<1, true><<<grid, block, 0, s>>>(fc, f->width, f->height, tiles_x, d_linear, d_rgba8, f->d_rays);
        else frag_kernel<1, false><<<grid, block, 0, s>>>(fc, f->width, f->height, tiles_x, d_linear, d_rgba8, f->d_rays);
    } else {
        if (srgb) frag_kernel<2, true><<<grid, block, 0, s>>>(fc, f->width, f->height, tiles_x, d_linear, d_rgba8, f->d_rays);
        else frag_kernel<2, false><<<grid, block, 0, s>>>(fc, f->width, f->height, tiles_x, d_linear, d_rgba8, f->d_rays);
    }
    FRAG_TRY(cudaGetLastError());
    FRAG_TRY(cudaEventRecord(f->ev_k1, s));
    f->stats.gpu_launches = 1;
    f->stats_pending = true;
    return FORMA_RT_OK;
}

}  // namespace

extern "C" {

int forma_frag_create(int device, uint32_t width, uint32_t height, forma_frag** out) {
    if (!out) return set_last_error(FORMA_RT_EINVAL, "forma_frag_create: out is null");
    *out = nullptr;
    if (!size_ok(width, height)) return set_last_error(FORMA_RT_EINVAL, "forma_frag_create: width/height out of range");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return set_last_error(FORMA_RT_ECUDA, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return set_last_error(FORMA_RT_EINVAL, "forma_frag_create: device ordinal out of range");
    FRAG_TRY(cudaSetDevice(device));
    forma_frag* f = new (std::nothrow) forma_frag();
    if (!f) return set_last_error(FORMA_RT_ENOMEM, "forma_frag_create: out of host memory");
    f->device = device; f->width = width; f->height = height;
    auto bail = [&](int rc) { delete f; return rc; };
    if ((e = cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking)) != cudaSuccess)
        return bail(set_last_error(FORMA_RT_ECUDA, std::string("cudaStreamCreate: ") + cudaGetErrorString(e)));
    for (cudaEvent_t* ev : {&f->ev_begin, &f->ev_k0, &f->ev_k1, &f->ev_end})
        if ((e = cudaEventCreate(ev)) != cudaSuccess)
            return bail(set_last_error(FORMA_RT_ECUDA, std::string("cudaEventCreate: ") + cudaGetErrorString(e)));
    if ((e = cudaMalloc((void**)&f->d_rays, sizeof(unsigned long long))) != cudaSuccess)
        return bail(set_last_error(FORMA_RT_ECUDA, std::string("cudaMalloc: ") + cudaGetErrorString(e)));
    if (int rc = alloc_frames(f)) return bail(rc);
    *out = f;
    return FORMA_RT_OK;
}

void forma_frag_destroy(forma_frag* fr) { delete fr; }

int forma_frag_resize(forma_frag* fr, uint32_t width, uint32_t height) {
    NEED_FRAG(fr);
    // wnt.rs:244 ignores a zero-sized request; a binding sees the same: nothing changes
    if (width == 0 || height == 0) return FORMA_RT_OK;
    if (!size_ok(width, height)) return set_last_error(FORMA_RT_EINVAL, "forma_frag_resize: width/height out of range");
    FRAG_TRY(cudaStreamSynchronize(fr->stream));
    fr->width = width; fr->height = height;
    return alloc_frames(fr);
}

int forma_frag_render(forma_frag* fr, int shader, const forma_frag_params* params, uint8_t* rgba8, float* linear_rgba) {
    NEED_FRAG(fr);
    cudaStream_t s = fr->stream;
    FRAG_TRY(cudaEventRecord(fr->ev_begin, s));
    if (int rc = launch(fr, shader, params, rgba8 ? fr->d_rgba8 : nullptr, linear_rgba ? fr->d_linear : nullptr, s)) return rc;
    if (rgba8) FRAG_TRY(cudaMemcpyAsync(rgba8, fr->d_rgba8, fr->pixels() * 4, cudaMemcpyDeviceToHost, s));
    if (linear_rgba) FRAG_TRY(cudaMemcpyAsync(linear_rgba, fr->d_linear, fr->pixels() * 16, cudaMemcpyDeviceToHost, s));
    FRAG_TRY(cudaEventRecord(fr->ev_end, s));
    FRAG_TRY(cudaStreamSynchronize(s));
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, fr->ev_begin, fr->ev_end) == cudaSuccess) fr->stats.total_ms = ms;
    return FORMA_RT_OK;
}

int forma_frag_render_device(forma_frag* fr, int shader, const forma_frag_params* params, void* d_rgba8,
                             void* d_linear_rgba, void* stream) {
    NEED_FRAG(fr);
    cudaStream_t s = stream ? (cudaStream_t)stream : fr->stream;
    return launch(fr, shader, params, (uint32_t*)d_rgba8, (float4*)d_linear_rgba, s);
}

int forma_frag_light_direction(float time, float out[3]) {
    if (!out) return set_last_error(FORMA_RT_EINVAL, "forma_frag_light_direction: out is null");
    V3 l;
    light_direction(time, &l);
    out[0] = l.x; out[1] = l.y; out[2] = l.z;
    return FORMA_RT_OK;
}

int forma_frag_synchronize(forma_frag* fr) {
    NEED_FRAG(fr);
    FRAG_TRY(cudaStreamSynchronize(fr->stream));
    return FORMA_RT_OK;
}

int forma_frag_stats_get(forma_frag* fr, forma_frag_stats* out) {
    NEED_FRAG(fr);
    if (!out) return set_last_error(FORMA_RT_EINVAL, "forma_frag_stats_get: out is null");
    if (fr->stats_pending) {
        FRAG_TRY(cudaEventSynchronize(fr->ev_k1));
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, fr->ev_k0, fr->ev_k1) == cudaSuccess) fr->stats.kernel_ms = ms;
        unsigned long long n = 0;
        FRAG_TRY(cudaMemcpy(&n, fr->d_rays, sizeof n, cudaMemcpyDeviceToHost));
        fr->stats.rays = n;
        fr->stats_pending = false;
    }
    *out = fr->stats;
    return FORMA_RT_OK;
}

}  // extern "C"
