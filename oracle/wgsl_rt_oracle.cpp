// wgsl_rt_oracle.cpp — CPU restatement of the reference's two fragment-shader ray tracers,
// src/shaders/rt.wgsl (`radiance`, rt.wgsl:127-158) and src/shaders/rt2.wgsl (`radiance`, rt2.wgsl:140-173),
// SURVEY.md 8(f) row f-4.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under fo-rma_b200/ may include, link or call this file; only tests/,
// __graft_entry__.smoke() and the cpu legs of the perf tools load it, as the checker for the CUDA path.
//
// PARITY UNPINNED: the reference runs these as wgpu fragment shaders (frontend/wnt.rs:145-148 loads rt.wgsl; rt2.wgsl
// is not loaded by anything), which is unavailable on this box (no graphics stack), and the repository holds no
// golden image of either.  WGSL leaves the precision of pow / normalize / the float -> texel conversion to the
// driver, so "the reference's result" is only defined up to those; what is restated here is the shader arithmetic as
// written, under these stated conventions:
//   * every f32 operation is IEEE round-to-nearest in the order the shader writes it, left-associative, never
//     contracted (-ffp-contract=off); dot(a, b) = a.x*b.x + a.y*b.y + a.z*b.z; length = sqrt(dot); normalize(v) =
//     v / length(v) (three divisions); reflect(e1, e2) = e1 - 2*dot(e2, e1)*e2; clamp = min(max(x, lo), hi) with IEEE
//     minNum / maxNum; mix(a, b, t) = a*(1-t) + b*t; smoothstep(lo, hi, x): t = clamp((x-lo)/(hi-lo), 0, 1),
//     t*t*(3 - 2*t) — the forms the WGSL specification lists first;
//   * pow(x, y) = exp2(y * log2(x)) — the WGSL specification's definition of its accuracy — with exp2 and log2 the
//     fixed f32 routines below (about 1 ulp; only + - * / and bit moves, so a second implementation of the same
//     recipe is bit-identical).  One exception, W1: pow(e, 2.0) with the literal exponent 2.0 (rt2.wgsl:58) is e*e.
//     Its base is a dot product that is negative for every sphere behind a ray; exp2(2*log2(negative)) is NaN, which
//     rt2's `trace` would count as a hit of that sphere (NaN < 0 is false), and every shader compiler strength-
//     reduces the literal square;
//   * module-scope and `let` constants are evaluated in f32 at run-time precision (normalize(vec3(1.0, 0.5, 0.5)),
//     vec3(0.6, 0.8, 1.0) * INTENSITY / GAMMA, 1.0 / GAMMA), not in abstract-float f64;
//   * cos / sin of rt2's `uniforms.time / 10f` (rt2.wgsl:144) are computed on the host in double and rounded to f32;
//   * the fragment position of pixel (x, y) is (x + 0.5, y + 0.5), row 0 on top; `resolution` and the aspect ratio are
//     the shaders' constants (1600 x 1200, 4/3) unless the caller passes others (added mode: a frame of another size
//     with the constants as written is a crop of the 1600 x 1200 image);
//   * outputs: the fragment's vec4 as f32 RGBA, and RGBA8 by the float -> unorm rule (NaN -> 0, clamp, x*255 + 0.5
//     truncated), optionally through the linear -> sRGB curve first (wnt.rs:124-132 prefers an sRGB surface):
//     c <= 0.0031308 ? 12.92*c : 1.055*pow(c, 1/2.4) - 0.055, alpha untouched.
// Each function cites the shader lines it follows.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {

inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// ---- fixed f32 log2 / exp2 (the recipe both sides implement) -------------------------------------------------------
// log2: argument reduction to [sqrt(1/2), sqrt(2)), s = f/(2+f), even/odd polynomial in s^2 (the classic fdlibm
// e_log2f scheme and constants), hi/lo split of the product with 1/ln 2.
float log2_fixed(float x) {
    uint32_t ix = f2u(x);
    int k = 0;
    if (ix < 0x00800000u) {                         // +0 or positive subnormal
        if (ix == 0) return -std::numeric_limits<float>::infinity();
        k -= 25; x *= 33554432.0f; ix = f2u(x);     // * 2^25, exact
    } else if (ix >= 0x7f800000u) {
        if (ix == 0x7f800000u) return x;            // +inf
        if (ix == 0x80000000u) return -std::numeric_limits<float>::infinity();  // -0
        return std::numeric_limits<float>::quiet_NaN();  // negative, or NaN
    }
    ix += 0x3f800000u - 0x3f3504f3u;
    k += (int)(ix >> 23) - 0x7f;
    ix = (ix & 0x007fffffu) + 0x3f3504f3u;
    x = u2f(ix);
    const float f = x - 1.0f;
    const float s = f / (2.0f + f);
    const float z = s * s;
    const float w = z * z;
    const float t1 = w * (0.40000972152f + w * 0.24279078841f);
    const float t2 = z * (0.66666662693f + w * 0.28498786688f);
    const float R = t2 + t1;
    const float hfsq = 0.5f * f * f;
    float hi = f - hfsq;
    hi = u2f(f2u(hi) & 0xfffff000u);
    const float lo = (f - hi) - hfsq + s * (hfsq + R);
    const float ivln2hi = 1.4428710938f, ivln2lo = -1.7605285393e-04f;
    return (lo + hi) * ivln2lo + lo * ivln2hi + hi * ivln2hi + (float)k;
}

// exp2: k = floor(x + 1/2), r = x - k (exact), e^(r ln 2) by a degree-7 Taylor polynomial in Horner form, scaled by
// 2^k in two exact-or-once-rounded steps so that subnormal results round once.
float exp2_fixed(float x) {
    if (!(x == x)) return x;
    if (x >= 128.0f) return std::numeric_limits<float>::infinity();
    if (x < -150.0f) return 0.0f;
    const float k = std::floor(x + 0.5f);
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
    return p * u2f((uint32_t)(k1 + 127) << 23) * u2f((uint32_t)(k2 + 127) << 23);
}

inline float pow_wgsl(float x, float y) { return exp2_fixed(y * log2_fixed(x)); }

// ---- WGSL value types and builtins ---------------------------------------------------------------------------------
struct V3 { float x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3 operator-(float s, V3 a) { return {s - a.x, s - a.y, s - a.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 splat(float s) { return {s, s, s}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 normalize(V3 v) { return v / std::sqrt(dot(v, v)); }
inline V3 reflect(V3 e1, V3 e2) { return e1 - (2.0f * dot(e2, e1)) * e2; }
inline float clampf(float x, float lo, float hi) { return std::fmin(std::fmax(x, lo), hi); }
inline float smoothstep(float lo, float hi, float x) {
    const float t = clampf((x - lo) / (hi - lo), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
inline V3 mix(V3 a, V3 b, float t) { return a * (1.0f - t) + b * t; }

struct Ray { V3 origin, direction; };

struct Frame {  // the per-frame values a host computes once
    float res_x, res_y, aspect;
    V3 light2;  // rt2.wgsl:144
};

// ==== rt.wgsl =======================================================================================================
namespace rt1 {
constexpr float EPSILON = 4e-4f;  // rt.wgsl:7
constexpr int SAMPLES = 1000;     // rt.wgsl:8
struct Material { V3 color; float f0; };                    // rt.wgsl:10-13
struct Light { V3 direction, color; };                      // rt.wgsl:15-18
struct Hit { float t; V3 normal; Material material; };      // rt.wgsl:25-29
struct Plane { float d; V3 normal; Material material; };    // rt.wgsl:31-35
struct Sphere { float radius; V3 center; Material material; };  // rt.wgsl:37-41
const Hit no_hit = {1e10f, {0, 0, 0}, {{-1, -1, -1}, -1}};

// rt.wgsl:43-52
Hit intersect_plane(const Plane& p, const Ray& r) {
    const float dotnd = dot(p.normal, r.direction);
    if (dotnd > 0.0f) return no_hit;
    const float t = -(dot(r.origin, p.normal) + p.d) / dotnd;
    return {t, p.normal, p.material};
}

// rt.wgsl:54-73
Hit intersect_sphere(const Sphere& s, const Ray& r) {
    const V3 op = s.center - r.origin;
    const float b = dot(op, r.direction);
    float det = b * b - dot(op, op) + s.radius * s.radius;
    if (det < 0.0f) return no_hit;
    det = std::sqrt(det);
    float t = b - det;
    if (t < 0.0f) t = b + det;
    if (t < 0.0f) return no_hit;
    return {t, (r.origin + t * r.direction - s.center) / s.radius, s.material};
}

// rt.wgsl:77-85
inline bool compare(Hit& a, const Hit& b) {
    if (b.material.f0 >= 0.0f && b.t < a.t) { a = b; return true; }
    return false;
}

// rt.wgsl:88-102
Hit intersect_scene(const Ray& ray, uint64_t& rays) {
    ++rays;
    const Sphere s1 = {1.0f, {-2, 1, 0}, {{1, 0, 0}, 0.04f}};
    const Sphere s2 = {1.0f, {0, 1, 0}, {{0, 1, 0}, 0.1f}};
    const Sphere s3 = {1.0f, {2, 1, 0}, {{0, 0, 1}, 0.5f}};
    const Plane p = {0.0f, {0, 1, 0}, {{0.5f, 0.4f, 0.3f}, 0.04f}};
    Hit hit = no_hit;
    compare(hit, intersect_plane(p, ray));
    compare(hit, intersect_sphere(s1, ray));
    compare(hit, intersect_sphere(s2, ray));
    compare(hit, intersect_sphere(s3, ray));
    return hit;
}

// rt.wgsl:104-111
V3 get_sky_color(V3 d) {
    const Light sun = {normalize({1.0f, 0.5f, 0.5f}), splat(1.0f)};
    const float transition = pow_wgsl(smoothstep(0.02f, 0.5f, d.y), 0.4f);
    const V3 sky = 2e0f * mix({0.52f, 0.77f, 1.0f}, {0.12f, 0.43f, 1.0f}, transition);
    const V3 sunc = sun.color * pow_wgsl(std::fabs(dot(d, sun.direction)), 5000.0f);
    return sky + sunc;
}

// rt.wgsl:113-115
inline float pow5(float x) { return x * x * x * x * x; }

// rt.wgsl:117-119
inline float fresnel(V3 h, V3 v, float f0) { return pow5(1.0f - clampf(dot(h, v), 0.0f, 1.0f)) * (1.0f - f0) + f0; }

// rt.wgsl:121-126
V3 account_for_directional_light(V3 p, V3 n, const Light& l, uint64_t& rays) {
    if (intersect_scene({p + EPSILON * l.direction, l.direction}, rays).material.f0 < 0.0f)
        return clampf(dot(n, l.direction), 0.0f, 1.0f) * l.color;
    return splat(0.0f);
}

// rt.wgsl:128-158
V3 radiance(Ray ray, uint64_t& rays, uint32_t& bounces) {
    V3 accum = splat(0.0f), attenuation = splat(1.0f);
    const Light sun_light = {normalize({1.0f, 0.5f, 0.0f}), splat(1e3f)};
    for (int i = 0; i < SAMPLES; i++) {
        const Hit hit = intersect_scene(ray, rays);
        if (hit.material.f0 >= 0.0f) {
            ++bounces;
            const float f = fresnel(hit.normal, -ray.direction, hit.material.f0);
            const V3 hit_pos = ray.origin + hit.t * ray.direction;
            V3 incoming = splat(0.0f);
            incoming = incoming + account_for_directional_light(hit_pos, hit.normal, sun_light, rays);
            accum = accum + hit.material.color;
            accum = accum + (1.0f - f) * attenuation * hit.material.color * incoming;
            attenuation = attenuation * f;
            const V3 d = reflect(ray.direction, hit.normal);
            ray = {ray.origin + hit.t * ray.direction + EPSILON * d, d};
        } else {
            accum = accum + get_sky_color(ray.direction);
            break;
        }
    }
    return accum / (float)SAMPLES;
}

// rt.wgsl:161-176
V3 fs_main(float px, float py, const Frame& fr, uint64_t& rays, uint32_t& bounces) {
    float ux = 2.0f * px / fr.res_x - 1.0f, uy = 2.0f * py / fr.res_y - 1.0f;
    uy = -uy;
    const V3 origin = {0.0f, 1.0f, 4.0f};
    const V3 direction = normalize({fr.aspect * ux, uy, -1.5f});
    const V3 c = radiance({origin, direction}, rays, bounces);
    return c / 1.0f;
}
}  // namespace rt1

// ==== rt2.wgsl ======================================================================================================
namespace rt2 {
constexpr float EPSILON = 1e-3f;     // rt2.wgsl:49
constexpr float EXPOSURE = 1e-2f;    // rt2.wgsl:52
constexpr float GAMMA = 2.2f;        // rt2.wgsl:53
constexpr float INTENSITY = 100.0f;  // rt2.wgsl:54
struct Material { V3 color; float diffuse, specular; };     // rt2.wgsl:25-29
struct Intersect { float len; V3 normal; Material material; };  // rt2.wgsl:31-35
struct Sphere { float radius; V3 position; Material material; };  // rt2.wgsl:37-41
struct Plane { V3 normal; Material material; };             // rt2.wgsl:43-46
const Intersect miss = {0.0f, {0, 0, 0}, {{0, 0, 0}, 0.0f, 0.0f}};

// rt2.wgsl:56-74; pow(e, 2.0) = e*e (W1)
Intersect intersect_sphere(const Ray& ray, const Sphere& s) {
    const V3 oc = s.position - ray.origin;
    const float l = dot(ray.direction, oc);
    const float det = l * l - dot(oc, oc) + s.radius * s.radius;
    if (det < 0.0f) return miss;
    float len = l - std::sqrt(det);
    if (len < 0.0f) len = l + std::sqrt(det);
    if (len < 0.0f) return miss;
    return {len, (ray.origin + len * ray.direction - s.position) / s.radius, s.material};
}

// rt2.wgsl:76-85
Intersect intersect_plane(const Ray& ray, const Plane& p) {
    const float len = -dot(ray.origin, p.normal) / dot(ray.direction, p.normal);
    if (len < 0.0f) return miss;
    return {len, p.normal, p.material};
}

inline bool lit(const Material& m) { return m.diffuse > 0.0f || m.specular > 0.0f; }

// rt2.wgsl:87-119 — no distance comparison: the last primitive hit, in the order plane, s1, s2, s4, s3, wins
Intersect trace(const Ray& ray, uint64_t& rays) {
    ++rays;
    const Sphere s1 = {2.0f, {-4.0f, 3.0f, 0.0f}, {{1.0f, 0.0f, 0.2f}, 1.0f, 0.001f}};
    const Sphere s2 = {3.0f, {4.0f, 3.0f, 0.0f}, {{0.0f, 0.2f, 1.0f}, 1.0f, 0.0f}};
    const Sphere s3 = {1.0f, {0.5f, 1.0f, 6.0f}, {{1.0f, 1.0f, 1.0f}, 0.5f, 0.25f}};
    const Sphere s4 = {1.0f, {6.0f, 1.0f, 4.0f}, {{0.0f, 1.0f, 0.2f}, 0.5f, 0.1f}};
    const Plane p1 = {{0.0f, 1.0f, 0.0f}, {{1.0f, 1.0f, 1.0f}, 1.0f, 0.0f}};
    Intersect intersection = miss;
    const Intersect plane = intersect_plane(ray, p1);
    if (lit(plane.material)) intersection = plane;
    Intersect sphere = intersect_sphere(ray, s1);
    if (lit(sphere.material)) intersection = sphere;
    sphere = intersect_sphere(ray, s2);
    if (lit(sphere.material)) intersection = sphere;
    sphere = intersect_sphere(ray, s4);
    if (lit(sphere.material)) intersection = sphere;
    sphere = intersect_sphere(ray, s3);
    if (lit(sphere.material)) intersection = sphere;
    return intersection;
}

// rt2.wgsl:121-138
inline bool is_vec_zero(V3 v) { return v.x == 0.0f && v.y == 0.0f && v.z == 0.0f; }
inline bool is_miss(const Intersect& i) {
    return i.len == 0.0f && is_vec_zero(i.normal) && is_vec_zero(i.material.color) && i.material.diffuse == 0.0f &&
           i.material.specular == 0.0f;
}

// rt2.wgsl:140-173
V3 radiance(Ray ray, const Frame& fr, uint64_t& rays, uint32_t& bounces) {
    const V3 ambient = V3{0.6f, 0.8f, 1.0f} * INTENSITY / GAMMA;
    const V3 light_color = splat(1.0f) * INTENSITY;
    const V3 light_dir = fr.light2;
    V3 color = splat(0.0f), fresnel = splat(0.0f), mask = splat(1.0f);
    for (int i = 0; i <= 16; i += 1) {
        const Intersect hit = trace(ray, rays);
        if (lit(hit.material)) {
            ++bounces;
            const V3 r0 = hit.material.color * hit.material.specular;
            const float hv = clampf(dot(hit.normal, -ray.direction), 0.0f, 1.0f);
            fresnel = r0 + (1.0f - r0) * pow_wgsl(1.0f - hv, 5.0f);
            mask = mask * fresnel;
            const Intersect result =
                trace({ray.origin + hit.len * ray.direction + EPSILON * light_dir, light_dir}, rays);
            if (is_miss(result)) {
                color = color + clampf(dot(hit.normal, light_dir), 0.0f, 1.0f) * light_color * hit.material.color *
                                    hit.material.diffuse * (1.0f - fresnel) * mask / fresnel;
            }
            const V3 reflection = reflect(ray.direction, hit.normal);
            ray = {ray.origin + hit.len * ray.direction + EPSILON * reflection, reflection};
        } else {
            const V3 spotlight = splat(1e6f) * pow_wgsl(std::fabs(dot(ray.direction, light_dir)), 250.0f);
            color = color + mask * (ambient + spotlight);
            break;
        }
    }
    return color;
}

// rt2.wgsl:175-188
V3 fs_main(float px, float py, const Frame& fr, uint64_t& rays, uint32_t& bounces) {
    float ux = 2.0f * px / fr.res_x - 1.0f, uy = 2.0f * py / fr.res_y - 1.0f;
    ux *= fr.aspect;
    uy = -uy;
    const Ray ray = {{0.0f, 2.5f, 12.0f}, normalize({ux, uy, -1.0f})};
    const V3 c = radiance(ray, fr, rays, bounces) * EXPOSURE;
    const float g = 1.0f / GAMMA;
    return {pow_wgsl(c.x, g), pow_wgsl(c.y, g), pow_wgsl(c.z, g)};
}
}  // namespace rt2

inline uint8_t to_unorm8(float x) {
    if (!(x == x)) return 0;
    if (x < 0.0f) x = 0.0f;
    if (x > 1.0f) x = 1.0f;
    return (uint8_t)(x * 255.0f + 0.5f);
}

inline float srgb_curve(float c) {
    if (c <= 0.0031308f) return 12.92f * c;   // NaN takes the other branch and stays NaN
    return 1.055f * pow_wgsl(c, 1.0f / 2.4f) - 0.055f;
}

}  // namespace

extern "C" {

float wgo_log2(float x) { return log2_fixed(x); }
float wgo_exp2(float x) { return exp2_fixed(x); }
float wgo_pow(float x, float y) { return pow_wgsl(x, y); }

// rt2.wgsl:144 — normalize(vec3(cos(time / 10f), 0.75, sin(time / 10f)))
void wgo_light2(float time, float out[3]) {
    const float a = time / 10.0f;
    const V3 l = normalize({(float)std::cos((double)a), 0.75f, (float)std::sin((double)a)});
    out[0] = l.x; out[1] = l.y; out[2] = l.z;
}

// shader: 1 = rt.wgsl, 2 = rt2.wgsl.  res_x / res_y <= 0 take the shaders' 1600 x 1200; aspect <= 0 takes 4/3
// (rt.wgsl:164) resp. res_x / res_y (rt2.wgsl:179).  linear_rgba (w*h*4 f32), rgba8 (w*h*4), bounces (w*h u32: hits
// along the pixel's path), rays (one u64: intersect_scene / trace calls) are each optional.
int wgo_render(int shader, uint32_t w, uint32_t h, float res_x, float res_y, float aspect, float time, int srgb,
               float* linear_rgba, uint8_t* rgba8, uint32_t* bounces, uint64_t* rays, uint32_t threads) {
    if (shader != 1 && shader != 2) return -1;
    Frame fr;
    fr.res_x = res_x > 0.0f ? res_x : 1600.0f;
    fr.res_y = res_y > 0.0f ? res_y : 1200.0f;
    fr.aspect = aspect > 0.0f ? aspect : (shader == 1 ? 4.0f / 3.0f : fr.res_x / fr.res_y);
    float l2[3];
    wgo_light2(time, l2);
    fr.light2 = {l2[0], l2[1], l2[2]};
    if (threads == 0) threads = 1;
    std::vector<uint64_t> ray_sums(threads, 0);
    auto band = [&](uint32_t tid) {
        uint64_t nr = 0;
        for (uint32_t y = tid; y < h; y += threads)
            for (uint32_t x = 0; x < w; ++x) {
                uint32_t nb = 0;
                const float px = (float)x + 0.5f, py = (float)y + 0.5f;
                const V3 c = shader == 1 ? rt1::fs_main(px, py, fr, nr, nb) : rt2::fs_main(px, py, fr, nr, nb);
                const size_t i = (size_t)y * w + x;
                if (linear_rgba) { float* q = linear_rgba + i * 4; q[0] = c.x; q[1] = c.y; q[2] = c.z; q[3] = 1.0f; }
                if (rgba8) {
                    uint8_t* q = rgba8 + i * 4;
                    if (srgb) { q[0] = to_unorm8(srgb_curve(c.x)); q[1] = to_unorm8(srgb_curve(c.y)); q[2] = to_unorm8(srgb_curve(c.z)); }
                    else { q[0] = to_unorm8(c.x); q[1] = to_unorm8(c.y); q[2] = to_unorm8(c.z); }
                    q[3] = 255;
                }
                if (bounces) bounces[i] = nb;
            }
        ray_sums[tid] = nr;
    };
    std::vector<std::thread> pool;
    for (uint32_t t = 1; t < threads; ++t) pool.emplace_back(band, t);
    band(0);
    for (auto& t : pool) t.join();
    if (rays) { uint64_t s = 0; for (uint64_t v : ray_sums) s += v; *rays = s; }
    return 0;
}

}  // extern "C"
