// forma_oracle.cpp — CPU ORACLE for the fo-rma `src/cpu_ray_tracer` hot path.
//
// *** TEST INFRASTRUCTURE ONLY ***  Nothing under fo-rma_b200/ (the product) may include, link or
// call this file.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs load liboracle.so, and only as the checker / CPU baseline.
//
// What it is: a bug-for-bug C++ restatement of the reference CPU path tracer (the reference is
// Rust and cannot be built here: no cargo/rustc, un-vendored `../kopek` dependency).  Every
// function cites the reference file:line (relative to /root/reference) it restates.  All f32
// arithmetic keeps the reference's evaluation order; build with  -O2 -ffp-contract=off  because
// rustc never contracts a*b+c.
//
// PARITY PIN STATUS
//   * Vec3 (primitives.rs:4-156): PINNED by the reference's own 7 known-answer tests
//     (primitives.rs:163-254), carried verbatim in tests/test_oracle_vec3_kat.py.
//   * Everything else (Sphere::hit, Plane::hit, scatter, camera, get_color, write-out):
//     **PARITY UNPINNED** — the reference has no test, fixture or golden image for it and its RNG
//     (rand::thread_rng, OS-seeded) is not reproducible.  This restatement, reviewed against the
//     cited lines, plus analytic tests, is the only pin.
//   * Triangles and the JSON->ray-traceable mapping are an EXTENSION with no reference behaviour
//     (the reference tracer never reads scenes/*.json and has no triangle intersector).
//
// RNG: the reference draws from rand::thread_rng() (third-party `rand` 0.8/0.9, ChaCha12, OS
// seed; Cargo.lock:2724-2738).  Only its distribution can be matched: uniform f32 in [0,1) with
// 24-bit resolution, (u32 >> 8) * 2^-24.  Two generators are offered:
//   mode 0  counter-based Philox4x32-7 keyed (seed) with counter (pixel, sample, slot, block) —
//           the generator the CUDA path uses, so same-RNG comparisons are meaningful;
//   mode 1  a sequential per-thread xoshiro128++ stream consumed in the reference's order (Q14) —
//           the stand-in for thread_rng in statistical (RMSE/PSNR) comparisons.

#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>
#include <algorithm>
#include <atomic>

namespace {

// ----------------------------------------------------------------------------------------------
// Vec3 — cpu_ray_tracer/primitives.rs:4-156
// ----------------------------------------------------------------------------------------------
struct Vec3 {
    float x, y, z;
    static Vec3 zero() { return {0.0f, 0.0f, 0.0f}; }                 // primitives.rs:18-24
    static Vec3 one() { return {1.0f, 1.0f, 1.0f}; }                  // primitives.rs:26-32
    float r() const { return x; }                                     // primitives.rs:38-48
    float g() const { return y; }
    float b() const { return z; }
    float length_squared() const { return x * x + y * y + z * z; }    // primitives.rs:50-52
    float length() const { return std::sqrt(length_squared()); }      // primitives.rs:54-56
    static float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // :58-60
    static Vec3 cross(Vec3 a, Vec3 b) {                               // primitives.rs:62-68 (Q13)
        return {a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x};
    }
    Vec3 unit_vector() const;                                         // primitives.rs:70-72
    Vec3 sqrt() const { return {std::sqrt(x), std::sqrt(y), std::sqrt(z)}; }  // :74-76
};
inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }  // :79-89
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }  // :91-101
inline Vec3 operator*(Vec3 a, Vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }  // :103-113
inline Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }       // :115-125
inline Vec3 operator*(float s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }       // :127-137
inline Vec3 operator/(Vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }       // :139-149
inline Vec3 Vec3::unit_vector() const { return *this / length(); }

// ----------------------------------------------------------------------------------------------
// Ray / HitRecord / ReflectRecord — cpu_ray_tracer/ray.rs:4-65
// ----------------------------------------------------------------------------------------------
struct Ray {
    Vec3 from, to;
    Vec3 origin() const { return from; }                              // ray.rs:21-23
    Vec3 direction() const { return to; }                             // ray.rs:25-27 (Q2: never normalised)
    Vec3 point_at(float t) const { return from + t * to; }            // ray.rs:30-32
};
struct HitRecord {                                                    // ray.rs:35-50
    float t = 0.0f;
    Vec3 p = Vec3::zero();
    Vec3 normal = Vec3::zero();
};
struct ReflectRecord {                                                // ray.rs:52-65
    Ray scattered;
    Vec3 attenuation;
};

// ----------------------------------------------------------------------------------------------
// RNG (stand-in for rand::thread_rng; see header).  Philox4x32-R is the published Random123
// algorithm (Salmon et al., SC'11): R rounds, multipliers 0xD2511F53 / 0xCD9E8D57, Weyl key
// increments 0x9E3779B9 / 0xBB67AE85.  The path draws from Philox4x32-7: the smallest round count the paper
// reports as Crush-resistant (it passes BigCrush; 10 is Random123's default safety margin), pinned like the
// 10-round form to Random123's published known-answer vectors (tests/test_oracle_analytic.py).  The choice of
// generator is this build's, not the reference's (its thread_rng is OS-seeded ChaCha12): only the distribution of
// the f32 draws is the reference's.  The CUDA kernels use the same round count (trace.cuh kPhiloxRounds).
// ----------------------------------------------------------------------------------------------
constexpr int kPathRounds = 7;
inline void philox4x32(int rounds, uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                       uint32_t out[4]) {
    for (int r = 0; r < rounds; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t n0 = hi1 ^ c1 ^ k0;
        uint32_t n1 = lo1;
        uint32_t n2 = hi0 ^ c3 ^ k1;
        uint32_t n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

inline float u32_to_unit_f32(uint32_t v) {  // rand's Standard f32: 24 mantissa bits, [0,1)
    return (float)(v >> 8) * (1.0f / 16777216.0f);
}

struct RngCtx {
    int mode = 0;             // 0 = counter Philox, 1 = sequential xoshiro128++
    uint32_t k0 = 0, k1 = 0;  // Philox key = seed
    uint32_t pixel = 0, sample = 0, slot = 0;
    uint32_t s[4] = {1, 2, 3, 4};  // xoshiro state (mode 1)
    uint64_t rays = 0;        // closest-hit queries (get_color calls)

    void block(uint32_t blk, float out[4]) const {
        uint32_t o[4];
        philox4x32(kPathRounds, k0, k1, pixel, sample, slot, blk, o);
        for (int i = 0; i < 4; ++i) out[i] = u32_to_unit_f32(o[i]);
    }
    static uint32_t rotl(uint32_t x, int k) { return (x << k) | (x >> (32 - k)); }
    float next_seq() {  // xoshiro128++ (Blackman & Vigna), public domain algorithm
        uint32_t result = rotl(s[0] + s[3], 7) + s[0];
        uint32_t t = s[1] << 9;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
        s[2] ^= t; s[3] = rotl(s[3], 11);
        return u32_to_unit_f32(result);
    }
    void seed_seq(uint64_t seed) {
        uint64_t z = seed;
        for (int i = 0; i < 4; ++i) {  // splitmix64
            z += 0x9E3779B97F4A7C15ull;
            uint64_t v = z;
            v = (v ^ (v >> 30)) * 0xBF58476D1CE4E5B9ull;
            v = (v ^ (v >> 27)) * 0x94D049BB133111EBull;
            v ^= v >> 31;
            s[i] = (uint32_t)(v >> 16) | 1u;
        }
    }
};

thread_local RngCtx* g_rng = nullptr;  // plays the role of thread_rng()'s TLS handle

// utility.rs:4-13  random_in_unit_circle.  Counter mode: try 0 = lanes 2,3 of camera block 0;
// tries 2j-1 / 2j = lanes (0,1) / (2,3) of block j>=1.
Vec3 random_in_unit_circle() {
    RngCtx& g = *g_rng;
    if (g.mode == 1) {
        float a = g.next_seq(), b = g.next_seq();
        Vec3 p = 2.0f * Vec3{a, b, 0.0f} - Vec3{1.0f, 1.0f, 0.0f};
        while (Vec3::dot(p, p) >= 1.0f) {
            a = g.next_seq(); b = g.next_seq();
            p = 2.0f * Vec3{a, b, 0.0f} - Vec3{1.0f, 1.0f, 0.0f};
        }
        return p;
    }
    float r[4];
    g.block(0, r);
    Vec3 p = 2.0f * Vec3{r[2], r[3], 0.0f} - Vec3{1.0f, 1.0f, 0.0f};
    uint32_t blk = 0;
    int half = 1;  // which half of the current block was just used (1 = lanes 2,3)
    while (Vec3::dot(p, p) >= 1.0f) {
        if (half == 1) { ++blk; g.block(blk, r); half = 0; } else { half = 1; }
        p = 2.0f * Vec3{r[2 * half], r[2 * half + 1], 0.0f} - Vec3{1.0f, 1.0f, 0.0f};
    }
    return p;
}

// utility.rs:15-25  random_in_unit_sphere.  Counter mode: try k = lanes 0..2 of scatter block k.
Vec3 random_in_unit_sphere() {
    RngCtx& g = *g_rng;
    if (g.mode == 1) {
        float a = g.next_seq(), b = g.next_seq(), c = g.next_seq();
        Vec3 p = 2.0f * Vec3{a, b, c} - Vec3{1.0f, 1.0f, 1.0f};
        while (Vec3::dot(p, p) >= 1.0f) {
            a = g.next_seq(); b = g.next_seq(); c = g.next_seq();
            p = 2.0f * Vec3{a, b, c} - Vec3{1.0f, 1.0f, 1.0f};
        }
        return p;
    }
    float r[4];
    uint32_t blk = 0;
    g.block(blk, r);
    Vec3 p = 2.0f * Vec3{r[0], r[1], r[2]} - Vec3{1.0f, 1.0f, 1.0f};
    while (Vec3::dot(p, p) >= 1.0f) {
        ++blk;
        g.block(blk, r);
        p = 2.0f * Vec3{r[0], r[1], r[2]} - Vec3{1.0f, 1.0f, 1.0f};
    }
    return p;
}

// utility.rs:27-30
Vec3 reflect(Vec3 v, Vec3 n) { return v - 2.0f * Vec3::dot(v, n) * n; }

// utility.rs:32-35  get_random.  Counter mode: lane 3 of scatter block 0.
float get_random() {
    RngCtx& g = *g_rng;
    if (g.mode == 1) return g.next_seq();
    float r[4];
    g.block(0, r);
    return r[3];
}

// utility.rs:37-47
bool refract(Vec3 v, Vec3 n, float ni_over_nt, Vec3* refracted) {
    Vec3 uv = v.unit_vector();
    float dt = Vec3::dot(uv, n);
    float discriminant = 1.0f - ni_over_nt * ni_over_nt * (1.0f - dt * dt);
    if (discriminant > 0.0f) {
        *refracted = ni_over_nt * (uv - dt * n) - std::sqrt(discriminant) * n;
        return true;
    }
    return false;
}

// f32::powi(5): compiler-rt __powisf2 / LLVM's expansion both give x * ((x*x)*(x*x)).
inline float powi5(float a) {
    float a2 = a * a;
    float a4 = a2 * a2;
    return a * a4;
}

// utility.rs:49-54
float schlick(float cosine, float ref_idx) {
    float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
    r0 = r0 * r0;
    return r0 + (1.0f - r0) * powi5(1.0f - cosine);
}

// ----------------------------------------------------------------------------------------------
// Hitable objects — shapes/hitable.rs:4-14, shapes/sphere.rs, shapes/plane.rs
// One tagged record stands in for Box<dyn Hitable>.  kind 2 (triangle) is the documented
// extension used by the JSON mapping (no reference behaviour).
// Layout is the C ABI of this oracle (mirrors, but is not shared with, forma_rt_object).
// ----------------------------------------------------------------------------------------------
struct Object {
    int32_t kind;      // 0 sphere, 1 plane, 2 triangle
    int32_t material;  // u8 in the reference: 0 lambertian, 1 metal, 2 dielectric, 3 "light"
    int32_t id;        // reported by primary_hits
    float fuzz;
    float color[3];
    float a[3];  // sphere center | plane position    | triangle v0
    float b[3];  // sphere radius,0,0 | plane orientation | triangle v1
    float c[3];  // unused        | plane size        | triangle v2
};
inline Vec3 v3(const float* f) { return {f[0], f[1], f[2]}; }

// shapes/sphere.rs:23-51
bool sphere_hit(const Object& s, Ray ray, float t_min, float t_max, HitRecord* hit_record) {
    Vec3 center = v3(s.a);
    float radius = s.b[0];
    Vec3 origin_to_center = ray.origin() - center;
    float a = Vec3::dot(ray.direction(), ray.direction());
    float b = Vec3::dot(origin_to_center, ray.direction());
    float c = Vec3::dot(origin_to_center, origin_to_center) - radius * radius;
    float discriminant = b * b - a * c;
    if (discriminant > 0.0f) {
        float root_one = (-b - std::sqrt(discriminant)) / a;
        if (root_one > t_min && root_one < t_max) {
            hit_record->t = root_one;
            hit_record->p = ray.point_at(root_one);
            hit_record->normal = (hit_record->p - center) / radius;
            return true;
        }
        float root_two = (-b + std::sqrt(discriminant)) / a;
        if (root_two > t_min && root_two < t_max) {
            hit_record->t = root_two;
            hit_record->p = ray.point_at(root_two);
            hit_record->normal = (hit_record->p - center) / radius;
            return true;
        }
    }
    return false;
}

// shapes/plane.rs:24-44  (Q4: the range test is on denom; Q5: t and p are written before the
// extent test and stay written on reject)
bool plane_hit(const Object& pl, Ray ray, float t_min, float t_max, HitRecord* hit_record) {
    Vec3 position = v3(pl.a), orientation = v3(pl.b), size = v3(pl.c);
    float denom = Vec3::dot(orientation, ray.direction());
    if (denom > t_min && denom < t_max) {
        Vec3 plane_to_ray = position - ray.origin();
        hit_record->t = Vec3::dot(plane_to_ray, orientation) / denom;
        hit_record->p = ray.point_at(hit_record->t);
        if (hit_record->p.x > position.x - size.x && hit_record->p.x < position.x + size.x &&
            hit_record->p.y > position.y - size.y && hit_record->p.y < position.y + size.y &&
            hit_record->p.z > position.z - size.z && hit_record->p.z < position.z + size.z) {
            hit_record->normal = orientation * -1.0f;
            return true;
        }
        return false;
    }
    return false;
}

// EXTENSION (no reference counterpart): Moeller-Trumbore, double sided, inclusive edges, strict
// (t_min, t_max) like sphere.rs:32; normal = unit(cross(e1,e2)) flipped to oppose the ray.
// Specified in DESIGN.md "Triangle extension".
bool triangle_hit(const Object& tr, Ray ray, float t_min, float t_max, HitRecord* hit_record) {
    Vec3 v0 = v3(tr.a), v1 = v3(tr.b), v2 = v3(tr.c);
    Vec3 e1 = v1 - v0, e2 = v2 - v0;
    Vec3 d = ray.direction();
    Vec3 pvec = Vec3::cross(d, e2);
    float det = Vec3::dot(e1, pvec);
    if (!(det != 0.0f)) return false;  // zero or NaN determinant: no hit
    Vec3 tvec = ray.origin() - v0;
    float uu = Vec3::dot(tvec, pvec);  // barycentric u scaled by det
    Vec3 qvec = Vec3::cross(tvec, e1);
    float vv = Vec3::dot(d, qvec);     // barycentric v scaled by det
    bool inside = det > 0.0f ? (uu >= 0.0f && vv >= 0.0f && uu + vv <= det)
                             : (uu <= 0.0f && vv <= 0.0f && uu + vv >= det);
    if (!inside) return false;
    float t = Vec3::dot(e2, qvec) / det;  // the only division, taken by rays inside the triangle
    if (t > t_min && t < t_max) {
        hit_record->t = t;
        hit_record->p = ray.point_at(t);
        Vec3 n = Vec3::cross(e1, e2).unit_vector();
        if (Vec3::dot(d, n) > 0.0f) n = n * -1.0f;
        hit_record->normal = n;
        return true;
    }
    return false;
}

bool object_hit(const Object& o, Ray ray, float t_min, float t_max, HitRecord* rec) {
    switch (o.kind) {
        case 0: return sphere_hit(o, ray, t_min, t_max, rec);
        case 1: return plane_hit(o, ray, t_min, t_max, rec);
        default: return triangle_hit(o, ray, t_min, t_max, rec);
    }
}

// sphere.rs:84-89 == plane.rs:101-106  (Q6: (p + n + r) - p in that order)
bool lambertian(const Object& o, HitRecord* hit_record, ReflectRecord* reflect_record) {
    Vec3 target = hit_record->p + hit_record->normal + random_in_unit_sphere();
    reflect_record->scattered = Ray{hit_record->p, target - hit_record->p};
    reflect_record->attenuation = v3(o.color);
    return true;
}

// sphere.rs:91-105 == plane.rs:108-122  (Q7: the sphere sample is drawn even when fuzz == 0)
bool metal(const Object& o, Ray ray, HitRecord* hit_record, ReflectRecord* reflect_record) {
    Vec3 reflected = reflect(ray.direction().unit_vector(), hit_record->normal);
    reflect_record->scattered = Ray{hit_record->p, reflected + o.fuzz * random_in_unit_sphere()};
    reflect_record->attenuation = v3(o.color);
    return Vec3::dot(reflect_record->scattered.direction(), hit_record->normal) > 0.0f;
}

// sphere.rs:107-145  (Q8)
bool dielectric(const Object& o, Ray ray, HitRecord* hit_record, ReflectRecord* reflect_record) {
    const float ref_idx = 1.3f;
    Vec3 outward_normal = Vec3::zero();
    Vec3 reflected = reflect(ray.direction(), hit_record->normal);
    float ni_over_nt;
    reflect_record->attenuation = v3(o.color);
    Vec3 refracted = Vec3::zero();
    float reflect_prob;
    float cosine;
    if (Vec3::dot(ray.direction(), hit_record->normal) > 0.0f) {
        outward_normal = outward_normal - hit_record->normal;
        ni_over_nt = ref_idx;
        cosine = ref_idx * Vec3::dot(ray.direction(), hit_record->normal) / ray.direction().length();
    } else {
        outward_normal = hit_record->normal;
        ni_over_nt = 1.0f / ref_idx;
        cosine = -Vec3::dot(ray.direction(), hit_record->normal) / ray.direction().length();
    }
    if (refract(ray.direction(), outward_normal, ni_over_nt, &refracted)) {
        reflect_prob = schlick(cosine, ref_idx);
    } else {
        reflect_prob = 1.0f;
    }
    if (get_random() < reflect_prob) {
        reflect_record->scattered = Ray{hit_record->p, reflected};
    } else {
        reflect_record->scattered = Ray{hit_record->p, refracted};
    }
    return true;
}

// sphere.rs:147-152  ("light" = lambertian with attenuation one; not emissive)
bool light(const Object&, HitRecord* hit_record, ReflectRecord* reflect_record) {
    Vec3 target = hit_record->p + hit_record->normal + random_in_unit_sphere();
    reflect_record->scattered = Ray{hit_record->p, target - hit_record->p};
    reflect_record->attenuation = Vec3::one();
    return true;
}

// sphere.rs:53-71 / plane.rs:46-60.  Triangles (extension) dispatch like Plane.
bool object_scatter(const Object& o, Ray ray, HitRecord* rec, ReflectRecord* refl) {
    if (o.kind == 0) {
        if (o.material == 0) return lambertian(o, rec, refl);
        else if (o.material == 1) return metal(o, ray, rec, refl);
        else if (o.material == 2) return dielectric(o, ray, rec, refl);
        else if (o.material == 3) return light(o, rec, refl);
        else return lambertian(o, rec, refl);
    }
    if (o.material == 0) return lambertian(o, rec, refl);
    else if (o.material == 1) return metal(o, ray, rec, refl);
    else return lambertian(o, rec, refl);
}

// ----------------------------------------------------------------------------------------------
// Camera — cpu_ray_tracer/camera.rs
// ----------------------------------------------------------------------------------------------
const float PI = 3.14159265359f;  // camera.rs:5

struct Camera {  // camera.rs:7-21 (C ABI of this oracle: 26 floats)
    Vec3 position, lower_left_corner;
    float aspect;
    Vec3 horizontal, vertical, u, v, w;
    float lens_radius, focus_dist, radius, rotation;
};

// camera.rs:24-60 generalised over (look_from, look_at, v_up, v_fov): Camera::new is the special
// case (0,0,1) -> (0,0,0), up +y, 60 deg.  The general form is the EXTENSION used for the JSON
// camera; arithmetic order is that of Camera::new.
Camera camera_look(Vec3 look_from, Vec3 look_at, Vec3 v_up, float v_fov, float aspect) {
    float aperture = 0.1f;
    float focus_dist = (look_from - look_at).length();
    float lens_radius = aperture / 2.0f;
    float theta = v_fov * PI / 180.0f;
    float half_height = std::tan(theta / 2.0f);
    float half_width = aspect * half_height;
    Vec3 position = look_from;
    Vec3 w = (look_from - look_at).unit_vector();
    Vec3 u = Vec3::cross(v_up, w).unit_vector();
    Vec3 v = Vec3::cross(w, u);
    Vec3 lower_left_corner =
        position - half_width * focus_dist * u - half_height * focus_dist * v - focus_dist * w;
    Vec3 horizontal = 2.0f * half_width * focus_dist * u;
    Vec3 vertical = 2.0f * half_height * focus_dist * v;
    Camera c;
    c.position = position; c.lower_left_corner = lower_left_corner; c.aspect = aspect;
    c.horizontal = horizontal; c.vertical = vertical; c.u = u; c.v = v; c.w = w;
    c.lens_radius = lens_radius;
    c.focus_dist = 2.0f; c.radius = 5.0f; c.rotation = 0.0f;  // camera.rs:56-58 (Q12)
    return c;
}

Camera camera_new(uint32_t width, uint32_t height) {  // camera.rs:24-60
    float aspect = (float)width / (float)height;
    return camera_look(Vec3{0.0f, 0.0f, 1.0f}, Vec3{0.0f, 0.0f, 0.0f}, Vec3{0.0f, 1.0f, 0.0f}, 60.0f, aspect);
}

// camera.rs:62-72
Ray camera_get_ray(const Camera& c, float s, float t, bool pinhole) {
    Vec3 offset = Vec3::zero();
    if (!pinhole) {
        Vec3 rd = c.lens_radius * random_in_unit_circle();
        offset = rd.x * c.u + rd.y * c.v;
    }
    return Ray{c.position + offset,
               c.lower_left_corner + s * c.horizontal + t * c.vertical - c.position - offset};
}

void camera_rebuild(Camera& c, Vec3 look_at) {  // shared tail of camera.rs:76-95 and :103-121
    Vec3 v_up{0.0f, 1.0f, 0.0f};
    float v_fov = 60.0f;
    float aperture = 0.1f;
    float focus_dist = (c.position - look_at).length();
    c.lens_radius = aperture / 2.0f;
    float theta = v_fov * PI / 180.0f;
    float half_height = std::tan(theta / 2.0f);
    float half_width = c.aspect * half_height;
    c.w = (c.position - look_at).unit_vector();
    c.u = Vec3::cross(v_up, c.w).unit_vector();
    c.v = Vec3::cross(c.w, c.u);
    c.lower_left_corner =
        c.position - half_width * focus_dist * c.u - half_height * focus_dist * c.v - focus_dist * c.w;
    c.horizontal = 2.0f * half_width * focus_dist * c.u;
    c.vertical = 2.0f * half_height * focus_dist * c.v;
}

void camera_translate(Camera& c, Vec3 delta) {  // camera.rs:74-95
    c.position = c.position + delta;
    camera_rebuild(c, Vec3{0.0f, 0.0f, -1.0f});
}

void camera_orbit(Camera& c, Vec3 delta) {  // camera.rs:97-122
    c.rotation += delta.x;
    c.radius += delta.z;
    c.position.x = c.radius * std::cos(c.rotation);
    c.position.y += delta.y;
    c.position.z = c.radius * std::sin(c.rotation);
    camera_rebuild(c, Vec3{0.0f, 0.0f, 0.0f});
}

// ----------------------------------------------------------------------------------------------
// get_color — cpu_ray_tracer/tracer.rs:189-219 (recursive, shared HitRecord across the scan)
// ----------------------------------------------------------------------------------------------
struct OBvh;
struct Trace {
    const Object* objects;
    size_t n;
    uint32_t max_depth;  // tracer.rs:10 MAX_DEPTH = 50 (u8)
    const OBvh* bvh = nullptr;  // checker-side accelerator (below); nullptr = the reference's scan
};

// ----------------------------------------------------------------------------------------------
// CHECKER-SIDE ACCELERATOR (no reference counterpart: tracer.rs:195-200 scans every object; shapes/aabb.rs is a
// stub).  Exists so that config D (100 000 spheres) can be checked at its full frame size: the brute-force scan
// costs O(n) per ray.  It returns EXACTLY what the scan returns and shares nothing with the product's LBVH:
//   * own tree: top-down median split of the centroid bounds, leaves of <= 4 primitives, boxes in double;
//   * the scan accepts object i iff its candidate t (sphere.rs:30-41: root_one if it exceeds t_min, else root_two;
//     the triangle's t) lies in (t_min, closest_so_far), so it ends with the lexicographic minimum of (t, i) over
//     all objects that have a candidate - which is what the walk keeps, in any visiting order - and with that
//     object's HitRecord, since a sphere / triangle only writes the record when it accepts;
//   * a node is skipped only when its box, grown by more than the distance at which sphere.rs:25-29's f32
//     discriminant can still turn positive for a true miss (~16 u |o - c|^2 / r, u = 2^-24; here 2^-18 (dist to the
//     box + its diagonal)^2 / r_min of the node, evaluated per node in double) plus 2^-12 of the scene scale for the
//     triangle test, is missed by the ray or entered beyond the best t so far.
// Only used when the list holds no plane (Plane::hit is order-sensitive, Q4 / Q5) and every coordinate is finite;
// tests/test_oracle_bvh.py checks it against the scan bit for bit before anything relies on it.
// ----------------------------------------------------------------------------------------------
struct OBvhNode {
    double lo[3], hi[3];
    double inv_rmin;       // largest 1 / |radius| below this node (0: triangles only)
    int left, right;       // children, or (first, -count) of a leaf's slice of `order`
};
struct OBvh {
    std::vector<OBvhNode> nodes;
    std::vector<uint32_t> order;  // primitive (list) indices, leaf slices contiguous
    double scale = 0.0;
};

inline void prim_box(const Object& o, double lo[3], double hi[3], double* inv_r) {
    if (o.kind == 0) {
        const double r = std::fabs((double)o.b[0]);
        for (int k = 0; k < 3; ++k) { lo[k] = (double)o.a[k] - r; hi[k] = (double)o.a[k] + r; }
        *inv_r = 1.0 / r;
    } else {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min({(double)o.a[k], (double)o.b[k], (double)o.c[k]});
            hi[k] = std::max({(double)o.a[k], (double)o.b[k], (double)o.c[k]});
        }
        *inv_r = 0.0;
    }
}

int obvh_build(OBvh& B, const Object* objs, uint32_t first, uint32_t count) {
    OBvhNode nd;
    double clo[3] = {1e300, 1e300, 1e300}, chi[3] = {-1e300, -1e300, -1e300};
    for (int k = 0; k < 3; ++k) { nd.lo[k] = 1e300; nd.hi[k] = -1e300; }
    nd.inv_rmin = 0.0;
    for (uint32_t i = first; i < first + count; ++i) {
        double lo[3], hi[3], q;
        prim_box(objs[B.order[i]], lo, hi, &q);
        nd.inv_rmin = std::max(nd.inv_rmin, q);
        for (int k = 0; k < 3; ++k) {
            nd.lo[k] = std::min(nd.lo[k], lo[k]); nd.hi[k] = std::max(nd.hi[k], hi[k]);
            const double c = 0.5 * (lo[k] + hi[k]);
            clo[k] = std::min(clo[k], c); chi[k] = std::max(chi[k], c);
        }
    }
    const int id = (int)B.nodes.size();
    B.nodes.push_back(nd);
    int axis = 0;
    for (int k = 1; k < 3; ++k) if (chi[k] - clo[k] > chi[axis] - clo[axis]) axis = k;
    if (count <= 4 || !(chi[axis] > clo[axis])) {
        B.nodes[id].left = (int)first; B.nodes[id].right = -(int)count;
        return id;
    }
    const uint32_t mid = first + count / 2;
    auto centre = [&](uint32_t p) {
        double lo[3], hi[3], q;
        prim_box(objs[p], lo, hi, &q);
        return lo[axis] + hi[axis];
    };
    std::nth_element(B.order.begin() + first, B.order.begin() + mid, B.order.begin() + first + count,
                     [&](uint32_t x, uint32_t y) { const double cx = centre(x), cy = centre(y); return cx < cy || (cx == cy && x < y); });
    const int l = obvh_build(B, objs, first, mid - first);
    const int r = obvh_build(B, objs, mid, first + count - mid);
    B.nodes[id].left = l; B.nodes[id].right = r;
    return id;
}

// false: the list cannot be accelerated exactly (a plane, or non-finite data) - the caller scans
bool obvh_make(OBvh& B, const Object* objs, size_t n) {
    if (n < 2 || n > 0x7fffffffu) return false;
    B.order.resize(n);
    for (size_t i = 0; i < n; ++i) {
        const Object& o = objs[i];
        if (o.kind != 0 && o.kind != 2) return false;
        double lo[3], hi[3], q;
        prim_box(o, lo, hi, &q);
        for (int k = 0; k < 3; ++k) {
            if (!std::isfinite(lo[k]) || !std::isfinite(hi[k])) return false;
            B.scale = std::max({B.scale, std::fabs(lo[k]), std::fabs(hi[k])});
        }
        if (o.kind == 0 && !std::isfinite(q)) return false;
        B.order[i] = (uint32_t)i;
    }
    B.nodes.reserve(n);
    obvh_build(B, objs, 0, (uint32_t)n);
    return true;
}

// entry parameter of the ray into the padded box, or -1 when it misses it (or enters beyond `best`)
inline double obvh_entry(const OBvh& B, const OBvhNode& nd, const double o[3], const double d[3], double best) {
    double dist2 = 0.0, diag2 = 0.0;
    for (int k = 0; k < 3; ++k) {
        const double c = 0.5 * (nd.lo[k] + nd.hi[k]), h = 0.5 * (nd.hi[k] - nd.lo[k]);
        dist2 += (o[k] - c) * (o[k] - c); diag2 += h * h;
    }
    const double reach = std::sqrt(dist2) + std::sqrt(diag2);
    const double pad = 3.814697265625e-06 * reach * reach * nd.inv_rmin + 2.44140625e-4 * B.scale + 1e-30;  // 2^-18, 2^-12
    double t0 = 0.0, t1 = best;
    for (int k = 0; k < 3; ++k) {
        const double lo = nd.lo[k] - pad, hi = nd.hi[k] + pad;
        if (d[k] == 0.0) { if (o[k] < lo || o[k] > hi) return -1.0; continue; }
        double a = (lo - o[k]) / d[k], b = (hi - o[k]) / d[k];
        if (a > b) std::swap(a, b);
        t0 = std::max(t0, a); t1 = std::min(t1, b);
    }
    if (!(t0 <= t1 * (1.0 + 1e-9) + 1e-9)) return -1.0;
    return t0;
}

// the scan of tracer.rs:195-200 over a plane-free list: winner index (or -1), its record and t
bool object_hit(const Object& o, Ray ray, float t_min, float t_max, HitRecord* rec);
int obvh_closest(const OBvh& B, const Object* objs, Ray ray, float t_min, HitRecord* rec_out, float* t_out) {
    const double o[3] = {ray.from.x, ray.from.y, ray.from.z}, d[3] = {ray.to.x, ray.to.y, ray.to.z};
    int best = -1;
    float best_t = std::numeric_limits<float>::max();
    int stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        const OBvhNode& nd = B.nodes[stack[--sp]];
        if (obvh_entry(B, nd, o, d, (double)best_t) < 0.0) continue;
        if (nd.right < 0) {
            for (int i = nd.left; i < nd.left - nd.right; ++i) {
                const int p = (int)B.order[i];
                HitRecord rec;
                // the candidate does not depend on t_max (see above): ask with the widest range, compare here
                if (object_hit(objs[p], ray, t_min, std::numeric_limits<float>::max(), &rec) &&
                    (rec.t < best_t || (rec.t == best_t && p < best))) {
                    best_t = rec.t; best = p; *rec_out = rec;
                }
            }
            continue;
        }
        // nearer child last on the stack = first off it
        const double el = obvh_entry(B, B.nodes[nd.left], o, d, (double)best_t);
        const double er = obvh_entry(B, B.nodes[nd.right], o, d, (double)best_t);
        if (sp + 2 > 128) return -2;  // cannot happen for a median-split tree over < 2^31 primitives
        if (el >= 0.0 && er >= 0.0) {
            if (el <= er) { stack[sp++] = nd.right; stack[sp++] = nd.left; }
            else { stack[sp++] = nd.left; stack[sp++] = nd.right; }
        } else if (el >= 0.0) stack[sp++] = nd.left;
        else if (er >= 0.0) stack[sp++] = nd.right;
    }
    *t_out = best_t;
    return best;
}

Vec3 get_color(Ray ray, const Trace& tr, uint32_t depth) {
    g_rng->rays++;
    HitRecord hit_record;
    const float t_min = 0.001f;
    float closest_so_far = std::numeric_limits<float>::max();
    const Object* temp_obj = nullptr;
    if (tr.bvh) {  // checker-side accelerator: the same winner and record as the scan below
        const int w = obvh_closest(*tr.bvh, tr.objects, ray, t_min, &hit_record, &closest_so_far);
        if (w >= 0) temp_obj = &tr.objects[w];
    } else
    for (size_t i = 0; i < tr.n; ++i) {
        const Object& obj = tr.objects[i];
        if (object_hit(obj, ray, t_min, closest_so_far, &hit_record)) {
            closest_so_far = hit_record.t;
            temp_obj = &obj;
        }
    }
    if (temp_obj) {
        ReflectRecord reflect_record{Ray{Vec3::zero(), Vec3::zero()}, Vec3::zero()};
        g_rng->slot = 1 + depth;  // counter RNG: one slot per scatter depth
        if (depth < tr.max_depth && object_scatter(*temp_obj, ray, &hit_record, &reflect_record)) {
            return reflect_record.attenuation * get_color(reflect_record.scattered, tr, depth + 1);
        }
        return Vec3::zero();
    }
    Vec3 unit_direction = ray.direction().unit_vector();
    float t = 0.5f * (unit_direction.y + 1.0f);
    return (1.0f - t) * Vec3{1.0f, 1.0f, 1.0f} + t * Vec3{0.5f, 0.7f, 1.0f};
}

// Rust `as u8` on f32: truncate toward zero, saturate, NaN -> 0 (tracer.rs:74-76, :181-183)
inline uint8_t as_u8(float f) {
    if (!(f == f)) return 0;
    if (f <= 0.0f) return 0;
    if (f >= 255.0f) return 255;
    return (uint8_t)(int)f;
}

enum : uint32_t { FLAG_NO_JITTER = 1, FLAG_PINHOLE = 2, FLAG_SEQ_RNG = 4, FLAG_MT_ROWS = 8, FLAG_BVH = 16 };

}  // namespace

// ================================================================================================
// C ABI (ctypes).  Struct layouts: `Object` (64 B) and `Camera` (26 floats) above.
// ================================================================================================
extern "C" {

int oracle_object_size() { return (int)sizeof(Object); }
int oracle_camera_floats() { return (int)(sizeof(Camera) / sizeof(float)); }

// ---- Vec3 known-answer surface (primitives.rs:163-254) ----
float oracle_vec3_length_squared(const float* a) { return v3(a).length_squared(); }
float oracle_vec3_length(const float* a) { return v3(a).length(); }
float oracle_vec3_dot(const float* a, const float* b) { return Vec3::dot(v3(a), v3(b)); }
void oracle_vec3_cross(const float* a, const float* b, float* o) { Vec3 r = Vec3::cross(v3(a), v3(b)); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void oracle_vec3_add(const float* a, const float* b, float* o) { Vec3 r = v3(a) + v3(b); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void oracle_vec3_sub(const float* a, const float* b, float* o) { Vec3 r = v3(a) - v3(b); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void oracle_vec3_mul(const float* a, const float* b, float* o) { Vec3 r = v3(a) * v3(b); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void oracle_vec3_scale(const float* a, float s, float* o) { Vec3 r = v3(a) * s; o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void oracle_vec3_lscale(float s, const float* a, float* o) { Vec3 r = s * v3(a); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void oracle_vec3_div(const float* a, float s, float* o) { Vec3 r = v3(a) / s; o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void oracle_vec3_unit(const float* a, float* o) { Vec3 r = v3(a).unit_vector(); o[0] = r.x; o[1] = r.y; o[2] = r.z; }

// ---- RNG ----
void oracle_philox4x32_10(uint32_t k0, uint32_t k1, const uint32_t* ctr, uint32_t* out) {
    philox4x32(10, k0, k1, ctr[0], ctr[1], ctr[2], ctr[3], out);
}
void oracle_philox4x32_r(int rounds, uint32_t k0, uint32_t k1, const uint32_t* ctr, uint32_t* out) {
    philox4x32(rounds, k0, k1, ctr[0], ctr[1], ctr[2], ctr[3], out);
}
int oracle_path_rounds() { return kPathRounds; }
float oracle_u32_to_f32(uint32_t v) { return u32_to_unit_f32(v); }

// ---- optics helpers ----
void oracle_reflect(const float* v, const float* n, float* o) { Vec3 r = reflect(v3(v), v3(n)); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
int oracle_refract(const float* v, const float* n, float ni_over_nt, float* o) {
    Vec3 r = Vec3::zero();
    bool ok = refract(v3(v), v3(n), ni_over_nt, &r);
    o[0] = r.x; o[1] = r.y; o[2] = r.z;
    return ok ? 1 : 0;
}
float oracle_schlick(float cosine, float ref_idx) { return schlick(cosine, ref_idx); }
int oracle_as_u8(float f) { return as_u8(f); }

// ---- single-object hit (unit tests): rec = t, p[3], normal[3]; pre-filled by the caller ----
int oracle_object_hit(const Object* o, const float* origin, const float* dir, float t_min, float t_max, float* rec) {
    HitRecord h;
    h.t = rec[0]; h.p = v3(rec + 1); h.normal = v3(rec + 4);
    bool hit = object_hit(*o, Ray{v3(origin), v3(dir)}, t_min, t_max, &h);
    rec[0] = h.t; rec[1] = h.p.x; rec[2] = h.p.y; rec[3] = h.p.z;
    rec[4] = h.normal.x; rec[5] = h.normal.y; rec[6] = h.normal.z;
    return hit ? 1 : 0;
}

// ---- camera ----
void oracle_camera_new(uint32_t w, uint32_t h, Camera* out) { *out = camera_new(w, h); }
void oracle_camera_look(const float* from, const float* at, const float* up, float v_fov, float aspect, Camera* out) {
    *out = camera_look(v3(from), v3(at), v3(up), v_fov, aspect);
}
void oracle_camera_orbit(Camera* c, const float* delta) { camera_orbit(*c, v3(delta)); }
void oracle_camera_translate(Camera* c, const float* delta) { camera_translate(*c, v3(delta)); }

// tracer.rs:30-52: key bits 00EQADWS -> delta -> camera.orbit(delta)
void oracle_update_camera(Camera* c, uint8_t keys, float delta_time) {
    Vec3 delta = Vec3::zero();
    if ((keys & 0b100000) == 0b100000) delta = delta + Vec3{0.0f, -1.0f, 0.0f} * delta_time;
    if ((keys & 0b10000) == 0b10000) delta = delta + Vec3{0.0f, 1.0f, 0.0f} * delta_time;
    if ((keys & 0b1000) == 0b1000) delta = delta + Vec3{1.0f, 0.0f, 0.0f} * delta_time;
    if ((keys & 0b100) == 0b100) delta = delta + Vec3{-1.0f, 0.0f, 0.0f} * delta_time;
    if ((keys & 0b10) == 0b10) delta = delta + Vec3{0.0f, 0.0f, -1.0f} * delta_time;
    if ((keys & 0b1) == 0b1) delta = delta + Vec3{0.0f, 0.0f, 1.0f} * delta_time;
    camera_orbit(*c, delta);
}

// ---- built-in scenes (cpu_ray_tracer/scenes.rs) ----
static Object mk_sphere(int id, Vec3 c, float r, int mat, Vec3 col, float fuzz) {
    Object o; std::memset(&o, 0, sizeof o);
    o.kind = 0; o.material = mat; o.id = id; o.fuzz = fuzz;
    o.color[0] = col.x; o.color[1] = col.y; o.color[2] = col.z;
    o.a[0] = c.x; o.a[1] = c.y; o.a[2] = c.z; o.b[0] = r;
    return o;
}
static Object mk_plane(int id, Vec3 pos, Vec3 ori, Vec3 size, int mat, Vec3 col, float fuzz) {
    Object o; std::memset(&o, 0, sizeof o);
    o.kind = 1; o.material = mat; o.id = id; o.fuzz = fuzz;
    o.color[0] = col.x; o.color[1] = col.y; o.color[2] = col.z;
    o.a[0] = pos.x; o.a[1] = pos.y; o.a[2] = pos.z;
    o.b[0] = ori.x; o.b[1] = ori.y; o.b[2] = ori.z;
    o.c[0] = size.x; o.c[1] = size.y; o.c[2] = size.z;
    return o;
}
// which: 0 get_simple_scene (scenes.rs:6-39), 1 get_plane_scene (:41-113), 2 get_objects (:115-156)
int oracle_builtin_scene(int which, Object* out, int cap) {
    std::vector<Object> v;
    if (which == 0) {
        v.push_back(mk_plane(0, {-1.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}, Vec3::one() * 5.0f, 1, {1.0f, 0.3f, 0.3f}, 0.05f));
        v.push_back(mk_sphere(1, {-0.5f, 0.0f, 0.0f}, 0.5f, 0, {0.0f, 0.66f, 0.13f}, 0.0f));
        v.push_back(mk_sphere(2, {0.5f, 0.0f, 0.0f}, 0.5f, 0, {0.7f, 0.43f, 0.0f}, 0.0f));
        v.push_back(mk_sphere(3, {0.0f, -1000.5f, 0.0f}, 1000.0f, 0, {0.3f, 0.3f, 0.3f}, 1.0f));
    } else if (which == 1) {
        v.push_back(mk_plane(0, {-1.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}, Vec3::one() * 100.0f, 1, {0.1f, 0.9f, 0.1f}, 0.0f));
    } else if (which == 2) {
        v.push_back(mk_sphere(0, {0.0f, 0.0f, -1.0f}, 0.5f, 0, {0.5f, 0.1f, 0.1f}, 0.0f));
        v.push_back(mk_sphere(1, {1.0f, 0.0f, -1.0f}, 0.5f, 1, {0.9f, 0.9f, 0.9f}, 0.2f));
        v.push_back(mk_sphere(2, {1.0f, 0.0f, -3.0f}, 0.5f, 1, {1.0f, 1.0f, 1.0f}, 1.0f));
        v.push_back(mk_sphere(3, {-1.0f, -0.0f, -1.0f}, 0.5f, 2, Vec3{0.1f, 0.5f, 0.1f}.sqrt().sqrt().sqrt(), 0.2f));
        v.push_back(mk_sphere(4, {0.0f, 0.0f, 1.0f}, 0.5f, 2, Vec3{0.5f, 0.5f, 0.3f}.sqrt().sqrt().sqrt(), 0.2f));
        v.push_back(mk_sphere(5, {0.0f, -100.5f, -1.0f}, 100.0f, 0, {0.1f, 0.3f, 0.9f}, 0.0f));
    } else {
        return -1;
    }
    int n = (int)v.size();
    for (int i = 0; i < n && i < cap; ++i) out[i] = v[i];
    return n;
}

// ---- primary-ray dump (config B): xi_u = xi_v = 0.5, pinhole; id = Object.id or -1; t = closest ----
void oracle_primary_hits(const Object* objs, uint64_t n, const Camera* cam, uint32_t width, uint32_t height,
                         int32_t* ids, float* ts) {
    RngCtx ctx;
    g_rng = &ctx;
    for (uint32_t y = 0; y < height; ++y) {
        for (uint32_t x = 0; x < width; ++x) {
            float u = ((float)x + 0.5f) / (float)width;
            float v = ((float)(height - y) + 0.5f) / (float)height;
            Ray ray = camera_get_ray(*cam, u, v, true);
            HitRecord rec;
            float closest = std::numeric_limits<float>::max();
            int32_t id = -1;
            for (uint64_t i = 0; i < n; ++i) {
                if (object_hit(objs[i], ray, 0.001f, closest, &rec)) {
                    closest = rec.t;
                    id = objs[i].id;
                }
            }
            ids[(size_t)x + (size_t)y * width] = id;
            ts[(size_t)x + (size_t)y * width] = closest;
        }
    }
    g_rng = nullptr;
}

// ---- render: tracer.rs:160-187 save_image's pixel/sample loops (and :57-81 render when spp = 1).
// Rows [y0, y1) only; samples get indices sample_offset .. sample_offset+spp-1 (sample-split).
// rgb (w*h*3, nullable): sqrt-gamma u8 of sum/spp.  linear_sum (w*h*3 f32, nullable): the raw
// per-pixel f32 sum in sample order, NOT divided.  Returns rays (get_color calls).
// nthreads > 1 interleaves rows over std::threads (the "all host cores" baseline); results are
// independent of nthreads in counter mode.
uint64_t oracle_render(const Object* objs, uint64_t n, const Camera* cam, uint32_t width, uint32_t height,
                       uint32_t spp, uint32_t sample_offset, uint32_t max_depth, uint64_t seed, uint32_t flags,
                       uint32_t y0, uint32_t y1, uint32_t nthreads, uint8_t* rgb, float* linear_sum) {
    if (nthreads == 0) nthreads = 1;
    if (y1 > height) y1 = height;
    std::atomic<uint64_t> total_rays{0};
    Trace tr{objs, (size_t)n, max_depth};
    OBvh bvh;  // FLAG_BVH: the checker-side accelerator when the list allows it, else the scan as always
    if ((flags & FLAG_BVH) && obvh_make(bvh, objs, (size_t)n)) tr.bvh = &bvh;
    auto worker = [&](uint32_t tid) {
        RngCtx ctx;
        ctx.mode = (flags & FLAG_SEQ_RNG) ? 1 : 0;
        ctx.k0 = (uint32_t)seed;
        ctx.k1 = (uint32_t)(seed >> 32);
        ctx.seed_seq(seed * 0x9E3779B97F4A7C15ull + tid);
        g_rng = &ctx;
        for (uint32_t y = y0 + tid; y < y1; y += nthreads) {
            for (uint32_t x = 0; x < width; ++x) {
                Vec3 col = Vec3::zero();
                for (uint32_t s = 0; s < spp; ++s) {
                    ctx.pixel = x + y * width;
                    ctx.sample = sample_offset + s;
                    ctx.slot = 0;
                    float xi_u, xi_v;
                    if (flags & FLAG_NO_JITTER) { xi_u = 0.5f; xi_v = 0.5f; }
                    else if (ctx.mode == 1) { xi_u = ctx.next_seq(); xi_v = ctx.next_seq(); }
                    else { float r[4]; ctx.block(0, r); xi_u = r[0]; xi_v = r[1]; }
                    float u = ((float)x + xi_u) / (float)width;                 // tracer.rs:171
                    float v = ((float)(height - y) + xi_v) / (float)height;     // tracer.rs:172 (Q1)
                    if (flags & FLAG_MT_ROWS) {
                        // render_mt (tracer.rs:83-134): four bands of t_height = height / 4 rows, each traced by its
                        // own thread with v = ((t_height - y) + xi) / height + t_id / 4 (:99-103), the bands appended
                        // in DESCENDING t_id (:126-131): output row y is row y % t_height of band 3 - y / t_height.
                        // Rows past 4 * t_height do not exist in its result (left zero here).
                        const uint32_t t_height = height / 4;
                        if (y >= 4 * t_height) continue;
                        const uint32_t t_id = 3 - y / t_height, yb = y % t_height;
                        v = ((float)(t_height - yb) + xi_v) / (float)height;
                        v += (float)t_id * (1.0f / 4.0f);
                    }
                    Ray ray = camera_get_ray(*cam, u, v, (flags & FLAG_PINHOLE) != 0);
                    col = col + get_color(ray, tr, 0);                          // tracer.rs:174
                }
                size_t idx = ((size_t)x + (size_t)y * width) * 3;
                if (linear_sum) { linear_sum[idx] = col.x; linear_sum[idx + 1] = col.y; linear_sum[idx + 2] = col.z; }
                if (rgb) {
                    Vec3 c = col / (float)spp;                                  // tracer.rs:177
                    c = Vec3{std::sqrt(c.r()), std::sqrt(c.g()), std::sqrt(c.b())};  // :178
                    rgb[idx] = as_u8(c.r() * 255.0f);                           // :180-184
                    rgb[idx + 1] = as_u8(c.g() * 255.0f);
                    rgb[idx + 2] = as_u8(c.b() * 255.0f);
                }
            }
        }
        total_rays += ctx.rays;
        g_rng = nullptr;
    };
    if (nthreads == 1) {
        worker(0);
    } else {
        std::vector<std::thread> th;
        for (uint32_t t = 0; t < nthreads; ++t) th.emplace_back(worker, t);
        for (auto& t : th) t.join();
    }
    return total_rays.load();
}

// tone-map a summed buffer exactly like tracer.rs:177-184 (used for multi-rank reduce checks)
void oracle_tonemap(const float* linear_sum, uint64_t npix, uint32_t spp, uint8_t* rgb) {
    for (uint64_t i = 0; i < npix; ++i) {
        Vec3 c = Vec3{linear_sum[3 * i], linear_sum[3 * i + 1], linear_sum[3 * i + 2]} / (float)spp;
        c = c.sqrt();
        rgb[3 * i] = as_u8(c.r() * 255.0f);
        rgb[3 * i + 1] = as_u8(c.g() * 255.0f);
        rgb[3 * i + 2] = as_u8(c.b() * 255.0f);
    }
}

}  // extern "C"
