// scene.cpp — see scene.hpp.  Host-side scene model of the tracer path.
#include "scene.hpp"

#include <cmath>
#include <cstring>
#include <stdexcept>

#include "json.hpp"

namespace forma {

// ---- Vec3 (cpu_ray_tracer/primitives.rs:50-76) ---------------------------------------------------
float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x}; }
float length(V3 a) { return std::sqrt(dot(a, a)); }
V3 unit(V3 a) { return a / length(a); }

static void put(float* dst, V3 v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }
static V3 get(const float* p) { return {p[0], p[1], p[2]}; }

// ---- Camera (cpu_ray_tracer/camera.rs) -----------------------------------------------------------
static const float kPi = 3.14159265359f;  // camera.rs:5
static const float kAperture = 0.1f;      // camera.rs:30,80,107
static const float kVFov = 60.0f;         // camera.rs:28,79,106

// camera.rs:31-44 == :81-94 == :108-121: basis + frustum from (position, look_at, up, fov, aspect)
static void rebuild(forma_rt_camera& c, V3 look_at, V3 v_up, float v_fov) {
    V3 position = get(c.position);
    float focus_dist = length(position - look_at);
    c.lens_radius = kAperture / 2.0f;
    float theta = v_fov * kPi / 180.0f;
    float half_height = std::tan(theta / 2.0f);
    float half_width = c.aspect * half_height;
    V3 w = unit(position - look_at);
    V3 u = unit(cross(v_up, w));
    V3 v = cross(w, u);
    V3 llc = position - half_width * focus_dist * u - half_height * focus_dist * v - focus_dist * w;
    put(c.lower_left_corner, llc);
    put(c.horizontal, 2.0f * half_width * focus_dist * u);
    put(c.vertical, 2.0f * half_height * focus_dist * v);
    put(c.u, u);
    put(c.v, v);
    put(c.w, w);
}

Camera Camera::look(V3 from, V3 at, V3 up, float v_fov, float aspect) {
    Camera cam;
    put(cam.c.position, from);
    cam.c.aspect = aspect;
    rebuild(cam.c, at, up, v_fov);
    cam.c.focus_dist = 2.0f;  // camera.rs:56-58: stored constants unrelated to the basis (Q12)
    cam.c.radius = 5.0f;
    cam.c.rotation = 0.0f;
    return cam;
}

Camera Camera::create(uint32_t width, uint32_t height) {
    return look({0.0f, 0.0f, 1.0f}, {0.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, kVFov, (float)width / (float)height);
}

void Camera::translate(V3 delta) {
    put(c.position, get(c.position) + delta);
    rebuild(c, {0.0f, 0.0f, -1.0f}, {0.0f, 1.0f, 0.0f}, kVFov);
}

void Camera::orbit(V3 delta) {
    c.rotation += delta.x;
    c.radius += delta.z;
    c.position[0] = c.radius * std::cos(c.rotation);
    c.position[1] += delta.y;
    c.position[2] = c.radius * std::sin(c.rotation);
    rebuild(c, {0.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, kVFov);
}

V3 keys_to_delta(uint8_t keys, float dt) {
    static const struct { uint8_t bit; V3 dir; } table[6] = {
        {0x20, {0.0f, -1.0f, 0.0f}},  // E
        {0x10, {0.0f, 1.0f, 0.0f}},   // Q
        {0x08, {1.0f, 0.0f, 0.0f}},   // A
        {0x04, {-1.0f, 0.0f, 0.0f}},  // D
        {0x02, {0.0f, 0.0f, -1.0f}},  // W
        {0x01, {0.0f, 0.0f, 1.0f}},   // S
    };
    V3 delta{0.0f, 0.0f, 0.0f};
    for (const auto& e : table)
        if (keys & e.bit) delta = delta + e.dir * dt;
    return delta;
}

// ---- built-in scenes (cpu_ray_tracer/scenes.rs) --------------------------------------------------
static forma_rt_object sphere(int id, V3 center, float radius, int material, V3 color, float fuzz) {
    forma_rt_object o;
    std::memset(&o, 0, sizeof o);
    o.kind = FORMA_RT_SPHERE; o.material = material; o.id = id; o.fuzz = fuzz;
    put(o.color, color); put(o.a, center); o.b[0] = radius;
    return o;
}
static forma_rt_object plane(int id, V3 position, V3 orientation, V3 size, int material, V3 color, float fuzz) {
    forma_rt_object o;
    std::memset(&o, 0, sizeof o);
    o.kind = FORMA_RT_PLANE; o.material = material; o.id = id; o.fuzz = fuzz;
    put(o.color, color); put(o.a, position); put(o.b, orientation); put(o.c, size);
    return o;
}
static V3 sqrt3x(V3 v) {  // Vec3::sqrt().sqrt().sqrt(), scenes.rs:141,148
    for (int i = 0; i < 3; ++i) v = {std::sqrt(v.x), std::sqrt(v.y), std::sqrt(v.z)};
    return v;
}

std::vector<forma_rt_object> builtin_scene(int which) {
    std::vector<forma_rt_object> s;
    switch (which) {
        case 0:  // get_simple_scene, scenes.rs:6-39
            s.push_back(plane(0, {-1.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}, V3{1.0f, 1.0f, 1.0f} * 5.0f, 1, {1.0f, 0.3f, 0.3f}, 0.05f));
            s.push_back(sphere(1, {-0.5f, 0.0f, 0.0f}, 0.5f, 0, {0.0f, 0.66f, 0.13f}, 0.0f));
            s.push_back(sphere(2, {0.5f, 0.0f, 0.0f}, 0.5f, 0, {0.7f, 0.43f, 0.0f}, 0.0f));
            s.push_back(sphere(3, {0.0f, -1000.5f, 0.0f}, 1000.0f, 0, {0.3f, 0.3f, 0.3f}, 1.0f));
            break;
        case 1:  // get_plane_scene, scenes.rs:41-113 (one live plane; the rest is commented out)
            s.push_back(plane(0, {-1.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}, V3{1.0f, 1.0f, 1.0f} * 100.0f, 1, {0.1f, 0.9f, 0.1f}, 0.0f));
            break;
        case 2:  // get_objects, scenes.rs:115-156
            s.push_back(sphere(0, {0.0f, 0.0f, -1.0f}, 0.5f, 0, {0.5f, 0.1f, 0.1f}, 0.0f));
            s.push_back(sphere(1, {1.0f, 0.0f, -1.0f}, 0.5f, 1, {0.9f, 0.9f, 0.9f}, 0.2f));
            s.push_back(sphere(2, {1.0f, 0.0f, -3.0f}, 0.5f, 1, {1.0f, 1.0f, 1.0f}, 1.0f));
            s.push_back(sphere(3, {-1.0f, -0.0f, -1.0f}, 0.5f, 2, sqrt3x({0.1f, 0.5f, 0.1f}), 0.2f));
            s.push_back(sphere(4, {0.0f, 0.0f, 1.0f}, 0.5f, 2, sqrt3x({0.5f, 0.5f, 0.3f}), 0.2f));
            s.push_back(sphere(5, {0.0f, -100.5f, -1.0f}, 100.0f, 0, {0.1f, 0.3f, 0.9f}, 0.0f));
            break;
        default:
            throw std::invalid_argument("unknown built-in scene (0 simple, 1 plane, 2 objects)");
    }
    return s;
}

// color_utils.rs:49-66: linear = (c / 255).clamp(0, 1); srgb = linear <= 0.0031308 ? 12.92 linear
// : 1.055 linear^(1/2.4) - 0.055; (srgb * 255).round() as u8 - all in f32, round half away from zero.
void srgb_encode_table(uint8_t table[256]) {
    const float threshold = 0.0031308f, linear_scale = 12.92f, gamma = 1.0f / 2.4f, a = 1.055f, b = 0.055f;
    for (int i = 0; i < 256; ++i) {
        float linear = (float)i / 255.0f;
        linear = linear < 0.0f ? 0.0f : (linear > 1.0f ? 1.0f : linear);
        const float srgb = linear <= threshold ? linear * linear_scale : a * std::pow(linear, gamma) - b;
        const float r = std::round(srgb * 255.0f);
        table[i] = (uint8_t)(r < 0.0f ? 0.0f : (r > 255.0f ? 255.0f : r));
    }
}

}  // namespace forma
