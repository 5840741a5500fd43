// scene.hpp — host-side mirror of the reference's scene model for the tracer path:
// Scene{camera, objects} (cpu_ray_tracer/scene.rs:4-7), Camera (camera.rs), the built-in scenes
// (scenes.rs) and the JSON loader (basics/scene_loader.rs) with the JSON -> ray-traceable mapping.
// All f32 arithmetic follows the reference's evaluation order (no FMA contraction on the host:
// this file is compiled with -ffp-contract=off).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/forma_rt.h"

namespace forma {

struct V3 {
    float x, y, z;
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
float dot(V3 a, V3 b);
V3 cross(V3 a, V3 b);
float length(V3 a);
V3 unit(V3 a);

// Camera (camera.rs:7-21).  `c` is the wire form handed to the kernels.
struct Camera {
    forma_rt_camera c{};
    static Camera create(uint32_t width, uint32_t height);                       // Camera::new, camera.rs:24-60
    static Camera look(V3 from, V3 at, V3 up, float v_fov, float aspect);        // same arithmetic, any pose
    void translate(V3 delta);                                                    // camera.rs:74-95
    void orbit(V3 delta);                                                        // camera.rs:97-122
};

struct JsonCamera {
    bool present = false;
    V3 position{0, 0, 0};
    float rotation[4]{0, 0, 0, 1};  // quat xyzw
    float fov = 60.0f;
};

// Scene (scene.rs:4-7): the camera plus the object list the tracer scans in order.
struct Scene {
    Camera camera;
    std::vector<forma_rt_object> objects;
    JsonCamera json_camera;
};

// scenes.rs: 0 get_simple_scene, 1 get_plane_scene, 2 get_objects.  Throws std::invalid_argument.
std::vector<forma_rt_object> builtin_scene(int which);

// basics/scene_loader.rs:3-66 + the mapping to ray-traceable objects.  Throws json::ParseError /
// std::runtime_error with a field name on schema errors.
void load_json_scene(const char* text, size_t len, std::vector<forma_rt_object>& objects, JsonCamera& cam);

// glam quat (xyzw) -> rotation matrix columns (Mat4::from_scale_rotation_translation)
void quat_axes(const float q[4], V3& ax, V3& ay, V3& az);

// tracer.rs:31-51: 00EQADWS key bits -> orbit delta
V3 keys_to_delta(uint8_t keys, float delta_time);

// color_utils::srgb_encode (color_utils.rs:49-66) for every u8 input: the function is a u8 -> u8 map
void srgb_encode_table(uint8_t table[256]);

}  // namespace forma
