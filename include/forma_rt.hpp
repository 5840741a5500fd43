// forma_rt.hpp — C++17 host-side mirror of the reference's tracer module over the C ABI (forma_rt.h).
//
// The reference is compiled code (Rust) and there is no Rust toolchain in this image, so this header is what the
// `forma_rt` crate of INTEGRATION.md would be, in the other compiled language the boundary speaks: the same module
// layout, names, argument meaning and error behaviour as src/cpu_ray_tracer/, every call going straight to
// libforma_rt.so.  Header-only; no CUDA, no torch.  Where the reference panics (`unwrap`), these throw forma::Error.
//
//   forma::primitives::Vec3                                   primitives.rs:4-156 (the fields; arithmetic stays on the GPU)
//   forma::scene::Scene { camera, objects }                   scene.rs:4-7
//   forma::scene::Object::translate / rotate                  shapes/hitable.rs:12-13 (plane.rs:62-68 assign; others no-op)
//   forma::camera::Camera::orbit / translate                  camera.rs:74-122
//   forma::tracer::TraceModel { scene, width, height, pixels } tracer.rs:12-17
//   forma::tracer::create_model(width, height)                tracer.rs:19-28
//   forma::tracer::update(model, keys, delta_time)            tracer.rs:30-55
//   forma::tracer::save_image(model, sample)                  tracer.rs:160-187 (writes out/basic.png)
//   forma::tracer::save_image_mt(model, sample)               tracer.rs:136-158 (writes out/basic_mt.png)
//   forma::scene_loader::construct_scene_from_json(model, s)  basics/scene_loader.rs:3-7 (+ the RT mapping)
#ifndef FORMA_RT_HPP
#define FORMA_RT_HPP

#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "forma_rt.h"

namespace forma {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error("forma_rt error " + std::to_string(c) + ": " + what), code(c) {}
};

inline void check(int rc) {
    if (rc != FORMA_RT_OK) throw Error(rc, forma_rt_last_error());
}

namespace primitives {
struct Vec3 {  // primitives.rs:4-9
    float x = 0.0f, y = 0.0f, z = 0.0f;
    Vec3() = default;
    Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}  // Vec3::new, primitives.rs:17-19
    static Vec3 zero() { return Vec3(); }                         // primitives.rs:25-31
};
}  // namespace primitives

namespace camera {
class Camera {  // camera.rs:7-21; lives in the handle, read back on demand
  public:
    explicit Camera(forma_rt* h) : h_(h) {}
    forma_rt_camera get() const { forma_rt_camera c; check(forma_rt_camera_get(h_, &c)); return c; }
    void orbit(primitives::Vec3 delta) { const float d[3] = {delta.x, delta.y, delta.z}; check(forma_rt_camera_orbit(h_, d)); }
    void translate(primitives::Vec3 delta) { const float d[3] = {delta.x, delta.y, delta.z}; check(forma_rt_camera_translate(h_, d)); }
    void use_json() { check(forma_rt_camera_json(h_)); }  // the camera of the last construct_scene_from_json (added)
  private:
    forma_rt* h_;
};
}  // namespace camera

namespace scene {
class Object {  // one `Box<dyn Hitable>` of Scene::objects
  public:
    Object(forma_rt* h, uint32_t i) : h_(h), i_(i) {}
    void translate(primitives::Vec3 v) { const float p[3] = {v.x, v.y, v.z}; check(forma_rt_set_object(h_, i_, p, nullptr)); }
    void rotate(primitives::Vec3 v) { const float o[3] = {v.x, v.y, v.z}; check(forma_rt_set_object(h_, i_, nullptr, o)); }
  private:
    forma_rt* h_;
    uint32_t i_;
};

class Objects {
  public:
    explicit Objects(forma_rt* h) : h_(h) {}
    size_t len() const { size_t n = 0; check(forma_rt_get_objects(h_, nullptr, 0, &n)); return n; }
    Object operator[](size_t i) const {
        if (i >= len()) throw std::out_of_range("index out of bounds: the len is " + std::to_string(len()) + " but the index is " + std::to_string(i));
        return Object(h_, (uint32_t)i);
    }
    std::vector<forma_rt_object> records() const {
        std::vector<forma_rt_object> v(len());
        size_t n = 0;
        check(forma_rt_get_objects(h_, v.data(), v.size(), &n));
        return v;
    }
  private:
    forma_rt* h_;
};

struct Scene {  // scene.rs:4-7
    camera::Camera camera;
    Objects objects;
    explicit Scene(forma_rt* h) : camera(h), objects(h) {}
};
}  // namespace scene

namespace tracer {
constexpr size_t CHANNEL_COUNT = 3;  // tracer.rs:9
constexpr uint8_t MAX_DEPTH = 50;    // tracer.rs:10

class TraceModel {  // tracer.rs:12-17
  public:
    scene::Scene scene;
    uint32_t width, height;
    std::vector<uint8_t> pixels;  // RGB8, row-major, row 0 = top

    TraceModel(uint32_t w, uint32_t h, const forma_rt_options* opts = nullptr)
        : scene(nullptr), width(w), height(h), pixels((size_t)w * h * CHANNEL_COUNT, 0) {
        check(forma_rt_create(opts, w, h, &h_));
        scene = scene::Scene(h_);
    }
    ~TraceModel() { if (h_) forma_rt_destroy(h_); }
    TraceModel(TraceModel&& o) noexcept : scene(o.h_), width(o.width), height(o.height), pixels(std::move(o.pixels)), h_(o.h_) { o.h_ = nullptr; }
    TraceModel(const TraceModel&) = delete;
    TraceModel& operator=(const TraceModel&) = delete;
    forma_rt* handle() const { return h_; }

  private:
    forma_rt* h_ = nullptr;
};

// tracer.rs:19-28.  `opts` (added): device ordinal, seed, flags; nullptr = device 0, seed 0.
inline TraceModel create_model(uint32_t width, uint32_t height, const forma_rt_options* opts = nullptr) {
    return TraceModel(width, height, opts);
}

// tracer.rs:30-55: keys = 00EQADWS; orbits the camera and replaces model.pixels with a fresh 1-spp frame
inline void update(TraceModel& model, uint8_t keys, float delta_time) {
    std::vector<uint8_t> px((size_t)model.width * model.height * CHANNEL_COUNT);
    check(forma_rt_update(model.handle(), keys, delta_time, px.data()));
    model.pixels = std::move(px);
}

// tracer.rs:160-187: `sample` samples per pixel, depth 50, written to out/basic.png; throws where the reference
// panics (tracer.rs:186 `unwrap`, e.g. ./out missing)
inline void save_image(const TraceModel& model, uint32_t sample, const char* path = "out/basic.png") {
    check(forma_rt_save_image(model.handle(), sample, path));
}

// tracer.rs:136-158: the 4-band variant, written to out/basic_mt.png (a fresh get_simple_scene(), post-gamma average)
inline void save_image_mt(const TraceModel& model, uint32_t sample, const char* path = "out/basic_mt.png") {
    check(forma_rt_save_image_mt(model.handle(), sample, path));
}

// save_image's loops without the PNG (added): the sqrt-gamma RGB8 frame of `sample` samples per pixel
inline std::vector<uint8_t> render(const TraceModel& model, uint32_t sample, uint32_t max_depth = MAX_DEPTH) {
    std::vector<uint8_t> px((size_t)model.width * model.height * CHANNEL_COUNT);
    check(forma_rt_render(model.handle(), sample, 0, max_depth, px.data(), nullptr));
    return px;
}
}  // namespace tracer

namespace scene_loader {
// basics/scene_loader.rs:3-7 `construct_scene_from_json(json: &str) -> SceneData`, here straight into the model's
// scene (the JSON -> ray-traceable mapping of DESIGN.md 4); throws where the reference's `unwrap` panics
inline void construct_scene_from_json(tracer::TraceModel& model, const std::string& json) {
    check(forma_rt_load_json(model.handle(), json.data(), json.size()));
}
}  // namespace scene_loader

}  // namespace forma
#endif  // FORMA_RT_HPP
