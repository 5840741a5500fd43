// scene_json.cpp — scenes/*.json (basics/scene_loader.rs:9-66 schema) -> Scene.objects.
//
// The reference tracer never reads these files (cpu_ray_tracer/scene.rs:10-15) and cannot trace
// triangles (shapes/aabb.rs, rectangle.rs are stubs), so this mapping is an EXTENSION, specified in
// DESIGN.md "Scene mapping":
//   * mesh "sphere"  -> analytic Sphere, center = position, radius = 0.5 * scale.x
//                       (primitives/sphere.rs:7 RADIUS = 0.5)
//   * any other mesh -> its triangle list from the reference's own mesh tables
//                       (primitives/{cube,quad,tetrahedron,triangle,circle,cylinder}.rs; unknown
//                       names fall back to the quad, basics/scene.rs:82-98), each vertex moved by
//                       glam's Mat4::from_scale_rotation_translation (primitives/primitive.rs:65-69)
//   * material       -> lambertian, colour CP0.palette[0] (color_utils.rs:101-109 via
//                       basics/scene.rs:292-294), unless the object carries the optional block
//                       "rt": {"material": "lambertian"|"metal"|"dielectric"|"light",
//                              "color": [r,g,b], "fuzz": f}
//   * id             -> (object index << 16) | triangle index within the mesh
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "json.hpp"
#include "scene.hpp"

namespace forma {
namespace {

struct Mesh {
    std::vector<V3> pos;
    std::vector<uint16_t> idx;
};

Mesh make_cube() {  // cube.rs:7-47: six faces x four corners, two triangles per face
    static const float P[24][3] = {
        {-.5f, -.5f, .5f},  {.5f, -.5f, .5f},  {.5f, .5f, .5f},  {-.5f, .5f, .5f},
        {-.5f, -.5f, -.5f}, {.5f, -.5f, -.5f}, {.5f, .5f, -.5f}, {-.5f, .5f, -.5f},
        {.5f, -.5f, -.5f},  {.5f, .5f, -.5f},  {.5f, .5f, .5f},  {.5f, -.5f, .5f},
        {-.5f, -.5f, -.5f}, {-.5f, .5f, -.5f}, {-.5f, .5f, .5f}, {-.5f, -.5f, .5f},
        {-.5f, .5f, -.5f},  {.5f, .5f, -.5f},  {.5f, .5f, .5f},  {-.5f, .5f, .5f},
        {-.5f, -.5f, -.5f}, {.5f, -.5f, -.5f}, {.5f, -.5f, .5f}, {-.5f, -.5f, .5f}};
    static const uint16_t I[36] = {0,  3,  1,  1,  3,  2,  4,  5,  6,  6,  7,  4,  8,  11, 10, 10, 9,  8,
                                   12, 13, 15, 13, 14, 15, 16, 17, 18, 16, 18, 19, 20, 23, 21, 21, 23, 22};
    Mesh m;
    for (auto& p : P) m.pos.push_back({p[0], p[1], p[2]});
    m.idx.assign(I, I + 36);
    return m;
}

Mesh make_quad() {  // quad.rs:6-17
    Mesh m;
    m.pos = {{-.5f, -.5f, 0.f}, {.5f, -.5f, 0.f}, {.5f, .5f, 0.f}, {-.5f, .5f, 0.f}};
    m.idx = {0, 1, 2, 2, 3, 0};
    return m;
}

Mesh make_tetrahedron() {  // tetrahedron.rs:10-32, X = 1
    Mesh m;
    m.pos = {{1, 1, -1},  {1, -1, 1},   {-1, 1, 1},  {-1, 1, 1}, {-1, -1, -1}, {1, 1, -1},
             {-1, 1, 1},  {1, -1, 1},   {-1, -1, -1}, {1, 1, -1}, {-1, -1, -1}, {1, -1, 1}};
    for (uint16_t i = 0; i < 12; ++i) m.idx.push_back(i);
    return m;
}

Mesh make_triangle() {  // triangle.rs:5-13
    Mesh m;
    m.pos = {{0.f, .5f, 0.f}, {-.5f, -.5f, 0.f}, {.5f, -.5f, 0.f}};
    m.idx = {0, 1, 2};
    return m;
}

// `i as f32 * 2.0 * PI / n as f32`, then RADIUS * cos / sin (circle.rs:68-70, cylinder.rs:72-74)
void sector(int i, int n, float& c, float& s) {
    const float pi = 3.14159265358979323846f;  // std::f32::consts::PI
    float ang = (float)i * 2.0f * pi / (float)n;
    c = 0.5f * std::cos(ang);
    s = 0.5f * std::sin(ang);
}

Mesh make_circle() {  // circle.rs:10-12,65-107
    const int n = 36;
    Mesh m;
    for (int i = 0; i < n; ++i) {
        float c, s;
        sector(i, n, c, s);
        m.pos.push_back({c, s, 0.f});
    }
    m.pos.push_back({0.f, 0.f, 0.f});
    for (int i = 0; i < n; ++i) {
        m.idx.push_back((uint16_t)i);
        m.idx.push_back((uint16_t)((i + 1) % n));
        m.idx.push_back((uint16_t)n);
    }
    return m;
}

Mesh make_cylinder() {  // cylinder.rs:68-190, sector_count = 30 (basics/scene.rs:90)
    const int n = 30;
    Mesh m;
    auto ring = [&](float y, bool with_centre) {
        for (int i = 0; i < n; ++i) {
            float c, s;
            sector(i, n, c, s);
            m.pos.push_back({c, y, s});
        }
        if (with_centre) m.pos.push_back({0.f, y, 0.f});
    };
    ring(.5f, true);    // top cap
    ring(-.5f, true);   // bottom cap
    ring(.5f, false);   // side, upper ring
    ring(-.5f, false);  // side, lower ring
    auto tri = [&](int a, int b, int c) {
        m.idx.push_back((uint16_t)a); m.idx.push_back((uint16_t)b); m.idx.push_back((uint16_t)c);
    };
    for (int i = 0; i < n; ++i) tri(i, (i + 1) % n, n);
    int off = n + 1;
    for (int i = 0; i < n; ++i) tri((i + 1) % n + off, i + off, n + off);
    off = 2 * (n + 1);
    for (int i = 0; i < n; ++i) {
        tri((i + 1) % n + off, i + off, (i + 1) % n + off + n);
        tri(i + off + n, (i + 1) % n + off + n, i + off);
    }
    return m;
}

const Mesh& mesh_for(const std::string& name) {
    static const Mesh cube = make_cube(), quad = make_quad(), tet = make_tetrahedron(), tri = make_triangle(),
                      circle = make_circle(), cyl = make_cylinder();
    if (name == "cube") return cube;
    if (name == "triangle") return tri;
    if (name == "circle") return circle;
    if (name == "cylinder") return cyl;
    if (name == "tetrahedron") return tet;
    return quad;  // basics/scene.rs:96-98
}

}  // namespace

// glam quat_to_axes (Mat4::from_scale_rotation_translation / Mat3::from_quat)
void quat_axes(const float q[4], V3& ax, V3& ay, V3& az) {
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float x2 = x + x, y2 = y + y, z2 = z + z;
    float xx = x * x2, xy = x * y2, xz = x * z2;
    float yy = y * y2, yz = y * z2, zz = z * z2;
    float wx = w * x2, wy = w * y2, wz = w * z2;
    ax = {1.0f - (yy + zz), xy + wz, xz - wy};
    ay = {xy - wz, 1.0f - (xx + zz), yz + wx};
    az = {xz + wy, yz - wx, 1.0f - (xx + yy)};
}

namespace {

using json::Value;

const Value& need(const Value& o, const char* key, Value::Type type, const char* ctx) {
    const Value* v = o.find(key);
    if (!v || v->type != type) throw std::runtime_error(std::string("scene JSON: missing or mistyped field `") + key + "` in " + ctx);
    return *v;
}
float num(const Value& o, const char* key, const char* ctx) { return (float)need(o, key, Value::Number, ctx).num; }
V3 vec3(const Value& o, const char* key, const char* ctx) {
    const Value& v = need(o, key, Value::Object, ctx);
    return {num(v, "x", key), num(v, "y", key), num(v, "z", key)};
}
void quat(const Value& o, const char* key, const char* ctx, float q[4]) {
    const Value& v = need(o, key, Value::Object, ctx);
    q[0] = num(v, "x", key); q[1] = num(v, "y", key); q[2] = num(v, "z", key); q[3] = num(v, "w", key);
}

}  // namespace

void load_json_scene(const char* text, size_t len, std::vector<forma_rt_object>& objects, JsonCamera& cam) {
    std::vector<forma_rt_object> out;
    size_t oi = 0;
    // objects[] entries are mapped as they are parsed (json.hpp parse_streaming): no DOM of the whole array
    auto on_object = [&](Value&& jo) {
        if (jo.type != Value::Object) throw std::runtime_error("scene JSON: objects[] entry is not an object");
        const std::string& mesh = need(jo, "mesh", Value::String, "object").str();
        need(jo, "material", Value::String, "object");  // rasteriser material name: required, not used here
        V3 position = vec3(jo, "position", "object");
        float rot[4];
        quat(jo, "rotation", "object", rot);
        V3 scale = vec3(jo, "scale", "object");

        forma_rt_object proto;
        std::memset(&proto, 0, sizeof proto);
        proto.material = 0;
        proto.color[0] = 0.263f; proto.color[1] = 0.208f; proto.color[2] = 0.655f;
        if (const Value* rt = jo.find("rt"); rt && rt->type == Value::Object) {
            if (const Value* m = rt->find("material"); m && m->type == Value::String) {
                if (m->str() == "metal") proto.material = 1;
                else if (m->str() == "dielectric") proto.material = 2;
                else if (m->str() == "light") proto.material = 3;
            }
            if (const Value* col = rt->find("color"); col && col->type == Value::Array && col->arr().size() == 3)
                for (int k = 0; k < 3; ++k) proto.color[k] = (float)col->arr()[k].num;
            if (const Value* f = rt->find("fuzz"); f && f->type == Value::Number) proto.fuzz = (float)f->num;
        }
        // ids pack the object index into the upper 16 bits; past 65535 objects it wraps (mod 2^16)
        int32_t id_base = (int32_t)((uint32_t)oi << 16);

        if (mesh == "sphere") {
            forma_rt_object o = proto;
            o.kind = FORMA_RT_SPHERE;
            o.id = id_base;
            o.a[0] = position.x; o.a[1] = position.y; o.a[2] = position.z;
            o.b[0] = 0.5f * scale.x;
            out.push_back(o);
        } else {
            const Mesh& m = mesh_for(mesh);
            V3 ax, ay, az;
            quat_axes(rot, ax, ay, az);
            V3 cx = ax * scale.x, cy = ay * scale.y, cz = az * scale.z;  // scaled matrix columns
            std::vector<V3> world(m.pos.size());
            for (size_t i = 0; i < m.pos.size(); ++i) {
                V3 p = m.pos[i];
                world[i] = {((cx.x * p.x + cy.x * p.y) + cz.x * p.z) + position.x,
                            ((cx.y * p.x + cy.y * p.y) + cz.y * p.z) + position.y,
                            ((cx.z * p.x + cy.z * p.y) + cz.z * p.z) + position.z};
            }
            for (size_t t = 0; t * 3 < m.idx.size(); ++t) {
                forma_rt_object o = proto;
                o.kind = FORMA_RT_TRIANGLE;
                o.id = id_base | (int32_t)t;
                V3 v0 = world[m.idx[3 * t]], v1 = world[m.idx[3 * t + 1]], v2 = world[m.idx[3 * t + 2]];
                o.a[0] = v0.x; o.a[1] = v0.y; o.a[2] = v0.z;
                o.b[0] = v1.x; o.b[1] = v1.y; o.b[2] = v1.z;
                o.c[0] = v2.x; o.c[1] = v2.y; o.c[2] = v2.z;
                out.push_back(o);
            }
        }
        ++oi;
    };
    Value root = json::Parser(text, len).parse_streaming("objects", on_object);
    if (root.type != Value::Object) throw std::runtime_error("scene JSON: top level is not an object");
    // SceneData { camera, lights, objects } — all three are required by serde (scene_loader.rs:10-14)
    const Value& jcam = need(root, "camera", Value::Object, "scene");
    need(root, "lights", Value::Array, "scene");
    need(root, "objects", Value::Array, "scene");

    JsonCamera c;
    c.present = true;
    c.position = vec3(jcam, "position", "camera");
    quat(jcam, "rotation", "camera", c.rotation);
    c.fov = num(jcam, "fov", "camera");

    objects.swap(out);
    cam = c;
}

}  // namespace forma
