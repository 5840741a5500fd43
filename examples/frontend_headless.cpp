// frontend_headless.cpp — the reference's interactive loop (src/frontend/macroquad.rs:4-98 `run`) without a window,
// written against include/forma_rt.hpp: create_model, then per frame objects[0].translate / rotate (the two egui
// sliders, macroquad.rs:12-13,66-67), tracer::update with the key bits (macroquad.rs:38-57,68), and the `R` key's
// save_image(&mut model, 50) (macroquad.rs:59-61) after the last frame.  The frame the reference copies into a
// texture (macroquad.rs:69-86) is summarised as an FNV-1a hash per frame on stdout.
//
//   frontend_headless <width> <height> <frames> [--keys 0x0a] [--dt 0.016] [--scene file.json] [--png out/basic.png]
//                     [--spp 50] [--seed 0]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>

#include "forma_rt.hpp"

using forma::primitives::Vec3;
namespace tracer = forma::tracer;

static uint64_t fnv1a(const std::vector<uint8_t>& v) {
    uint64_t h = 1469598103934665603ull;
    for (uint8_t b : v) { h ^= b; h *= 1099511628211ull; }
    return h;
}

int main(int argc, char** argv) {
    if (argc < 4) {
        std::fprintf(stderr, "usage: %s width height frames [--keys K] [--dt S] [--scene file.json] [--png path] [--spp N] [--seed N]\n", argv[0]);
        return 2;
    }
    const uint32_t width = (uint32_t)std::atoi(argv[1]), height = (uint32_t)std::atoi(argv[2]);
    const int frames = std::atoi(argv[3]);
    uint8_t keys = 0;
    float dt = 0.016f;
    uint32_t spp = 50;
    uint64_t seed = 0;
    std::string scene_file, png;
    for (int i = 4; i + 1 < argc; i += 2) {
        const std::string k = argv[i];
        if (k == "--keys") keys = (uint8_t)std::strtoul(argv[i + 1], nullptr, 0);
        else if (k == "--dt") dt = (float)std::atof(argv[i + 1]);
        else if (k == "--scene") scene_file = argv[i + 1];
        else if (k == "--png") png = argv[i + 1];
        else if (k == "--spp") spp = (uint32_t)std::atoi(argv[i + 1]);
        else if (k == "--seed") seed = std::strtoull(argv[i + 1], nullptr, 0);
        else { std::fprintf(stderr, "unknown option %s\n", k.c_str()); return 2; }
    }
    try {
        forma_rt_options opts{};
        opts.struct_size = sizeof opts;
        opts.seed = seed;
        tracer::TraceModel model = tracer::create_model(width, height, &opts);  // macroquad.rs:7
        if (!scene_file.empty()) {
            std::ifstream f(scene_file);
            if (!f) { std::fprintf(stderr, "cannot read %s\n", scene_file.c_str()); return 2; }
            std::stringstream ss;
            ss << f.rdbuf();
            forma::scene_loader::construct_scene_from_json(model, ss.str());
            model.scene.camera.use_json();
        }
        Vec3 translation(-1.0f, 0.0f, 0.0f), orientation(-1.0f, 0.0f, 0.0f);  // macroquad.rs:12-13
        for (int frame = 0; frame < frames; ++frame) {
            if (scene_file.empty()) {  // the sliders act on the built-in scene's plane (objects[0])
                orientation.y = 0.25f * (float)frame;  // a slider moved a notch per frame
                model.scene.objects[0].translate(translation);  // macroquad.rs:66
                model.scene.objects[0].rotate(orientation);     // macroquad.rs:67
            }
            const auto t0 = std::chrono::steady_clock::now();
            tracer::update(model, keys, dt);                    // macroquad.rs:68
            const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            std::printf("frame %d: %zu bytes, fnv1a %016llx, %.3f ms\n", frame, model.pixels.size(),
                        (unsigned long long)fnv1a(model.pixels), ms);
        }
        if (!png.empty()) {
            tracer::save_image(model, spp, png.c_str());        // macroquad.rs:59-61
            std::printf("saved %s (%u spp)\n", png.c_str(), spp);
        }
    } catch (const forma::Error& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
