// postfx_oracle.cpp — CPU restatement of the reference's post-processing compute filters.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under fo-rma_b200/ may include, link or call this file; only tests/,
// __graft_entry__.smoke() and the cpu legs of the perf tools load it, as the checker for the CUDA path.
//
// PARITY UNPINNED: the reference runs these filters as WGSL compute shaders through wgpu, which is unavailable on
// this box (no graphics stack), and its repository holds no golden image for them.  What is restated here is the
// shader arithmetic as written, under these stated conventions:
//   * textures are Rgba8Unorm (rendering_utils.rs:168-193): textureLoad yields c / 255 (IEEE division, f32);
//     textureStore converts with the D3D / Vulkan float -> unorm rule: NaN -> 0, clamp to [0, 1], x * 255 + 0.5,
//     fraction dropped;
//   * every f32 operation is IEEE round-to-nearest in the order the shader writes it, never contracted
//     (-ffp-contract=off); fract(x) = x - floor(x); length(v) = sqrt(x*x + y*y);
//   * a textureLoad outside the texture yields (0, 0, 0, 0) (naga's ReadZeroSkipWrite policy); a textureStore outside
//     the texture is dropped;
//   * a chain follows PostProcessor::update_effects' ping-pong (post_processor.rs:133-197) on the state of the
//     reference's first frame: both intermediate textures and the output start zeroed.
// Each function cites the shader it follows (src/shaders/compute/<name>.comp.wgsl).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

// shader_utils.rs:58-71
enum Effect : int {
    kNone = 0, kNoise, kPixelate, kInvertColor, kWave, kInterlace, kFlipAxis, kGrayscale, kStep, kWatercolor,
    kChromostereopsis, kAnaglyph, kEffectCount
};

struct Px { float r, g, b, a; };

inline float decode(uint8_t c) { return (float)c / 255.0f; }

inline uint8_t encode(float x) {
    if (!(x == x)) return 0;
    if (x < 0.0f) x = 0.0f;
    if (x > 1.0f) x = 1.0f;
    return (uint8_t)(x * 255.0f + 0.5f);
}

struct Tex {
    const uint8_t* p; uint32_t w, h;
    Px load(int64_t x, int64_t y) const {
        if (x < 0 || y < 0 || x >= (int64_t)w || y >= (int64_t)h) return Px{0.0f, 0.0f, 0.0f, 0.0f};
        const uint8_t* q = p + ((size_t)y * w + (size_t)x) * 4;
        return Px{decode(q[0]), decode(q[1]), decode(q[2]), decode(q[3])};
    }
};

inline void store(uint8_t* dst, uint32_t w, uint32_t h, int64_t x, int64_t y, Px c) {
    if (x < 0 || y < 0 || x >= (int64_t)w || y >= (int64_t)h) return;
    uint8_t* q = dst + ((size_t)y * w + (size_t)x) * 4;
    q[0] = encode(c.r); q[1] = encode(c.g); q[2] = encode(c.b); q[3] = encode(c.a);
}

inline float fractf(float x) { return x - std::floor(x); }

// grayscale.comp.wgsl:15, step.comp.wgsl:15, wave.comp.wgsl:16 — "NTSC formula"
inline float luma(Px c) { return c.r * 0.299f + c.g * 0.587f + c.b * 0.114f; }

// noise.comp.wgsl:28-33
inline float hash2(uint32_t px, uint32_t py, float time) {
    const float fx = (float)px / 10.0f, fy = (float)py / 10.0f;
    const float t = time * 0.05f;
    const float vx = fx * 0.3183099f + t, vy = fy * 0.3678794f + t;
    return fractf(23.0f * fractf(vx * vy * (vx + vy)));
}

// watercolor.comp.wgsl:7-10
inline float hash_wc(float px, float py) {
    const float x = fractf(px * 0.1031f), y = fractf(py * 0.1031f), z = fractf(px * 0.1031f);
    const float d = x * (y + 33.333f) + y * (z + 33.333f) + z * (x + 33.333f);
    return fractf((x + y) * d);
}

Px shade(int effect, const Tex& src, uint32_t x, uint32_t y, float time) {
    switch (effect) {
    case kNone:  // none.comp.wgsl:14-15
        return src.load(x, y);
    case kNoise: {  // noise.comp.wgsl:22-25
        const Px c = src.load(x, y);
        const float n = hash2(x, y, time) / 20.0f;
        return Px{c.r + n, c.g + n, c.b + n, 1.0f};
    }
    case kPixelate:  // pixelate.comp.wgsl:16-22: PIXEL_SIZE 8, one sample per block
        return src.load((int64_t)(x / 8u) * 8, (int64_t)(y / 8u) * 8);
    case kInvertColor: {  // invert_color.comp.wgsl:14-17
        const Px c = src.load(x, y);
        return Px{1.0f - c.r, 1.0f - c.g, 1.0f - c.b, 1.0f};
    }
    case kWave: {  // wave.comp.wgsl:15-26
        const Px c = src.load(x, y);
        const float l = luma(c);
        Px s{l, l, l, 1.0f};
        if (c.r > 0.4f) s = Px{1.0f, c.g, 0.1f * c.b, 1.0f};
        if (c.g > 0.4f) s = Px{0.1f, c.g, 0.1f * c.b, 1.0f};
        if (c.b > 0.4f) s = Px{0.1f * c.r, c.g, 1.0f, 1.0f};
        return s;
    }
    case kInterlace: {  // interlace.comp.wgsl:14-16: select(1.0, 0.0, even row)
        const Px c = src.load(x, y);
        const float f = (y % 2u) == 0u ? 0.0f : 1.0f;
        return Px{c.r * f, c.g * f, c.b * f, c.a};
    }
    case kGrayscale: {  // grayscale.comp.wgsl:14-17
        const Px c = src.load(x, y);
        const float l = luma(c);
        return Px{l, l, l, c.a};
    }
    case kStep: {  // step.comp.wgsl:14-19
        const Px c = src.load(x, y);
        const float step_size = 0.2f;
        const float b = std::floor(luma(c) / step_size) * step_size;
        return Px{b, b, b, c.a};
    }
    case kWatercolor: {  // watercolor.comp.wgsl:20-53
        Px c = src.load(x, y);
        const float posx = (float)x, posy = (float)y;
        for (int i = 0; i < 50; ++i) {
            const float fi = (float)i;
            const float csx = fi * 123.45f, csy = 67.89f;
            const float rsx = fi * 234.56f, rsy = 78.9f;
            const float ksx = fi * 345.67f, ksy = 89.01f;
            const float qw = (float)src.w / 4.0f, qh = (float)src.h / 4.0f;
            const float cx = qw + hash_wc(csx, csy) * qw * 2.0f;
            const float cy = qh + hash_wc(csx + 1.0f, csy + 0.0f) * qh * 2.0f;
            const float radius = 10.0f + hash_wc(rsx, rsy) * 200.0f;
            const float red = hash_wc(ksx, ksy);
            const float dx = posx - cx, dy = posy - cy;
            const float distance = std::sqrt(dx * dx + dy * dy);
            if (distance <= radius) {
                c.r += red * 0.05f; c.g += 0.0f * 0.05f; c.b += 0.0f * 0.05f; c.a += 1.0f * 0.05f;
            }
        }
        return c;
    }
    case kChromostereopsis: {  // chromostereopsis.comp.wgsl:14-17: saturate(sign(r - b))
        const Px c = src.load(x, y);
        const float diff = c.r - c.b;
        const float sg = diff > 0.0f ? 1.0f : (diff < 0.0f ? -1.0f : 0.0f);
        const float r = sg < 0.0f ? 0.0f : (sg > 1.0f ? 1.0f : sg);
        return Px{r, 0.0f, 1.0f - r, 1.0f};
    }
    case kAnaglyph: {  // anaglyph.comp.wgsl:14-17: offset 10 pixels
        const float r = src.load((int64_t)x - 10, y).r;
        const float b = src.load((int64_t)x + 10, y).b;
        return Px{r, 0.0f, b, 1.0f};
    }
    default:
        return Px{0.0f, 0.0f, 0.0f, 0.0f};
    }
}

void apply_rows(int effect, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h, float time, uint32_t y0,
                uint32_t y1) {
    const Tex t{src, w, h};
    for (uint32_t y = y0; y < y1; ++y)
        for (uint32_t x = 0; x < w; ++x) {
            if (effect == kFlipAxis) store(dst, w, h, y, x, t.load(x, y));  // flip_axis.comp.wgsl:14-20: (x, y) -> (y, x)
            else store(dst, w, h, x, y, shade(effect, t, x, y, time));
        }
}

void apply(int effect, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h, float time, uint32_t threads) {
    if (threads <= 1 || h < threads) { apply_rows(effect, src, dst, w, h, time, 0, h); return; }
    std::vector<std::thread> pool;
    for (uint32_t k = 0; k < threads; ++k) {
        const uint32_t y0 = (uint32_t)((uint64_t)h * k / threads), y1 = (uint32_t)((uint64_t)h * (k + 1) / threads);
        pool.emplace_back(apply_rows, effect, src, dst, w, h, time, y0, y1);
    }
    for (auto& t : pool) t.join();
}

}  // namespace

extern "C" {

// One dispatch of one filter (post_processor.rs:111-125).  Texels of dst the filter does not write keep their value.
int fxo_apply(int effect, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h, float time, uint32_t threads) {
    if (effect < 0 || effect >= kEffectCount || !src || !dst || w == 0 || h == 0) return -1;
    apply(effect, src, dst, w, h, time, threads);
    return 0;
}

// The active effects in order, wired as PostProcessor::update_effects wires them (post_processor.rs:133-197):
//   effect 0 reads the input and writes the output if it is the only one, else intermediate 1;
//   the last effect writes the output and reads intermediate 1 if its index is odd, else intermediate 2;
//   a middle effect with an odd index reads intermediate 1 and writes intermediate 2, an even one the reverse.
// Intermediates and output start zeroed.  n == 0 runs nothing and leaves dst untouched.
int fxo_chain(const uint8_t* effects, uint32_t n, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h, float time,
              uint32_t threads) {
    if ((!effects && n) || !src || !dst || w == 0 || h == 0) return -1;
    for (uint32_t i = 0; i < n; ++i) if (effects[i] >= kEffectCount) return -1;
    if (n == 0) return 0;
    const size_t bytes = (size_t)w * h * 4;
    std::vector<uint8_t> t1(n > 1 ? bytes : 0, 0), t2(n > 2 ? bytes : 0, 0);
    std::memset(dst, 0, bytes);
    for (uint32_t i = 0; i < n; ++i) {
        const uint8_t* rd; uint8_t* wr;
        if (i == 0) { rd = src; wr = n == 1 ? dst : t1.data(); }
        else if (i == n - 1) { wr = dst; rd = (i % 2 == 1) ? t1.data() : t2.data(); }
        else if (i % 2 == 1) { rd = t1.data(); wr = t2.data(); }
        else { rd = t2.data(); wr = t1.data(); }
        apply(effects[i], rd, wr, w, h, time, threads);
    }
    return 0;
}

}  // extern "C"
