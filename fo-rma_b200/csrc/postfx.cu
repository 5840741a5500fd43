// postfx.cu — the reference's post-processing compute filters (src/shaders/compute/*.comp.wgsl, chained by
// rendering/post_processor.rs) as CUDA kernels for sm_100a, behind include/forma_fx.h.
//
// Layout: a frame is width*height RGBA8 texels (4 B, the wgpu Rgba8Unorm texture), row-major, row 0 on top.
// Every filter reads 4 B and writes 4 B per texel, so a pass is bound by memory traffic: 8 B per texel.  What this file
// does about that:
//   * a run of per-texel filters is ONE pass: the chain is cut into segments [any filter][per-texel filter]*, each
//     segment one launch of fx_segment_kernel that keeps the texel in registers and re-quantises it to unorm8
//     between filters exactly as the reference's textureStore / textureLoad pair does, so the bytes are the same as
//     one dispatch per filter (FORMA_FX_UNFUSED runs that form; tests compare the two);
//   * a thread owns four neighbouring texels (one 16-byte load, one 16-byte store; a warp covers 512 contiguous
//     bytes) and keeps kInFlight such quads in flight: all loads of an iteration are issued before the first is
//     used, which is what it takes to cover the HBM latency (one quad per thread left the pass at 35 % of the copy
//     peak); grids are a multiple of the SM count with a grid-stride loop;
//   * flip_axis is a 32x32 shared-memory tile transpose, so both its loads and its stores are coalesced;
//   * watercolor's 50 circles depend on the frame size only: a 64-thread kernel evaluates them once per resize
//     with the shader's own arithmetic, the pass reads them from shared memory.  The shader's
//     `sqrt(d2) <= radius` is evaluated as `d2 <= T` with T the largest float whose IEEE square root is <= radius
//     (sqrt is monotonic, so the two predicates are the same set), and a circle is skipped for a whole row when
//     dy*dy alone exceeds T.
// Arithmetic conventions are stated in include/forma_fx.h; the build has -fmad=false, IEEE division and sqrt.
#include "postfx.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/forma_fx.h"
#include "divconst.cuh"
#include "kernels.hpp"

namespace {

constexpr int kFxBlock = 256;
constexpr uint32_t kMaxOps = 16;       // per-texel filters fused into one pass
constexpr int kCircles = 50;           // watercolor.comp.wgsl:26
constexpr int kInFlight = 4;           // quads (16 B) a thread loads before it touches the first

struct FxPass {
    const uint8_t* src;
    uint8_t* dst;
    uint32_t w, h, quads_per_row;
    float time;
    uint32_t n_ops;
    uint32_t fast_div;          // n / 10 and n / 20 as multiply + correction (verified when the handle was created)
    uint32_t ops[kMaxOps / 4];  // one filter per byte
    const float4* circles;  // (centre x, centre y, T: d2 <= T <=> sqrt(d2) <= radius, red * 0.05)
};

__device__ __forceinline__ float fractf_(float x) { return x - floorf(x); }

// textureStore to Rgba8Unorm: NaN -> 0, clamp, x * 255 + 0.5, fraction dropped.  The unorm8 value is the LOW BYTE of
// the result (the other bytes are not zero): the truncation is an add of 2^23 rounded toward -inf, which for
// 0 <= v < 2^23 leaves floor(v) in the low mantissa bits (full-rate FADD.RM instead of a quarter-rate F2I).
__device__ __forceinline__ uint32_t enc(float x) {
    const float v = __fadd_rn(__fmul_rn(__saturatef(x), 255.0f), 0.5f);
    return __float_as_uint(__fadd_rd(v, 8388608.0f));
}
// textureLoad from Rgba8Unorm: channel k of texel t as c / 255 (lut[c] holds that quotient)
__device__ __forceinline__ float chan(uint32_t t, int k, const float* lut) { return lut[(t >> (8 * k)) & 255u]; }

// A filter maps packed texels to packed texels: it decodes only the channels it reads and encodes only the channels
// it computes.  A channel it passes through keeps its byte, which is what textureStore(textureLoad(c)) gives for every
// c (enc(c / 255) == c for all 256 values; checked in tests/); so two filters in a row need no conversion between them.
__device__ __forceinline__ uint32_t gray_texel(uint32_t e, uint32_t t) { return __byte_perm(e, t, 0x7000); }  // e,e,e,t.a
__device__ __forceinline__ uint32_t rgb1_texel(uint32_t er, uint32_t eg, uint32_t eb) {                       // r,g,b,255
    return __byte_perm(__byte_perm(er, eg, 0x0040), eb, 0x4410) | 0xFF000000u;
}

__device__ __forceinline__ float luma(float r, float g, float b) { return r * 0.299f + g * 0.587f + b * 0.114f; }

// noise.comp.wgsl:28-33
__device__ __forceinline__ float hash2(uint32_t px, uint32_t py, float time, bool fast) {
    const float fx = forma::div_const((float)px, 10.0f, 0.1f, fast), fy = forma::div_const((float)py, 10.0f, 0.1f, fast);
    const float t = time * 0.05f;
    const float vx = fx * 0.3183099f + t, vy = fy * 0.3678794f + t;
    return fractf_(23.0f * fractf_(vx * vy * (vx + vy)));
}

// watercolor.comp.wgsl:7-10
__device__ __forceinline__ float hash_wc(float px, float py) {
    const float x = fractf_(px * 0.1031f), y = fractf_(py * 0.1031f), z = fractf_(px * 0.1031f);
    const float d = x * (y + 33.333f) + y * (z + 33.333f) + z * (x + 33.333f);
    return fractf_((x + y) * d);
}

// watercolor.comp.wgsl:26-47 — everything in the loop body that does not depend on the texel
__global__ void watercolor_circles_kernel(uint32_t w, uint32_t h, float4* out) {
    const int i = threadIdx.x;
    if (i >= kCircles) return;
    const float fi = (float)i;
    const float csx = fi * 123.45f, csy = 67.89f;
    const float rsx = fi * 234.56f, rsy = 78.9f;
    const float ksx = fi * 345.67f, ksy = 89.01f;
    const float qw = (float)w / 4.0f, qh = (float)h / 4.0f;
    const float cx = qw + hash_wc(csx, csy) * qw * 2.0f;
    const float cy = qh + hash_wc(csx + 1.0f, csy + 0.0f) * qh * 2.0f;
    const float radius = 10.0f + hash_wc(rsx, rsy) * 200.0f;
    const float red = hash_wc(ksx, ksy);
    // T = max { t : sqrtf(t) <= radius }  (radius >= 10; IEEE sqrt is monotonic, so the search is a few ulps)
    float t = radius * radius;
    while (sqrtf(t) > radius) t = __uint_as_float(__float_as_uint(t) - 1u);
    while (sqrtf(__uint_as_float(__float_as_uint(t) + 1u)) <= radius) t = __uint_as_float(__float_as_uint(t) + 1u);
    out[i] = make_float4(cx, cy, t, red * 0.05f);
}

// one per-texel filter on the four packed texels of a thread; `op` is uniform over the grid
__device__ __forceinline__ void apply_op(int op, uint32_t (&t)[4], uint32_t x0, uint32_t y, float time, bool fast,
                                         const float* lut, const float4* circ) {
    switch (op) {
    case FORMA_FX_NOISE:  // noise.comp.wgsl:22-25: (rgb + hash / 20, 1)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // hash2 is 0 over large parts of a big frame (the product has no fraction bits left); 0 / 20 would take
            // the division's slow path
            const float hsh = hash2(x0 + i, y, time, fast);
            const float n = fast && hsh >= 5.9604645e-08f ? forma::div_const(hsh, 20.0f, 0.05f, true)
                                                          : (hsh == 0.0f ? 0.0f : hsh / 20.0f);
            t[i] = rgb1_texel(enc(chan(t[i], 0, lut) + n), enc(chan(t[i], 1, lut) + n), enc(chan(t[i], 2, lut) + n));
        }
        break;
    case FORMA_FX_INVERT_COLOR:  // invert_color.comp.wgsl:14-17: (1 - rgb, 1); enc(1 - c / 255) == 255 - c for every c
#pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = ~t[i] | 0xFF000000u;
        break;
    case FORMA_FX_WAVE:  // wave.comp.wgsl:15-26
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float r = chan(t[i], 0, lut), g = chan(t[i], 1, lut), b = chan(t[i], 2, lut);
            const uint32_t el = enc(luma(r, g, b));
            uint32_t er = el, eb = el;
            bool keep_g = false;
            if (r > 0.4f) { er = 255u; eb = enc(0.1f * b); keep_g = true; }
            if (g > 0.4f) { er = enc(0.1f); eb = enc(0.1f * b); keep_g = true; }
            if (b > 0.4f) { er = enc(0.1f * r); eb = 255u; keep_g = true; }
            t[i] = rgb1_texel(er, keep_g ? t[i] >> 8 : el, eb);
        }
        break;
    case FORMA_FX_INTERLACE:  // interlace.comp.wgsl:14-16: rgb * 0 on even rows, rgb * 1 on odd ones, alpha stays
        if ((y & 1u) == 0u) {
#pragma unroll
            for (int i = 0; i < 4; ++i) t[i] &= 0xFF000000u;
        }
        break;
    case FORMA_FX_GRAYSCALE:  // grayscale.comp.wgsl:14-17
#pragma unroll
        for (int i = 0; i < 4; ++i)
            t[i] = gray_texel(enc(luma(chan(t[i], 0, lut), chan(t[i], 1, lut), chan(t[i], 2, lut))), t[i]);
        break;
    case FORMA_FX_STEP:  // step.comp.wgsl:14-19
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float l = luma(chan(t[i], 0, lut), chan(t[i], 1, lut), chan(t[i], 2, lut));
            t[i] = gray_texel(enc(floorf(l / 0.2f) * 0.2f), t[i]);
        }
        break;
    case FORMA_FX_WATERCOLOR: {  // watercolor.comp.wgsl:26-52: 50 circles, each adds (red, 0, 0, 1) * 0.05 inside
        const float py = (float)y, px = (float)x0;
        float r[4], a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { r[i] = chan(t[i], 0, lut); a[i] = chan(t[i], 3, lut); }
        for (int k = 0; k < kCircles; ++k) {
            const float4 q = circ[k];
            const float dy = py - q.y;
            const float dy2 = dy * dy;
            if (dy2 > q.z) continue;  // RN(dx*dx + dy2) >= dy2: no texel of this row is inside
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float dx = (px + (float)i) - q.x;  // x0 + i < 2^24: the sum is exact, = f32(x)
                if (dx * dx + dy2 <= q.z) { r[i] += q.w; a[i] += 0.05f; }
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)  // (enc r, g, b, enc a)
            t[i] = __byte_perm(__byte_perm(enc(r[i]), t[i], 0x7650), enc(a[i]), 0x4210);
        break;
    }
    case FORMA_FX_CHROMOSTEREOPSIS:  // chromostereopsis.comp.wgsl:14-17: r = saturate(sign(r - b)) -> (r, 0, 1 - r, 1);
        // c / 255 is strictly increasing in c, so r - b > 0 exactly when the red byte is the larger one
#pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = (t[i] & 255u) > ((t[i] >> 16) & 255u) ? 0xFF0000FFu : 0xFFFF0000u;
        break;
    default:  // none (and pixelate / anaglyph, whose work is the gather in front of the pass)
        break;
    }
}

__shared__ float s_lut[256];          // c / 255
__shared__ float4 s_circ[kCircles];   // watercolor's circles
__shared__ uint32_t s_ops[kMaxOps];   // the filters of a fused pass

// The filters of a fused pass on one quad.  Not inlined: the pass keeps kInFlight quads in flight, and four copies of
// every filter's code did not fit the instruction cache (no_instruction was the top stall).
__device__ __noinline__ uint4 run_filters(uint4 tv, uint32_t x0, uint32_t y, float time, uint32_t n_ops, bool fast) {
    uint32_t t[4] = {tv.x, tv.y, tv.z, tv.w};
    for (uint32_t k = 0; k < n_ops; ++k) {
        // the coordinates go through an opaque move so that no filter's texel-independent part (noise's hash) is
        // hoisted out of this loop and evaluated for passes that do not contain the filter
        uint32_t xo = x0, yo = y;
        asm volatile("" : "+r"(xo), "+r"(yo));
        apply_op((int)s_ops[k], t, xo, yo, time, fast, s_lut, s_circ);
    }
    return make_uint4(t[0], t[1], t[2], t[3]);
}

// ONE >= 0: the pass is that single filter, compiled as straight-line code; ONE < 0: the filter list of a fused pass.
template <bool VEC, int ONE>
__global__ void __launch_bounds__(kFxBlock, 4) fx_segment_kernel(const FxPass P) {
    float* const lut = s_lut;
    float4* const circ = s_circ;
    uint32_t* const ops = s_ops;
    lut[threadIdx.x] = (float)threadIdx.x / 255.0f;
    if (threadIdx.x < kMaxOps) {
        const uint32_t word = threadIdx.x < 4 ? P.ops[0] : threadIdx.x < 8 ? P.ops[1] : threadIdx.x < 12 ? P.ops[2] : P.ops[3];
        ops[threadIdx.x] = (word >> ((threadIdx.x & 3u) * 8u)) & 255u;
    }
    __syncthreads();
    bool has_wc = ONE == FORMA_FX_WATERCOLOR, arithmetic = false;
    if (ONE < 0) {
        for (uint32_t k = 0; k < P.n_ops; ++k) {
            const int op = (int)ops[k];
            has_wc |= op == FORMA_FX_WATERCOLOR;
            arithmetic |= !(op == FORMA_FX_NONE || op == FORMA_FX_PIXELATE || op == FORMA_FX_ANAGLYPH);
        }
    } else {
        arithmetic = !(ONE == FORMA_FX_NONE || ONE == FORMA_FX_PIXELATE || ONE == FORMA_FX_ANAGLYPH);
    }
    if (has_wc && threadIdx.x < kCircles) circ[threadIdx.x] = P.circles[threadIdx.x];
    __syncthreads();

    const uint32_t* src = reinterpret_cast<const uint32_t*>(P.src);
    uint32_t* dst = reinterpret_cast<uint32_t*>(P.dst);
    const uint32_t quads = P.quads_per_row * P.h;  // <= 2^29
    const int first = ONE >= 0 ? ONE : (int)(P.ops[0] & 255u);
    const uint32_t span = gridDim.x * (kFxBlock * kInFlight);
    for (uint32_t base = blockIdx.x * (kFxBlock * kInFlight) + threadIdx.x; base < quads; base += span) {
        uint32_t t[kInFlight][4];
        uint32_t xs[kInFlight], ys[kInFlight];
#pragma unroll
        for (int u = 0; u < kInFlight; ++u) {  // every load of the iteration goes out before the first use
            const uint32_t q = base + u * kFxBlock;
            if (q >= quads) { xs[u] = 0xFFFFFFFFu; continue; }
            const uint32_t y = q / P.quads_per_row;
            const uint32_t x0 = (q - y * P.quads_per_row) * 4u;
            xs[u] = x0; ys[u] = y;
            const size_t row = (size_t)y * P.w;
            if (first == FORMA_FX_PIXELATE) {  // pixelate.comp.wgsl:16-22: x0 % 4 == 0, so the four share a block
                const uint32_t v = __ldg(src + (size_t)(y & ~7u) * P.w + (x0 & ~7u));
                t[u][0] = t[u][1] = t[u][2] = t[u][3] = v;
            } else if (first == FORMA_FX_ANAGLYPH && VEC) {
                // anaglyph.comp.wgsl:14-17: red from x - 10, blue from x + 10.  x0 % 4 == 0 and w % 4 == 0, so the
                // texels come as aligned pairs (x0 - 10, x0 - 9), (x0 - 8, x0 - 7), (x0 + 10, ..), (x0 + 12, ..), each
                // pair wholly inside or wholly outside the row
                const uint2* row2 = reinterpret_cast<const uint2*>(src + row);
                const uint2 z = make_uint2(0u, 0u);
                const uint2 r0 = x0 >= 12u ? __ldg(row2 + (x0 - 10u) / 2u) : z;
                const uint2 r1 = x0 >= 8u ? __ldg(row2 + (x0 - 8u) / 2u) : z;
                const uint2 b0 = x0 + 10u < P.w ? __ldg(row2 + (x0 + 10u) / 2u) : z;
                const uint2 b1 = x0 + 12u < P.w ? __ldg(row2 + (x0 + 12u) / 2u) : z;
                t[u][0] = (r0.x & 255u) | (b0.x & 0x00FF0000u) | 0xFF000000u;
                t[u][1] = (r0.y & 255u) | (b0.y & 0x00FF0000u) | 0xFF000000u;
                t[u][2] = (r1.x & 255u) | (b1.x & 0x00FF0000u) | 0xFF000000u;
                t[u][3] = (r1.y & 255u) | (b1.y & 0x00FF0000u) | 0xFF000000u;
            } else if (first == FORMA_FX_ANAGLYPH) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t x = x0 + i;
                    const uint32_t r = x >= 10u && x - 10u < P.w ? __ldg(src + row + x - 10u) & 255u : 0u;
                    const uint32_t b = x + 10u < P.w ? __ldg(src + row + x + 10u) & 0x00FF0000u : 0u;
                    t[u][i] = r | b | 0xFF000000u;
                }
            } else if (VEC) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + row + x0));
                t[u][0] = v.x; t[u][1] = v.y; t[u][2] = v.z; t[u][3] = v.w;
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) t[u][i] = x0 + i < P.w ? __ldg(src + row + x0 + i) : 0u;
            }
        }
#pragma unroll
        for (int u = 0; u < kInFlight; ++u) {
            if (xs[u] == 0xFFFFFFFFu) continue;
            const uint32_t x0 = xs[u], y = ys[u];
            if (ONE >= 0) {
                apply_op(ONE, t[u], x0, y, P.time, P.fast_div != 0u, lut, circ);
            } else if (arithmetic) {  // else only none / pixelate / anaglyph: the gather in front was all the work
                const uint4 r = run_filters(make_uint4(t[u][0], t[u][1], t[u][2], t[u][3]), x0, y, P.time, P.n_ops,
                                            P.fast_div != 0u);
                t[u][0] = r.x; t[u][1] = r.y; t[u][2] = r.z; t[u][3] = r.w;
            }
            const size_t row = (size_t)y * P.w;
            if (VEC) {
                *reinterpret_cast<uint4*>(dst + row + x0) = make_uint4(t[u][0], t[u][1], t[u][2], t[u][3]);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) if (x0 + i < P.w) dst[row + x0 + i] = t[u][i];
            }
        }
    }
}

// flip_axis.comp.wgsl:14-20: texel (x, y) goes to (y, x); targets outside the frame are dropped.
// A block moves a 2x2 group of 32x32 tiles per iteration (16 loads per thread in flight before the barrier; rows of
// 256 contiguous bytes on both the read and the write side).
constexpr uint32_t kFlipTiles = 4;
__global__ void __launch_bounds__(256) fx_flip_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst,
                                                     uint32_t w, uint32_t h, uint32_t groups_x, uint32_t groups) {
    __shared__ uint32_t tile[kFlipTiles][32][33];
    const uint32_t lx = threadIdx.x & 31u, ly = threadIdx.x >> 5;
    for (uint32_t g = blockIdx.x; g < groups; g += gridDim.x) {
        const uint32_t gx = (g % groups_x) * 64u, gy = (g / groups_x) * 64u;
#pragma unroll
        for (uint32_t u = 0; u < kFlipTiles; ++u) {
            const uint32_t tx = gx + (u & 1u) * 32u, ty = gy + (u >> 1) * 32u;
#pragma unroll
            for (uint32_t j = ly; j < 32u; j += 8u) {
                const uint32_t x = tx + lx, y = ty + j;
                if (x < w && y < h) tile[u][j][lx] = __ldg(src + (size_t)y * w + x);
            }
        }
        __syncthreads();
#pragma unroll
        for (uint32_t u = 0; u < kFlipTiles; ++u) {
            const uint32_t tx = gx + (u >> 1) * 32u, ty = gy + (u & 1u) * 32u;  // u & 1 walks along the destination row
            const uint32_t v = ((u & 1u) << 1) | (u >> 1);                        // the tile loaded from (tx, ty)
#pragma unroll
            for (uint32_t j = ly; j < 32u; j += 8u) {
                // source texel (sx, sy) = (tx + j, ty + lx) lands at column sy, row sx
                const uint32_t sx = tx + j, sy = ty + lx;
                if (sx < w && sy < h && sy < w && sx < h) dst[(size_t)sx * w + sy] = tile[v][lx][j];
            }
        }
        __syncthreads();
    }
}

// RGB8 (the tracer's frame, tracer.rs:177-184) -> RGBA8 with alpha 255, four texels per thread
__global__ void __launch_bounds__(256) fx_expand_kernel(const uint8_t* __restrict__ rgb, uint32_t* __restrict__ rgba,
                                                       uint64_t npix) {
    const uint64_t quads = npix / 4u;
    const uint32_t* in = reinterpret_cast<const uint32_t*>(rgb);
    for (uint64_t q = (uint64_t)blockIdx.x * 256u + threadIdx.x; q < quads; q += (uint64_t)gridDim.x * 256u) {
        const uint32_t a = __ldg(in + 3 * q), b = __ldg(in + 3 * q + 1), c = __ldg(in + 3 * q + 2);
        const uint4 o = make_uint4((a & 0x00FFFFFFu) | 0xFF000000u, (a >> 24) | ((b & 0xFFFFu) << 8) | 0xFF000000u,
                                   (b >> 16) | ((c & 0xFFu) << 16) | 0xFF000000u, (c >> 8) | 0xFF000000u);
        *reinterpret_cast<uint4*>(rgba + 4 * q) = o;
    }
    if (blockIdx.x == 0 && threadIdx.x < (npix & 3u)) {
        const uint64_t p = quads * 4u + threadIdx.x;
        rgba[p] = (uint32_t)rgb[3 * p] | ((uint32_t)rgb[3 * p + 1] << 8) | ((uint32_t)rgb[3 * p + 2] << 16) | 0xFF000000u;
    }
}

inline bool per_texel(int e) {
    return e != FORMA_FX_FLIP_AXIS && e != FORMA_FX_PIXELATE && e != FORMA_FX_ANAGLYPH;
}

}  // namespace

// =====================================================================================================
// handle
// =====================================================================================================
struct forma_fx {
    int device = 0;
    uint32_t width = 0, height = 0, flags = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_k0 = nullptr, ev_k1 = nullptr, ev_end = nullptr;
    std::vector<uint8_t> effects;
    // a pass = effects [first, first + count) in one launch; rd / wr: 0 input, 1 / 2 intermediates, 3 output
    struct Pass { uint32_t first, count; int rd, wr; bool zero_rest; };
    std::vector<Pass> plan;
    bool fast_div = false;    // noise's divisions by 10 and 20 verified against IEEE division on this device
    bool ragged_flip = false;  // flip_axis on a non-square frame: the chain runs unfused on the reference's wiring
    uint8_t *d_t1 = nullptr, *d_t2 = nullptr, *d_in = nullptr, *d_out = nullptr;
    float4* d_circles = nullptr;
    forma_fx_stats stats{};
    bool stats_pending = false;

    size_t bytes() const { return (size_t)width * height * 4; }
    void release_frames() {
        for (uint8_t** p : {&d_t1, &d_t2, &d_in, &d_out}) { if (*p) cudaFree(*p); *p = nullptr; }
    }
    ~forma_fx() {
        cudaSetDevice(device);
        if (stream) cudaStreamSynchronize(stream);
        release_frames();
        if (d_circles) cudaFree(d_circles);
        for (cudaEvent_t e : {ev_begin, ev_k0, ev_k1, ev_end}) if (e) cudaEventDestroy(e);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {

using forma::set_last_error;

#define FX_TRY(expr)                                                                                    \
    do {                                                                                                \
        cudaError_t e__ = (expr);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return set_last_error(FORMA_RT_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e__)); \
    } while (0)
#define NEED_FX(f) do { if (!(f)) return set_last_error(FORMA_RT_EINVAL, "null forma_fx handle"); cudaSetDevice((f)->device); } while (0)

int alloc_frames(forma_fx* f) {
    f->release_frames();
    for (uint8_t** p : {&f->d_t1, &f->d_t2, &f->d_in, &f->d_out}) {
        FX_TRY(cudaMalloc((void**)p, f->bytes()));
        FX_TRY(cudaMemsetAsync(*p, 0, f->bytes(), f->stream));  // wgpu textures start zeroed
    }
    watercolor_circles_kernel<<<1, 64, 0, f->stream>>>(f->width, f->height, f->d_circles);
    FX_TRY(cudaGetLastError());
    FX_TRY(cudaStreamSynchronize(f->stream));  // later runs may come in on a caller's stream
    return FORMA_RT_OK;
}

// update_effects' wiring (post_processor.rs:142-194) when every effect is its own pass, else the fused segments
void make_plan(forma_fx* f) {
    f->plan.clear();
    const uint32_t n = (uint32_t)f->effects.size();
    bool has_flip = false;
    for (uint8_t e : f->effects) has_flip |= e == FORMA_FX_FLIP_AXIS;
    f->ragged_flip = has_flip && f->width != f->height;
    if ((f->flags & FORMA_FX_UNFUSED) || f->ragged_flip) {
        bool written[4] = {false, false, false, false};  // this run has filled the texture
        for (uint32_t i = 0; i < n; ++i) {
            int rd, wr;
            if (i == 0) { rd = 0; wr = n == 1 ? 3 : 1; }
            else if (i == n - 1) { wr = 3; rd = (i % 2 == 1) ? 1 : 2; }
            else if (i % 2 == 1) { rd = 1; wr = 2; }
            else { rd = 2; wr = 1; }
            // flip_axis on a non-square frame leaves texels unwritten.  They keep the texture's content: an earlier
            // pass's output if this run has written the texture, else the zeros of the reference's first frame.
            const bool zero_rest = f->ragged_flip && f->effects[i] == FORMA_FX_FLIP_AXIS && !written[wr];
            written[wr] = true;
            f->plan.push_back({i, 1, rd, wr, zero_rest});
        }
    } else {
        uint32_t i = 0;
        while (i < n) {
            uint32_t cnt = 1;
            if (f->effects[i] != FORMA_FX_FLIP_AXIS)
                while (i + cnt < n && cnt < kMaxOps && per_texel(f->effects[i + cnt])) ++cnt;
            f->plan.push_back({i, cnt, 0, 0, false});
            i += cnt;
        }
        const size_t m = f->plan.size();
        for (size_t k = 0; k < m; ++k) {
            f->plan[k].rd = k == 0 ? 0 : f->plan[k - 1].wr;
            f->plan[k].wr = k == m - 1 ? 3 : (k % 2 == 0 ? 1 : 2);
        }
    }
    f->stats.n_effects = n;
    f->stats.n_passes = (uint32_t)f->plan.size();
}

template <int ONE>
void launch_one(bool vec, uint32_t grid, cudaStream_t s, const FxPass& P) {
    if (vec) fx_segment_kernel<true, ONE><<<grid, kFxBlock, 0, s>>>(P);
    else fx_segment_kernel<false, ONE><<<grid, kFxBlock, 0, s>>>(P);
}

void launch_segment(int one, bool vec, uint32_t grid, cudaStream_t s, const FxPass& P) {
    switch (one) {
    case FORMA_FX_NONE: launch_one<FORMA_FX_NONE>(vec, grid, s, P); break;
    case FORMA_FX_NOISE: launch_one<FORMA_FX_NOISE>(vec, grid, s, P); break;
    case FORMA_FX_PIXELATE: launch_one<FORMA_FX_PIXELATE>(vec, grid, s, P); break;
    case FORMA_FX_INVERT_COLOR: launch_one<FORMA_FX_INVERT_COLOR>(vec, grid, s, P); break;
    case FORMA_FX_WAVE: launch_one<FORMA_FX_WAVE>(vec, grid, s, P); break;
    case FORMA_FX_INTERLACE: launch_one<FORMA_FX_INTERLACE>(vec, grid, s, P); break;
    case FORMA_FX_GRAYSCALE: launch_one<FORMA_FX_GRAYSCALE>(vec, grid, s, P); break;
    case FORMA_FX_STEP: launch_one<FORMA_FX_STEP>(vec, grid, s, P); break;
    case FORMA_FX_WATERCOLOR: launch_one<FORMA_FX_WATERCOLOR>(vec, grid, s, P); break;
    case FORMA_FX_CHROMOSTEREOPSIS: launch_one<FORMA_FX_CHROMOSTEREOPSIS>(vec, grid, s, P); break;
    case FORMA_FX_ANAGLYPH: launch_one<FORMA_FX_ANAGLYPH>(vec, grid, s, P); break;
    default: launch_one<-1>(vec, grid, s, P); break;
    }
}

int launch_pass(forma_fx* f, const uint8_t* effects, uint32_t count, const uint8_t* src, uint8_t* dst, float time,
                cudaStream_t s) {
    const uint32_t w = f->width, h = f->height;
    if (effects[0] == FORMA_FX_FLIP_AXIS) {
        const uint32_t groups_x = (w + 63) / 64, groups = groups_x * ((h + 63) / 64);
        const uint32_t grid = std::min<uint32_t>(groups, (uint32_t)f->sm_count * 8u);
        fx_flip_kernel<<<grid, 256, 0, s>>>((const uint32_t*)src, (uint32_t*)dst, w, h, groups_x, groups);
    } else {
        FxPass P{};
        P.src = src; P.dst = dst; P.w = w; P.h = h; P.quads_per_row = (w + 3) / 4; P.time = time;
        P.n_ops = count;
        for (uint32_t k = 0; k < count; ++k) P.ops[k / 4] |= (uint32_t)effects[k] << ((k % 4) * 8);
        P.circles = f->d_circles;
        P.fast_div = f->fast_div && w <= (1u << 24) && h <= (1u << 24) ? 1u : 0u;
        const uint64_t quads = (uint64_t)P.quads_per_row * h;
        const uint64_t per_block = (uint64_t)kFxBlock * kInFlight;
        const uint64_t want = (quads + per_block - 1) / per_block;
        const uint32_t grid = (uint32_t)std::min<uint64_t>(want, (uint64_t)f->sm_count * 8u);
        const bool vec = w % 4 == 0 && (((uintptr_t)src | (uintptr_t)dst) & 15u) == 0;
        launch_segment(count == 1 ? (int)effects[0] : -1, vec, grid, s, P);
    }
    FX_TRY(cudaGetLastError());
    return FORMA_RT_OK;
}

// the chain on device frames; events ev_k0 / ev_k1 bracket the passes when s is the handle's stream
int run_chain(forma_fx* f, const uint8_t* d_src, uint8_t* d_dst, float time, cudaStream_t s) {
    const bool own = s == f->stream;
    if (own) FX_TRY(cudaEventRecord(f->ev_k0, s));
    uint32_t launches = 0;
    for (const forma_fx::Pass& p : f->plan) {
        const uint8_t* rd = p.rd == 0 ? d_src : (p.rd == 1 ? f->d_t1 : f->d_t2);
        uint8_t* wr = p.wr == 3 ? d_dst : (p.wr == 1 ? f->d_t1 : f->d_t2);
        if (p.zero_rest) {  // the texels outside the min(w, h) square, which flip_axis cannot reach
            const size_t w = f->width, h = f->height;
            if (w > h) FX_TRY(cudaMemset2DAsync(wr + h * 4, w * 4, 0, (w - h) * 4, h, s));
            else FX_TRY(cudaMemsetAsync(wr + w * w * 4, 0, (h - w) * w * 4, s));
        }
        if (int rc = launch_pass(f, f->effects.data() + p.first, p.count, rd, wr, time, s)) return rc;
        ++launches;
    }
    if (own) { FX_TRY(cudaEventRecord(f->ev_k1, s)); f->stats_pending = true; }
    f->stats.gpu_launches = launches;
    f->stats.bytes_moved = (uint64_t)launches * f->bytes() * 2u;
    return FORMA_RT_OK;
}

}  // namespace

namespace forma {

int fx_device(const forma_fx* fx) { return fx->device; }
uint32_t fx_width(const forma_fx* fx) { return fx->width; }
uint32_t fx_height(const forma_fx* fx) { return fx->height; }

int fx_run_from_rgb8(forma_fx* f, const uint8_t* d_rgb, float time, cudaStream_t s, const uint8_t** d_out) {
    const uint64_t npix = (uint64_t)f->width * f->height;
    const uint32_t grid = (uint32_t)std::min<uint64_t>((npix / 4 + 255) / 256 + 1, (uint64_t)f->sm_count * 16u);
    fx_expand_kernel<<<grid, 256, 0, s>>>(d_rgb, (uint32_t*)f->d_in, npix);
    FX_TRY(cudaGetLastError());
    if (f->plan.empty()) { *d_out = f->d_in; return FORMA_RT_OK; }  // no active effect: the frame as it is
    if (int rc = run_chain(f, f->d_in, f->d_out, time, s)) return rc;
    f->stats.gpu_launches += 1;
    *d_out = f->d_out;
    return FORMA_RT_OK;
}

}  // namespace forma

// =====================================================================================================
// C ABI (include/forma_fx.h)
// =====================================================================================================
extern "C" {

int forma_fx_create(int device, uint32_t width, uint32_t height, uint32_t flags, forma_fx** out) {
    if (!out) return set_last_error(FORMA_RT_EINVAL, "forma_fx_create: out is null");
    *out = nullptr;
    if (width == 0 || height == 0 || (uint64_t)width * height > (1ull << 31))
        return set_last_error(FORMA_RT_EINVAL, "forma_fx_create: width/height out of range");
    if (flags & ~FORMA_FX_UNFUSED) return set_last_error(FORMA_RT_EINVAL, "forma_fx_create: unknown flag");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return set_last_error(FORMA_RT_ECUDA, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return set_last_error(FORMA_RT_EINVAL, "forma_fx_create: device ordinal out of range");
    FX_TRY(cudaSetDevice(device));
    forma_fx* f = new (std::nothrow) forma_fx();
    if (!f) return set_last_error(FORMA_RT_ENOMEM, "forma_fx_create: out of host memory");
    f->device = device; f->width = width; f->height = height; f->flags = flags;
    auto bail = [&](int rc) { delete f; return rc; };
    if ((e = cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking)) != cudaSuccess)
        return bail(set_last_error(FORMA_RT_ECUDA, std::string("cudaStreamCreate: ") + cudaGetErrorString(e)));
    for (cudaEvent_t* ev : {&f->ev_begin, &f->ev_k0, &f->ev_k1, &f->ev_end})
        if ((e = cudaEventCreate(ev)) != cudaSuccess)
            return bail(set_last_error(FORMA_RT_ECUDA, std::string("cudaEventCreate: ") + cudaGetErrorString(e)));
    if ((e = cudaMalloc((void**)&f->d_circles, 64 * sizeof(float4))) != cudaSuccess)
        return bail(set_last_error(FORMA_RT_ECUDA, std::string("cudaMalloc: ") + cudaGetErrorString(e)));
    cudaDeviceGetAttribute(&f->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (int rc = alloc_frames(f)) return bail(rc);
    {   // noise divides texel coordinates (integers < 2^24) by 10 and a hash in [0, 1) by 20: compare the multiply form
        // with IEEE division for every f32 numerator in those ranges once; any mismatch keeps the division
        unsigned int* d_bad = reinterpret_cast<unsigned int*>(f->d_circles + 60);
        unsigned int bad = 1;
        const char* env = std::getenv("FORMA_RT_FAST_DIV");
        if (!env || std::atoi(env) != 0) {
            if (cudaMemsetAsync(d_bad, 0, sizeof bad, f->stream) == cudaSuccess &&
                forma::launch_div_verify(10.0f, 16777216.0f, d_bad, f->sm_count, f->stream) == cudaSuccess &&
                forma::launch_div_verify(20.0f, 1.0f, d_bad, f->sm_count, f->stream) == cudaSuccess &&
                cudaMemcpyAsync(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost, f->stream) == cudaSuccess &&
                cudaStreamSynchronize(f->stream) == cudaSuccess)
                f->fast_div = bad == 0;
        }
    }
    make_plan(f);
    *out = f;
    return FORMA_RT_OK;
}

void forma_fx_destroy(forma_fx* fx) { delete fx; }

int forma_fx_resize(forma_fx* fx, uint32_t width, uint32_t height) {
    NEED_FX(fx);
    if (width == 0 || height == 0 || (uint64_t)width * height > (1ull << 31))
        return set_last_error(FORMA_RT_EINVAL, "forma_fx_resize: width/height out of range");
    FX_TRY(cudaStreamSynchronize(fx->stream));
    fx->width = width; fx->height = height;
    if (int rc = alloc_frames(fx)) return rc;
    make_plan(fx);
    return FORMA_RT_OK;
}

int forma_fx_update_effects(forma_fx* fx, const uint8_t* effects, uint32_t n) {
    NEED_FX(fx);
    if ((!effects && n) || n > FORMA_FX_MAX_EFFECTS)
        return set_last_error(FORMA_RT_EINVAL, "forma_fx_update_effects: bad effect list");
    for (uint32_t i = 0; i < n; ++i)
        if (effects[i] >= FORMA_FX_EFFECT_COUNT)
            return set_last_error(FORMA_RT_EINVAL, "forma_fx_update_effects: unknown effect " + std::to_string(effects[i]));
    fx->effects.assign(effects, effects + n);
    make_plan(fx);
    return FORMA_RT_OK;
}

int forma_fx_run(forma_fx* fx, const uint8_t* src_rgba, uint8_t* dst_rgba, float time) {
    NEED_FX(fx);
    if (!src_rgba || !dst_rgba) return set_last_error(FORMA_RT_EINVAL, "forma_fx_run: null frame");
    if (fx->plan.empty()) { fx->stats.gpu_launches = 0; fx->stats.bytes_moved = 0; return FORMA_RT_OK; }
    cudaStream_t s = fx->stream;
    FX_TRY(cudaEventRecord(fx->ev_begin, s));
    FX_TRY(cudaMemcpyAsync(fx->d_in, src_rgba, fx->bytes(), cudaMemcpyHostToDevice, s));
    if (int rc = run_chain(fx, fx->d_in, fx->d_out, time, s)) return rc;
    FX_TRY(cudaMemcpyAsync(dst_rgba, fx->d_out, fx->bytes(), cudaMemcpyDeviceToHost, s));
    FX_TRY(cudaEventRecord(fx->ev_end, s));
    FX_TRY(cudaStreamSynchronize(s));
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, fx->ev_begin, fx->ev_end) == cudaSuccess) fx->stats.total_ms = ms;
    return FORMA_RT_OK;
}

int forma_fx_run_device(forma_fx* fx, const void* d_src_rgba, void* d_dst_rgba, float time, void* stream) {
    NEED_FX(fx);
    if (!d_src_rgba || !d_dst_rgba || d_src_rgba == d_dst_rgba)
        return set_last_error(FORMA_RT_EINVAL, "forma_fx_run_device: frames must be two distinct device buffers");
    if (fx->plan.empty()) { fx->stats.gpu_launches = 0; fx->stats.bytes_moved = 0; return FORMA_RT_OK; }
    cudaStream_t s = stream ? (cudaStream_t)stream : fx->stream;
    return run_chain(fx, (const uint8_t*)d_src_rgba, (uint8_t*)d_dst_rgba, time, s);
}

int forma_fx_apply_device(forma_fx* fx, int effect, const void* d_src_rgba, void* d_dst_rgba, float time, void* stream) {
    NEED_FX(fx);
    if (effect < 0 || effect >= FORMA_FX_EFFECT_COUNT) return set_last_error(FORMA_RT_EINVAL, "forma_fx_apply_device: unknown effect");
    if (!d_src_rgba || !d_dst_rgba || d_src_rgba == d_dst_rgba)
        return set_last_error(FORMA_RT_EINVAL, "forma_fx_apply_device: frames must be two distinct device buffers");
    cudaStream_t s = stream ? (cudaStream_t)stream : fx->stream;
    const uint8_t e = (uint8_t)effect;
    return launch_pass(fx, &e, 1, (const uint8_t*)d_src_rgba, (uint8_t*)d_dst_rgba, time, s);
}

const char* forma_fx_effect_name(int effect) {
    static const char* const names[FORMA_FX_EFFECT_COUNT] = {
        "none", "noise", "pixelate", "invert_color", "wave", "interlace", "flipaxis", "grayscale", "step",
        "watercolor", "chromostereopsis", "anaglyph"};
    return effect >= 0 && effect < FORMA_FX_EFFECT_COUNT ? names[effect] : nullptr;
}

int forma_fx_synchronize(forma_fx* fx) {
    NEED_FX(fx);
    FX_TRY(cudaStreamSynchronize(fx->stream));
    return FORMA_RT_OK;
}

int forma_fx_stats_get(forma_fx* fx, forma_fx_stats* out) {
    NEED_FX(fx);
    if (!out) return set_last_error(FORMA_RT_EINVAL, "forma_fx_stats_get: out is null");
    if (fx->stats_pending) {
        FX_TRY(cudaEventSynchronize(fx->ev_k1));
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, fx->ev_k0, fx->ev_k1) == cudaSuccess) fx->stats.kernel_ms = ms;
        fx->stats_pending = false;
    }
    *out = fx->stats;
    return FORMA_RT_OK;
}

}  // extern "C"
