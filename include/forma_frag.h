/* forma_frag.h — C ABI of the reference's two fragment-shader ray tracers as CUDA passes.
 *
 * src/shaders/rt.wgsl and src/shaders/rt2.wgsl each trace a fixed scene (three resp. four spheres over a ground
 * plane, mirror bounces, one shadow ray per hit) in `fs_main`, one invocation per pixel of a full-screen quad;
 * frontend/wnt.rs:145-148 builds the render pipeline around rt.wgsl, rt2.wgsl is a second shader of the same shape
 * that nothing loads.  The reference has no FFI for this either: the entry points below are what a binding of
 * wnt.rs's `State` (new / resize / render) would bind.  They live in libforma_rt.so and follow forma_rt.h's
 * conventions: 0 = ok, < 0 = error with the message in forma_rt_last_error(); caller owns every host buffer; one
 * handle per host thread at a time; no CPU path — a box without a CUDA device gets FORMA_RT_ECUDA.  There is no
 * window, surface or swapchain: a frame is a headless buffer.
 *
 * Arithmetic (the conventions WGSL leaves to the driver, fixed here and in oracle/wgsl_rt_oracle.cpp): every f32
 * operation IEEE, in the order the shader writes it, never contracted; pow(x, y) = exp2(y * log2(x)) with fixed f32
 * exp2 / log2 routines, except the literal square pow(e, 2.0) of rt2.wgsl:58 = e * e; constants evaluated in f32.
 * DESIGN.md 8d lists them all.
 *
 * Frames: row-major, row 0 on top; pixel (x, y) is the fragment at position (x + 0.5, y + 0.5).
 *   linear: 4 x f32 per pixel, fs_main's return value (alpha 1);
 *   rgba8:  4 bytes per pixel, the float -> unorm conversion of a render target (NaN -> 0, clamp, x*255 + 0.5
 *           truncated), through the linear -> sRGB curve first when FORMA_FRAG_SRGB_SURFACE is set (wnt.rs:124-132
 *           picks an sRGB surface format when there is one), alpha 255.
 */
#ifndef FORMA_FRAG_H
#define FORMA_FRAG_H

#include <stddef.h>
#include <stdint.h>

#include "forma_rt.h"

#ifdef __cplusplus
extern "C" {
#endif

#define FORMA_FRAG_API FORMA_RT_API

enum forma_frag_shader {
    FORMA_FRAG_RT = 1, /* src/shaders/rt.wgsl  */
    FORMA_FRAG_RT2 = 2 /* src/shaders/rt2.wgsl */
};

/* forma_frag_params.flags */
#define FORMA_FRAG_SRGB_SURFACE 1u /* rgba8 is an ...UnormSrgb target: linear -> sRGB before the unorm conversion */

typedef struct forma_frag_params {
    float time;       /* rt2.wgsl:8-13 `uniforms.time` (seconds); rt.wgsl has no uniform */
    float res_x;      /* `resolution` of fs_main (rt.wgsl:163, rt2.wgsl:177-179); <= 0 = the shaders' 1600 x 1200 */
    float res_y;
    float aspect;     /* <= 0 = the shaders' own: 4/3 (rt.wgsl:164) resp. res_x / res_y (rt2.wgsl:180) */
    uint32_t flags;
} forma_frag_params;

typedef struct forma_frag forma_frag;

typedef struct forma_frag_stats {
    uint64_t rays;         /* intersect_scene (rt.wgsl:88) / trace (rt2.wgsl:87) calls of the last frame */
    uint32_t gpu_launches; /* kernel launches of the last frame */
    float kernel_ms;       /* device time of the last frame's kernel (CUDA events on the launching stream) */
    float total_ms;        /* last forma_frag_render: kernel + D2H of the requested frames */
} forma_frag_stats;

/* wnt.rs:81-237 `State::new`: a width x height target on `device`. */
FORMA_FRAG_API int forma_frag_create(int device, uint32_t width, uint32_t height, forma_frag** out);
FORMA_FRAG_API void forma_frag_destroy(forma_frag* fr);

/* wnt.rs:243-250 `State::resize`. */
FORMA_FRAG_API int forma_frag_resize(forma_frag* fr, uint32_t width, uint32_t height);

/* wnt.rs:259-302 `State::render`: one full-screen pass of `shader`'s fs_main; params may be NULL (time 0, the
 * shaders' constants, unorm target).  rgba8 (width*height*4 bytes) and linear_rgba (width*height*4 floats) are host
 * buffers, each nullable. */
FORMA_FRAG_API int forma_frag_render(forma_frag* fr, int shader, const forma_frag_params* params, uint8_t* rgba8,
                                     float* linear_rgba);

/* The same into device buffers (each nullable, 16-byte aligned), asynchronous on `stream` (NULL = the handle's). */
FORMA_FRAG_API int forma_frag_render_device(forma_frag* fr, int shader, const forma_frag_params* params,
                                            void* d_rgba8, void* d_linear_rgba, void* stream);

/* rt2.wgsl:144 — the light direction fs_main derives from `time` (computed on the host, once per frame). */
FORMA_FRAG_API int forma_frag_light_direction(float time, float out[3]);

FORMA_FRAG_API int forma_frag_synchronize(forma_frag* fr);
FORMA_FRAG_API int forma_frag_stats_get(forma_frag* fr, forma_frag_stats* out);

#ifdef __cplusplus
}
#endif
#endif /* FORMA_FRAG_H */
