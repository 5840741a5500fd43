/* forma_fx.h — C ABI of the post-processing filters that follow a traced frame.
 *
 * The reference runs twelve WGSL compute filters (src/shaders/compute/<name>.comp.wgsl, 8x8 workgroups) over an
 * Rgba8Unorm texture, chained through two intermediate textures by rendering/post_processor.rs.  The reference has no
 * FFI for this either; the entry points below are what a `forma_fx` binding of `PostProcessor` would bind, one per
 * method.  They live in the same shared library as forma_rt.h (libforma_rt.so) and follow its conventions: 0 = ok,
 * < 0 = error with the message in forma_rt_last_error(); caller owns every host buffer; one handle per host thread at
 * a time; there is no CPU path — a box without a CUDA device gets FORMA_RT_ECUDA.
 *
 * Frames are RGBA8, row-major, row 0 on top, 4 bytes per texel (the layout of wgpu's Rgba8Unorm texture).
 * Arithmetic: textureLoad = c / 255 in f32, textureStore = NaN -> 0, clamp, x * 255 + 0.5 truncated; every f32
 * operation IEEE in the order the shader writes it (the library is built with -fmad=false); loads outside the frame
 * read (0,0,0,0), stores outside it are dropped.
 */
#ifndef FORMA_FX_H
#define FORMA_FX_H

#include <stddef.h>
#include <stdint.h>

#include "forma_rt.h"

#ifdef __cplusplus
extern "C" {
#endif

#define FORMA_FX_API FORMA_RT_API

/* shader_utils.rs:58-71 — enum Effect, same order */
enum forma_fx_effect {
    FORMA_FX_NONE = 0,
    FORMA_FX_NOISE = 1,
    FORMA_FX_PIXELATE = 2,
    FORMA_FX_INVERT_COLOR = 3,
    FORMA_FX_WAVE = 4,
    FORMA_FX_INTERLACE = 5,
    FORMA_FX_FLIP_AXIS = 6,
    FORMA_FX_GRAYSCALE = 7,
    FORMA_FX_STEP = 8,
    FORMA_FX_WATERCOLOR = 9,
    FORMA_FX_CHROMOSTEREOPSIS = 10,
    FORMA_FX_ANAGLYPH = 11,
    FORMA_FX_EFFECT_COUNT = 12
};

/* forma_fx_create flags */
#define FORMA_FX_UNFUSED 1u /* one launch per effect on the reference's ping-pong textures (the default fuses runs of
                               per-texel effects into one pass; both give the same bytes) */

#define FORMA_FX_MAX_EFFECTS 64u

typedef struct forma_fx forma_fx;

typedef struct forma_fx_stats {
    uint32_t n_effects;    /* active effects */
    uint32_t n_passes;     /* kernel launches per run (fused passes; = n_effects when unfused) */
    uint32_t gpu_launches; /* launches of the last run */
    float kernel_ms;       /* device time of the last run's passes (CUDA events on the handle's stream) */
    float total_ms;        /* last forma_fx_run: H2D + passes + D2H */
    uint64_t bytes_moved;  /* algorithmic bytes of the last run: 8 B per texel and pass */
} forma_fx_stats;

/* PostProcessor::new (post_processor.rs:42-99): intermediate frames for width x height, no active effect. */
FORMA_FX_API int forma_fx_create(int device, uint32_t width, uint32_t height, uint32_t flags, forma_fx** out);
FORMA_FX_API void forma_fx_destroy(forma_fx* fx);

/* PostProcessor::resize (post_processor.rs:199-262): new frame size, same active effects. */
FORMA_FX_API int forma_fx_resize(forma_fx* fx, uint32_t width, uint32_t height);

/* PostProcessor::update_effects (post_processor.rs:133-197): the active effects, in the order they run
 * (forma_fx_effect values, n <= FORMA_FX_MAX_EFFECTS; n == 0 = no pass, the output is left untouched). */
FORMA_FX_API int forma_fx_update_effects(forma_fx* fx, const uint8_t* effects, uint32_t n);

/* PostProcessor::run (post_processor.rs:101-129) on host frames: src_rgba -> device, the chain, -> dst_rgba.
 * `time` is control_uniform.values[0] (seconds since PostProcessor::new in the reference; only `noise` reads it). */
FORMA_FX_API int forma_fx_run(forma_fx* fx, const uint8_t* src_rgba, uint8_t* dst_rgba, float time);

/* The same on device frames (width*height*4 bytes each, src != dst), asynchronous on `stream` (NULL = the handle's). */
FORMA_FX_API int forma_fx_run_device(forma_fx* fx, const void* d_src_rgba, void* d_dst_rgba, float time, void* stream);

/* One filter, one dispatch (post_processor.rs:111-125), device frames.  Texels the filter does not write (flip_axis
 * on a non-square frame) keep what d_dst_rgba held. */
FORMA_FX_API int forma_fx_apply_device(forma_fx* fx, int effect, const void* d_src_rgba, void* d_dst_rgba, float time,
                                       void* stream);

/* Trace a frame with `rt` (forma_rt_render's arguments) and run the chain over it without leaving the device:
 * RGB8 frame -> RGBA8 (alpha 255) -> effects -> dst_rgba (host, width*height*4).  rt and fx must have the same size
 * and device.  rgb (host, width*height*3, nullable) receives the unfiltered frame. */
FORMA_FX_API int forma_fx_run_traced(forma_fx* fx, forma_rt* rt, uint32_t spp, uint32_t sample_offset,
                                     uint32_t max_depth, float time, uint8_t* rgb, uint8_t* dst_rgba);

/* shader_utils.rs:73-88 — effect_to_name; NULL for an unknown effect. */
FORMA_FX_API const char* forma_fx_effect_name(int effect);

FORMA_FX_API int forma_fx_synchronize(forma_fx* fx);
FORMA_FX_API int forma_fx_stats_get(forma_fx* fx, forma_fx_stats* out);

#ifdef __cplusplus
}
#endif
#endif /* FORMA_FX_H */
