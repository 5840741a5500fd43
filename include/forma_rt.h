/* forma_rt.h — C ABI of libforma_rt.so, the B200 (sm_100a) drop-in for the per-pixel path-tracing
 * loop of fo-rma's `src/cpu_ray_tracer` (reference paths below are relative to the fo-rma tree).
 *
 * The reference has no FFI layer of its own (no `extern`, `#[no_mangle]` or cdylib): its boundary
 * is the four public Rust functions of cpu_ray_tracer/tracer.rs plus Scene / Camera / Hitable and
 * the JSON loader.  Each entry point below names the reference interface it replaces; a thin Rust
 * crate keeps the Rust names and signatures and forwards to these symbols.
 *
 * Conventions: every call returns 0 on success and a negative FORMA_RT_E* code on failure;
 * forma_rt_last_error() returns a thread-local NUL-terminated message for the last failure on the
 * calling thread (the reference panics instead: tracer.rs:157,186, scene_loader.rs:4).  The caller
 * owns every host buffer; the library owns all device memory.  A handle is used from one host
 * thread at a time (the reference has a single caller, frontend/macroquad.rs:14-92).  There is no
 * CPU fallback: without a usable CUDA device forma_rt_create fails with FORMA_RT_ECUDA.
 */
#ifndef FORMA_RT_H
#define FORMA_RT_H

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define FORMA_RT_API __attribute__((visibility("default")))
#else
#define FORMA_RT_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define FORMA_RT_ABI_VERSION 3

/* error codes */
#define FORMA_RT_OK 0
#define FORMA_RT_EINVAL (-1)  /* bad argument / state (e.g. render before a camera exists) */
#define FORMA_RT_EJSON (-2)   /* scene JSON does not parse or lacks a required field */
#define FORMA_RT_ECUDA (-3)   /* CUDA runtime error, or no device */
#define FORMA_RT_EIO (-4)     /* file could not be written */
#define FORMA_RT_ENOMEM (-5)

/* forma_rt_options.flags */
#define FORMA_RT_NO_JITTER 1u /* xi_u = xi_v = 0.5 instead of tracer.rs:66-67's two draws (added mode) */
#define FORMA_RT_PINHOLE 2u   /* skip camera.rs:63-64's lens sample: offset = 0 (added mode) */
#define FORMA_RT_HOST_ONLY 4u /* no device: scene/camera bookkeeping only; every render call fails with FORMA_RT_ECUDA */
#define FORMA_RT_SRGB_OUT 8u  /* RGB8 outputs pass through color_utils::srgb_encode (color_utils.rs:49-66) on the GPU (added mode) */

/* forma_rt_options.accel */
#define FORMA_RT_ACCEL_AUTO 0u  /* brute-force scan for small scenes, LBVH once spheres + 2 x triangles >= 32 (the measured crossover) */
#define FORMA_RT_ACCEL_BRUTE 1u /* tracer.rs:195-200's linear scan, as-is */
#define FORMA_RT_ACCEL_LBVH 2u  /* on-GPU Morton-code LBVH over the spheres (reference has none: aabb.rs is a stub) */

/* forma_rt_object.kind — one record per entry of Scene.objects (cpu_ray_tracer/scene.rs:4-7) */
#define FORMA_RT_SPHERE 0   /* shapes/sphere.rs:8-14 */
#define FORMA_RT_PLANE 1    /* shapes/plane.rs:8-15  */
#define FORMA_RT_TRIANGLE 2 /* extension for the JSON mesh names; no reference counterpart */

typedef struct forma_rt forma_rt; /* opaque: TraceModel (tracer.rs:12-17) + its device mirror */

typedef struct forma_rt_options {
    uint32_t struct_size; /* = sizeof(forma_rt_options) */
    int32_t device;       /* CUDA device ordinal */
    uint64_t seed;        /* Philox key; the reference's thread_rng is OS-seeded (utility.rs:5) */
    uint32_t flags;       /* FORMA_RT_NO_JITTER | FORMA_RT_PINHOLE | FORMA_RT_HOST_ONLY | FORMA_RT_SRGB_OUT */
    uint32_t accel;       /* FORMA_RT_ACCEL_* */
} forma_rt_options;

/* 64 bytes; the flat form of a `Box<dyn Hitable>` (shapes/hitable.rs:4-14). */
typedef struct forma_rt_object {
    int32_t kind;     /* FORMA_RT_SPHERE | PLANE | TRIANGLE */
    int32_t material; /* sphere.rs:60-70: 0 lambertian, 1 metal, 2 dielectric, 3 "light", other lambertian */
    int32_t id;       /* what forma_rt_primary_hits reports for this object */
    float fuzz;
    float color[3];
    float a[3]; /* sphere center        | plane position    | triangle v0 */
    float b[3]; /* sphere (radius,0,0)  | plane orientation | triangle v1 */
    float c[3]; /* unused               | plane size        | triangle v2 */
} forma_rt_object;

/* cpu_ray_tracer/camera.rs:7-21, same field order; 26 floats */
typedef struct forma_rt_camera {
    float position[3];
    float lower_left_corner[3];
    float aspect;
    float horizontal[3];
    float vertical[3];
    float u[3], v[3], w[3];
    float lens_radius;
    float focus_dist;
    float radius;
    float rotation;
} forma_rt_camera;

typedef struct forma_rt_stats {
    uint64_t rays;          /* closest-hit queries (get_color calls, tracer.rs:189) of the last render call */
    uint64_t paths;         /* (pixel, sample) pairs of the last render call */
    float kernel_ms;        /* CUDA-event time of the trace kernel(s) of the last render call */
    float total_ms;         /* CUDA-event time of the whole call on the device stream, copies included */
    uint32_t gpu_launches;  /* kernels launched by the last render call */
    uint32_t n_objects, n_spheres, n_planes, n_triangles;
    uint32_t bvh_nodes;     /* 0 when the brute-force scan is in use */
    uint32_t accel_in_use;  /* FORMA_RT_ACCEL_BRUTE | FORMA_RT_ACCEL_LBVH */
    uint32_t sm_count;
    uint32_t sm_clock_khz;  /* cudaDevAttrClockRate */
    uint32_t reserved;
    uint64_t culled_paths;  /* paths of `rays` that ran NO intersection test: every sample of a pixel whose whole primary
                               beam misses the scene's bounding sphere goes camera ray -> sky (ABI 3) */
} forma_rt_stats;

/* forma_rt_walk_stats: what one frame's LBVH traversal touches (SURVEY 8(d), config D's L2 roofline) */
typedef struct forma_rt_walk_stats {
    uint64_t rays;          /* closest-hit queries of the instrumented frame */
    uint64_t node_visits;   /* internal nodes whose two child boxes were tested */
    uint64_t leaf_tests;    /* leaf primitives handed to the sphere / triangle test */
    uint32_t node_bytes;    /* bytes fetched per node visit (one 32-byte record) */
    uint32_t leaf_bytes;    /* bytes fetched per leaf test (a sphere: 16; a triangle: 48) */
    uint32_t bvh_nodes;     /* 0: the scene is scanned, there is no walk */
    float l2_read_gbs;      /* streaming-from-L2 read bandwidth measured on this device by the same call */
} forma_rt_walk_stats;

/* ---- life cycle ------------------------------------------------------------------------------ */
/* tracer::create_model (tracer.rs:19-28): width x height model with Scene::new's content
 * (scene.rs:10-15: Camera::new + scenes::get_simple_scene). */
FORMA_RT_API int forma_rt_create(const forma_rt_options* opts, uint32_t width, uint32_t height, forma_rt** out);
FORMA_RT_API void forma_rt_destroy(forma_rt* ctx);
FORMA_RT_API const char* forma_rt_last_error(void);
FORMA_RT_API uint32_t forma_rt_abi_version(void);

/* ---- scene: Scene.objects (scene.rs:4-7) ------------------------------------------------------ */
/* scenes.rs: 0 get_simple_scene (:6-39), 1 get_plane_scene (:41-113), 2 get_objects (:115-156) */
FORMA_RT_API int forma_rt_load_builtin(forma_rt* ctx, int which);
/* scene_loader::construct_scene_from_json (basics/scene_loader.rs:3-7), schema :9-66, plus the
 * JSON -> ray-traceable mapping (DESIGN.md "Scene mapping"; optional per-object "rt" block).
 * Also records the JSON camera for forma_rt_camera_json.  Unknown keys are ignored (serde default). */
FORMA_RT_API int forma_rt_load_json(forma_rt* ctx, const char* json, size_t len);
/* replace Scene.objects by an explicit list */
FORMA_RT_API int forma_rt_load_objects(forma_rt* ctx, const forma_rt_object* objs, size_t n);
/* read Scene.objects back: *n = count; copies min(*n, cap) records when objs != NULL */
FORMA_RT_API int forma_rt_get_objects(forma_rt* ctx, forma_rt_object* objs, size_t cap, size_t* n);
/* Hitable::translate / Hitable::rotate on objects[idx] (hitable.rs:12-13, plane.rs:62-68; the
 * frontend calls both on objects[0] every frame, macroquad.rs:66-67).  Planes ASSIGN position /
 * orientation; every other kind ignores the call, as in the reference.  NULL = leave unchanged. */
FORMA_RT_API int forma_rt_set_object(forma_rt* ctx, uint32_t idx, const float* position, const float* orientation);

/* ---- camera: Scene.camera (camera.rs) --------------------------------------------------------- */
FORMA_RT_API int forma_rt_camera_new(forma_rt* ctx);  /* Camera::new(width, height), camera.rs:24-60 */
/* Camera::new's arithmetic for an arbitrary look-at (added; used by the JSON camera) */
FORMA_RT_API int forma_rt_camera_look(forma_rt* ctx, const float from[3], const float at[3], const float up[3], float v_fov);
FORMA_RT_API int forma_rt_camera_json(forma_rt* ctx);                          /* camera of the last load_json (added) */
FORMA_RT_API int forma_rt_camera_orbit(forma_rt* ctx, const float delta[3]);     /* Camera::orbit, camera.rs:97-122 */
FORMA_RT_API int forma_rt_camera_translate(forma_rt* ctx, const float delta[3]); /* Camera::translate, camera.rs:74-95 */
FORMA_RT_API int forma_rt_camera_get(forma_rt* ctx, forma_rt_camera* out);
FORMA_RT_API int forma_rt_camera_set(forma_rt* ctx, const forma_rt_camera* cam);

/* ---- rendering -------------------------------------------------------------------------------- */
/* tracer::update (tracer.rs:30-55): keys = 00EQADWS, orbit the camera, then one 1-spp frame
 * (render, tracer.rs:57-81) into rgb[width*height*3] (RGB8, row-major, row 0 = top).
 * Frame k of a handle uses sample index k so successive frames are decorrelated. */
FORMA_RT_API int forma_rt_update(forma_rt* ctx, uint8_t keys, float delta_time, uint8_t* rgb);
/* Progressive refinement (added mode).  The handle keeps a per-pixel f32 running sum.
 * forma_rt_accumulate traces samples [n, n + spp_more) of every pixel ON TOP of the n accumulated so far and
 * writes the average of all n + spp_more: the sum continues tracer.rs:170-175's chain, so k calls are
 * bit-identical to one forma_rt_render of the total.  *spp_total (nullable) receives n + spp_more.
 * forma_rt_update_progressive is tracer::update (keys, orbit) with that buffer: one more sample per frame
 * while the camera and the scene stand still, a fresh start at sample 0 whenever either changed. */
FORMA_RT_API int forma_rt_accumulate(forma_rt* ctx, uint32_t spp_more, uint32_t max_depth, uint8_t* rgb,
                                    float* linear_rgba, uint32_t* spp_total);
FORMA_RT_API int forma_rt_accumulate_reset(forma_rt* ctx);
FORMA_RT_API int forma_rt_update_progressive(forma_rt* ctx, uint8_t keys, float delta_time, uint8_t* rgb,
                                            uint32_t* spp_total);
/* tracer::save_image's loops (tracer.rs:168-184) without the PNG write: spp samples per pixel,
 * sample indices sample_offset .. sample_offset+spp-1, paths up to max_depth scatters
 * (tracer.rs:10: 50).  rgb (nullable): width*height*3 sqrt-gamma bytes of sum/spp.
 * linear_rgba (nullable): width*height*4 floats, the per-pixel f32 SUM over the samples in sample
 * order (r,g,b) and the sample count in .a. */
FORMA_RT_API int forma_rt_render(forma_rt* ctx, uint32_t spp, uint32_t sample_offset, uint32_t max_depth, uint8_t* rgb,
                    float* linear_rgba);
/* tracer::save_image (tracer.rs:160-187): forma_rt_render + PNG to `path` ("out/basic.png" in the
 * reference).  Fails with FORMA_RT_EIO where the reference panics. */
FORMA_RT_API int forma_rt_save_image(forma_rt* ctx, uint32_t sample, const char* path);
/* tracer::render_mt (tracer.rs:83-134), the 4-thread variant `update` has commented out (tracer.rs:53): one 1-spp
 * frame in four bands of height / 4 rows - v = ((t_height - y) + xi) / height + t_id / 4, bands appended in descending
 * t_id - of a FRESH get_simple_scene() (tracer.rs:105: not the model's scene, whatever it holds) through the model's
 * camera, sqrt gamma and `as u8` per sample.  rgb receives what its Vec holds: width * 4 * (height / 4) * 3 bytes
 * (nothing when height < 4).  `frame` is the sample index of the counter RNG.
 * forma_rt_render_mt_average is tracer::save_image_mt's loops (tracer.rs:139-154): frames 0 .. sample-1 averaged
 * AFTER gamma in f32 (acc += p as f32 / sample as f32), `as u8`, into a full width * height * 3 frame whose rows past
 * 4 * (height / 4) stay 0; forma_rt_save_image_mt writes that frame as a PNG ("out/basic_mt.png" in the reference)
 * and fails with FORMA_RT_EIO where the reference panics. */
FORMA_RT_API int forma_rt_render_mt(forma_rt* ctx, uint32_t frame, uint8_t* rgb);
FORMA_RT_API int forma_rt_render_mt_average(forma_rt* ctx, uint32_t sample, uint8_t* rgb);
FORMA_RT_API int forma_rt_save_image_mt(forma_rt* ctx, uint32_t sample, const char* path);
/* primary-ray dump (added; config B): xi = 0.5, pinhole.  prim_id[width*height] = object id of the
 * closest hit or -1; t[width*height] = its distance (FLT_MAX on a miss). */
FORMA_RT_API int forma_rt_primary_hits(forma_rt* ctx, int32_t* prim_id, float* t);

/* Tile-interleaved multi-GPU split: after forma_rt_set_partition(rank, world) every render call of
 * this handle traces only the 32x4-pixel tiles (tx, ty) with (ty * pitch + tx) % world == rank - pitch = the smallest
 * integer >= tiles per row that is coprime with world, so the deal rotates from row to row - and leaves the
 * other pixels of its output buffers untouched, so `world` handles - one per GPU - can write one frame (each
 * rank stores straight into rank 0's buffer through a peer pointer).  Every pixel is traced by exactly one rank
 * with the same sample order: bit-exact against a single GPU.  (0, 1) restores the whole frame. */
FORMA_RT_API int forma_rt_set_partition(forma_rt* ctx, uint32_t rank, uint32_t world);

/* ---- device-resident entry points (one process per GPU; torch.distributed moves the buffers) --- */
/* Same as forma_rt_render but the outputs are DEVICE pointers on ctx's device and nothing is copied
 * to the host; runs on `stream` (a cudaStream_t, NULL = the handle's own stream) without
 * synchronising it.  d_sum_rgba: width*height float4 (nullable).  d_rgb: width*height*3 bytes (nullable). */
FORMA_RT_API int forma_rt_render_device(forma_rt* ctx, uint32_t spp, uint32_t sample_offset, uint32_t max_depth, void* d_rgb,
                           void* d_sum_rgba, void* stream);
/* forma_rt_render_device WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same results (bit-identical sums, bytes and ray count).  The scratch of the
 * job-queue kernels (per-path radiance, cross-chunk sum, job counter) lives in a file-scope static device array of
 * 8 MiB + 256 B that every GPU loading the library holds until exit, so
 *   - width * height must not exceed FORMA_RT_NOALLOC_MAX_PIXELS (else FORMA_RT_EINVAL; any spp is fine);
 *   - the scene must already be resident: one forma_rt_render / forma_rt_render_device after every scene change
 *     (the upload and the LBVH build allocate), else FORMA_RT_EINVAL;
 *   - all calls on a GPU share the array, which keeps the last call's contents (everything a launch reads it has
 *     cleared or written first): calls that OVERLAP on the GPU - two streams, two handles, two threads - race.
 * The call itself makes no cudaMalloc / cudaFree / allocating library call; the caller owns d_rgb / d_sum_rgba. */
#define FORMA_RT_NOALLOC_MAX_PIXELS 262144 /* 2^18, e.g. 512 x 512 or 640 x 400 */
FORMA_RT_API int forma_rt_render_device_noalloc(forma_rt* ctx, uint32_t spp, uint32_t sample_offset, uint32_t max_depth,
                                               void* d_rgb, void* d_sum_rgba, void* stream);
/* tracer.rs:177-184 on a summed buffer: rgb = as_u8(sqrt(sum / spp_total) * 255), pixels
 * [pixel_begin, pixel_end).  Used after the cross-GPU reduce of the per-rank sums. */
FORMA_RT_API int forma_rt_tonemap_device(forma_rt* ctx, const void* d_sum_rgba, uint32_t spp_total, uint64_t pixel_begin,
                            uint64_t pixel_end, void* d_rgb, void* stream);
/* Fused exchange step for the sample-split: for pixels [pixel_begin, pixel_end) sum the n_peers
 * float4 buffers (device pointers that may live on peer GPUs, mapped through CUDA IPC / P2P) in rank
 * order and tone-map in the same kernel — the reduce and the write-out never round-trip HBM. */
FORMA_RT_API int forma_rt_reduce_tonemap_device(forma_rt* ctx, const void* const* d_peer_sums, uint32_t n_peers,
                                   uint32_t spp_total, uint64_t pixel_begin, uint64_t pixel_end, void* d_rgb,
                                   void* d_sum_out, void* stream);
/* CUDA IPC plumbing for the above: export a handle (64 bytes) of a device allocation of this
 * process, open a peer's handle, close it. */
/* Stream-ordered flags in peer-visible device memory (added): how the sample-split exchange orders its kernels
 * across GPUs without host-side barriers.  A flag is a zero-initialised uint32 in a forma_rt_device_alloc'ed buffer that
 * the peers map through forma_rt_ipc_open.  set: once everything queued on `stream` before the call is complete and
 * visible system-wide, *d_flag = value.  wait: work queued on `stream` after the call starts only when every
 * *d_flags[i] has reached `value` (n <= 16; values only grow; comparison is wrap-safe).
 * A wait never spins for ever: after FORMA_RT_FLAG_TIMEOUT_MS (environment, default 30 000) it gives up, lets the
 * stream go on and records the flag; the next forma_rt_flag_wait_device / forma_rt_synchronize / forma_rt_flag_error of
 * the handle fails with FORMA_RT_ECUDA and names it (a peer rank died between frames, or two ranks share one GPU -
 * `world` must not exceed the number of GPUs).  forma_rt_flag_error reports 0 or 1 + the index of that flag. */
FORMA_RT_API int forma_rt_flag_set_device(forma_rt* ctx, void* d_flag, uint32_t value, void* stream);
FORMA_RT_API int forma_rt_flag_wait_device(forma_rt* ctx, const void* const* d_flags, uint32_t n, uint32_t value,
                                           void* stream);
FORMA_RT_API int forma_rt_flag_error(forma_rt* ctx, uint32_t* flag_index_plus_one);

FORMA_RT_API int forma_rt_ipc_export(forma_rt* ctx, void* d_ptr, uint8_t handle[64]);
FORMA_RT_API int forma_rt_ipc_open(forma_rt* ctx, const uint8_t handle[64], void** d_ptr);
FORMA_RT_API int forma_rt_ipc_close(forma_rt* ctx, void* d_ptr);
/* scratch device allocations owned by the handle (freed by forma_rt_destroy) */
FORMA_RT_API int forma_rt_device_alloc(forma_rt* ctx, size_t bytes, void** d_ptr); /* zero-filled */
FORMA_RT_API int forma_rt_device_free(forma_rt* ctx, void* d_ptr);
FORMA_RT_API int forma_rt_synchronize(forma_rt* ctx);
/* page-locked host memory owned by the handle (freed by forma_rt_destroy): a `pixels` buffer (TraceModel.pixels,
 * tracer.rs:16) allocated here makes forma_rt_update / forma_rt_render copy the frame at the pinned D2H rate - the
 * rate bench.py's e2e figure is quoted at; any other host buffer works too, through a staged copy. */
FORMA_RT_API int forma_rt_host_alloc(forma_rt* ctx, size_t bytes, void** h_ptr);
FORMA_RT_API int forma_rt_host_free(forma_rt* ctx, void* h_ptr);

/* ---- misc -------------------------------------------------------------------------------------- */
/* Instrumented (untimed) frame: the same traversal as forma_rt_render(spp, 0, max_depth) with per-lane counters,
 * through a second instantiation of the tree kernel; no frame is returned.  Also runs a streaming read of a 32 MiB
 * buffer out of L2 (one warm-up pass, then timed passes, CUDA events) and reports its bandwidth.  Added; the
 * reference has no acceleration structure. */
FORMA_RT_API int forma_rt_walk_stats_get(forma_rt* ctx, uint32_t spp, uint32_t max_depth, forma_rt_walk_stats* out);

FORMA_RT_API int forma_rt_stats_get(forma_rt* ctx, forma_rt_stats* out);
/* color_utils::srgb_encode (color_utils.rs:49-66) in place on nbytes of device RGB8 (4-byte aligned);
 * the function maps u8 -> u8, forma_rt_srgb_table returns that map (host, no device needed). */
FORMA_RT_API int forma_rt_srgb_encode_device(forma_rt* ctx, void* d_rgb, size_t nbytes, void* stream);
FORMA_RT_API int forma_rt_srgb_table(uint8_t table[256]);
/* RGB8 -> PNG file (the `image` crate's role at tracer.rs:186) */
FORMA_RT_API int forma_rt_write_png(const char* path, const uint8_t* rgb, uint32_t width, uint32_t height);

#ifdef __cplusplus
}
#endif
#endif /* FORMA_RT_H */
