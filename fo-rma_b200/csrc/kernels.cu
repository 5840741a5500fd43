// kernels.cu — sm_100a kernels of the tracer path and their launchers.
//   render_kernel        tracer.rs:160-187 (save_image's pixel x sample loops) / :57-81 (render)
//   primary_hits_kernel  config-B dump: closest hit of the un-jittered pinhole primary ray
//   tonemap_kernel       tracer.rs:177-184 on a summed f32 buffer
//   reduce_tonemap_kernel  the cross-GPU exchange step fused with the write-out (peer loads)
// Built with -fmad=false (see trace.cuh).
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "kernels.hpp"
#include "trace.cuh"

namespace forma {

// Pixel mapping: a 128-thread block covers a 32x4 pixel tile; warp w covers the 8x4 sub-tile at
// x offset 8w (compact primary-ray footprint for the LBVH walk), lane -> (lane & 7, lane >> 3).
constexpr int kTileW = 32, kTileH = 4, kBlock = 128;

template <int MODE>
__device__ __forceinline__ void stage_scene(const SceneView& S, const float4*& sph, const float4*& tri) {
    if constexpr (MODE == kScanSmem) {
        extern __shared__ float4 smem_f4[];
        const int ns = S.n_spheres, nt3 = 3 * S.n_tris;
        for (int i = threadIdx.x; i < ns; i += blockDim.x) smem_f4[i] = S.spheres[i];
        for (int i = threadIdx.x; i < nt3; i += blockDim.x) smem_f4[ns + i] = S.tris[i];
        __syncthreads();
        sph = smem_f4;
        tri = smem_f4 + ns;
    } else {
        sph = S.spheres;
        tri = S.tris;
    }
}

// tracer.rs:177-184 for one pixel per thread: f32 sum out, and / or average in linear space, sqrt
// gamma, (c * 255.0) as u8 — the RGB8 tile staged in shared memory and stored as coalesced words.
__device__ __forceinline__ void write_out(const RenderParams& P, float3 col, uint32_t x, uint32_t y, int lx, int ly,
                                          bool live, uint8_t (*tile_rgb)[kTileW * 3], uint32_t bx, uint32_t by) {
    const float n_spp = (float)(P.spp + P.spp_prior);
    if (P.sum && live) P.sum[(size_t)x + (size_t)y * P.width] = make_float4(col.x, col.y, col.z, n_spp);

    if (P.rgb) {
        // tracer.rs:177-184: average in linear space, sqrt gamma, (c * 255.0) as u8
        float3 c = col / n_spp;
        const uint32_t r = as_u8(sqrtf(c.x) * 255.0f), g = as_u8(sqrtf(c.y) * 255.0f), b = as_u8(sqrtf(c.z) * 255.0f);
        // word stores need a 4-byte aligned frame (a caller's device pointer may be a slice) and rows that are whole words
        const bool full_tile = (bx + 1) * kTileW <= P.width && (by + 1) * kTileH <= P.height &&
                               (P.width * 3u) % 4u == 0 && (reinterpret_cast<uintptr_t>(P.rgb) & 3u) == 0u;
        if (full_tile) {  // stage 4 rows x 96 bytes, write them as 96 coalesced 32-bit words
            tile_rgb[ly][lx * 3 + 0] = (uint8_t)r;
            tile_rgb[ly][lx * 3 + 1] = (uint8_t)g;
            tile_rgb[ly][lx * 3 + 2] = (uint8_t)b;
            __syncthreads();
            if (threadIdx.x < kTileH * 24) {
                const int row = threadIdx.x / 24, word = threadIdx.x % 24;
                const size_t byte = ((size_t)(by * kTileH + row) * P.width + (size_t)bx * kTileW) * 3;
                reinterpret_cast<uint32_t*>(P.rgb + byte)[word] = reinterpret_cast<const uint32_t*>(tile_rgb[row])[word];
            }
        } else if (live) {
            uint8_t* px = P.rgb + ((size_t)x + (size_t)y * P.width) * 3;
            px[0] = (uint8_t)r; px[1] = (uint8_t)g; px[2] = (uint8_t)b;
        }
    }
}

#ifndef FORMA_SCAN_MIN_BLOCKS
#define FORMA_SCAN_MIN_BLOCKS 7  // 72 registers, no spills, 28 resident warps / SM (profiles/r04i_occ.txt: C 119.8 -> 121.5 Grays/s)
#endif
template <int MODE>
__global__ void __launch_bounds__(kBlock, FORMA_SCAN_MIN_BLOCKS) render_kernel(const __grid_constant__ RenderParams P) {
    // The grid is 1-D over the tiles THIS launch owns: CTA j traces tile t = j * part_mod + part_rem of a part_pitch-wide
    // tile grid (forma_rt_set_partition; one GPU: t = j).  Launching every tile and letting the other GPUs' CTAs return
    // starved the SMs of working CTAs: config C split 8 ways took 1.63 ms per rank instead of 0.5 (profiles/r05v).
    const uint32_t t = blockIdx.x * P.part_mod + P.part_rem;
    const uint32_t bx = t % P.part_pitch, by = t / P.part_pitch;
    if (bx * kTileW >= P.width) return;  // a padding column of the pitch (uniform for the CTA)
    const float4 *sph, *tri;
    stage_scene<MODE>(P.scene, sph, tri);

    __shared__ __align__(16) uint8_t tile_rgb[kTileH][kTileW * 3];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lx = warp * 8 + (lane & 7), ly = lane >> 3;
    const uint32_t x = bx * kTileW + lx, y = by * kTileH + ly;
    const bool live = x < P.width && y < P.height;

    uint32_t rays = 0, culled = 0;
    float3 col = f3(0.0f, 0.0f, 0.0f);
    if (live) {
        if (P.prior) col = xyz(P.prior[(size_t)x + (size_t)y * P.width]);  // progressive: continue tracer.rs:170-175's sum
        PathRng rng{P.rk, x + y * P.width, 0u};
        const float fx = (float)x, fy = (float)(P.height - y);  // tracer.rs:171-172 (Q1: height - y)
        const float fw = (float)P.width, fh = (float)P.height;
        const bool no_jitter = P.flags & FORMA_RT_NO_JITTER, pinhole = P.flags & FORMA_RT_PINHOLE;
        const bool fast_div = P.fast_div != 0u;
        // tracer.rs:168-175's sample loop and get_color's recursion (tracer.rs:189-219) as ONE flat
        // loop of closest-hit queries: a lane whose path ended starts the next sample of its own pixel
        // in the same iteration the other lanes spend on their next bounce, so a warp runs for the
        // longest per-pixel SUM of path lengths, not for spp x the longest path.  Per pixel the samples
        // are still traced and added in order, so the f32 sums match the reference's loop exactly.
        int stack[kMaxDepthCap];  // list indices of the hit objects: attenuations multiply innermost first
        uint32_t s = 0, depth = 0;
        bool alive = false;
        float3 o = f3(0.0f, 0.0f, 0.0f), d = o;
        if (beam_culled(P.cam, P.scene, x, y, P.width, P.height)) {
            // No primary ray of this pixel can hit anything: every sample is camera ray -> sky.  One Philox
            // block per iteration and lane: block 0 of the lane's next sample, or - when its lens try was
            // rejected (utility.rs:4-13) - the next retry block of the SAME sample, so the retries of some
            // lanes ride along with the fresh samples of the others instead of costing the warp a pass of
            // their own.  Each lane still adds its samples in order.
            uint32_t blk = 0;
            float xu = 0.0f, xv = 0.0f;
            while (s < P.spp) {
                rng.sample = P.sample_offset + s;
                const float4 k = rng.block_k(0, blk);  // xi = k * 2^-24 exactly, so 2 xi - 1 = RN(k 2^-23 - 1) and x + xi = RN(k 2^-24 + x): one FFMA each
                // block 0 = (xi_u, xi_v, lens try 0); a retry block = two lens tries (x, y), (z, w) (utility.rs:4-13).
                // Both readings share the (z, w) try, so one branch-free evaluation serves fresh and retrying lanes.
                const bool retry = blk != 0u;
                const float s23 = 1.0f / 8388608.0f, s24 = 1.0f / 16777216.0f;
                const float ax = __fmaf_rn(k.x, s23, -1.0f), ay = __fmaf_rn(k.y, s23, -1.0f);
                const float bx = __fmaf_rn(k.z, s23, -1.0f), by = __fmaf_rn(k.w, s23, -1.0f);
                const bool use_a = retry && !(ax * ax + ay * ay >= 1.0f);
                const bool ok = use_a || pinhole || !(bx * bx + by * by >= 1.0f);
                const float px = use_a ? ax : bx, py = use_a ? ay : by;
                if (!retry) {  // tracer.rs:171-172: x + xi_u, (height - y) + xi_v
                    xu = no_jitter ? fx + 0.5f : __fmaf_rn(k.x, s24, fx);
                    xv = no_jitter ? fy + 0.5f : __fmaf_rn(k.y, s24, fy);
                }
                if (ok) {
                    const float u = div_const(xu, fw, P.inv_w, fast_div);  // tracer.rs:171
                    const float v = div_const(xv, fh, P.inv_h, fast_div);  // tracer.rs:172
                    camera_ray_at(P.cam, px, py, u, v, pinhole, o, d);
                    col = col + sky(d);  // tracer.rs:174 with get_color = tracer.rs:213-217
                    ++s; blk = 0;
                } else {
                    ++blk;
                }
            }
            rays = P.spp;
            culled = P.spp;  // reported beside the ray count: these "queries" ran no intersection test
        }
        for (;;) {
            if (!alive) {
                if (s >= P.spp) break;
                rng.sample = P.sample_offset + s;
                Rand4 r0 = rng.block(0, 0);
                if (no_jitter) { r0.x = 0.5f; r0.y = 0.5f; }
                const float u = div_const(fx + r0.x, fw, P.inv_w, fast_div);
                const float v = div_const(fy + r0.y, fh, P.inv_h, fast_div);
                camera_ray(P.cam, rng, r0, u, v, pinhole, o, d);
                depth = 0;
                alive = true;
            }
            HitState h;
            ++rays;
            closest_hit<MODE>(P.scene, sph, tri, o, d, h);
            float3 pc = f3(0.0f, 0.0f, 0.0f);
            bool done = true;
            if (h.obj < 0) pc = sky(d);                      // tracer.rs:213-217
            else if (depth < P.max_depth) {                  // tracer.rs:205
                float3 no, nd;
                if (scatter(P.scene, rng, 1u + depth, d, h, no, nd)) {
                    stack[depth] = h.obj;
                    ++depth;
                    o = no; d = nd;
                    done = false;
                }
            }
            if (done) {
                if (black_path(P.scene, pc)) depth = 0;
                while (depth > 0) {  // tracer.rs:206-207 on the way back up
                    --depth;
                    pc = attenuation(P.scene, stack[depth]) * pc;
                }
                col = col + pc;      // tracer.rs:174
                ++s;
                alive = false;
            }
        }
    }

    // closest-hit query count: one atomic per warp
    uint32_t wr = __reduce_add_sync(0xffffffffu, rays);
    if (lane == 0 && wr) atomicAdd(P.rays, (unsigned long long)wr);
    wr = __reduce_add_sync(0xffffffffu, culled);
    if (lane == 0 && wr) atomicAdd(P.rays + 1, (unsigned long long)wr);

    write_out(P, col, x, y, lx, ly, live, tile_rgb, bx, by);
}

// A batch is what a warp holds when the queue runs dry, i.e. the drain at the end of a launch: at 8 samples (256
// paths, ~4 ms of a warp's time on config D) the tail cost 4-5 ms per launch - nothing on a 230 ms launch, 15 % of the
// 30 ms launch each GPU of an 8-way tile split runs (tools/partition_probe.py: 36.3 -> 32.2 ms at 2 samples).
#ifndef FORMA_BATCH_SAMPLES
#define FORMA_BATCH_SAMPLES 2
#endif
constexpr int kBatchSamples = FORMA_BATCH_SAMPLES;

// Scan-mode path tracer for frames that BOUNCE (the reference's own scenes, scenes.rs:6-156: a plane or a
// radius-1000 ground sphere behind every pixel).  render_kernel above gives a pixel to a thread for all its samples:
// on such frames a thread next to a silhouette traces 5-10x the rays of its neighbours, a 128-thread CTA runs as
// long as its slowest pixel and the launch ends with a long tail (profiles/r05a_*: SMs busy 18 % of the elapsed
// cycles at 600x600x50 spp, 55 % at 1080p x 64 spp).  Here the (pixel, sample) pairs are JOBS: persistent warps draw
// them in batches (kBatchSamples samples of one 8x4 tile, one atomic per batch) and every lane takes its next job the
// moment its path ends, so every lane of every warp traces one ray per iteration until the queue is empty.  A finished
// path stores its radiance in rad[sample][pixel]; accumulate_kernel adds them per pixel in sample order (the f32
// chain of tracer.rs:170-175: bit-identical sums).  600x600x50 spp: get_simple_scene 5.49 -> 1.17 ms, get_objects
// 5.45 -> 1.31 ms.  Measured and dropped (profiles/r05m_perf.txt): a per-lane state machine that runs the Philox
// blocks of camera rays, scatter tries and rejection retries in common passes (1-3 per round) ahead of one trace pass -
// bit-exact, but 80 registers and a thinner trace pass: 21.4 ms against this loop's 18.4 ms at 1080p x 64 spp.
#ifndef FORMA_QUEUE_MIN_BLOCKS
#define FORMA_QUEUE_MIN_BLOCKS 6
#endif
template <int MODE>
__global__ void __launch_bounds__(kBlock, FORMA_QUEUE_MIN_BLOCKS) render_queue_kernel(const __grid_constant__ RenderParams P) {
    const float4 *sph, *tri;
    stage_scene<MODE>(P.scene, sph, tri);
    const SceneView& S = P.scene;
    const int lane = threadIdx.x & 31;
    constexpr unsigned kFull = 0xffffffffu;
    const uint32_t ctas = P.part_pitch * ((P.height + kTileH - 1) / kTileH);
    const uint32_t own = ctas > P.part_rem ? (ctas - P.part_rem + P.part_mod - 1) / P.part_mod : 0u;
    const uint32_t groups = (P.chunk_k + kBatchSamples - 1) / kBatchSamples;
    const uint32_t n_batches = own * 4u * groups;
    const uint32_t npix = P.width * P.height;
    const float fw = (float)P.width, fh = (float)P.height;
    const bool no_jitter = P.flags & FORMA_RT_NO_JITTER, pinhole = P.flags & FORMA_RT_PINHOLE;
    uint32_t batch_next = 0, batch_jobs = 0, batch_x = 0, batch_y = 0, batch_k0 = 0;  // warp-uniform

    uint32_t rays = 0;
    int astack[kMaxDepthCap];  // list indices of the hit objects: attenuations multiply innermost first
    PathRng rng{P.rk, 0u, 0u};
    uint32_t depth = 0;
    bool alive = false, lane_done = false;
    float3 o = f3(0.0f, 0.0f, 0.0f), d = o;
    for (;;) {
        // ---- hand the warp's next jobs to the lanes whose path has ended (tracer.rs:171-173 for each)
        for (;;) {
            const bool want = !alive && !lane_done;
            const unsigned need = __ballot_sync(kFull, want);
            if (need == 0u) break;
            if (batch_next >= batch_jobs) {  // warp-uniform: draw the next batch
                unsigned bi = 0;
                if (lane == 0) bi = atomicAdd(P.job_counter, 1u);
                bi = __shfl_sync(kFull, bi, 0);
                if (bi >= n_batches) { if (want) lane_done = true; break; }
                const uint32_t sub = bi / groups, grp = bi % groups, t = (sub >> 2) * P.part_mod + P.part_rem;
                batch_x = (t % P.part_pitch) * kTileW + (sub & 3u) * 8u; batch_y = (t / P.part_pitch) * kTileH;
                batch_k0 = grp * kBatchSamples;
                batch_jobs = batch_x < P.width ? min((uint32_t)kBatchSamples, P.chunk_k - batch_k0) * 32u : 0u;  // padding column
                batch_next = 0;
                if (batch_jobs == 0u) continue;
            }
            const uint32_t job = batch_next + __popc(need & ((1u << lane) - 1u));
            batch_next = min(batch_next + (uint32_t)__popc(need), batch_jobs);
            if (want && job < batch_jobs) {
                const uint32_t jx = batch_x + (job & 7u), jy = batch_y + ((job >> 3) & 3u), k = batch_k0 + (job >> 5);
                if (jx < P.width && jy < P.height) {  // jobs of pixels outside the image (edge tiles) are dropped
                    rng.pixel = jx + jy * P.width;
                    rng.sample = P.sample_offset + P.chunk_base + k;
                    Rand4 r0 = rng.block(0, 0);
                    if (no_jitter) { r0.x = 0.5f; r0.y = 0.5f; }
                    const float u = div_const((float)jx + r0.x, fw, P.inv_w, P.fast_div != 0u);                // tracer.rs:171
                    const float v = div_const((float)(P.height - jy) + r0.y, fh, P.inv_h, P.fast_div != 0u);   // tracer.rs:172 (Q1: height - y)
                    camera_ray(P.cam, rng, r0, u, v, pinhole, o, d);
                    depth = 0;
                    alive = true;
                }
            }
        }
        if (__ballot_sync(kFull, alive) == 0u) break;  // the queue is empty and every path of the warp has ended
        // ---- one closest-hit query per lane: get_color (tracer.rs:189-219), unrolled
        if (alive) {
            HitState h;
            ++rays;
            closest_hit<MODE>(S, sph, tri, o, d, h);
            float3 pc = f3(0.0f, 0.0f, 0.0f);
            bool done = true;
            if (h.obj < 0) pc = sky(d);                      // tracer.rs:213-217
            else if (depth < P.max_depth) {                  // tracer.rs:205
                float3 no, nd;
                if (scatter(S, rng, 1u + depth, d, h, no, nd)) {
                    astack[depth] = h.obj;
                    ++depth;
                    o = no; d = nd;
                    done = false;
                }
            }
            if (done) {
                if (black_path(S, pc)) depth = 0;
                while (depth > 0) {  // tracer.rs:206-207 on the way back up
                    --depth;
                    pc = attenuation(S, astack[depth]) * pc;
                }
                const uint32_t k = rng.sample - P.sample_offset - P.chunk_base;
                P.rad[(size_t)k * npix + rng.pixel] = make_float4(pc.x, pc.y, pc.z, 1.0f);
                alive = false;
            }
        }
    }
    const uint32_t wr = __reduce_add_sync(kFull, rays);
    if (lane == 0 && wr) atomicAdd(P.rays, (unsigned long long)wr);
}

// Tree-mode path tracer (persistent warps).  Two things make a pixel-per-thread loop slow on scenes
// with real bounces: rays of one warp need very different numbers of LBVH nodes, and pixels of one
// tile need very different numbers of rays (sky next to a cluster of spheres).  So here
//  * the (pixel, sample) pairs of a chunk of samples are JOBS.  A warp draws them in batches (8
//    samples of one 8x4-pixel tile) from one global queue and hands them to its lanes as their
//    paths end, so no lane idles while its warp has work and no warp idles while work is left
//    anywhere on the GPU, yet the rays of a warp stay inside one or two tiles;
//  * a finished path writes its radiance to rad[sample in chunk][pixel]; accumulate_kernel then
//    adds every pixel's samples IN SAMPLE ORDER - the same f32 chain as tracer.rs:170-175, so the
//    sums stay bit-identical to the flat loop and to the CPU restatement.  Cost: 32 B of HBM traffic per
//    path, nothing next to the thousands of instructions a bouncing path takes;
//  * every lane is a small state machine so that a warp never waits on its slowest ray:
//      kShade    winner -> HitRecord, trailing planes, scatter / sky / termination (tracer.rs:201-217)
//      kNewPath  draw a job, generate its camera ray                      (tracer.rs:171-173)
//      kStart    planes that precede the tree's run, the linear "big" primitives, walk_begin
//      kWalk     resumable LBVH walk
//    The warp walks until `refill` lanes have finished their walk (or none is walking), then those
//    lanes shade, scatter or draw their next job and rejoin.
enum : int { kNewPath = 0, kStart = 1, kWalk = 2, kShade = 3, kLaneDone = 4 };
#ifndef FORMA_TREE_MIN_BLOCKS
#define FORMA_TREE_MIN_BLOCKS 8  // 64 registers: 32 resident warps / SM (profiles/r01p_occ.txt: A 134 -> 122 ms, D 142 -> 138 ms)
#endif
#ifndef FORMA_STEPS_PER_VOTE
#define FORMA_STEPS_PER_VOTE 2   // node steps between two looks at the warp's state (one ballot each): A 52.5 -> 51.0 ms, D 100.3 -> 97.8 ms
#endif
// What a lane keeps IN REGISTERS across the walk is deliberately small - the ray's grid (9), closest_so_far, the
// winner (one unified primitive index), node, sp, the stack base and the phase.  Everything the node loop does not
// read - the ray itself (o, d, a), the job (pixel, sample) and the depth - waits in a per-thread record in shared
// memory (`cold`, 40 B per thread, conflict-free) and is loaded where a leaf is tested or a lane is shaded; the
// HitRecord is rebuilt when the walk is over; the warp's batch lives in shared memory; the ray counter is
// warp-uniform; the grid is parked in local memory between two entries of the walk loop so that the shading code
// in between cannot make the allocator spill it.  Round 1's kernel carried ~20 more live values through the loop
// and paid 6 LDL of spilled grid constants + ~16 address instructions per node visit at the 64-register cap
// (profiles/r04i_occ.txt r04q, r05a_A): 13.5 -> 17.6 lanes per instruction, 344 -> 301 warp instructions per ray.
// COUNT = true is the instrumented instantiation (forma_rt_walk_stats): the same traversal, plus two per-lane
// counters - internal nodes visited and leaf primitives tested - summed into P.walk_counters.  Never timed.
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, FORMA_TREE_MIN_BLOCKS) render_tree_kernel(const __grid_constant__ RenderParams P) {
    const SceneView& S = P.scene;
    const float4 *sph = S.spheres, *tri = S.tris;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr unsigned kFull = 0xffffffffu;
    // this launch owns the 32x4 tiles t = j * part_mod + part_rem of a part_pitch-wide tile grid (the scan kernel's
    // CTA tiles; columns >= ctas_x are padding); each is four 8x4 warp tiles
    const uint32_t ctas = P.part_pitch * ((P.height + kTileH - 1) / kTileH);
    const uint32_t own = ctas > P.part_rem ? (ctas - P.part_rem + P.part_mod - 1) / P.part_mod : 0u;
    const uint32_t groups = (P.chunk_k + kBatchSamples - 1) / kBatchSamples;
    const uint32_t n_batches = own * 4u * groups;  // < 2^32: checked by the launcher
    const uint32_t npix = P.width * P.height;
    // the warp's current batch (warp-uniform, touched only when lanes are refilled): next job, jobs, x, y, first sample
    __shared__ uint32_t batch_smem[kBlock / 32][8];
    volatile uint32_t* batch = batch_smem[warp];
    if (lane < 8) batch[lane] = 0u;
    __syncwarp();
    // the lane's cold state: [0] = (o, a), [1] = (d, depth), job = (pixel, sample)
    __shared__ float4 cold_smem[2][kBlock];
    __shared__ uint2 job_smem[kBlock];
    auto cold_o = [&](float3& o, float& a) { const float4 v = cold_smem[0][threadIdx.x]; o = xyz(v); a = v.w; };
    auto cold_d = [&](float3& d, uint32_t& depth) { const float4 v = cold_smem[1][threadIdx.x]; d = xyz(v); depth = __float_as_uint(v.w); };

    uint32_t rays = 0;          // warp-uniform: every lane counts the warp's closest-hit queries
    unsigned long long n_nodes = 0, n_leaves = 0;  // COUNT only
    int astack[kMaxDepthCap];   // list indices of the hit objects: attenuations multiply innermost first
    __shared__ int2 stack_smem[kSmemStack * kWalkBlock];
    int2 stack_local[kStackCap];
    int phase = kNewPath;
    float closest = FLT_MAX;
    int winner = -1;            // unified primitive index of the nearest accepted hit of the tree's run
    float grid_park[9];
    volatile float* gpark = grid_park;
    auto park = [&](const RayGrid& g) {
        gpark[0] = g.ax; gpark[1] = g.ay; gpark[2] = g.az; gpark[3] = g.blx; gpark[4] = g.bly; gpark[5] = g.blz;
        gpark[6] = g.bhx; gpark[7] = g.bhy; gpark[8] = g.bhz;
    };
    int node = kWalkDone, sp = 0;

    // planes that precede the tree's run in Scene.objects (tracer.rs:195-200 meets them first)
    auto planes_before_tree = [&](float3 o, float3 d, HitState& h) {
        h.t = 0.0f; h.p = f3(0.0f, 0.0f, 0.0f); h.n = h.p;
        h.closest = FLT_MAX; h.obj = -1;
        for (int sg = 0; sg < S.tree_seg; ++sg) {
            const Segment sgm = S.segments[sg];
            if (sgm.plane >= 0) plane_hit(S, sgm.plane, o, d, h);
        }
    };

    for (;;) {
        // ---- A. shade the rays whose walk has finished
        if (phase == kShade) {
            float3 o, d; float a; uint32_t depth;
            cold_o(o, a); cold_d(d, depth);
            const uint2 job = job_smem[threadIdx.x];
            const PathRng rng{P.rk, job.x, job.y};
            HitState h;
            planes_before_tree(o, d, h);   // the HitRecord as the scan left it in front of the run (not kept across the walk)
            h.closest = closest;
            for (int sg = S.tree_seg; sg < S.n_segments; ++sg) {
                Segment sgm = S.seg0;
                if (sg > 0) sgm = S.segments[sg];
                if (sg == S.tree_seg) finalize_winner(sph, tri, o, d, winner_of(S, winner), h);
                if (sgm.plane >= 0) plane_hit(S, sgm.plane, o, d, h);
            }
            float3 pc = f3(0.0f, 0.0f, 0.0f);
            bool done = true;
            if (h.obj < 0) pc = sky(d);                      // tracer.rs:213-217
            else if (depth < P.max_depth) {                  // tracer.rs:205
                float3 no, nd;
                if (scatter(S, rng, 1u + depth, d, h, no, nd)) {
                    astack[depth] = h.obj;
                    ++depth;
                    cold_smem[0][threadIdx.x] = make_float4(no.x, no.y, no.z, dot(nd, nd));  // a: sphere.rs:25
                    cold_smem[1][threadIdx.x] = make_float4(nd.x, nd.y, nd.z, __uint_as_float(depth));
                    done = false;
                }
            }
            phase = kStart;
            if (done) {
                if (black_path(S, pc)) depth = 0;
                while (depth > 0) {  // tracer.rs:206-207 on the way back up
                    --depth;
                    pc = attenuation(S, astack[depth]) * pc;
                }
                // where this path's radiance goes in `rad`: [sample in chunk][pixel]
                const uint32_t k = rng.sample - P.sample_offset - P.chunk_base;
                P.rad[(size_t)k * npix + rng.pixel] = make_float4(pc.x, pc.y, pc.z, 1.0f);
                phase = kNewPath;
            }
        }
        // ---- B. hand the warp's next jobs to the lanes that need one.  The warp owns one BATCH at a time:
        // kBatchSamples consecutive samples of one 8x4 tile, job j = (sample j >> 5, pixel j & 31), so
        // a refilled lane stays inside the tile its neighbours are tracing; one atomic per batch.
        for (;;) {
            const bool want = phase == kNewPath;
            const unsigned need = __ballot_sync(kFull, want);
            if (need == 0u) break;
            uint32_t batch_next = batch[0], batch_jobs = batch[1];
            if (batch_next >= batch_jobs) {  // warp-uniform: draw the next batch
                unsigned bi = 0;
                if (lane == 0) bi = atomicAdd(P.job_counter, 1u);
                bi = __shfl_sync(kFull, bi, 0);
                if (bi >= n_batches) { if (want) phase = kLaneDone; break; }
                const uint32_t sub = bi / groups, grp = bi % groups, t = (sub >> 2) * P.part_mod + P.part_rem;
                const uint32_t bx = (t % P.part_pitch) * kTileW + (sub & 3u) * 8u, by = (t / P.part_pitch) * kTileH;
                const uint32_t k0 = grp * kBatchSamples;
                batch_jobs = bx < P.width ? min((uint32_t)kBatchSamples, P.chunk_k - k0) * 32u : 0u;  // padding column
                batch_next = 0;
                __syncwarp();
                if (lane == 0) { batch[0] = 0u; batch[1] = batch_jobs; batch[2] = bx; batch[3] = by; batch[4] = k0; }
                __syncwarp();
                if (batch_jobs == 0u) continue;
            }
            const uint32_t job = batch_next + __popc(need & ((1u << lane) - 1u));
            const uint32_t bx = batch[2], by = batch[3], k0 = batch[4];
            __syncwarp();
            if (lane == 0) batch[0] = min(batch_next + (uint32_t)__popc(need), batch_jobs);
            __syncwarp();
            if (want && job < batch_jobs) {
                // ---- C. decode the job, generate its camera ray
                const uint32_t jx = bx + (job & 7u), jy = by + ((job >> 3) & 3u), k = k0 + (job >> 5);
                if (jx < P.width && jy < P.height) {  // jobs of pixels outside the image (edge tiles) are dropped
                    const float fw = (float)P.width, fh = (float)P.height;
                    const PathRng rng{P.rk, jx + jy * P.width, P.sample_offset + P.chunk_base + k};
                    Rand4 r0 = rng.block(0, 0);
                    if (P.flags & FORMA_RT_NO_JITTER) { r0.x = 0.5f; r0.y = 0.5f; }
                    const float u = div_const((float)jx + r0.x, fw, P.inv_w, P.fast_div != 0u);                // tracer.rs:171
                    const float v = div_const((float)(P.height - jy) + r0.y, fh, P.inv_h, P.fast_div != 0u);   // tracer.rs:172 (Q1: height - y)
                    float3 o, d;
                    camera_ray(P.cam, rng, r0, u, v, (P.flags & FORMA_RT_PINHOLE) != 0u, o, d);
                    cold_smem[0][threadIdx.x] = make_float4(o.x, o.y, o.z, dot(d, d));  // a: sphere.rs:25
                    cold_smem[1][threadIdx.x] = make_float4(d.x, d.y, d.z, __uint_as_float(0u));
                    job_smem[threadIdx.x] = make_uint2(rng.pixel, rng.sample);
                    phase = kStart;
                }
            }
        }
        // ---- D. tracer.rs:190-194, then everything the scan meets before the tree
        rays += __popc(__ballot_sync(kFull, phase == kStart));
        if (phase == kStart) {
            float3 o, d; float a; uint32_t depth;
            cold_o(o, a); cold_d(d, depth);
            closest = FLT_MAX;
            if (S.tree_seg > 0) { HitState h; planes_before_tree(o, d, h); closest = h.closest; }
            winner = -1;
            phase = kShade;
            node = kWalkDone;
            if (!scene_culled(S, o, d, a)) {
                for (int i = 0; i < S.big_spheres; ++i) try_prim(S, sph, tri, i, o, d, a, closest, winner);
                for (int i = 0; i < S.big_tris; ++i) try_prim(S, sph, tri, S.n_spheres + i, o, d, a, closest, winner);
                park(ray_grid(S, o, d, a, closest));
                node = 0; sp = 0;
                phase = kWalk;
            }
        }
        if (__ballot_sync(kFull, phase != kLaneDone) == 0u) break;  // the queue is empty and every lane has finished
        // ---- E. walk until enough lanes are idle to be worth refilling
        {
            // Convergent walk: per step every walking lane takes ONE stack entry (if it needs one) and tests ONE
            // internal node, so lanes never wait for each other inside a node or a pop loop.  Lanes that reach a
            // leaf, finish or idle stop progressing; the warp looks at them only once `leaf_batch` more lanes have
            // stopped (one ballot per FORMA_STEPS_PER_VOTE steps until then), tests the waiting leaves together,
            // and leaves the loop when `refill` lanes are ready to shade.  Any visiting order gives the scan's
            // result (the tie rule of try_prim).  Lanes that are not walking hold node == kWalkDone, so the loop
            // looks at `node` only.
            int wake_at = 0;
            WalkStack wstack;
            wstack.init(stack_smem + threadIdx.x, stack_local);
            const RayGrid g{gpark[0], gpark[1], gpark[2], gpark[3], gpark[4], gpark[5], gpark[6], gpark[7], gpark[8]};
            for (;;) {
#pragma unroll
                for (int step = 0; step < FORMA_STEPS_PER_VOTE; ++step) {
                    walk_pop(node, sp, wstack, closest);
                    if (COUNT && node >= 0) ++n_nodes;
                    if (node >= 0) walk_inner(S, g, node, sp, wstack, closest);
                }
                const bool busy = node >= 0 || node == kWalkPop;
                const int stopped = 32 - __popc(__ballot_sync(kFull, busy));
                if (stopped < wake_at) continue;
                if (phase == kWalk && node == kWalkDone) phase = kShade;
                const bool at_leaf = phase == kWalk && !busy;
                const unsigned leaves = __ballot_sync(kFull, at_leaf);
                const int idle = __popc(__ballot_sync(kFull, phase != kLaneDone && phase != kWalk));
                if (idle >= (int)P.refill || (stopped == 32 && leaves == 0u)) break;
                if (at_leaf) {
                    if (COUNT) ++n_leaves;
                    float3 o, d; float a; uint32_t depth;
                    cold_o(o, a); cold_d(d, depth);
                    try_prim(S, sph, tri, ~node, o, d, a, closest, winner);
                    node = kWalkPop;
                }
                wake_at = min(32, stopped - __popc(leaves) + (int)P.leaf_batch);
            }
        }
    }

    // closest-hit query count (warp-uniform)
    if (lane == 0 && rays) atomicAdd(P.rays, (unsigned long long)rays);
    if (COUNT) {
        for (int m = 16; m >= 1; m >>= 1) {
            n_nodes += __shfl_xor_sync(kFull, n_nodes, m);
            n_leaves += __shfl_xor_sync(kFull, n_leaves, m);
        }
        if (lane == 0) { atomicAdd(P.walk_counters, n_nodes); atomicAdd(P.walk_counters + 1, n_leaves); }
    }
}

// Four RGB8 pixels = 12 bytes = three aligned 32-bit words: the write-out kernels below give a thread four
// neighbouring pixels and store words, not bytes (the frame may be a peer GPU's, written over NVLink).
__device__ __forceinline__ void store_rgb8x4(uint8_t* rgb, size_t first_pixel, const uint32_t r[4], const uint32_t g[4],
                                             const uint32_t b[4]) {
    uint32_t* w = reinterpret_cast<uint32_t*>(rgb + 3 * first_pixel);  // first_pixel % 4 == 0 and rgb 4-byte aligned
    w[0] = r[0] | g[0] << 8 | b[0] << 16 | r[1] << 24;
    w[1] = g[1] | b[1] << 8 | r[2] << 16 | g[2] << 24;
    w[2] = b[2] | r[3] << 8 | g[3] << 16 | b[3] << 24;
}
// tracer.rs:177-184: average in linear space, sqrt gamma, (c * 255.0) as u8
__device__ __forceinline__ void tone_rgb(float3 col, float n_spp, uint32_t& r, uint32_t& g, uint32_t& b) {
    const float3 c = col / n_spp;
    r = as_u8(sqrtf(c.x) * 255.0f); g = as_u8(sqrtf(c.y) * 255.0f); b = as_u8(sqrtf(c.z) * 255.0f);
}

// tracer.rs:170-184 for the job-queue modes: acc[p] += rad[0][p], rad[1][p], ... in sample order; after the last
// chunk the f32 sum and / or the sqrt-gamma RGB8 pixel are written out.  A thread owns four neighbouring pixels
// of a row (never more than one 32x4 tile, so the multi-GPU tile split keeps or skips them together).
__global__ void accumulate_kernel(const __grid_constant__ RenderParams P, int first, int last) {
    const size_t npix = (size_t)P.width * P.height;
    const uint32_t gpr = (P.width + 3u) / 4u;  // groups per row
    const size_t ngroups = (size_t)gpr * P.height;
    const bool words = P.width % 4u == 0u && (reinterpret_cast<uintptr_t>(P.rgb) & 3u) == 0u;
    for (size_t gi = blockIdx.x * (size_t)blockDim.x + threadIdx.x; gi < ngroups; gi += (size_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(gi / gpr), x0 = (uint32_t)(gi % gpr) * 4u;
        const uint32_t n = min(4u, P.width - x0);
        if (P.part_mod > 1u && ((y / kTileH) * P.part_pitch + x0 / kTileW) % P.part_mod != P.part_rem) continue;  // another GPU's tile
        const size_t p0 = (size_t)y * P.width + x0;
        float3 col[4];
#pragma unroll
        for (uint32_t j = 0; j < 4u; ++j) {
            col[j] = f3(0.0f, 0.0f, 0.0f);
            if (j < n) {
                if (!first) { const float4 c = P.acc[p0 + j]; col[j] = f3(c.x, c.y, c.z); }
                else if (P.prior) col[j] = xyz(P.prior[p0 + j]);
            }
        }
        for (uint32_t k = 0; k < P.chunk_k; ++k) {
            float4 r[4];
#pragma unroll
            for (uint32_t j = 0; j < 4u; ++j) r[j] = j < n ? P.rad[k * npix + p0 + j] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
            for (uint32_t j = 0; j < 4u; ++j) col[j] = col[j] + f3(r[j].x, r[j].y, r[j].z);  // tracer.rs:174
        }
        if (!last) {
#pragma unroll
            for (uint32_t j = 0; j < 4u; ++j) if (j < n) P.acc[p0 + j] = make_float4(col[j].x, col[j].y, col[j].z, 0.0f);
            continue;
        }
        const float n_spp = (float)(P.spp + P.spp_prior);
        if (P.sum) {
#pragma unroll
            for (uint32_t j = 0; j < 4u; ++j) if (j < n) P.sum[p0 + j] = make_float4(col[j].x, col[j].y, col[j].z, n_spp);
        }
        if (P.rgb) {
            uint32_t r[4], g[4], b[4];
#pragma unroll
            for (uint32_t j = 0; j < 4u; ++j) tone_rgb(col[j], n_spp, r[j], g[j], b[j]);
            if (words && n == 4u) store_rgb8x4(P.rgb, p0, r, g, b);
            else
                for (uint32_t j = 0; j < n; ++j) { uint8_t* px = P.rgb + 3 * (p0 + j); px[0] = (uint8_t)r[j]; px[1] = (uint8_t)g[j]; px[2] = (uint8_t)b[j]; }
        }
    }
}

// tracer::render_mt (tracer.rs:83-134): four bands of t_height = height / 4 rows, each traced by
// its own thread in the reference, v = ((t_height - y) + xi) / height + t_id / 4 (:99-103), one sample per pixel, sqrt
// gamma and `as u8` per sample (:107-112); the bands are appended in DESCENDING t_id (:126-131), so output row r is row
// r % t_height of band 3 - r / t_height.  Rows past 4 * t_height do not exist in its result.  One thread per pixel;
// the path is get_color (tracer.rs:189-219) unrolled into a loop with the attenuations multiplied innermost first.
__global__ void __launch_bounds__(kBlock) render_mt_kernel(const __grid_constant__ MtParams P) {
    const SceneView& S = P.scene;
    const float4 *sph = S.spheres, *tri = S.tris;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t x = blockIdx.x * kTileW + warp * 8 + (lane & 7), r = blockIdx.y * kTileH + (lane >> 3);
    const uint32_t t_height = P.height / 4u;
    uint32_t rays = 0;
    if (x < P.width && r < 4u * t_height) {
        const uint32_t t_id = 3u - r / t_height, yb = r % t_height;
        PathRng rng{P.rk, x + r * P.width, P.frame};
        const Rand4 r0 = rng.block(0, 0);
        const float u = ((float)x + r0.x) / (float)P.width;                    // tracer.rs:99
        float v = ((float)(t_height - yb) + r0.y) / (float)P.height;           // tracer.rs:100
        v += (float)t_id * (1.0f / 4.0f);                                      // tracer.rs:101
        float3 o, d;
        camera_ray(P.cam, rng, r0, u, v, false, o, d);                         // tracer.rs:102
        int astack[kMaxDepthCap];
        uint32_t depth = 0;
        float3 pc = f3(0.0f, 0.0f, 0.0f);
        for (;;) {                                                             // tracer.rs:103 get_color(ray, scene, 0)
            HitState h;
            closest_hit<kScanGlobal>(S, sph, tri, o, d, h);
            ++rays;
            if (h.obj < 0) { pc = sky(d); break; }
            float3 no, nd;
            if (depth < P.max_depth && scatter(S, rng, 1u + depth, d, h, no, nd)) {
                astack[depth++] = h.obj;
                o = no; d = nd;
                continue;
            }
            break;  // black
        }
        while (depth > 0) { --depth; pc = attenuation(S, astack[depth]) * pc; }
        uint8_t* px = P.rgb + ((size_t)x + (size_t)r * P.width) * 3;
        px[0] = (uint8_t)as_u8(sqrtf(pc.x) * 255.0f);                          // tracer.rs:107-112
        px[1] = (uint8_t)as_u8(sqrtf(pc.y) * 255.0f);
        px[2] = (uint8_t)as_u8(sqrtf(pc.z) * 255.0f);
    }
    const uint32_t wr = __reduce_add_sync(0xffffffffu, rays);
    if (lane == 0 && wr) atomicAdd(P.rays, (unsigned long long)wr);
}

// tracer::save_image_mt (tracer.rs:139-154): acc[i] += pixels[i] as f32 / sample as f32, then acc[i] as u8
__global__ void mt_accumulate_kernel(const uint8_t* __restrict__ frame, float* __restrict__ acc, float sample, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        acc[i] = acc[i] + (float)frame[i] / sample;
}
__global__ void mt_finish_kernel(const float* __restrict__ acc, uint8_t* __restrict__ out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = (uint8_t)as_u8(acc[i]);
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) primary_hits_kernel(const __grid_constant__ PrimaryParams P) {
    const float4 *sph, *tri;
    stage_scene<MODE>(P.scene, sph, tri);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t x = blockIdx.x * kTileW + warp * 8 + (lane & 7), y = blockIdx.y * kTileH + (lane >> 3);
    if (x >= P.width || y >= P.height) return;
    const float u = ((float)x + 0.5f) / (float)P.width;
    const float v = ((float)(P.height - y) + 0.5f) / (float)P.height;
    const RoundKeys no_keys{};
    PathRng rng{no_keys, 0u, 0u};
    Rand4 r0{0.5f, 0.5f, 0.0f, 0.0f};
    float3 o, d;
    camera_ray(P.cam, rng, r0, u, v, true, o, d);
    HitState h;
    closest_hit<MODE>(P.scene, sph, tri, o, d, h);
    const size_t i = (size_t)x + (size_t)y * P.width;
    P.prim_id[i] = h.obj >= 0 ? P.scene.ids[h.obj] : -1;
    P.t[i] = h.closest;
}

// tracer.rs:177-184 over pixels [begin, end) of a summed float4 buffer.  A thread owns the aligned group of four
// pixels 4 g .. 4 g + 3: whole groups inside the range leave as three 32-bit words, the ragged ends as bytes.
__global__ void tonemap_kernel(const float4* __restrict__ sum, float spp, uint64_t begin, uint64_t end,
                               uint8_t* __restrict__ rgb) {
    const bool words = (reinterpret_cast<uintptr_t>(rgb) & 3u) == 0u;
    const uint64_t g0 = begin / 4u, g1 = (end + 3u) / 4u;
    for (uint64_t gi = g0 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; gi < g1; gi += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t p0 = 4u * gi;
        uint32_t r[4], g[4], b[4];
        bool in[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            in[j] = p0 + j >= begin && p0 + j < end;
            const float4 s = in[j] ? sum[p0 + j] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            tone_rgb(f3(s.x, s.y, s.z), spp, r[j], g[j], b[j]);
        }
        if (words && in[0] && in[3]) store_rgb8x4(rgb, p0, r, g, b);
        else
            for (int j = 0; j < 4; ++j) if (in[j]) { uint8_t* px = rgb + 3 * (p0 + j); px[0] = (uint8_t)r[j]; px[1] = (uint8_t)g[j]; px[2] = (uint8_t)b[j]; }
    }
}

// Sample-split exchange fused with the write-out: each rank owns pixels [begin, end); it loads the
// n_peers partial sums of those pixels straight from the peers' HBM over NVLink (P2P-mapped
// pointers), adds them in rank order, and tone-maps — no intermediate reduced buffer in HBM.  Four pixels per
// thread: 64 contiguous bytes from every peer in flight, and the RGB8 result - usually into rank 0's peer-mapped
// frame - leaves as three 32-bit words instead of twelve single bytes.
constexpr int kMaxPeers = 16;
struct PeerPtrs {
    const float4* p[kMaxPeers];
};
__global__ void reduce_tonemap_kernel(const __grid_constant__ PeerPtrs peers, int n_peers, float spp, uint64_t begin,
                                      uint64_t end, uint8_t* __restrict__ rgb, float4* __restrict__ sum_out) {
    const bool words = (reinterpret_cast<uintptr_t>(rgb) & 3u) == 0u;
    const uint64_t g0 = begin / 4u, g1 = (end + 3u) / 4u;
    for (uint64_t gi = g0 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; gi < g1; gi += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t p0 = 4u * gi;
        bool in[4];
        float4 acc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) { in[j] = p0 + j >= begin && p0 + j < end; acc[j] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); }
        for (int r0 = 0; r0 < n_peers; r0 += 4) {  // 4 peers x 4 pixels = 16 loads in flight before the adds
            float4 v[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    v[r][j] = (r0 + r < n_peers && in[j]) ? __ldg(peers.p[r0 + r] + p0 + j) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
            for (int r = 0; r < 4; ++r)
                if (r0 + r < n_peers) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (r0 + r == 0) acc[j] = v[r][j];  // the sum STARTS with rank 0's partial (not 0 + it: -0 stays -0)
                        else { acc[j].x += v[r][j].x; acc[j].y += v[r][j].y; acc[j].z += v[r][j].z; acc[j].w += v[r][j].w; }
                    }
                }
        }
        if (sum_out) {
#pragma unroll
            for (int j = 0; j < 4; ++j) if (in[j]) sum_out[p0 + j] = acc[j];
        }
        if (rgb) {
            uint32_t r[4], g[4], b[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) tone_rgb(f3(acc[j].x, acc[j].y, acc[j].z), spp, r[j], g[j], b[j]);
            if (words && in[0] && in[3]) store_rgb8x4(rgb, p0, r, g, b);
            else
                for (int j = 0; j < 4; ++j) if (in[j]) { uint8_t* px = rgb + 3 * (p0 + j); px[0] = (uint8_t)r[j]; px[1] = (uint8_t)g[j]; px[2] = (uint8_t)b[j]; }
        }
    }
}

static int stream_grid(uint64_t n, int sm_count);

// color_utils::srgb_encode (color_utils.rs:49-66) on an RGB8 buffer: a u8 -> u8 map, so a 256-entry table
// (built on the host, context.cu) applied to four bytes per thread-iteration.  HBM-bound: 1 B in + 1 B out.
struct SrgbLut {
    uint8_t v[256];
};
__global__ void srgb_encode_kernel(uint8_t* __restrict__ rgb, size_t nbytes, const __grid_constant__ SrgbLut lut) {
    __shared__ uint8_t t[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) t[i] = lut.v[i];
    __syncthreads();
    const size_t nwords = nbytes / 4;
    uint32_t* w = reinterpret_cast<uint32_t*>(rgb);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t x = w[i];
        w[i] = (uint32_t)t[x & 0xff] | ((uint32_t)t[(x >> 8) & 0xff] << 8) | ((uint32_t)t[(x >> 16) & 0xff] << 16) |
               ((uint32_t)t[x >> 24] << 24);
    }
    if (blockIdx.x == 0 && threadIdx.x < (nbytes & 3)) { const size_t j = nwords * 4 + threadIdx.x; rgb[j] = t[rgb[j]]; }
}

// Exhaustive check of div_const for divisor b: every f32 numerator in {0} U [2^-24, hi] (the kernels form
// (x + xi) and ((height - y) + xi) with xi a multiple of 2^-24 in [0, 1), so nothing else can occur).
__global__ void div_verify_kernel(float b, float y, uint32_t hi_bits, unsigned int* __restrict__ mismatches) {
    const uint32_t lo_bits = 0x33800000u;  // 2^-24
    unsigned int bad = 0;
    for (uint64_t i = lo_bits + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i <= hi_bits;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const float n = __uint_as_float((uint32_t)i);
        bad += __float_as_uint(div_const(n, b, y, true)) != __float_as_uint(n / b);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) bad += __float_as_uint(div_const(0.0f, b, y, true)) != __float_as_uint(0.0f / b);
    if (bad) atomicAdd(mismatches, bad);
}

cudaError_t launch_srgb_encode(uint8_t* rgb, size_t nbytes, const uint8_t lut[256], int sm_count, cudaStream_t stream) {
    if (nbytes == 0) return cudaSuccess;
    if (reinterpret_cast<uintptr_t>(rgb) & 3) return cudaErrorMisalignedAddress;
    SrgbLut L;
    for (int i = 0; i < 256; ++i) L.v[i] = lut[i];
    srgb_encode_kernel<<<stream_grid(nbytes / 4 + 1, sm_count), 256, 0, stream>>>(rgb, nbytes, L);
    return cudaGetLastError();
}

cudaError_t launch_div_verify(float b, float hi, unsigned int* d_mismatches, int sm_count, cudaStream_t stream) {
    div_verify_kernel<<<sm_count * 8, 256, 0, stream>>>(b, 1.0f / b, __builtin_bit_cast(uint32_t, hi), d_mismatches);
    return cudaGetLastError();
}

// ---- launchers -------------------------------------------------------------------------------------
static dim3 tile_grid(uint32_t w, uint32_t h) { return dim3((w + kTileW - 1) / kTileW, (h + kTileH - 1) / kTileH); }


// Job-queue modes (tree walk, or the scan with a queue): chunks of at most P.rad_chunk samples; per chunk one
// persistent trace launch (as many CTAs as fit on the device at once) and one in-order accumulate launch.
template <class K>
static cudaError_t launch_queue_chunks(K kernel, size_t smem, RenderParams P, int sm_count, cudaStream_t stream, uint32_t* launches) {
    if (!P.rad || !P.acc || !P.job_counter || P.rad_chunk == 0) return cudaErrorInvalidValue;
    const uint64_t tiles = (uint64_t)((P.width + 7) / 8) * ((P.height + 3) / 4);
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBlock, smem);
    if (e != cudaSuccess) return e;
    const size_t npix = (size_t)P.width * P.height;
    for (uint32_t base = 0; base < P.spp; base += P.rad_chunk) {
        uint32_t k = P.spp - base < P.rad_chunk ? P.spp - base : P.rad_chunk;
        while (tiles * k >= (1ull << 32)) k = (k + 1) / 2;  // batch indices are 32-bit
        P.chunk_base = base; P.chunk_k = k;
        // no more CTAs than there are batches to draw (a 1-spp 600x600 frame has 11 250)
        const uint64_t batches = tiles * ((k + kBatchSamples - 1) / kBatchSamples);
        const uint64_t want = (batches + kBlock / 32 - 1) / (kBlock / 32);
        const int grid = (int)std::min<uint64_t>((uint64_t)sm_count * (per_sm > 0 ? per_sm : 1), want ? want : 1);
        if ((e = cudaMemsetAsync(P.job_counter, 0, sizeof(unsigned int), stream)) != cudaSuccess) return e;
        kernel<<<grid, kBlock, smem, stream>>>(P);
        accumulate_kernel<<<stream_grid((npix + 3) / 4 + P.height, sm_count), 256, 0, stream>>>(P, base == 0, base + k >= P.spp);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        *launches += 2;
        if (k < P.rad_chunk && base + k < P.spp) base -= P.rad_chunk - k;  // a halved chunk: continue right after it
    }
    return cudaSuccess;
}

cudaError_t launch_render(const RenderParams& P, int sm_count, cudaStream_t stream, uint32_t* launches) {
    *launches = 0;
    if (P.scene.bvh_nodes) {
        if (P.walk_counters) return launch_queue_chunks(render_tree_kernel<true>, 0, P, sm_count, stream, launches);
        return launch_queue_chunks(render_tree_kernel<false>, 0, P, sm_count, stream, launches);
    }
    if (P.queue) {
        const auto k_smem = render_queue_kernel<kScanSmem>;
        const auto k_glob = render_queue_kernel<kScanGlobal>;
        if (P.scene.smem_bytes > 0) {
            cudaError_t e = cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, P.scene.smem_bytes);
            if (e != cudaSuccess) return e;
            return launch_queue_chunks(k_smem, (size_t)P.scene.smem_bytes, P, sm_count, stream, launches);
        }
        return launch_queue_chunks(k_glob, 0, P, sm_count, stream, launches);
    }
    // the tiles this launch owns: t = j * part_mod + part_rem < part_pitch * tile rows
    const uint32_t n_tiles = P.part_pitch * ((P.height + kTileH - 1) / kTileH);
    const dim3 grid(n_tiles > P.part_rem ? (n_tiles - P.part_rem + P.part_mod - 1) / P.part_mod : 0u);
    if (grid.x == 0) return cudaSuccess;
    if (P.scene.smem_bytes > 0) {
        cudaError_t e = cudaFuncSetAttribute(render_kernel<kScanSmem>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             P.scene.smem_bytes);
        if (e != cudaSuccess) return e;
        render_kernel<kScanSmem><<<grid, kBlock, P.scene.smem_bytes, stream>>>(P);
    } else {
        render_kernel<kScanGlobal><<<grid, kBlock, 0, stream>>>(P);
    }
    *launches = 1;
    return cudaGetLastError();
}

cudaError_t launch_primary_hits(const PrimaryParams& P, cudaStream_t stream) {
    const dim3 grid = tile_grid(P.width, P.height);
    if (P.scene.bvh_nodes) {
        primary_hits_kernel<kTree><<<grid, kBlock, 0, stream>>>(P);
    } else if (P.scene.smem_bytes > 0) {
        cudaError_t e = cudaFuncSetAttribute(primary_hits_kernel<kScanSmem>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             P.scene.smem_bytes);
        if (e != cudaSuccess) return e;
        primary_hits_kernel<kScanSmem><<<grid, kBlock, P.scene.smem_bytes, stream>>>(P);
    } else {
        primary_hits_kernel<kScanGlobal><<<grid, kBlock, 0, stream>>>(P);
    }
    return cudaGetLastError();
}

// Streaming-from-L2 read microbenchmark (the denominator of config D's L2 roofline is measured on the
// box, not assumed): per pass the grid reads the buffer once, 16 bytes per lane and load (ld.global.cg: L2, not L1),
// 8 loads in flight per lane; the XOR of everything keeps the loads alive.  The caller runs an untimed pass first, so
// the timed passes hit L2.
__global__ void __launch_bounds__(256) l2_read_kernel(const uint4* __restrict__ buf, size_t n16, uint32_t passes,
                                                      uint32_t* __restrict__ sink) {
    uint4 acc = make_uint4(0u, 0u, 0u, 0u);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (uint32_t p = 0; p < passes; ++p) {
        // a different starting phase per CTA and pass, so that CTAs do not march through the buffer in lock step
        size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x + (size_t)p * 977u * blockDim.x) % n16;
        const size_t per_thread = n16 / stride;
        for (size_t k = 0; k + 8 <= per_thread; k += 8) {
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) { v[u] = __ldcg(buf + i); i += stride; if (i >= n16) i -= n16; }
#pragma unroll
            for (int u = 0; u < 8; ++u) { acc.x ^= v[u].x; acc.y ^= v[u].y; acc.z ^= v[u].z; acc.w ^= v[u].w; }
        }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x9e3779b9u) *sink = acc.x;  // practically never: keeps the loads live
}

cudaError_t launch_l2_read(const void* buf, size_t bytes, uint32_t passes, uint32_t* sink, int sm_count, cudaStream_t stream,
                           size_t* bytes_per_pass) {
    const size_t n16 = bytes / 16;
    const int grid = sm_count * 2;  // 16 warps / SM, 8 x 16 B in flight per lane
    const size_t stride = (size_t)grid * 256, rounds = n16 / stride / 8;
    if (rounds == 0) return cudaErrorInvalidValue;
    if (bytes_per_pass) *bytes_per_pass = rounds * 8 * stride * 16;
    l2_read_kernel<<<grid, 256, 0, stream>>>((const uint4*)buf, n16, passes, sink);
    return cudaGetLastError();
}

static int stream_grid(uint64_t n, int sm_count) {
    uint64_t blocks = (n + 255) / 256;
    uint64_t cap = (uint64_t)sm_count * 8;  // a multiple of the SM count; grid-stride covers the rest
    return (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
}

cudaError_t launch_tonemap(const float4* sum, uint32_t spp, uint64_t begin, uint64_t end, uint8_t* rgb, int sm_count,
                           cudaStream_t stream) {
    if (end <= begin) return cudaSuccess;
    tonemap_kernel<<<stream_grid((end - begin) / 4 + 2, sm_count), 256, 0, stream>>>(sum, (float)spp, begin, end, rgb);
    return cudaGetLastError();
}

cudaError_t launch_reduce_tonemap(const float4* const* peers, int n_peers, uint32_t spp, uint64_t begin, uint64_t end,
                                  uint8_t* rgb, float4* sum_out, int sm_count, cudaStream_t stream) {
    if (n_peers < 1 || n_peers > kMaxPeers) return cudaErrorInvalidValue;
    if (end <= begin) return cudaSuccess;
    PeerPtrs pp{};
    for (int i = 0; i < n_peers; ++i) pp.p[i] = peers[i];
    reduce_tonemap_kernel<<<stream_grid((end - begin) / 4 + 2, sm_count), 256, 0, stream>>>(pp, n_peers, (float)spp, begin, end,
                                                                                  rgb, sum_out);
    return cudaGetLastError();
}

cudaError_t launch_render_mt(const MtParams& P, cudaStream_t stream) {
    if (P.height / 4u == 0u) return cudaSuccess;  // t_height = 0: render_mt returns an empty Vec
    const dim3 grid((P.width + kTileW - 1) / kTileW, (4u * (P.height / 4u) + kTileH - 1) / kTileH);
    render_mt_kernel<<<grid, kBlock, 0, stream>>>(P);
    return cudaGetLastError();
}

static int stream_grid(uint64_t n, int sm_count);
cudaError_t launch_mt_accumulate(const uint8_t* frame, float* acc, uint32_t sample, size_t n, int sm_count, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    mt_accumulate_kernel<<<stream_grid(n, sm_count), 256, 0, stream>>>(frame, acc, (float)sample, n);
    return cudaGetLastError();
}

cudaError_t launch_mt_finish(const float* acc, uint8_t* out, size_t n, int sm_count, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    mt_finish_kernel<<<stream_grid(n, sm_count), 256, 0, stream>>>(acc, out, n);
    return cudaGetLastError();
}

// ---- stream-ordered flags in peer-visible memory ------------------------------------------------------------
// The sample-split exchange orders its kernels ACROSS GPUs with these instead of host-side barriers: a rank raises
// its flag once its partial sums are complete, and the peers' reduce kernels are queued behind a wait on it.
// Release / acquire at system scope; the flag lives in memory the peers map through CUDA IPC (NVLink).
__global__ void flag_set_kernel(uint32_t* flag, uint32_t value) {
    __threadfence_system();  // everything earlier in this stream is visible system-wide before the flag moves
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}

struct FlagPtrs { const uint32_t* f[kMaxPeers]; };

// one thread per flag spins until its flag has reached `value` (wrap-safe signed distance) - or until timeout_ns have
// passed (%globaltimer): a peer that died between frames must not leave this GPU spinning for ever.  A timed-out
// wait records 1 + the index of its flag in *err (host-mapped) and lets the stream go on; the host side reports it.
__global__ void flag_wait_kernel(const __grid_constant__ FlagPtrs p, int n, uint32_t value, unsigned long long timeout_ns,
                                 unsigned int* err) {
    if ((int)threadIdx.x < n) {
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        uint32_t x;
        for (uint32_t spin = 0;; ++spin) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(x) : "l"(p.f[threadIdx.x]) : "memory");
            if ((int32_t)(x - value) >= 0) break;
            if ((spin & 1023u) == 1023u) {
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > timeout_ns) {  // plain stores: the word is host-mapped, no atomics needed to report ONE late flag
                    if (err && *(volatile unsigned int*)err == 0u) *(volatile unsigned int*)err = 1u + threadIdx.x;
                    break;
                }
            }
        }
    }
    __syncthreads();
    __threadfence_system();
}

cudaError_t launch_flag_set(uint32_t* flag, uint32_t value, cudaStream_t stream) {
    flag_set_kernel<<<1, 1, 0, stream>>>(flag, value);
    return cudaGetLastError();
}

cudaError_t launch_flag_wait(const uint32_t* const* flags, int n, uint32_t value, unsigned long long timeout_ns,
                             unsigned int* d_err, cudaStream_t stream) {
    if (n < 1 || n > kMaxPeers) return cudaErrorInvalidValue;
    FlagPtrs pp{};
    for (int i = 0; i < n; ++i) pp.f[i] = flags[i];
    flag_wait_kernel<<<1, kMaxPeers, 0, stream>>>(pp, n, value, timeout_ns, d_err);
    return cudaGetLastError();
}

// ---- no-allocation variant of the job-queue launch (forma_rt_render_device_noalloc; no counterpart in the
// reference: added for callers that cannot call cudaMalloc - code under CUDA
// graph capture, tight serving loops).  The scratch of render_tree_kernel / render_queue_kernel + accumulate_kernel
// - the job counter, the per-path radiance `rad`, the cross-chunk sum `acc` - lives in ONE file-scope device array
// of a fixed, modest size instead of the handle's cudaMalloc'ed buffers:
//     [0, 256)                         job counter (cleared per chunk by the launcher)
//     [256, 256 + 4 MiB)               rad: kNoAllocSlots float4 path slots
//     [256 + 4 MiB, 256 + 8 MiB)       acc: kNoAllocSlots float4 pixels
// = 8 MiB + 256 B held by every GPU that loads the library, until exit.  Every call shares it and it keeps the
// last call's contents: the job counter is cleared before every launch, every rad / acc slot a launch reads has
// been written by that launch first, so nothing else needs to start at zero - but two calls that overlap on the
// GPU (two streams, two handles, two threads) RACE on it.  The array's name is only a host-side stand-in: the device
// address comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr size_t kNoAllocSlots = FORMA_RT_NOALLOC_MAX_PIXELS;
constexpr size_t kNoAllocBytes = 256 + 2 * kNoAllocSlots * sizeof(float4);
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

cudaError_t launch_render_noalloc(RenderParams P, int device, int sm_count, cudaStream_t stream, uint32_t* launches) {
    *launches = 0;
    static void* base_of_device[64] = {};
    static std::mutex base_lock;  // handles of several host threads may meet here for the first time
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    {
        std::lock_guard<std::mutex> guard(base_lock);
        if (!base_of_device[device]) {  // set-up, once per device; the caller has made `device` current
            void* p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess) return e;
            base_of_device[device] = p;
        }
    }
    const size_t npix = (size_t)P.width * P.height;
    if (npix == 0 || npix > kNoAllocSlots) return cudaErrorInvalidValue;  // the entry point reports this with the limit
    unsigned char* base = static_cast<unsigned char*>(base_of_device[device]);
    P.job_counter = reinterpret_cast<unsigned int*>(base);
    P.rad = reinterpret_cast<float4*>(base + 256);
    P.acc = reinterpret_cast<float4*>(base + 256 + kNoAllocSlots * sizeof(float4));
    const size_t chunk = kNoAllocSlots / npix;  // samples of the frame that fit the radiance slots (>= 1)
    P.rad_chunk = (uint32_t)(chunk < 64 ? chunk : 64);
    if (P.rad_chunk > P.spp) P.rad_chunk = P.spp;
    if (!P.scene.bvh_nodes) P.queue = 1u;  // scanned scenes: the job-queue scan kernel is the one with scratch
    return launch_render(P, sm_count, stream, launches);
}

}  // namespace forma
