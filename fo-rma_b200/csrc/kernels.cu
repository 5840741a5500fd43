// kernels.cu — sm_100a kernels of the tracer path and their launchers.
//   render_kernel        tracer.rs:160-187 (save_image's pixel x sample loops) / :57-81 (render)
//   primary_hits_kernel  config-B dump: closest hit of the un-jittered pinhole primary ray
//   tonemap_kernel       tracer.rs:177-184 on a summed f32 buffer
//   reduce_tonemap_kernel  the cross-GPU exchange step fused with the write-out (peer loads)
// Built with -fmad=false (see trace.cuh).
#include <cstdio>

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

template <int MODE>
__global__ void __launch_bounds__(kBlock) render_kernel(const __grid_constant__ RenderParams P) {
    const float4 *sph, *tri;
    stage_scene<MODE>(P.scene, sph, tri);

    __shared__ __align__(16) uint8_t tile_rgb[kTileH][kTileW * 3];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lx = warp * 8 + (lane & 7), ly = lane >> 3;
    const uint32_t x = blockIdx.x * kTileW + lx, y = blockIdx.y * kTileH + ly;
    const bool live = x < P.width && y < P.height;

    uint32_t rays = 0;
    float3 col = f3(0.0f, 0.0f, 0.0f);
    if (live) {
        PathRng rng{P.rk, x + y * P.width, 0u};
        const float fx = (float)x, fy = (float)(P.height - y);  // tracer.rs:171-172 (Q1: height - y)
        const float fw = (float)P.width, fh = (float)P.height;
        const bool no_jitter = P.flags & FORMA_RT_NO_JITTER, pinhole = P.flags & FORMA_RT_PINHOLE;
        // tracer.rs:168-175's sample loop and get_color's recursion (tracer.rs:189-219) as ONE flat
        // loop of closest-hit queries: a lane whose path ended starts the next sample of its own pixel
        // in the same iteration the other lanes spend on their next bounce, so a warp runs for the
        // longest per-pixel SUM of path lengths, not for spp x the longest path.  Per pixel the samples
        // are still traced and added in order, so the f32 sums match the reference's loop exactly.
        int stack[kMaxDepthCap];  // list indices of the hit objects: attenuations multiply innermost first
        uint32_t s = 0, depth = 0;
        bool alive = false;
        float3 o = f3(0.0f, 0.0f, 0.0f), d = o;
        for (;;) {
            if (!alive) {
                if (s >= P.spp) break;
                rng.sample = P.sample_offset + s;
                Rand4 r0 = rng.block(0, 0);
                if (no_jitter) { r0.x = 0.5f; r0.y = 0.5f; }
                const float u = (fx + r0.x) / fw;
                const float v = (fy + r0.y) / fh;
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

    if (P.sum && live) P.sum[(size_t)x + (size_t)y * P.width] = make_float4(col.x, col.y, col.z, (float)P.spp);

    if (P.rgb) {
        // tracer.rs:177-184: average in linear space, sqrt gamma, (c * 255.0) as u8
        float3 c = col / (float)P.spp;
        const uint32_t r = as_u8(sqrtf(c.x) * 255.0f), g = as_u8(sqrtf(c.y) * 255.0f), b = as_u8(sqrtf(c.z) * 255.0f);
        const bool full_tile = (blockIdx.x + 1) * kTileW <= P.width && (blockIdx.y + 1) * kTileH <= P.height &&
                               (P.width * 3u) % 4u == 0;
        if (full_tile) {  // stage 4 rows x 96 bytes, write them as 96 coalesced 32-bit words
            tile_rgb[ly][lx * 3 + 0] = (uint8_t)r;
            tile_rgb[ly][lx * 3 + 1] = (uint8_t)g;
            tile_rgb[ly][lx * 3 + 2] = (uint8_t)b;
            __syncthreads();
            if (threadIdx.x < kTileH * 24) {
                const int row = threadIdx.x / 24, word = threadIdx.x % 24;
                const size_t byte = ((size_t)(blockIdx.y * kTileH + row) * P.width + (size_t)blockIdx.x * kTileW) * 3;
                reinterpret_cast<uint32_t*>(P.rgb + byte)[word] = reinterpret_cast<const uint32_t*>(tile_rgb[row])[word];
            }
        } else if (live) {
            uint8_t* px = P.rgb + ((size_t)x + (size_t)y * P.width) * 3;
            px[0] = (uint8_t)r; px[1] = (uint8_t)g; px[2] = (uint8_t)b;
        }
    }
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

// tracer.rs:177-184 over pixels [begin, end) of a summed float4 buffer
__global__ void tonemap_kernel(const float4* __restrict__ sum, float spp, uint64_t begin, uint64_t end,
                               uint8_t* __restrict__ rgb) {
    for (uint64_t i = begin + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < end;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const float4 s = sum[i];
        float3 c = f3(s.x, s.y, s.z) / spp;
        rgb[3 * i + 0] = (uint8_t)as_u8(sqrtf(c.x) * 255.0f);
        rgb[3 * i + 1] = (uint8_t)as_u8(sqrtf(c.y) * 255.0f);
        rgb[3 * i + 2] = (uint8_t)as_u8(sqrtf(c.z) * 255.0f);
    }
}

// Sample-split exchange fused with the write-out: each rank owns pixels [begin, end); it loads the
// n_peers partial sums of those pixels straight from the peers' HBM over NVLink (P2P-mapped
// pointers), adds them in rank order, and tone-maps — no intermediate reduced buffer in HBM.
constexpr int kMaxPeers = 16;
struct PeerPtrs {
    const float4* p[kMaxPeers];
};
__global__ void reduce_tonemap_kernel(const __grid_constant__ PeerPtrs peers, int n_peers, float spp, uint64_t begin,
                                      uint64_t end, uint8_t* __restrict__ rgb, float4* __restrict__ sum_out) {
    for (uint64_t i = begin + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < end;
         i += (uint64_t)gridDim.x * blockDim.x) {
        float4 v[kMaxPeers];
#pragma unroll
        for (int r = 0; r < kMaxPeers; ++r)
            if (r < n_peers) v[r] = __ldg(peers.p[r] + i);  // all peer loads in flight before the adds
        float4 acc = v[0];
#pragma unroll
        for (int r = 1; r < kMaxPeers; ++r)
            if (r < n_peers) { acc.x += v[r].x; acc.y += v[r].y; acc.z += v[r].z; acc.w += v[r].w; }
        if (sum_out) sum_out[i] = acc;
        if (rgb) {
            float3 c = f3(acc.x, acc.y, acc.z) / spp;
            rgb[3 * i + 0] = (uint8_t)as_u8(sqrtf(c.x) * 255.0f);
            rgb[3 * i + 1] = (uint8_t)as_u8(sqrtf(c.y) * 255.0f);
            rgb[3 * i + 2] = (uint8_t)as_u8(sqrtf(c.z) * 255.0f);
        }
    }
}

// ---- launchers -------------------------------------------------------------------------------------
static dim3 tile_grid(uint32_t w, uint32_t h) { return dim3((w + kTileW - 1) / kTileW, (h + kTileH - 1) / kTileH); }

cudaError_t launch_render(const RenderParams& P, cudaStream_t stream) {
    const dim3 grid = tile_grid(P.width, P.height);
    if (P.scene.bvh_nodes) {
        render_kernel<kTree><<<grid, kBlock, 0, stream>>>(P);
    } else if (P.scene.smem_bytes > 0) {
        cudaError_t e = cudaFuncSetAttribute(render_kernel<kScanSmem>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             P.scene.smem_bytes);
        if (e != cudaSuccess) return e;
        render_kernel<kScanSmem><<<grid, kBlock, P.scene.smem_bytes, stream>>>(P);
    } else {
        render_kernel<kScanGlobal><<<grid, kBlock, 0, stream>>>(P);
    }
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

static int stream_grid(uint64_t n, int sm_count) {
    uint64_t blocks = (n + 255) / 256;
    uint64_t cap = (uint64_t)sm_count * 8;  // a multiple of the SM count; grid-stride covers the rest
    return (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
}

cudaError_t launch_tonemap(const float4* sum, uint32_t spp, uint64_t begin, uint64_t end, uint8_t* rgb, int sm_count,
                           cudaStream_t stream) {
    if (end <= begin) return cudaSuccess;
    tonemap_kernel<<<stream_grid(end - begin, sm_count), 256, 0, stream>>>(sum, (float)spp, begin, end, rgb);
    return cudaGetLastError();
}

cudaError_t launch_reduce_tonemap(const float4* const* peers, int n_peers, uint32_t spp, uint64_t begin, uint64_t end,
                                  uint8_t* rgb, float4* sum_out, int sm_count, cudaStream_t stream) {
    if (n_peers < 1 || n_peers > kMaxPeers) return cudaErrorInvalidValue;
    if (end <= begin) return cudaSuccess;
    PeerPtrs pp{};
    for (int i = 0; i < n_peers; ++i) pp.p[i] = peers[i];
    reduce_tonemap_kernel<<<stream_grid(end - begin, sm_count), 256, 0, stream>>>(pp, n_peers, (float)spp, begin, end,
                                                                                  rgb, sum_out);
    return cudaGetLastError();
}

}  // namespace forma
