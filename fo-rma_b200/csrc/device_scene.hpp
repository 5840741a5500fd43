// device_scene.hpp — how Scene.objects (cpu_ray_tracer/scene.rs:4-7) is laid out in HBM, and the
// kernel parameter blocks.  Shared by the host packer (context.cu) and the kernels (trace.cuh).
//
// The reference scans a Vec<Box<dyn Hitable>> in list order with one shared HitRecord
// (tracer.rs:190-200).  Only Plane::hit is order-sensitive (it ignores t_max for t and clobbers the
// record, plane.rs:25-41), so the list is cut into SEGMENTS at every plane: inside a segment the
// spheres and triangles are order-independent up to the tie rule "equal t keeps the earlier list
// entry", and are stored type-sorted as float4 SoA:
//
//   spheres[i]      = (cx, cy, cz, r)                                  16 B
//   tris[3i..3i+2]  = (v0 | n.x), (e1 | n.y), (e2 | n.z)               48 B   e1 = v1-v0, e2 = v2-v0,
//                                                                      n = unit(cross(e1,e2)), host f32
//   planes[3k..3k+2]= position, orientation, size (w unused)            48 B
//   *_obj[i]        = list index of packed primitive i (ties, material lookup)
//   material[j]     = (r, g, b, fuzz) of list entry j                  16 B
//   kindmat[j]      = kind | material << 8;  ids[j] = reported id
//   segments[s]     = (sph_begin, sph_end, tri_begin, tri_end, plane)  plane = -1 when none follows
//
// With accel = LBVH the spheres of the (single) segment are additionally indexed by a binary
// radix tree (lbvh.cu): bvh_nodes[2k], [2k+1] = child boxes of internal node k.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/forma_rt.h"

namespace forma {

constexpr int kMaxDepthCap = 64;  // size of the per-path attenuation stack (reference: MAX_DEPTH 50)

struct Segment {
    int sph_begin, sph_end, tri_begin, tri_end, plane;
};

struct SceneView {
    const float4* spheres;
    const int* sphere_obj;
    const float4* tris;
    const int* tri_obj;
    const float4* planes;
    const int* plane_obj;
    const Segment* segments;
    Segment seg0;              // = segments[0], as a kernel-parameter (constant-bank) copy
    const float4* material;
    const int* kindmat;
    const int* ids;
    // LBVH over spheres[0 .. n_spheres) (nullptr = brute force)
    const uint4* bvh_nodes;    // 2 uint4 (32 B) per internal node: quantised child boxes + child codes, see lbvh.cu
    const int* bvh_prim;       // unified primitive index (< n_spheres: sphere) of sorted leaf l
    int big_spheres, big_tris; // with a tree: the first big_* packed primitives are scanned linearly
    int finite_materials;      // 1: every colour component of every list entry is finite (see black_path)
    int tree_seg;              // with a tree: index of the one segment that holds spheres / triangles
    float bvh_center[3];       // bounding sphere of everything the tree indexes
    float bvh_r2;              // its radius squared
    float bvh_eps;             // absolute box pad: 2^-16 x scene scale
    float bvh_qmin[3];         // the tree's 16-bit grid: plane q on axis k lies at bvh_qmin[k] + q * bvh_qstep[k]
    float bvh_qstep[3];
    float bvh_q;               // max 1/|radius| over the spheres of the tree (0 without spheres): scales the sphere pad
    float bvh_rmax;            // largest |radius| in the tree
    float cull_center[3];      // bounding sphere of every finite sphere and triangle (closest_hit's cull);
    float cull_r2;             // radius squared; +inf = never cull
    int n_segments;
    int n_spheres;
    int n_tris;
    int n_planes;
    int smem_bytes;            // > 0: spheres+tris are staged in shared memory by the kernel
};

// Philox4x32-10 round keys: k[2r], k[2r+1] = (seed_lo + r*0x9E3779B9, seed_hi + r*0xBB67AE85)
struct RoundKeys {
    uint32_t k[20];
    void expand(uint64_t seed) {
        uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
        for (int r = 0; r < 10; ++r) { k[2 * r] = k0; k[2 * r + 1] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
    }
};

struct RenderParams {
    forma_rt_camera cam;
    SceneView scene;
    uint32_t width, height;
    uint32_t spp, sample_offset, max_depth, flags;
    RoundKeys rk;              // Philox key schedule of the seed
    uint8_t* rgb;              // nullable
    float4* sum;               // nullable
    unsigned long long* rays;  // device counter
    const float4* prior;       // nullable: per-pixel f32 sum of the spp_prior samples traced by earlier calls (progressive mode)
    uint32_t spp_prior;        // the new samples continue that sum; the write-out averages over spp_prior + spp
    // tile-interleaved multi-GPU split: this launch owns the 32x4 tiles (tx, ty) with (ty * part_pitch + tx) % part_mod == part_rem;
    // part_pitch >= tiles per row is coprime with part_mod, so the deal rotates from row to row
    uint32_t part_mod, part_rem, part_pitch;
    float inv_w, inv_h;        // RN(1 / width), RN(1 / height)
    uint32_t fast_div;         // 1: div_const's multiply form was verified exhaustively for both divisors
    uint32_t queue;            // scan mode: 1 = render_queue_kernel ((pixel, sample) jobs), 0 = render_kernel (a pixel per thread)
    // ---- job-queue modes (render_tree_kernel / render_queue_kernel + accumulate_kernel, kernels.cu) ----
    float4* rad;               // [chunk sample][pixel] radiance of the finished paths of one chunk
    float4* acc;               // [pixel] running in-order sum across chunks
    unsigned int* job_counter; // next (pixel, sample) job of the chunk
    uint32_t rad_chunk;        // capacity of `rad` in samples
    uint32_t chunk_base, chunk_k; // samples [chunk_base, chunk_base + chunk_k) are traced by this launch
    uint32_t leaf_batch;       // lanes that must stop before the warp looks at its waiting leaves
    uint32_t refill;           // refill a warp's finished lanes once this many are idle
    unsigned long long* walk_counters;  // nullable: [0] += internal nodes visited, [1] += leaf primitives tested (the
                                        // instrumented instantiation of render_tree_kernel; forma_rt_walk_stats)
};

// tracer::render_mt (tracer.rs:83-134): one 1-spp frame in four bands (kernels.cu render_mt_kernel)
struct MtParams {
    forma_rt_camera cam;       // the model's camera (tracer.rs:90 clones it)
    SceneView scene;           // a fresh get_simple_scene() (tracer.rs:105), never the model's scene
    uint32_t width, height, frame, max_depth;
    RoundKeys rk;
    uint8_t* rgb;              // width * 4 * (height / 4) * 3 bytes
    unsigned long long* rays;
};

struct PrimaryParams {
    forma_rt_camera cam;
    SceneView scene;
    uint32_t width, height;
    int32_t* prim_id;
    float* t;
};

}  // namespace forma
