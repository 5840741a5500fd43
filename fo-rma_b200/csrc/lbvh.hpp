// lbvh.hpp — on-GPU linear BVH over the scene's finite primitives (the reference has no
// acceleration structure: shapes/aabb.rs is a stub and tracer.rs:195-200 scans every object).
#pragma once
#include <cuda_runtime.h>

namespace forma {

// What the builder indexes.  Primitives use a unified index: p < n_spheres -> spheres[p], else
// triangle p - n_spheres.  The first big_spheres / big_tris entries of each packed array are the
// "big" primitives the tree leaves out (scanned linearly instead).
struct LbvhInput {
    const float4* spheres;  // (cx, cy, cz, r)
    const float4* tris;     // 3 float4 per triangle: (v0|nx), (e1|ny), (e2|nz)
    int n_spheres, n_tris;
    int big_spheres, big_tris;
    int n_bvh_spheres;      // n_spheres - big_spheres
    int count;              // primitives in the tree: n_bvh_spheres + (n_tris - big_tris)
    float cmin[3];          // centroid bounds -> 10-bit Morton grid: cell = (c - cmin) * cscale
    float cscale[3];
    float qmin[3];          // the 16-bit grid the node boxes are stored on: plane = qmin + q / qinv
    float qinv[3];
};

struct LbvhDevice {
    uint4* nodes = nullptr;   // 2 uint4 per internal node (layout in lbvh.cu)
    int* prim = nullptr;      // unified primitive index of sorted leaf l
    int n_leaves = 0;
    int capacity = 0;
    cudaError_t build(const LbvhInput& in, cudaStream_t stream);
    void release();
};

}  // namespace forma
