// lbvh.hpp — on-GPU linear BVH over the scene's spheres (the reference has no acceleration
// structure: shapes/aabb.rs is a stub and tracer.rs:195-200 scans every object).
#pragma once
#include <cuda_runtime.h>

namespace forma {

struct LbvhDevice {
    float4* nodes = nullptr;      // 4 float4 per internal node (see lbvh.cu)
    int n_leaves = 0;
    // Builds the tree over spheres[0..n) (device pointer, (cx,cy,cz,r)) on `stream`.
    cudaError_t build(const float4* spheres, int n, cudaStream_t stream);
    void release();
};

}  // namespace forma
