// lbvh.cu — on-GPU linear BVH (Morton codes -> radix sort -> Karras 2012 hierarchy -> bottom-up
// refit) over the finite primitives of the scene.  The reference has no acceleration structure
// (shapes/aabb.rs is a stub; tracer.rs:195-200 scans every object), so this only has to return what
// the scan returns; see trace.cuh `bvh_closest` for how the walk stays exact.
//
// Node record (2 x uint4 = 32 B per internal node k, children L and R), read with one 256-bit load:
//   w0 = Llo.x | Llo.y << 16   w1 = Llo.z | Lhi.x << 16   w2 = Lhi.y | Lhi.z << 16
//   w3 = Rlo.x | Rlo.y << 16   w4 = Rlo.z | Rhi.x << 16   w5 = Rhi.y | Rhi.z << 16
//   w6 = code(L)  w7 = code(R)          code >= 0: internal node; code < 0: ~(unified primitive index)
// Box planes are u16 on a grid over the tree's bounding box (qmin + q / qinv per axis), lo rounded DOWN and
// hi rounded UP by a whole extra cell, so a stored box always contains the f32 box: the walk (trace.cuh)
// stays conservative and pads every box by coef(ray) * q_tree + eps so that the f32 sphere test of
// sphere.rs:23-51 can never report a hit that the box test culled.  prim[l] = primitive of sorted leaf l is
// kept for tooling only.
#include <cfloat>
#include <cstdint>

#include <cub/device/device_radix_sort.cuh>

#include "lbvh.hpp"

namespace forma {
namespace {

__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// Leaf bounds, 1/r and the 30-bit Morton code of the centroid of primitive `first + i`
// (unified index: < n_spheres -> sphere, else triangle).
__global__ void leaf_kernel(LbvhInput in, float4* __restrict__ leaf_lo, float4* __restrict__ leaf_hi,
                            uint32_t* __restrict__ keys, int* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= in.count) return;
    int prim;
    float3 lo, hi;
    float q = 0.0f;
    if (i < in.n_bvh_spheres) {
        const int s = in.big_spheres + i;
        const float4 sp = in.spheres[s];
        const float ar = fabsf(sp.w);  // sphere.rs:27 squares the radius: r and -r are the same sphere
        lo = make_float3(sp.x - ar, sp.y - ar, sp.z - ar);
        hi = make_float3(sp.x + ar, sp.y + ar, sp.z + ar);
        q = 1.0f / ar;
        prim = s;
    } else {
        const int t = in.big_tris + (i - in.n_bvh_spheres);
        const float4 a = in.tris[3 * t], e1 = in.tris[3 * t + 1], e2 = in.tris[3 * t + 2];
        const float3 v0 = make_float3(a.x, a.y, a.z);
        const float3 v1 = make_float3(a.x + e1.x, a.y + e1.y, a.z + e1.z);
        const float3 v2 = make_float3(a.x + e2.x, a.y + e2.y, a.z + e2.z);
        lo = make_float3(fminf(v0.x, fminf(v1.x, v2.x)), fminf(v0.y, fminf(v1.y, v2.y)), fminf(v0.z, fminf(v1.z, v2.z)));
        hi = make_float3(fmaxf(v0.x, fmaxf(v1.x, v2.x)), fmaxf(v0.y, fmaxf(v1.y, v2.y)), fmaxf(v0.z, fmaxf(v1.z, v2.z)));
        prim = in.n_spheres + t;
    }
    leaf_lo[i] = make_float4(lo.x, lo.y, lo.z, q);
    leaf_hi[i] = make_float4(hi.x, hi.y, hi.z, 0.0f);
    const float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
    const float fx = fminf(fmaxf((cx - in.cmin[0]) * in.cscale[0], 0.0f), 1023.0f);
    const float fy = fminf(fmaxf((cy - in.cmin[1]) * in.cscale[1], 0.0f), 1023.0f);
    const float fz = fminf(fmaxf((cz - in.cmin[2]) * in.cscale[2], 0.0f), 1023.0f);
    keys[i] = (expand_bits((uint32_t)fx) << 2) | (expand_bits((uint32_t)fy) << 1) | expand_bits((uint32_t)fz);
    vals[i] = i;  // value = leaf id before sorting
    (void)prim;
}

// common-prefix length of keys i and j, index as tie-break (Karras 2012, section 4)
__device__ __forceinline__ int delta(const uint32_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = keys[i], b = keys[j];
    return a == b ? 32 + __clz((uint32_t)i ^ (uint32_t)j) : __clz(a ^ b);
}

__global__ void hierarchy_kernel(const uint32_t* __restrict__ keys, int n, int2* __restrict__ children,
                                 int* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = (lo == gamma) ? ~gamma : gamma;              // leaf ids are stored complemented
    const int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    children[i] = make_int2(left, right);
    // parent slots: internal nodes [0, n-1), leaves [n-1, 2n-1)
    parent[left < 0 ? (n - 1) + ~left : left] = i;
    parent[right < 0 ? (n - 1) + ~right : right] = i;
    if (i == 0) parent[0] = -1;
}

struct Box {
    float3 lo, hi;
    float q;
};

__device__ __forceinline__ Box child_box(int c, const float4* __restrict__ leaf_lo, const float4* __restrict__ leaf_hi,
                                         const int* __restrict__ sorted_leaf, const float4* __restrict__ node_lo,
                                         const float4* __restrict__ node_hi) {
    float4 a, b;
    if (c < 0) { const int l = sorted_leaf[~c]; a = leaf_lo[l]; b = leaf_hi[l]; }
    else { a = __ldcg(node_lo + c); b = __ldcg(node_hi + c); }  // written by another SM in this launch: bypass L1
    return Box{make_float3(a.x, a.y, a.z), make_float3(b.x, b.y, b.z), a.w};
}

// Bottom-up: each leaf climbs; the second thread to reach a node (atomic flag) sees both children
// complete, writes the node record and the node's own box, and continues upward.
__device__ __forceinline__ int leaf_to_prim(const LbvhInput& in, int i) {  // leaf id (pre-sort) -> unified primitive index
    return i < in.n_bvh_spheres ? in.big_spheres + i : in.n_spheres + in.big_tris + (i - in.n_bvh_spheres);
}

__global__ void refit_kernel(LbvhInput in, int n, const int2* __restrict__ children, const int* __restrict__ parent,
                             const float4* __restrict__ leaf_lo, const float4* __restrict__ leaf_hi,
                             const int* __restrict__ sorted_leaf, float4* node_lo, float4* node_hi,
                             unsigned int* __restrict__ flags, uint4* __restrict__ nodes) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= n) return;
    int node = parent[(n - 1) + leaf];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&flags[node], 1u) == 0u) return;  // first arrival: the sibling subtree is not done yet
        __threadfence();
        const int2 ch = children[node];
        const Box L = child_box(ch.x, leaf_lo, leaf_hi, sorted_leaf, node_lo, node_hi);
        const Box R = child_box(ch.y, leaf_lo, leaf_hi, sorted_leaf, node_lo, node_hi);
        const int cl = ch.x < 0 ? ~leaf_to_prim(in, sorted_leaf[~ch.x]) : ch.x;
        const int cr = ch.y < 0 ? ~leaf_to_prim(in, sorted_leaf[~ch.y]) : ch.y;
        // (v - qmin) * qinv errs by < 0.02 cells (65535 x a few ulp), so floor - 1 / ceil + 1 are safely outside
        auto qlo = [&](float v, int k) { return (uint32_t)fminf(fmaxf(floorf((v - in.qmin[k]) * in.qinv[k]) - 1.0f, 0.0f), 65535.0f); };
        auto qhi = [&](float v, int k) { return (uint32_t)fminf(fmaxf(ceilf((v - in.qmin[k]) * in.qinv[k]) + 1.0f, 0.0f), 65535.0f); };
        nodes[2 * node + 0] = make_uint4(qlo(L.lo.x, 0) | qlo(L.lo.y, 1) << 16, qlo(L.lo.z, 2) | qhi(L.hi.x, 0) << 16,
                                         qhi(L.hi.y, 1) | qhi(L.hi.z, 2) << 16, qlo(R.lo.x, 0) | qlo(R.lo.y, 1) << 16);
        nodes[2 * node + 1] = make_uint4(qlo(R.lo.z, 2) | qhi(R.hi.x, 0) << 16, qhi(R.hi.y, 1) | qhi(R.hi.z, 2) << 16,
                                         (uint32_t)cl, (uint32_t)cr);
        __stcg(node_lo + node, make_float4(fminf(L.lo.x, R.lo.x), fminf(L.lo.y, R.lo.y), fminf(L.lo.z, R.lo.z), fmaxf(L.q, R.q)));
        __stcg(node_hi + node, make_float4(fmaxf(L.hi.x, R.hi.x), fmaxf(L.hi.y, R.hi.y), fmaxf(L.hi.z, R.hi.z), 0.0f));
        node = parent[node];
    }
}

// prim[l] = unified primitive index of sorted leaf l
__global__ void leaf_prim_kernel(LbvhInput in, const int* __restrict__ sorted_leaf, int* __restrict__ prim) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= in.count) return;
    const int i = sorted_leaf[l];
    prim[l] = leaf_to_prim(in, i);
}

}  // namespace

#define LB_TRY(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { free_tmp(); return e__; } } while (0)

cudaError_t LbvhDevice::build(const LbvhInput& in, cudaStream_t stream) {
    const int n = in.count;
    if (n < 2) return cudaErrorInvalidValue;
    void* tmp[12] = {};
    int ntmp = 0;
    auto free_tmp = [&]() { for (int i = 0; i < ntmp; ++i) cudaFree(tmp[i]); ntmp = 0; };
    auto alloc = [&](void** p, size_t bytes) { cudaError_t e = cudaMalloc(p, bytes); if (e == cudaSuccess) tmp[ntmp++] = *p; return e; };

    if (n > capacity) {  // persistent outputs
        release();
        cudaError_t e = cudaMalloc((void**)&nodes, sizeof(uint4) * 2 * (size_t)(n - 1));
        if (e == cudaSuccess) e = cudaMalloc((void**)&prim, sizeof(int) * (size_t)n);
        if (e != cudaSuccess) { release(); return e; }
        capacity = n;
    }
    n_leaves = n;

    float4 *leaf_lo, *leaf_hi, *node_lo, *node_hi;
    uint32_t *keys, *keys_alt;
    int *vals, *vals_alt, *parent;
    int2* children;
    unsigned int* flags;
    LB_TRY(alloc((void**)&leaf_lo, sizeof(float4) * n));
    LB_TRY(alloc((void**)&leaf_hi, sizeof(float4) * n));
    LB_TRY(alloc((void**)&node_lo, sizeof(float4) * n));
    LB_TRY(alloc((void**)&node_hi, sizeof(float4) * n));
    LB_TRY(alloc((void**)&keys, sizeof(uint32_t) * n));
    LB_TRY(alloc((void**)&keys_alt, sizeof(uint32_t) * n));
    LB_TRY(alloc((void**)&vals, sizeof(int) * n));
    LB_TRY(alloc((void**)&vals_alt, sizeof(int) * n));
    LB_TRY(alloc((void**)&parent, sizeof(int) * (2 * (size_t)n - 1)));
    LB_TRY(alloc((void**)&children, sizeof(int2) * (n - 1)));
    LB_TRY(alloc((void**)&flags, sizeof(unsigned int) * (n - 1)));

    const int B = 256, G = (n + B - 1) / B;
    leaf_kernel<<<G, B, 0, stream>>>(in, leaf_lo, leaf_hi, keys, vals);
    LB_TRY(cudaGetLastError());

    cub::DoubleBuffer<uint32_t> dk(keys, keys_alt);
    cub::DoubleBuffer<int> dv(vals, vals_alt);
    size_t sort_bytes = 0;
    LB_TRY(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, dk, dv, n, 0, 30, stream));
    void* sort_tmp = nullptr;
    LB_TRY(alloc(&sort_tmp, sort_bytes ? sort_bytes : 1));
    LB_TRY(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, dk, dv, n, 0, 30, stream));
    const uint32_t* sorted_keys = dk.Current();
    const int* sorted_leaf = dv.Current();

    hierarchy_kernel<<<G, B, 0, stream>>>(sorted_keys, n, children, parent);
    LB_TRY(cudaGetLastError());
    LB_TRY(cudaMemsetAsync(flags, 0, sizeof(unsigned int) * (n - 1), stream));
    refit_kernel<<<G, B, 0, stream>>>(in, n, children, parent, leaf_lo, leaf_hi, sorted_leaf, node_lo, node_hi, flags, nodes);
    LB_TRY(cudaGetLastError());
    leaf_prim_kernel<<<G, B, 0, stream>>>(in, sorted_leaf, prim);
    LB_TRY(cudaGetLastError());
    LB_TRY(cudaStreamSynchronize(stream));
    free_tmp();
    return cudaSuccess;
}

void LbvhDevice::release() {
    if (nodes) cudaFree(nodes);
    if (prim) cudaFree(prim);
    nodes = nullptr; prim = nullptr;
    n_leaves = 0; capacity = 0;
}

}  // namespace forma
