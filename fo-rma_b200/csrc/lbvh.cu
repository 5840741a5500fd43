// lbvh.cu — placeholder until the Morton/Karras builder lands (next milestone).
#include "lbvh.hpp"

namespace forma {

cudaError_t LbvhDevice::build(const float4*, int, cudaStream_t) { return cudaErrorNotSupported; }
void LbvhDevice::release() {
    if (nodes) cudaFree(nodes);
    nodes = nullptr;
    n_leaves = 0;
}

}  // namespace forma
