// divconst.cuh — exact division by a launch-constant divisor without the IEEE division sequence.
#pragma once

namespace forma {

// n / b for a launch-constant divisor b with y = RN(1 / b): q0 = RN(n y), q = RN(q0 + RN(n - q0 b) y)
// (Markstein's correction).  The host enables `fast` only after div_verify_kernel has compared this
// with IEEE division for EVERY f32 numerator the kernels can form for this b (kernels.cu).
__device__ __forceinline__ float div_const(float n, float b, float y, bool fast) {
    if (!fast) return n / b;
    const float q0 = n * y;
    return __fmaf_rn(__fmaf_rn(-q0, b, n), y, q0);
}

}  // namespace forma
