// kernels.hpp — launchers exported by kernels.cu / lbvh.cu to the context (context.cu).
#pragma once
#include <cuda_runtime.h>

#include "device_scene.hpp"

namespace forma {

// *launches = kernels launched (tree mode: a trace + an accumulate launch per chunk of samples)
cudaError_t launch_render(const RenderParams& P, int sm_count, cudaStream_t stream, uint32_t* launches);
// *d_mismatches += numerators in {0} U [2^-24, hi] for which div_const(n, b) != n / b
cudaError_t launch_div_verify(float b, float hi, unsigned int* d_mismatches, int sm_count, cudaStream_t stream);
// color_utils::srgb_encode (color_utils.rs:49-66) in place on nbytes of RGB8; lut = the u8 -> u8 map
cudaError_t launch_srgb_encode(uint8_t* rgb, size_t nbytes, const uint8_t lut[256], int sm_count, cudaStream_t stream);
cudaError_t launch_primary_hits(const PrimaryParams& P, cudaStream_t stream);
// launch_render's job-queue path with its scratch in a file-scope static device array (kernels.cu): allocates nothing.
// width * height <= FORMA_RT_NOALLOC_MAX_PIXELS; `device` = the current device (per-device symbol address cache)
cudaError_t launch_render_noalloc(RenderParams P, int device, int sm_count, cudaStream_t stream, uint32_t* launches);
cudaError_t launch_tonemap(const float4* sum, uint32_t spp, uint64_t begin, uint64_t end, uint8_t* rgb, int sm_count,
                           cudaStream_t stream);
cudaError_t launch_reduce_tonemap(const float4* const* peers, int n_peers, uint32_t spp, uint64_t begin, uint64_t end,
                                  uint8_t* rgb, float4* sum_out, int sm_count, cudaStream_t stream);

// the grid reads buf `passes` times through L2; *bytes_per_pass = what one pass reads (bytes rounded down to whole
// rounds of 8 loads per thread)
cudaError_t launch_l2_read(const void* buf, size_t bytes, uint32_t passes, uint32_t* sink, int sm_count, cudaStream_t stream,
                           size_t* bytes_per_pass);

// tracer::render_mt / save_image_mt (tracer.rs:83-158)
cudaError_t launch_render_mt(const MtParams& P, cudaStream_t stream);
cudaError_t launch_mt_accumulate(const uint8_t* frame, float* acc, uint32_t sample, size_t n, int sm_count, cudaStream_t stream);
cudaError_t launch_mt_finish(const float* acc, uint8_t* out, size_t n, int sm_count, cudaStream_t stream);
// stream-ordered flags in peer-visible memory (kernels.cu): *flag = value after everything earlier on `stream` is visible
// system-wide; later work on `stream` starts once every *flags[i] has reached value (n <= 16)
cudaError_t launch_flag_set(uint32_t* flag, uint32_t value, cudaStream_t stream);
// a wait that times out stores 1 + the flag's index in *d_err (host-mapped, may be null) and lets the stream continue
cudaError_t launch_flag_wait(const uint32_t* const* flags, int n, uint32_t value, unsigned long long timeout_ns,
                             unsigned int* d_err, cudaStream_t stream);

}  // namespace forma
