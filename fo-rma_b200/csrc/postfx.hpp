// postfx.hpp — internal interface between postfx.cu (filters + forma_fx handle) and context.cu.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

struct forma_fx;

namespace forma {

// sets forma_rt_last_error() for the calling thread and returns `code` (context.cu)
int set_last_error(int code, const std::string& msg);

// forma_fx internals context.cu needs for forma_fx_run_traced
int fx_device(const forma_fx* fx);
uint32_t fx_width(const forma_fx* fx);
uint32_t fx_height(const forma_fx* fx);
// RGB8 frame on the device -> the handle's RGBA8 input frame (alpha 255) -> chain -> *d_out (the handle's output
// frame), all on `stream`; returns a forma_rt status
int fx_run_from_rgb8(forma_fx* fx, const uint8_t* d_rgb, float time, cudaStream_t stream, const uint8_t** d_out);

}  // namespace forma
