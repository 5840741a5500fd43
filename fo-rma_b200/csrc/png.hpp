// png.hpp — RGB8 -> PNG file; the role of the `image` crate at tracer.rs:186 (img_buf.save).
#pragma once
#include <cstdint>
#include <string>

namespace forma {
// Returns false and fills `err` when the file cannot be written.
bool write_png_rgb8(const char* path, const uint8_t* rgb, uint32_t width, uint32_t height, std::string& err);
}
