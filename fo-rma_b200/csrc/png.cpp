// png.cpp — minimal PNG encoder (8-bit RGB, filter 0, zlib "stored" deflate blocks: no compression,
// no third-party code).  Stands in for `image::ImageBuffer::save` (tracer.rs:186).
#include "png.hpp"

#include <cstdio>
#include <cstring>
#include <vector>

namespace forma {
namespace {

uint32_t crc_table[256];
bool crc_ready = false;
void crc_init() {
    for (uint32_t n = 0; n < 256; ++n) {
        uint32_t c = n;
        for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        crc_table[n] = c;
    }
    crc_ready = true;
}
uint32_t crc32(uint32_t crc, const uint8_t* p, size_t n) {
    if (!crc_ready) crc_init();
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = crc_table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}
void be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}
void chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
    be32(out, (uint32_t)data.size());
    size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    be32(out, crc32(0, out.data() + start, out.size() - start));
}

}  // namespace

bool write_png_rgb8(const char* path, const uint8_t* rgb, uint32_t width, uint32_t height, std::string& err) {
    if (!path || !rgb || width == 0 || height == 0) { err = "write_png: bad argument"; return false; }
    // raw scanlines, each prefixed by filter type 0
    const size_t stride = (size_t)width * 3;
    std::vector<uint8_t> raw((stride + 1) * height);
    for (uint32_t y = 0; y < height; ++y) {
        raw[(stride + 1) * y] = 0;
        std::memcpy(&raw[(stride + 1) * y + 1], rgb + stride * y, stride);
    }
    // zlib stream of stored blocks
    std::vector<uint8_t> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;  // adler32
    size_t pos = 0;
    while (pos < raw.size()) {
        size_t n = raw.size() - pos;
        if (n > 65535) n = 65535;
        z.push_back(pos + n == raw.size() ? 1 : 0);
        z.push_back((uint8_t)(n & 0xff)); z.push_back((uint8_t)(n >> 8));
        z.push_back((uint8_t)(~n & 0xff)); z.push_back((uint8_t)((~n >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        for (size_t i = 0; i < n; ++i) {
            a += raw[pos + i]; if (a >= 65521) a -= 65521;
            b += a; if (b >= 65521) b -= 65521;
        }
        pos += n;
    }
    be32(z, (b << 16) | a);

    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr;
    be32(ihdr, width); be32(ihdr, height);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk(out, "IHDR", ihdr);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", {});

    FILE* f = std::fopen(path, "wb");
    if (!f) { err = std::string("write_png: cannot open ") + path; return false; }
    size_t w = std::fwrite(out.data(), 1, out.size(), f);
    bool ok = (w == out.size());
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) err = std::string("write_png: short write to ") + path;
    return ok;
}

}  // namespace forma
