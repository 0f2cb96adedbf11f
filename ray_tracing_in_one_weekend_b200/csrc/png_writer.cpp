// Minimal PNG writer for the output stage (replaces image::save_buffer at src/main.rs:128; the reference
// links the `image` crate, this image has zlib but no libpng).  8-bit RGB, filter type 0 on every row,
// one IDAT chunk compressed with zlib's deflate.
#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/rtgpu.h"

namespace rtb {
int fail_png(const std::string& msg);  // rtgpu.cu: sets rt_last_error, returns RT_ERR_IO
}

namespace {
void put_u32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24));
    v.push_back((uint8_t)(x >> 16));
    v.push_back((uint8_t)(x >> 8));
    v.push_back((uint8_t)x);
}
void chunk(std::vector<uint8_t>& out, const char type[4], const uint8_t* data, size_t n) {
    put_u32(out, (uint32_t)n);
    size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    if (n) out.insert(out.end(), data, data + n);
    put_u32(out, (uint32_t)crc32(0L, out.data() + at, (uInt)(n + 4)));
}
}  // namespace

extern "C" int rt_write_png(const char* path, const uint8_t* rgb8, uint32_t width, uint32_t height) {
    if (!path || !rgb8 || width == 0 || height == 0) return rtb::fail_png("rt_write_png: NULL argument or empty image");
    if ((uint64_t)width * height * 3u + height >= (1ull << 31)) return rtb::fail_png("rt_write_png: image too large");
    std::vector<uint8_t> raw((size_t)height * (1 + (size_t)width * 3));
    for (uint32_t y = 0; y < height; y++) {
        uint8_t* row = raw.data() + (size_t)y * (1 + (size_t)width * 3);
        row[0] = 0;  // filter: none
        std::copy(rgb8 + (size_t)y * width * 3, rgb8 + (size_t)(y + 1) * width * 3, row + 1);
    }
    uLongf zn = compressBound((uLong)raw.size());
    std::vector<uint8_t> z(zn);
    if (compress2(z.data(), &zn, raw.data(), (uLong)raw.size(), 6) != Z_OK) return rtb::fail_png("rt_write_png: deflate failed");
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr;
    put_u32(ihdr, width);
    put_u32(ihdr, height);
    const uint8_t tail[5] = {8, 2, 0, 0, 0};  // 8 bits, colour type 2 (RGB), deflate, adaptive filtering, no interlace
    ihdr.insert(ihdr.end(), tail, tail + 5);
    chunk(out, "IHDR", ihdr.data(), ihdr.size());
    chunk(out, "IDAT", z.data(), zn);
    chunk(out, "IEND", nullptr, 0);
    FILE* f = std::fopen(path, "wb");
    if (!f) return rtb::fail_png(std::string("rt_write_png: cannot open ") + path);
    size_t w = std::fwrite(out.data(), 1, out.size(), f);
    if (std::fclose(f) != 0 || w != out.size()) return rtb::fail_png(std::string("rt_write_png: short write to ") + path);
    return RT_OK;
}
