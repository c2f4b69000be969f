// image_io.h — dependency-free PNG encoder for the headless output stage: 8-bit RGBA, rows written
// top-down from a bottom-up plane (row 0 of the framebuffer is the bottom scan-line, reference
// src/Global/RenderPin.cu:91-92), zlib stream made of STORED deflate blocks (no compression, no zlib needed).
// Host C++ only; header-only so that the test twin can encode without the CUDA library.
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

namespace png_detail {
inline uint32_t crc32(const uint8_t *p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        ready = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return ~crc;
}
inline void be32(std::vector<uint8_t> &o, uint32_t v) { o.push_back(v >> 24); o.push_back(v >> 16); o.push_back(v >> 8); o.push_back(v); }
inline void chunk(std::vector<uint8_t> &o, const char type[4], const std::vector<uint8_t> &data) {
    be32(o, (uint32_t) data.size());
    const size_t at = o.size();
    o.insert(o.end(), type, type + 4);
    o.insert(o.end(), data.begin(), data.end());
    be32(o, crc32(o.data() + at, o.size() - at));
}
} // namespace png_detail

// rgba: H rows of W pixels, row 0 = bottom scan-line when bottom_up.  out: the complete .png file.
inline void png_encode_rgba8(const uint8_t *rgba, uint32_t W, uint32_t H, bool bottom_up, std::vector<uint8_t> &out) {
    using namespace png_detail;
    out.clear();
    const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    out.insert(out.end(), sig, sig + 8);
    std::vector<uint8_t> ihdr;
    be32(ihdr, W); be32(ihdr, H);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);      // 8-bit RGBA, no interlace
    chunk(out, "IHDR", ihdr);
    // filtered scan-lines (filter type 0), top-down
    const size_t stride = (size_t) W * 4 + 1;
    std::vector<uint8_t> raw(stride * H);
    for (uint32_t y = 0; y < H; y++) {
        const uint8_t *src = rgba + (size_t) (bottom_up ? H - 1 - y : y) * W * 4;
        raw[y * stride] = 0;
        for (size_t k = 0; k < (size_t) W * 4; k++) raw[y * stride + 1 + k] = src[k];
    }
    // zlib container, stored blocks of at most 65535 bytes, Adler-32 of the raw data
    std::vector<uint8_t> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    size_t pos = 0;
    do {
        const size_t n = raw.size() - pos < 65535 ? raw.size() - pos : 65535;
        z.push_back(pos + n == raw.size() ? 1 : 0);
        z.push_back(n & 0xFF); z.push_back(n >> 8); z.push_back(~n & 0xFF); z.push_back((~n >> 8) & 0xFF);
        for (size_t k = 0; k < n; k++) {
            z.push_back(raw[pos + k]);
            a = (a + raw[pos + k]) % 65521u;
            b = (b + a) % 65521u;
        }
        pos += n;
    } while (pos < raw.size());
    be32(z, (b << 16) | a);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", std::vector<uint8_t>());
}
