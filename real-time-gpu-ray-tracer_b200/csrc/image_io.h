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


// ---- OpenEXR, uncompressed scan-line file, 32-bit float channels ---------------------------------------------------
// planes: n_channels interleaved floats per pixel (3: radiance RGB -> channels B, G, R; 1: depth -> channel Z), H rows of
// W pixels, row 0 = bottom scan-line when bottom_up.  EXR stores channels in alphabetical order, one block per
// scan-line (y, byte count, then each channel's W values), top-down (lineOrder INCREASING_Y), after a table of block
// offsets.  Everything little-endian.
namespace exr_detail {
inline void put(std::vector<uint8_t> &o, const void *p, size_t n) { const uint8_t *b = (const uint8_t *) p; o.insert(o.end(), b, b + n); }
inline void i32(std::vector<uint8_t> &o, int32_t v) { put(o, &v, 4); }
inline void f32(std::vector<uint8_t> &o, float v) { put(o, &v, 4); }
inline void str(std::vector<uint8_t> &o, const char *s) { while (*s) o.push_back((uint8_t) *s++); o.push_back(0); }
inline void attr(std::vector<uint8_t> &o, const char *name, const char *type, const std::vector<uint8_t> &value) {
    str(o, name); str(o, type); i32(o, (int32_t) value.size()); o.insert(o.end(), value.begin(), value.end());
}
} // namespace exr_detail

inline void exr_encode_float(const float *plane, uint32_t W, uint32_t H, int n_channels, bool bottom_up, std::vector<uint8_t> &out) {
    using namespace exr_detail;
    out.clear();
    i32(out, 20000630);                       // magic
    i32(out, 2);                              // version 2, single-part scan-line
    // channel c of the file <- component src[c] of the plane (alphabetical channel names)
    const char *names3[3] = {"B", "G", "R"};
    const int src3[3] = {2, 1, 0};
    const char *names1[1] = {"Z"};
    const int src1[1] = {0};
    const char **names = n_channels == 3 ? names3 : names1;
    const int *src = n_channels == 3 ? src3 : src1;
    std::vector<uint8_t> v;
    for (int c = 0; c < n_channels; c++) { str(v, names[c]); i32(v, 2 /*FLOAT*/); v.push_back(0); v.push_back(0); v.push_back(0); v.push_back(0); i32(v, 1); i32(v, 1); }
    v.push_back(0);
    attr(out, "channels", "chlist", v);
    v.assign(1, 0);
    attr(out, "compression", "compression", v);                 // NO_COMPRESSION
    v.clear(); i32(v, 0); i32(v, 0); i32(v, (int32_t) W - 1); i32(v, (int32_t) H - 1);
    attr(out, "dataWindow", "box2i", v);
    attr(out, "displayWindow", "box2i", v);
    v.assign(1, 0);
    attr(out, "lineOrder", "lineOrder", v);                     // INCREASING_Y
    v.clear(); f32(v, 1.0f);
    attr(out, "pixelAspectRatio", "float", v);
    v.clear(); f32(v, 0.0f); f32(v, 0.0f);
    attr(out, "screenWindowCenter", "v2f", v);
    v.clear(); f32(v, 1.0f);
    attr(out, "screenWindowWidth", "float", v);
    out.push_back(0);                         // end of header
    const size_t line_bytes = (size_t) W * 4 * n_channels, block = 8 + line_bytes;
    const uint64_t first = out.size() + (uint64_t) H * 8;
    for (uint32_t y = 0; y < H; y++) { const uint64_t off = first + (uint64_t) y * block; put(out, &off, 8); }
    std::vector<float> row(W);
    for (uint32_t y = 0; y < H; y++) {
        const float *line = plane + (size_t) (bottom_up ? H - 1 - y : y) * W * n_channels;
        i32(out, (int32_t) y);
        i32(out, (int32_t) line_bytes);
        for (int c = 0; c < n_channels; c++) {
            for (uint32_t x = 0; x < W; x++) row[x] = line[(size_t) x * n_channels + src[c]];
            put(out, row.data(), (size_t) W * 4);
        }
    }
}
