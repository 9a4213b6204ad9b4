// =====================================================================================
// imageio.cpp -- PNG (decode + encode) and PPM output for the drop-in executables.
//
// Replaces the capability the reference gets from the vendored LodePNG
// (Image::loadPNG / Image::savePNG, src/neuralUtils/image.cu:36-110) and from
// sdkSavePPM4ub (src/common/helper_image.h:252-329).  Own implementation written from
// RFC 1950/1951 (zlib/deflate) and the PNG specification: 8- and 16-bit, colour types
// 0/2/3/4/6, non-interlaced (every file in matcaps/ is of that kind).  The encoder emits
// stored (uncompressed) deflate blocks -- valid PNG, no compression.
// =====================================================================================
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/nsdf.h"
#include "../../include/nsdf_inflate.hpp"

namespace {

using nsdf::zl::zlib_inflate;


uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32(const uint8_t* d, size_t n, uint32_t crc = 0) {
    if (!crc_ready) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            crc_table[i] = c;
        }
        crc_ready = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = crc_table[(crc ^ d[i]) & 0xFF] ^ (crc >> 8);
    return ~crc;
}

uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
void put_be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}

int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool read_file(const char* path, std::vector<uint8_t>& out) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const size_t got = out.empty() ? 0 : std::fread(out.data(), 1, out.size(), f);
    std::fclose(f);
    return got == out.size();
}

// Decodes to RGBA8 rows.  Throws on anything unsupported.
void png_decode(const std::vector<uint8_t>& file, std::vector<uint8_t>& rgba, int& w, int& h) {
    static const uint8_t magic[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    if (file.size() < 33 || std::memcmp(file.data(), magic, 8) != 0) throw std::runtime_error("png: bad signature");
    size_t pos = 8;
    int depth = 0, ctype = 0, interlace = 0;
    bool have_ihdr = false;
    std::vector<uint8_t> idat, plte, trns;
    while (pos + 12 <= file.size()) {
        const uint32_t len = be32(&file[pos]);
        const char* type = (const char*)&file[pos + 4];
        if (len > file.size() - pos - 12) throw std::runtime_error("png: truncated chunk");
        const uint8_t* body = &file[pos + 8];
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len < 13) throw std::runtime_error("png: bad IHDR");
            w = (int)be32(body); h = (int)be32(body + 4);
            depth = body[8]; ctype = body[9]; interlace = body[12];
            have_ihdr = true;
        } else if (!std::memcmp(type, "PLTE", 4)) plte.assign(body, body + len);
        else if (!std::memcmp(type, "tRNS", 4)) trns.assign(body, body + len);
        else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!std::memcmp(type, "IEND", 4)) break;
        pos += 12 + len;
    }
    if (!have_ihdr || w <= 0 || h <= 0 || w > 32768 || h > 32768) throw std::runtime_error("png: bad dimensions");
    if (interlace) throw std::runtime_error("png: interlaced images are not supported");
    int channels;
    switch (ctype) {
        case 0: channels = 1; break;
        case 2: channels = 3; break;
        case 3: channels = 1; break;
        case 4: channels = 2; break;
        case 6: channels = 4; break;
        default: throw std::runtime_error("png: bad colour type");
    }
    if (!(depth == 8 || depth == 16 || (depth < 8 && (ctype == 0 || ctype == 3) && (depth == 1 || depth == 2 || depth == 4))))
        throw std::runtime_error("png: unsupported bit depth");
    const size_t bpp_bits = (size_t)channels * depth;
    const size_t stride = ((size_t)w * bpp_bits + 7) / 8;
    const size_t bpp = bpp_bits >= 8 ? bpp_bits / 8 : 1;
    std::vector<uint8_t> raw = zlib_inflate(idat.data(), idat.size(), (stride + 1) * h);
    if (raw.size() < (stride + 1) * (size_t)h) throw std::runtime_error("png: not enough image data");
    std::vector<uint8_t> prev(stride, 0), cur(stride);
    rgba.assign((size_t)w * h * 4, 255);
    for (int y = 0; y < h; ++y) {
        const uint8_t* src = &raw[(stride + 1) * y];
        const int ft = src[0];
        ++src;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int v = src[i];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: throw std::runtime_error("png: bad filter type");
            }
            cur[i] = (uint8_t)v;
        }
        uint8_t* dst = &rgba[(size_t)y * w * 4];
        for (int x = 0; x < w; ++x) {
            auto sample = [&](int ch) -> int {   // 8-bit value of channel ch at pixel x
                if (depth == 8) return cur[(size_t)x * channels + ch];
                if (depth == 16) return cur[((size_t)x * channels + ch) * 2];   // high byte (lodepng's 16->8 rule)
                const int per = 8 / depth, idx = x, byte = cur[idx / per];
                const int shift = 8 - depth * (idx % per + 1);
                return (byte >> shift) & ((1 << depth) - 1);
            };
            uint8_t* o = dst + 4 * x;
            if (ctype == 0) {
                int g = sample(0);
                if (depth < 8) g = g * 255 / ((1 << depth) - 1);
                o[0] = o[1] = o[2] = (uint8_t)g;
            } else if (ctype == 2) {
                o[0] = (uint8_t)sample(0); o[1] = (uint8_t)sample(1); o[2] = (uint8_t)sample(2);
            } else if (ctype == 3) {
                const size_t i = (size_t)sample(0);
                if (3 * i + 2 >= plte.size() + 0 && 3 * i + 2 >= plte.size()) throw std::runtime_error("png: palette index out of range");
                o[0] = plte[3 * i]; o[1] = plte[3 * i + 1]; o[2] = plte[3 * i + 2];
                if (i < trns.size()) o[3] = trns[i];
            } else if (ctype == 4) {
                o[0] = o[1] = o[2] = (uint8_t)sample(0); o[3] = (uint8_t)sample(1);
            } else {
                o[0] = (uint8_t)sample(0); o[1] = (uint8_t)sample(1); o[2] = (uint8_t)sample(2); o[3] = (uint8_t)sample(3);
            }
        }
        prev.swap(cur);
    }
}

void chunk(std::vector<uint8_t>& out, const char* type, const std::vector<uint8_t>& body) {
    put_be32(out, (uint32_t)body.size());
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), body.begin(), body.end());
    put_be32(out, crc32(&out[start], out.size() - start));
}

}  // namespace

extern "C" {

// Image::loadPNG, src/neuralUtils/image.cu:36-65 (decode to RGBA8, pack a<<24|b<<16|g<<8|r, :58).
int nsdf_png_load(const char* path, uint32_t* texels, size_t capacity, int* w, int* h) {
    if (!path || !w || !h) return NSDF_EINVAL;
    std::vector<uint8_t> file;
    if (!read_file(path, file)) return NSDF_EIO;
    std::vector<uint8_t> rgba;
    try {
        png_decode(file, rgba, *w, *h);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "nsdf_png_load(%s): %s\n", path, e.what());
        return NSDF_EIO;
    }
    if (!texels) return NSDF_OK;
    const size_t n = (size_t)*w * *h;
    if (capacity < n) return NSDF_EINVAL;
    for (size_t i = 0; i < n; ++i)
        texels[i] = ((uint32_t)rgba[4 * i + 3] << 24) | ((uint32_t)rgba[4 * i + 2] << 16) | ((uint32_t)rgba[4 * i + 1] << 8) |
                    rgba[4 * i];
    return NSDF_OK;
}

// Image::savePNG, src/neuralUtils/image.cu:67-110.  The reference's "flip" walks the RGBA byte
// stream backwards texel by texel (:84-98): output texel i = input texel n-1-i (180-degree turn),
// channel order kept.
int nsdf_png_save(const char* path, const uint32_t* texels, int w, int h, int flip) {
    if (!path || !texels || w <= 0 || h <= 0) return NSDF_EINVAL;
    const size_t n = (size_t)w * h, stride = (size_t)w * 4 + 1;
    std::vector<uint8_t> raw(stride * h);
    for (int y = 0; y < h; ++y) {
        uint8_t* row = &raw[stride * y];
        row[0] = 0;
        for (int x = 0; x < w; ++x) {
            const size_t i = (size_t)y * w + x;
            const uint32_t t = texels[flip ? n - 1 - i : i];
            row[1 + 4 * x] = (uint8_t)t; row[2 + 4 * x] = (uint8_t)(t >> 8);
            row[3 + 4 * x] = (uint8_t)(t >> 16); row[4 + 4 * x] = (uint8_t)(t >> 24);
        }
    }
    std::vector<uint8_t> z;
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    for (size_t pos = 0; pos < raw.size();) {
        const size_t len = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + len == raw.size() ? 1 : 0);
        z.push_back((uint8_t)len); z.push_back((uint8_t)(len >> 8));
        z.push_back((uint8_t)~len); z.push_back((uint8_t)(~len >> 8));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + len);
        pos += len;
    }
    for (size_t i = 0; i < raw.size();) {
        const size_t blk = std::min<size_t>(5552, raw.size() - i);
        for (size_t k = 0; k < blk; ++k) { a += raw[i + k]; b += a; }
        a %= 65521u; b %= 65521u;
        i += blk;
    }
    put_be32(z, (b << 16) | a);
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, (uint32_t)w); put_be32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk(out, "IHDR", ihdr);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", {});
    FILE* f = std::fopen(path, "wb");
    if (!f) return NSDF_EIO;
    const size_t wr = std::fwrite(out.data(), 1, out.size(), f);
    std::fclose(f);
    return wr == out.size() ? NSDF_OK : NSDF_EIO;
}

// __savePPM via sdkSavePPM4ub (src/common/helper_image.h:252-329): header "P6\n<W>\n<H>\n255\n",
// then RGB triplets (alpha dropped).
int nsdf_ppm_save(const char* path, const uint32_t* texels, int w, int h) {
    if (!path || !texels || w <= 0 || h <= 0) return NSDF_EINVAL;
    FILE* f = std::fopen(path, "wb");
    if (!f) return NSDF_EIO;
    std::fprintf(f, "P6\n%d\n%d\n255\n", w, h);
    std::vector<uint8_t> row((size_t)w * 3);
    bool ok = true;
    for (int y = 0; y < h && ok; ++y) {
        for (int x = 0; x < w; ++x) {
            const uint32_t t = texels[(size_t)y * w + x];
            row[3 * x] = (uint8_t)t; row[3 * x + 1] = (uint8_t)(t >> 8); row[3 * x + 2] = (uint8_t)(t >> 16);
        }
        ok = std::fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    std::fclose(f);
    return ok ? NSDF_OK : NSDF_EIO;
}

int nsdf_zlib_inflate(const unsigned char* in, size_t n, unsigned char* out, size_t capacity, size_t* n_out) {
    if (!in || (!out && capacity)) return NSDF_EINVAL;
    try {
        const std::vector<uint8_t> v = zlib_inflate(in, n, capacity);
        if (n_out) *n_out = v.size();
        if (v.size() > capacity) return NSDF_EINVAL;
        if (!v.empty()) std::memcpy(out, v.data(), v.size());
        return NSDF_OK;
    } catch (const std::exception&) {
        return NSDF_EIO;
    }
}

}  // extern "C"
