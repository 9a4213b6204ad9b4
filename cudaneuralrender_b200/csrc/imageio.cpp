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

namespace {

struct BitReader {
    const uint8_t* p;
    size_t n, pos = 0;
    uint32_t bitbuf = 0;
    int bitcnt = 0;
    BitReader(const uint8_t* d, size_t len) : p(d), n(len) {}
    uint32_t bits(int need) {
        uint32_t v = bitbuf;
        while (bitcnt < need) {
            if (pos >= n) throw std::runtime_error("deflate: out of input");
            v |= (uint32_t)p[pos++] << bitcnt;
            bitcnt += 8;
        }
        bitbuf = need < 32 ? (v >> need) : 0;
        bitcnt -= need;
        return need < 32 ? (v & ((1u << need) - 1u)) : v;
    }
    void align() { bitbuf = 0; bitcnt = 0; }
};

struct Huffman {
    uint16_t count[16];
    uint16_t symbol[288];
    void build(const uint8_t* lengths, int n) {
        std::memset(count, 0, sizeof(count));
        for (int i = 0; i < n; ++i) count[lengths[i]]++;
        count[0] = 0;
        uint16_t offs[16];
        offs[1] = 0;
        for (int l = 1; l < 15; ++l) offs[l + 1] = offs[l] + count[l];
        for (int i = 0; i < n; ++i)
            if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
    }
    int decode(BitReader& br) const {
        int code = 0, first = 0, index = 0;
        for (int len = 1; len <= 15; ++len) {
            code |= (int)br.bits(1);
            const int c = count[len];
            if (code - c < first) return symbol[index + (code - first)];
            index += c;
            first += c;
            first <<= 1;
            code <<= 1;
        }
        throw std::runtime_error("deflate: bad Huffman code");
    }
};

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

void inflate_codes(BitReader& br, const Huffman& lit, const Huffman& dist, std::vector<uint8_t>& out) {
    for (;;) {
        int sym = lit.decode(br);
        if (sym < 256) { out.push_back((uint8_t)sym); continue; }
        if (sym == 256) return;
        sym -= 257;
        if (sym >= 29) throw std::runtime_error("deflate: bad length symbol");
        const int len = kLenBase[sym] + (int)br.bits(kLenExtra[sym]);
        const int ds = dist.decode(br);
        if (ds >= 30) throw std::runtime_error("deflate: bad distance symbol");
        const size_t d = kDistBase[ds] + br.bits(kDistExtra[ds]);
        if (d > out.size()) throw std::runtime_error("deflate: distance too far back");
        const size_t start = out.size() - d;
        for (int i = 0; i < len; ++i) out.push_back(out[start + i]);
    }
}

// zlib stream (RFC 1950) -> bytes
std::vector<uint8_t> zlib_inflate(const uint8_t* data, size_t n, size_t expect) {
    if (n < 6) throw std::runtime_error("zlib: stream too short");
    if ((data[0] & 0x0F) != 8 || ((data[0] << 8) | data[1]) % 31 != 0) throw std::runtime_error("zlib: bad header");
    if (data[1] & 0x20) throw std::runtime_error("zlib: preset dictionary not supported");
    BitReader br(data + 2, n - 2);
    std::vector<uint8_t> out;
    out.reserve(expect);
    int last;
    do {
        last = (int)br.bits(1);
        const int type = (int)br.bits(2);
        if (type == 0) {
            br.align();
            if (br.pos + 4 > br.n) throw std::runtime_error("deflate: truncated stored block");
            const unsigned len = br.p[br.pos] | (br.p[br.pos + 1] << 8);
            const unsigned nlen = br.p[br.pos + 2] | (br.p[br.pos + 3] << 8);
            br.pos += 4;
            if ((len ^ 0xFFFFu) != nlen) throw std::runtime_error("deflate: stored block length mismatch");
            if (br.pos + len > br.n) throw std::runtime_error("deflate: truncated stored block");
            out.insert(out.end(), br.p + br.pos, br.p + br.pos + len);
            br.pos += len;
        } else if (type == 1) {
            uint8_t l[288];
            int i = 0;
            for (; i < 144; ++i) l[i] = 8;
            for (; i < 256; ++i) l[i] = 9;
            for (; i < 280; ++i) l[i] = 7;
            for (; i < 288; ++i) l[i] = 8;
            Huffman lit, dist;
            lit.build(l, 288);
            uint8_t dl[30];
            for (i = 0; i < 30; ++i) dl[i] = 5;
            dist.build(dl, 30);
            inflate_codes(br, lit, dist, out);
        } else if (type == 2) {
            const int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
            if (nlen > 286 || ndist > 30) throw std::runtime_error("deflate: bad code counts");
            static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            uint8_t cl[19] = {0};
            for (int i = 0; i < ncode; ++i) cl[order[i]] = (uint8_t)br.bits(3);
            Huffman clh;
            clh.build(cl, 19);
            uint8_t lengths[320] = {0};
            int idx = 0;
            while (idx < nlen + ndist) {
                const int sym = clh.decode(br);
                if (sym < 16) { lengths[idx++] = (uint8_t)sym; continue; }
                int rep, val = 0;
                if (sym == 16) {
                    if (idx == 0) throw std::runtime_error("deflate: repeat without previous length");
                    val = lengths[idx - 1];
                    rep = 3 + (int)br.bits(2);
                } else if (sym == 17) rep = 3 + (int)br.bits(3);
                else rep = 11 + (int)br.bits(7);
                if (idx + rep > nlen + ndist) throw std::runtime_error("deflate: too many code lengths");
                while (rep--) lengths[idx++] = (uint8_t)val;
            }
            Huffman lit, dist;
            lit.build(lengths, nlen);
            dist.build(lengths + nlen, ndist);
            inflate_codes(br, lit, dist, out);
        } else {
            throw std::runtime_error("deflate: reserved block type");
        }
    } while (!last);
    return out;
}

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

}  // extern "C"
