// =====================================================================================
// nsdf_inflate.hpp -- zlib / DEFLATE decoder (RFC 1950 / 1951), header-only.
//
// Shared by the PNG reader (csrc/imageio.cpp: the matcaps the reference loads through lodepng,
// src/neuralUtils/image.cu:36-65) and the HDF5 reader (nsdf_h5.hpp: datasets stored with the deflate filter, which
// the reference reads through libhdf5, src/neuralNetwork.cpp:85-151).  Stored, fixed-Huffman and dynamic-Huffman
// blocks; no preset dictionary.  Throws std::runtime_error on malformed input.
// =====================================================================================
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace nsdf {
namespace zl {

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

static const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

inline void inflate_codes(BitReader& br, const Huffman& lit, const Huffman& dist, std::vector<uint8_t>& out) {
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
inline std::vector<uint8_t> zlib_inflate(const uint8_t* data, size_t n, size_t expect) {
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

}  // namespace zl
}  // namespace nsdf
