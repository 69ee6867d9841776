// Host-side codecs that sit on the per-task latency path: pinned deflate (fixed Huffman, greedy
// single-probe LZ77) for the deterministic PNG writer, and SHA-256 for CIDv0 hashing of large
// outputs. Pure host code; it lives in the same shared library so the Python layer has one binding.
// The Python twin in arbius_b200/codecs/deflate.py implements the identical algorithm bit for bit.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "../include/arbius_b200.h"

namespace {

struct BitWriter {
    uint8_t* out;
    size_t cap, pos = 0;
    uint64_t acc = 0;
    int nbits = 0;
    bool overflow = false;
    BitWriter(uint8_t* o, size_t c) : out(o), cap(c) {}
    inline void put(uint32_t v, int n) {   // LSB-first
        acc |= static_cast<uint64_t>(v) << nbits;
        nbits += n;
        while (nbits >= 8) {
            if (pos < cap) out[pos] = static_cast<uint8_t>(acc); else overflow = true;
            ++pos; acc >>= 8; nbits -= 8;
        }
    }
    inline void flush() { if (nbits > 0) { if (pos < cap) out[pos] = static_cast<uint8_t>(acc); else overflow = true; ++pos; acc = 0; nbits = 0; } }
};

inline uint32_t rev(uint32_t v, int n) { uint32_t r = 0; for (int i = 0; i < n; ++i) { r = (r << 1) | (v & 1); v >>= 1; } return r; }

inline void put_litlen(BitWriter& w, int sym) {   // fixed Huffman, RFC 1951 3.2.6 (codes are MSB-first)
    if (sym < 144) w.put(rev(0x30 + sym, 8), 8);
    else if (sym < 256) w.put(rev(0x190 + sym - 144, 9), 9);
    else if (sym < 280) w.put(rev(sym - 256, 7), 7);
    else w.put(rev(0xC0 + sym - 280, 8), 8);
}

const int kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const int kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const int kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const int kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

}  // namespace

// Returns the number of bytes written, or (size_t)-1 if `cap` was too small. Worst case output is
// n * 9 / 8 + 16 bytes.
extern "C" size_t ab_deflate_fixed(const uint8_t* data, size_t n, uint8_t* out, size_t cap) {
    BitWriter w(out, cap);
    w.put(1, 1);   // BFINAL
    w.put(1, 2);   // BTYPE = 01 fixed Huffman
    std::vector<int64_t> table(1 << 15, -1);
    size_t i = 0;
    while (i < n) {
        size_t mlen = 0, mdist = 0;
        if (i + 2 < n) {
            const uint32_t key = (static_cast<uint32_t>(data[i]) << 16) | (static_cast<uint32_t>(data[i + 1]) << 8) | data[i + 2];
            const uint32_t h = (key * 2654435761u) >> 17;
            const int64_t cand = table[h];
            table[h] = static_cast<int64_t>(i);
            if (cand >= 0 && i - static_cast<size_t>(cand) <= 32768 && data[cand] == data[i] &&
                data[cand + 1] == data[i + 1] && data[cand + 2] == data[i + 2]) {
                const size_t maxl = n - i < 258 ? n - i : 258;
                size_t l = 3;
                while (l < maxl && data[cand + l] == data[i + l]) ++l;
                mlen = l; mdist = i - static_cast<size_t>(cand);
            }
        }
        if (mlen >= 3) {
            int li = 28;
            while (kLenBase[li] > static_cast<int>(mlen)) --li;
            put_litlen(w, 257 + li);
            if (kLenExtra[li]) w.put(static_cast<uint32_t>(mlen - kLenBase[li]), kLenExtra[li]);
            int di = 29;
            while (kDistBase[di] > static_cast<int>(mdist)) --di;
            w.put(rev(di, 5), 5);
            if (kDistExtra[di]) w.put(static_cast<uint32_t>(mdist - kDistBase[di]), kDistExtra[di]);
            i += mlen;
        } else {
            put_litlen(w, data[i]);
            ++i;
        }
    }
    put_litlen(w, 256);
    w.flush();
    return w.overflow ? static_cast<size_t>(-1) : w.pos;
}

// ---- SHA-256 (FIPS 180-4) -------------------------------------------------------------------------
namespace {
const uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01,
    0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc,
    0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147,
    0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08,
    0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208,
    0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
inline uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
void sha256_block(uint32_t* s, const uint8_t* p) {
    uint32_t w[64];
    for (int i = 0; i < 16; ++i) w[i] = (uint32_t(p[4 * i]) << 24) | (uint32_t(p[4 * i + 1]) << 16) | (uint32_t(p[4 * i + 2]) << 8) | p[4 * i + 3];
    for (int i = 16; i < 64; ++i) {
        const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3);
        const uint32_t s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
        w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4], f = s[5], g = s[6], h = s[7];
    for (int i = 0; i < 64; ++i) {
        const uint32_t S1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25);
        const uint32_t t1 = h + S1 + ((e & f) ^ (~e & g)) + K256[i] + w[i];
        const uint32_t S0 = rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22);
        const uint32_t t2 = S0 + ((a & b) ^ (a & c) ^ (b & c));
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e; s[5] += f; s[6] += g; s[7] += h;
}
}  // namespace

extern "C" void ab_sha256(const uint8_t* data, size_t n, uint8_t* digest32) {
    uint32_t s[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    size_t i = 0;
    for (; i + 64 <= n; i += 64) sha256_block(s, data + i);
    uint8_t tail[128] = {0};
    const size_t rem = n - i;
    memcpy(tail, data + i, rem);
    tail[rem] = 0x80;
    const size_t tl = rem + 9 <= 64 ? 64 : 128;
    const uint64_t bits = static_cast<uint64_t>(n) * 8;
    for (int k = 0; k < 8; ++k) tail[tl - 1 - k] = static_cast<uint8_t>(bits >> (8 * k));
    sha256_block(s, tail);
    if (tl == 128) sha256_block(s, tail + 64);
    for (int k = 0; k < 8; ++k) { digest32[4 * k] = s[k] >> 24; digest32[4 * k + 1] = s[k] >> 16; digest32[4 * k + 2] = s[k] >> 8; digest32[4 * k + 3] = s[k]; }
}
