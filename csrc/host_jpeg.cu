// Deterministic baseline JPEG encoder (integer-only) for the MJPEG-in-MP4 video writer.
// The solution CID of a video task covers the MP4 bytes (SURVEY §2.7 "Output file encoding is part of the CID"),
// so nothing here may depend on libjpeg/ffmpeg builds: fixed quantisation tables, the Annex K Huffman tables,
// 4:4:4 sampling, an integer separable DCT with 64-bit accumulators, no restart markers, no optimisation passes.
// Pure host code, part of the same shared library.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "../include/arbius_b200.h"

namespace {

const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
const uint8_t kQLumaBase[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57,
                                69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55,
                                64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100,
                                103, 99};
const uint8_t kQChromaBase[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                                  99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                  99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
// Annex K.3 typical Huffman tables
const uint8_t kDcLumBits[17] = {0, 0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
const uint8_t kDcVals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
const uint8_t kDcChrBits[17] = {0, 0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
const uint8_t kAcLumBits[17] = {0, 0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
const uint8_t kAcLumVals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14,
    0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09,
    0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a,
    0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65,
    0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88,
    0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9,
    0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca,
    0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea,
    0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
const uint8_t kAcChrBits[17] = {0, 0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
const uint8_t kAcChrVals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32,
    0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16,
    0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39,
    0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64,
    0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86,
    0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7,
    0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8,
    0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9,
    0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};

struct Huff { uint16_t code[256]; uint8_t len[256]; };
void build_huff(const uint8_t* bits, const uint8_t* vals, Huff& h) {
    memset(&h, 0, sizeof(h));
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
        for (int i = 0; i < bits[l]; ++i) { h.code[vals[k]] = static_cast<uint16_t>(code); h.len[vals[k]] = static_cast<uint8_t>(l); ++code; ++k; }
        code <<= 1;
    }
}

// cos((2x+1) u pi / 16) * sqrt(2)/2-normalised, scaled by 2^14 (computed once, exact integers below)
const int32_t kC[8][8] = {
    {5793, 5793, 5793, 5793, 5793, 5793, 5793, 5793},
    {8035, 6811, 4551, 1598, -1598, -4551, -6811, -8035},
    {7568, 3135, -3135, -7568, -7568, -3135, 3135, 7568},
    {6811, -1598, -8035, -4551, 4551, 8035, 1598, -6811},
    {5793, -5793, -5793, 5793, 5793, -5793, -5793, 5793},
    {4551, -8035, 1598, 6811, -6811, -1598, 8035, -4551},
    {3135, -7568, 7568, -3135, -3135, 7568, -7568, 3135},
    {1598, -4551, 6811, -8035, 8035, -6811, 4551, -1598}};   // = round(0.5 * c(u) * cos(...) * 2^14), c(0)=1/sqrt2

struct Out {
    std::vector<uint8_t>& v;
    uint32_t acc = 0; int n = 0;
    explicit Out(std::vector<uint8_t>& vv) : v(vv) {}
    void bits(uint32_t code, int len) {
        acc = (acc << len) | (code & ((1u << len) - 1)); n += len;
        while (n >= 8) { uint8_t b = static_cast<uint8_t>(acc >> (n - 8)); v.push_back(b); if (b == 0xFF) v.push_back(0); n -= 8; }
    }
    void flush() { if (n > 0) bits(0x7F, 8 - n); }
};

inline int bitlen(int v) { int a = v < 0 ? -v : v, l = 0; while (a) { ++l; a >>= 1; } return l; }

void encode_block(Out& o, const int16_t* q, int& pred, const Huff& dc, const Huff& ac) {
    const int diff = q[0] - pred; pred = q[0];
    int s = bitlen(diff);
    o.bits(dc.code[s], dc.len[s]);
    if (s) o.bits(static_cast<uint32_t>(diff < 0 ? diff - 1 : diff), s);
    int run = 0;
    for (int k = 1; k < 64; ++k) {
        const int v = q[kZigzag[k]];
        if (v == 0) { ++run; continue; }
        while (run > 15) { o.bits(ac.code[0xF0], ac.len[0xF0]); run -= 16; }
        s = bitlen(v);
        const int sym = (run << 4) | s;
        o.bits(ac.code[sym], ac.len[sym]);
        o.bits(static_cast<uint32_t>(v < 0 ? v - 1 : v), s);
        run = 0;
    }
    if (run) o.bits(ac.code[0], ac.len[0]);
}

void put16(std::vector<uint8_t>& v, int x) { v.push_back(static_cast<uint8_t>(x >> 8)); v.push_back(static_cast<uint8_t>(x)); }

}  // namespace

// rgb: H x W x 3 uint8. quality 1..100 (libjpeg scaling rule). Returns bytes written or (size_t)-1 if cap too small.
extern "C" size_t ab_jpeg_encode(const uint8_t* rgb, int w, int h, int quality, uint8_t* out, size_t cap) {
    if (!rgb || w <= 0 || h <= 0) return static_cast<size_t>(-1);
    quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
    const int scale = quality < 50 ? 5000 / quality : 200 - 2 * quality;
    uint8_t ql[64], qc[64];
    for (int i = 0; i < 64; ++i) {
        int a = (kQLumaBase[i] * scale + 50) / 100, b = (kQChromaBase[i] * scale + 50) / 100;
        ql[i] = static_cast<uint8_t>(a < 1 ? 1 : (a > 255 ? 255 : a));
        qc[i] = static_cast<uint8_t>(b < 1 ? 1 : (b > 255 ? 255 : b));
    }
    static Huff dcl, dcc, acl, acc;
    static bool built = false;
    if (!built) { build_huff(kDcLumBits, kDcVals, dcl); build_huff(kDcChrBits, kDcVals, dcc); build_huff(kAcLumBits, kAcLumVals, acl); build_huff(kAcChrBits, kAcChrVals, acc); built = true; }
    std::vector<uint8_t> v;
    v.reserve(static_cast<size_t>(w) * h / 2 + 1024);
    // SOI, JFIF APP0
    const uint8_t app0[] = {0xFF, 0xD8, 0xFF, 0xE0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
    v.insert(v.end(), app0, app0 + sizeof(app0));
    for (int t = 0; t < 2; ++t) {           // DQT (zigzag order)
        v.push_back(0xFF); v.push_back(0xDB); put16(v, 67); v.push_back(static_cast<uint8_t>(t));
        for (int i = 0; i < 64; ++i) v.push_back((t ? qc : ql)[kZigzag[i]]);
    }
    v.push_back(0xFF); v.push_back(0xC0); put16(v, 17); v.push_back(8); put16(v, h); put16(v, w); v.push_back(3);
    for (int c = 0; c < 3; ++c) { v.push_back(static_cast<uint8_t>(c + 1)); v.push_back(0x11); v.push_back(c ? 1 : 0); }
    struct { const uint8_t* bits; const uint8_t* vals; int n; int cls; } hts[4] = {
        {kDcLumBits, kDcVals, 12, 0x00}, {kAcLumBits, kAcLumVals, 162, 0x10}, {kDcChrBits, kDcVals, 12, 0x01}, {kAcChrBits, kAcChrVals, 162, 0x11}};
    for (auto& t : hts) {
        v.push_back(0xFF); v.push_back(0xC4); put16(v, 19 + t.n); v.push_back(static_cast<uint8_t>(t.cls));
        for (int i = 1; i <= 16; ++i) v.push_back(t.bits[i]);
        for (int i = 0; i < t.n; ++i) v.push_back(t.vals[i]);
    }
    const uint8_t sos[] = {0xFF, 0xDA, 0, 12, 3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0};
    v.insert(v.end(), sos, sos + sizeof(sos));

    Out o(v);
    int pred[3] = {0, 0, 0};
    for (int by = 0; by < h; by += 8) {
        for (int bx = 0; bx < w; bx += 8) {
            int32_t comp[3][64];
            for (int y = 0; y < 8; ++y) {
                const int yy = by + y < h ? by + y : h - 1;
                for (int x = 0; x < 8; ++x) {
                    const int xx = bx + x < w ? bx + x : w - 1;
                    const uint8_t* px = rgb + (static_cast<size_t>(yy) * w + xx) * 3;
                    const int r = px[0], g = px[1], b = px[2];
                    // BT.601 full range, 16-bit fixed point, round to nearest
                    comp[0][y * 8 + x] = ((19595 * r + 38470 * g + 7471 * b + 32768) >> 16) - 128;
                    comp[1][y * 8 + x] = ((-11059 * r - 21709 * g + 32768 * b + 32768) >> 16);
                    comp[2][y * 8 + x] = ((32768 * r - 27439 * g - 5329 * b + 32768) >> 16);
                }
            }
            for (int c = 0; c < 3; ++c) {
                int64_t tmp[64];
                for (int y = 0; y < 8; ++y)
                    for (int u = 0; u < 8; ++u) {
                        int64_t s = 0;
                        for (int x = 0; x < 8; ++x) s += static_cast<int64_t>(kC[u][x]) * comp[c][y * 8 + x];
                        tmp[y * 8 + u] = s;                       // scaled 2^14
                    }
                int16_t q[64];
                const uint8_t* qt = c ? qc : ql;
                for (int vv = 0; vv < 8; ++vv)
                    for (int u = 0; u < 8; ++u) {
                        int64_t s = 0;
                        for (int y = 0; y < 8; ++y) s += static_cast<int64_t>(kC[vv][y]) * tmp[y * 8 + u];
                        // s is the DCT coefficient scaled by 2^28; divide by quant step with round-half-away
                        const int64_t den = static_cast<int64_t>(qt[vv * 8 + u]) << 28;
                        const int64_t half = den >> 1;
                        q[vv * 8 + u] = static_cast<int16_t>(s >= 0 ? (s + half) / den : -((-s + half) / den));
                    }
                encode_block(o, q, pred[c], c ? dcc : dcl, c ? acc : acl);
            }
        }
    }
    o.flush();
    v.push_back(0xFF); v.push_back(0xD9);
    if (v.size() > cap) return static_cast<size_t>(-1);
    memcpy(out, v.data(), v.size());
    return v.size();
}
