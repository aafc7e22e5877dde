// image_write.hpp — BMP and baseline-JPEG writers for the host front end.
// The reference writes its result through stb_image_write v1.15 (src/Renderer.cpp:256-257:
// stbi_write_jpg(..., quality 60) and stbi_write_bmp; stb/stb_image_write.h:491-498, 1440-1560).  stb is a
// third-party tree that is not part of this repository, so the front end carries its own encoders — written to make
// the SAME FILES, byte for byte, as that stb version does for 3-component input (tests/test_host_frontend.py compares
// them with the files the reference's own stb object writes):
//   BMP   24 bpp, bottom-up, rows padded to 4 bytes, biSizeImage and the resolution fields left 0 as stb leaves them;
//   JPEG  baseline sequential (ITU-T T.81), Annex K tables scaled by the IJG quality rule, ONE DQT and ONE DHT segment
//         holding all tables, 2x2 chroma subsampling (16x16 MCUs, box filter) at quality <= 90 and 4:4:4 above,
//         the Arai-Agui-Nakajima float DCT with the scale factors folded into the quantiser, round-half-away-from-zero
//         by truncation of v +- 0.5, edge replication for partial blocks, one-padding of the last byte.
// Single-precision operation order is part of the contract (build with -ffp-contract=off, as the Makefile does).
// Input is u8 RGB, row-major, top row first.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace imgw {

inline bool write_file(const std::string& path, const std::vector<uint8_t>& bytes) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    std::fclose(f);
    return ok;
}

// ---- BMP: BITMAPFILEHEADER + BITMAPINFOHEADER, 24 bpp, bottom-up, rows padded to 4 bytes
inline std::vector<uint8_t> encode_bmp(const uint8_t* rgb, int w, int h) {
    const uint32_t stride = (uint32_t)((w * 3 + 3) & ~3);
    const uint32_t data = stride * (uint32_t)h, off = 14 + 40;
    std::vector<uint8_t> out(off + data, 0);
    auto put16 = [&](size_t p, uint32_t v) { out[p] = v & 255; out[p + 1] = (v >> 8) & 255; };
    auto put32 = [&](size_t p, uint32_t v) { put16(p, v & 0xFFFF); put16(p + 2, v >> 16); };
    out[0] = 'B'; out[1] = 'M';
    put32(2, off + data); put32(10, off);
    put32(14, 40); put32(18, (uint32_t)w); put32(22, (uint32_t)h);
    put16(26, 1); put16(28, 24);   // biSizeImage (offset 34) stays 0: allowed for BI_RGB, and what stb writes
    for (int y = 0; y < h; ++y) {
        uint8_t* row = out.data() + off + (size_t)(h - 1 - y) * stride;
        const uint8_t* src = rgb + (size_t)y * w * 3;
        for (int x = 0; x < w; ++x) {  // BGR
            row[3 * x] = src[3 * x + 2]; row[3 * x + 1] = src[3 * x + 1]; row[3 * x + 2] = src[3 * x];
        }
    }
    return out;
}

// ---- JPEG ------------------------------------------------------------------------
namespace detail {

static const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22,
                                    15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
// T.81 Annex K.1, natural (row-major) order
static const uint8_t kQLuma[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57,
                                   69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55,
                                   64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const uint8_t kQChroma[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                                     99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                     99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
// T.81 Annex K.3: BITS (codes per length 1..16) and HUFFVAL
static const uint8_t kDcLumaBits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const uint8_t kDcChromaBits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const uint8_t kDcVals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t kAcLumaBits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const uint8_t kAcLumaVals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32,
    0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16,
    0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45,
    0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69,
    0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94,
    0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6,
    0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8,
    0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};
static const uint8_t kAcChromaBits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const uint8_t kAcChromaVals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81,
    0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34,
    0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44,
    0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68,
    0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92,
    0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4,
    0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6,
    0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};

struct HuffTable {
    uint16_t code[256];
    uint8_t len[256];
    HuffTable(const uint8_t* bits, const uint8_t* vals) {
        for (int i = 0; i < 256; ++i) { code[i] = 0; len[i] = 0; }
        uint16_t c = 0;
        int k = 0;
        for (int l = 1; l <= 16; ++l) {          // T.81 Annex C: canonical code assignment
            for (int i = 0; i < bits[l - 1]; ++i) { code[vals[k]] = c++; len[vals[k]] = (uint8_t)l; ++k; }
            c <<= 1;
        }
    }
};

struct BitWriter {
    std::vector<uint8_t>& out;
    uint32_t acc = 0;
    int n = 0;
    explicit BitWriter(std::vector<uint8_t>& o) : out(o) {}
    void put(uint32_t bits, int len) {
        acc = (acc << len) | (bits & ((1u << len) - 1u));
        n += len;
        while (n >= 8) {
            const uint8_t b = (uint8_t)(acc >> (n - 8));
            out.push_back(b);
            if (b == 0xFF) out.push_back(0);     // byte stuffing
            n -= 8;
        }
    }
    void flush() { if (n > 0) put(0x7F, 8 - n); }  // pad with ones
};

// One 8-point Arai-Agui-Nakajima forward DCT (scaled: the missing factors live in the quantiser), in place on
// d[0], d[st], ..., d[7 st].  Five multiplies; the operation order below is what fixes the bytes.
inline void aan_dct8(float* d, int st) {
    const float a0 = d[0], a1 = d[st], a2 = d[2 * st], a3 = d[3 * st], a4 = d[4 * st], a5 = d[5 * st], a6 = d[6 * st], a7 = d[7 * st];
    const float s07 = a0 + a7, m07 = a0 - a7, s16 = a1 + a6, m16 = a1 - a6, s25 = a2 + a5, m25 = a2 - a5, s34 = a3 + a4, m34 = a3 - a4;
    // even half
    const float e0 = s07 + s34, e3 = s07 - s34, e1 = s16 + s25, e2 = s16 - s25;
    d[0] = e0 + e1;
    d[4 * st] = e0 - e1;
    const float r = (e2 + e3) * 0.707106781f;
    d[2 * st] = e3 + r;
    d[6 * st] = e3 - r;
    // odd half (rotator without extra negations)
    const float o0 = m34 + m25, o1 = m25 + m16, o2 = m16 + m07;
    const float z5 = (o0 - o2) * 0.382683433f;
    const float z2 = o0 * 0.541196100f + z5;
    const float z4 = o2 * 1.306562965f + z5;
    const float z3 = o1 * 0.707106781f;
    const float p = m07 + z3, m = m07 - z3;
    d[5 * st] = m + z2;
    d[3 * st] = m - z2;
    d[st] = p + z4;
    d[7 * st] = p - z4;
}

}  // namespace detail

inline std::vector<uint8_t> encode_jpeg(const uint8_t* rgb, int w, int h, int quality) {
    using namespace detail;
    quality = quality ? quality : 90;
    const bool subsample = quality <= 90;
    quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
    const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;  // IJG quality scaling
    uint8_t qz[2][64];        // quantisation tables in ZIGZAG order (as the DQT segment stores them)
    for (int z = 0; z < 64; ++z) {
        const int i = kZigzag[z];
        const int l = (kQLuma[i] * scale + 50) / 100, c = (kQChroma[i] * scale + 50) / 100;
        qz[0][z] = (uint8_t)(l < 1 ? 1 : (l > 255 ? 255 : l));
        qz[1][z] = (uint8_t)(c < 1 ? 1 : (c > 255 ? 255 : c));
    }
    // natural index -> zigzag position
    uint8_t zpos[64];
    for (int z = 0; z < 64; ++z) zpos[kZigzag[z]] = (uint8_t)z;
    // descale factors of the AAN DCT folded into the quantiser: 1 / (q * s[row] * s[col])
    static const float aan[8] = {1.0f * 2.828427125f, 1.387039845f * 2.828427125f, 1.306562965f * 2.828427125f, 1.175875602f * 2.828427125f,
                                 1.0f * 2.828427125f, 0.785694958f * 2.828427125f, 0.541196100f * 2.828427125f, 0.275899379f * 2.828427125f};
    float fq[2][64];
    for (int row = 0, k = 0; row < 8; ++row)
        for (int col = 0; col < 8; ++col, ++k)
            for (int t = 0; t < 2; ++t) fq[t][k] = 1 / (qz[t][zpos[k]] * aan[row] * aan[col]);

    std::vector<uint8_t> out;
    out.reserve((size_t)w * h / 4 + 1024);
    auto b8 = [&](int v) { out.push_back((uint8_t)v); };
    auto b16 = [&](int v) { b8(v >> 8); b8(v & 255); };
    b16(0xFFD8);                                                        // SOI
    b16(0xFFE0); b16(16); for (char ch : {'J', 'F', 'I', 'F', '\0'}) b8(ch);
    b16(0x0101); b8(0); b16(1); b16(1); b8(0); b8(0);                  // JFIF 1.01, aspect 1:1
    b16(0xFFDB); b16(2 + 65 + 65);                                      // ONE DQT segment, both tables
    for (int t = 0; t < 2; ++t) { b8(t); for (int z = 0; z < 64; ++z) b8(qz[t][z]); }
    b16(0xFFC0); b16(17); b8(8); b16(h); b16(w); b8(3);                // SOF0
    b8(1); b8(subsample ? 0x22 : 0x11); b8(0); b8(2); b8(0x11); b8(1); b8(3); b8(0x11); b8(1);
    b16(0xFFC4); b16(2 + (17 + 12) + (17 + 162) + (17 + 12) + (17 + 162));   // ONE DHT segment, four tables
    auto dht = [&](int cls_id, const uint8_t* bits, const uint8_t* vals, int nvals) {
        b8(cls_id);
        for (int i = 0; i < 16; ++i) b8(bits[i]);
        for (int i = 0; i < nvals; ++i) b8(vals[i]);
    };
    dht(0x00, kDcLumaBits, kDcVals, 12);
    dht(0x10, kAcLumaBits, kAcLumaVals, 162);
    dht(0x01, kDcChromaBits, kDcVals, 12);
    dht(0x11, kAcChromaBits, kAcChromaVals, 162);
    b16(0xFFDA); b16(12); b8(3); b8(1); b8(0x00); b8(2); b8(0x11); b8(3); b8(0x11); b8(0); b8(63); b8(0);  // SOS

    const HuffTable dcL(kDcLumaBits, kDcVals), acL(kAcLumaBits, kAcLumaVals);
    const HuffTable dcC(kDcChromaBits, kDcVals), acC(kAcChromaBits, kAcChromaVals);
    BitWriter bw(out);

    // one 8x8 data unit inside `du` (row stride `st`): DCT, quantise, entropy-code; returns the new DC predictor
    auto code_unit = [&](float* du, int st, const float* fqt, int pred, const HuffTable& dc, const HuffTable& ac) -> int {
        for (int y = 0; y < 8; ++y) aan_dct8(du + y * st, 1);          // rows
        for (int x = 0; x < 8; ++x) aan_dct8(du + x, st);              // columns
        int zz[64];
        for (int y = 0, j = 0; y < 8; ++y)
            for (int x = 0; x < 8; ++x, ++j) {
                const float v = du[y * st + x] * fqt[j];
                zz[zpos[j]] = (int)(v < 0 ? v - 0.5f : v + 0.5f);
            }
        auto magnitude = [](int v, int& nbits, uint32_t& bits) {
            int a = v < 0 ? -v : v;
            nbits = 0;
            while (a) { ++nbits; a >>= 1; }
            bits = (uint32_t)(v < 0 ? v + (1 << nbits) - 1 : v);
        };
        int nb; uint32_t bits;
        magnitude(zz[0] - pred, nb, bits);
        bw.put(dc.code[nb], dc.len[nb]);
        if (nb) bw.put(bits, nb);
        int run = 0, last = 63;
        while (last > 0 && zz[last] == 0) --last;
        for (int i = 1; i <= last; ++i) {
            if (zz[i] == 0) { ++run; continue; }
            while (run >= 16) { bw.put(ac.code[0xF0], ac.len[0xF0]); run -= 16; }  // ZRL
            magnitude(zz[i], nb, bits);
            bw.put(ac.code[(run << 4) | nb], ac.len[(run << 4) | nb]);
            bw.put(bits, nb);
            run = 0;
        }
        if (last < 63) bw.put(ac.code[0x00], ac.len[0x00]);                        // EOB
        return zz[0];
    };

    int pY = 0, pU = 0, pV = 0;
    const int B = subsample ? 16 : 8;                                   // MCU edge in pixels
    std::vector<float> Yb(B * B), Ub(B * B), Vb(B * B);
    for (int by = 0; by < h; by += B)
        for (int bx = 0; bx < w; bx += B) {
            for (int y = 0, pos = 0; y < B; ++y) {
                const int sy = (by + y < h) ? by + y : h - 1;          // edge replicate
                for (int x = 0; x < B; ++x, ++pos) {
                    const int sx = (bx + x < w) ? bx + x : w - 1;
                    const uint8_t* p = rgb + ((size_t)sy * w + sx) * 3;
                    const float r = p[0], g = p[1], b = p[2];
                    Yb[pos] = +0.29900f * r + 0.58700f * g + 0.11400f * b - 128;
                    Ub[pos] = -0.16874f * r - 0.33126f * g + 0.50000f * b;
                    Vb[pos] = +0.50000f * r - 0.41869f * g - 0.08131f * b;
                }
            }
            if (subsample) {
                pY = code_unit(Yb.data() + 0, 16, fq[0], pY, dcL, acL);
                pY = code_unit(Yb.data() + 8, 16, fq[0], pY, dcL, acL);
                pY = code_unit(Yb.data() + 128, 16, fq[0], pY, dcL, acL);
                pY = code_unit(Yb.data() + 136, 16, fq[0], pY, dcL, acL);
                float su[64], sv[64];
                for (int yy = 0, pos = 0; yy < 8; ++yy)
                    for (int xx = 0; xx < 8; ++xx, ++pos) {
                        const int j = yy * 32 + xx * 2;
                        su[pos] = (Ub[j] + Ub[j + 1] + Ub[j + 16] + Ub[j + 17]) * 0.25f;
                        sv[pos] = (Vb[j] + Vb[j + 1] + Vb[j + 16] + Vb[j + 17]) * 0.25f;
                    }
                pU = code_unit(su, 8, fq[1], pU, dcC, acC);
                pV = code_unit(sv, 8, fq[1], pV, dcC, acC);
            } else {
                pY = code_unit(Yb.data(), 8, fq[0], pY, dcL, acL);
                pU = code_unit(Ub.data(), 8, fq[1], pU, dcC, acC);
                pV = code_unit(Vb.data(), 8, fq[1], pV, dcC, acC);
            }
        }
    bw.flush();
    out.push_back(0xFF); out.push_back(0xD9);                                               // EOI
    return out;
}

inline bool write_bmp(const std::string& path, const uint8_t* rgb, int w, int h) { return write_file(path, encode_bmp(rgb, w, h)); }
inline bool write_jpg(const std::string& path, const uint8_t* rgb, int w, int h, int quality) {
    return write_file(path, encode_jpeg(rgb, w, h, quality));
}

}  // namespace imgw
