// rt2_ingest_codecs.cpp — the two image codecs the reference's asset path needs beyond PNG / ZIP:
//   * baseline JPEG, what `gltf::import` decodes through the `image` crate for glTF base-colour maps
//     (src/gltf/mod.rs:50-52, src/gltf/image.rs:57-109);
//   * OpenEXR PIZ, the OpenEXR default for photographs — `image::open("…skybox2.exr")`, src/main.rs:63.
// Both are restatements of the published formats (ITU-T T.81 with the JFIF colour transform; OpenEXR's PIZ: bitmap + LUT, 2-D
// Haar-like wavelet, canonical Huffman with run-length symbol).  The reference's decoders live in un-vendored crates
// (image 0.25.5 -> zune-jpeg, exr), so the arithmetic is pinned to the de-facto reference implementations the tests can run here:
// libjpeg-turbo's integer IDCT ("islow"), fancy chroma upsampling and 16-bit fixed-point colour conversion (PIL decodes with it),
// and OpenEXR itself (OpenCV writes the fixtures with it).  PIZ is lossless: bit-exact.  Host code, no CUDA.
#include "rt2_ingest.h"

#include <algorithm>
#include <cstring>

namespace rt2ingest {

// =====================================================================================================================
// OpenEXR PIZ
// =====================================================================================================================
namespace {

const int HUF_ENCBITS = 16, HUF_ENCSIZE = (1 << HUF_ENCBITS) + 1;
const int SHORT_ZEROCODE_RUN = 59, LONG_ZEROCODE_RUN = 63, SHORTEST_LONG_RUN = 2 + LONG_ZEROCODE_RUN - SHORT_ZEROCODE_RUN;

struct BitReader {
    const uint8_t* p; const uint8_t* end; uint64_t c = 0; int lc = 0;
    bool get(int n, uint32_t* out) {                    // MSB first
        while (lc < n) { if (p >= end) return false; c = (c << 8) | *p++; lc += 8; }
        lc -= n;
        *out = (uint32_t)((c >> lc) & ((1ull << n) - 1));
        return true;
    }
};

// ImfHuf: the packed table stores a code length (6 bits) per symbol im..iM with zero runs; codes are canonical, assigned from the
// longest length down (hufCanonicalCodeTable)
bool huf_uncompress(const uint8_t* src, size_t n, uint16_t* raw, size_t n_raw, std::string& err) {
    if (n_raw == 0) return true;
    if (n < 20) { err = "EXR PIZ: truncated Huffman header"; return false; }
    uint32_t im, iM, nbits;
    std::memcpy(&im, src, 4); std::memcpy(&iM, src + 4, 4); std::memcpy(&nbits, src + 12, 4);
    if (im >= (uint32_t)HUF_ENCSIZE || iM >= (uint32_t)HUF_ENCSIZE || im > iM) { err = "EXR PIZ: bad Huffman symbol range"; return false; }
    std::vector<uint8_t> len((size_t)HUF_ENCSIZE, 0);
    BitReader br{src + 20, src + n};
    for (uint32_t s = im; s <= iM; s++) {
        uint32_t l;
        if (!br.get(6, &l)) { err = "EXR PIZ: truncated Huffman table"; return false; }
        if (l == (uint32_t)LONG_ZEROCODE_RUN) {
            uint32_t z;
            if (!br.get(8, &z)) { err = "EXR PIZ: truncated Huffman table"; return false; }
            z += SHORTEST_LONG_RUN;
            if (s + z > iM + 1) { err = "EXR PIZ: bad zero run in the Huffman table"; return false; }
            s += z - 1;
        } else if (l >= (uint32_t)SHORT_ZEROCODE_RUN) {
            const uint32_t z = l - SHORT_ZEROCODE_RUN + 2;
            if (s + z > iM + 1) { err = "EXR PIZ: bad zero run in the Huffman table"; return false; }
            s += z - 1;
        } else len[s] = (uint8_t)l;
    }
    // canonical codes: count per length, first code per length from the longest length down
    uint64_t count[59] = {0}, first[59] = {0};
    for (uint32_t s = im; s <= iM; s++) count[len[s]]++;
    { uint64_t c = 0; for (int l = 58; l > 0; l--) { const uint64_t nc = (c + count[l]) >> 1; first[l] = c; c = nc; } }
    // symbols of each length in increasing symbol order get consecutive codes: sorted table per length for decoding
    std::vector<uint32_t> start(60, 0), sym;
    { uint32_t acc = 0; for (int l = 1; l <= 58; l++) { start[(size_t)l] = acc; acc += (uint32_t)count[l]; } start[59] = acc; sym.resize(acc); }
    { std::vector<uint32_t> fill(start.begin(), start.end()); for (uint32_t s = im; s <= iM; s++) if (len[s]) sym[fill[len[s]]++] = s; }
    // the bit stream follows the table at the next byte boundary
    const uint8_t* data = br.p;
    if ((uint64_t)(src + n - data) * 8 < nbits) { err = "EXR PIZ: truncated Huffman data"; return false; }
    BitReader in{data, src + n};
    const uint32_t rlc = iM;                             // run-length symbol
    size_t out = 0;
    uint64_t bits_left = nbits;
    while (bits_left > 0 && out <= n_raw) {
        uint64_t code = 0; int l = 0; uint32_t s = 0xFFFFFFFFu;
        while (l < 58) {
            uint32_t b;
            if (bits_left == 0 || !in.get(1, &b)) { l = 59; break; }
            bits_left--; code = (code << 1) | b; l++;
            if (count[l] && code >= first[l] && code - first[l] < count[l]) { s = sym[start[(size_t)l] + (uint32_t)(code - first[l])]; break; }
        }
        if (s == 0xFFFFFFFFu) {
            if (out == n_raw) break;                     // trailing pad bits
            err = "EXR PIZ: bad Huffman code"; return false;
        }
        if (s == rlc) {
            uint32_t cs;
            if (bits_left < 8 || !in.get(8, &cs)) { err = "EXR PIZ: truncated run length"; return false; }
            bits_left -= 8;
            if (out == 0 || out + cs > n_raw) { err = "EXR PIZ: bad run length"; return false; }
            const uint16_t v = raw[out - 1];
            for (uint32_t k = 0; k < cs; k++) raw[out++] = v;
        } else {
            if (out >= n_raw) { err = "EXR PIZ: too much Huffman data"; return false; }
            raw[out++] = (uint16_t)s;
        }
    }
    if (out != n_raw) { err = "EXR PIZ: Huffman data decodes to the wrong size"; return false; }
    return true;
}

// ImfWav: inverse of the 2-D wavelet step on a pair (l = average, h = difference)
inline void wdec14(uint16_t l, uint16_t h, uint16_t& a, uint16_t& b) {
    const int16_t ls = (int16_t)l, hs = (int16_t)h;
    const int hi = hs, ai = ls + (hi & 1) + (hi >> 1);
    a = (uint16_t)(int16_t)ai; b = (uint16_t)(int16_t)(ai - hi);
}
inline void wdec16(uint16_t l, uint16_t h, uint16_t& a, uint16_t& b) {
    const int m = l, d = h;
    const int bb = (m - (d >> 1)) & 0xFFFF, aa = (d + bb - 0x8000) & 0xFFFF;
    b = (uint16_t)bb; a = (uint16_t)aa;
}
void wav2_decode(uint16_t* in, int nx, int ox, int ny, int oy, uint16_t mx) {
    const bool w14 = mx < (1 << 14);
    const int n = nx > ny ? ny : nx;
    int p = 1, p2;
    while (p <= n) p <<= 1;
    p >>= 1; p2 = p; p >>= 1;
    while (p >= 1) {
        uint16_t* py = in;
        uint16_t* ey = in + (ptrdiff_t)oy * (ny - p2);
        const int oy1 = oy * p, oy2 = oy * p2, ox1 = ox * p, ox2 = ox * p2;
        uint16_t i00, i01, i10, i11;
        for (; py <= ey; py += oy2) {
            uint16_t* px = py;
            uint16_t* ex = py + (ptrdiff_t)ox * (nx - p2);
            for (; px <= ex; px += ox2) {
                uint16_t *p01 = px + ox1, *p10 = px + oy1, *p11 = p10 + ox1;
                if (w14) { wdec14(*px, *p10, i00, i10); wdec14(*p01, *p11, i01, i11); wdec14(i00, i01, *px, *p01); wdec14(i10, i11, *p10, *p11); }
                else { wdec16(*px, *p10, i00, i10); wdec16(*p01, *p11, i01, i11); wdec16(i00, i01, *px, *p01); wdec16(i10, i11, *p10, *p11); }
            }
            if (nx & p) {
                uint16_t* p10 = px + oy1;
                if (w14) wdec14(*px, *p10, i00, *p10); else wdec16(*px, *p10, i00, *p10);
                *px = i00;
            }
        }
        if (ny & p) {
            uint16_t* px = py;
            uint16_t* ex = py + (ptrdiff_t)ox * (nx - p2);
            for (; px <= ex; px += ox2) {
                uint16_t* p01 = px + ox1;
                if (w14) wdec14(*px, *p01, i00, *p01); else wdec16(*px, *p01, i00, *p01);
                *px = i00;
            }
        }
        p2 = p; p >>= 1;
    }
}

}  // namespace

// One PIZ block (ImfPizCompressor::uncompress): `chan_words[c]` = 16-bit words per pixel of channel c (1 = HALF, 2 = FLOAT / UINT);
// out = the block as it is stored uncompressed (per scan line, channels in file order).
bool exr_piz_decode(const uint8_t* src, size_t n, const std::vector<int>& chan_words, uint32_t width, uint32_t rows, std::vector<uint8_t>& out, std::string& err) {
    size_t words_per_line = 0;
    for (int cw : chan_words) words_per_line += (size_t)cw * width;
    const size_t total = words_per_line * rows;
    std::vector<uint16_t> tmp(total);
    if (n < 4) { err = "EXR PIZ: truncated block"; return false; }
    uint16_t min_nz, max_nz;
    std::memcpy(&min_nz, src, 2); std::memcpy(&max_nz, src + 2, 2);
    size_t pos = 4;
    std::vector<uint8_t> bitmap(8192, 0);
    if (min_nz <= max_nz) {
        const size_t len = (size_t)max_nz - min_nz + 1;
        if (max_nz >= 8192 || pos + len > n) { err = "EXR PIZ: bad bitmap"; return false; }
        std::memcpy(bitmap.data() + min_nz, src + pos, len); pos += len;
    }
    std::vector<uint16_t> lut(65536, 0);
    uint32_t k = 0;
    for (uint32_t i = 0; i < 65536; i++) if (i == 0 || (bitmap[i >> 3] & (1u << (i & 7)))) lut[k++] = (uint16_t)i;
    const uint16_t max_value = (uint16_t)(k - 1);
    if (pos + 4 > n) { err = "EXR PIZ: truncated block"; return false; }
    int32_t hlen; std::memcpy(&hlen, src + pos, 4); pos += 4;
    if (hlen < 0 || pos + (size_t)hlen > n) { err = "EXR PIZ: bad Huffman length"; return false; }
    if (!huf_uncompress(src + pos, (size_t)hlen, tmp.data(), total, err)) return false;
    // wavelet-decode each channel plane (channel-major in the temporary buffer), each 16-bit component of a 32-bit channel on its own
    uint16_t* plane = tmp.data();
    for (int cw : chan_words) {
        for (int j = 0; j < cw; j++) wav2_decode(plane + j, (int)width, cw, (int)rows, (int)width * cw, max_value);
        plane += (size_t)width * rows * cw;
    }
    for (size_t i = 0; i < total; i++) tmp[i] = lut[tmp[i]];
    // channel-major planes -> scan lines with the channels of a line back to back
    out.resize(total * 2);
    uint8_t* o = out.data();
    std::vector<const uint16_t*> row_ptr(chan_words.size());
    { const uint16_t* q = tmp.data(); for (size_t c = 0; c < chan_words.size(); c++) { row_ptr[c] = q; q += (size_t)width * rows * chan_words[c]; } }
    for (uint32_t y = 0; y < rows; y++)
        for (size_t c = 0; c < chan_words.size(); c++) {
            const size_t nw = (size_t)width * chan_words[c];
            std::memcpy(o, row_ptr[c], nw * 2); o += nw * 2; row_ptr[c] += nw;
        }
    return true;
}

// =====================================================================================================================
// baseline JPEG (ITU-T T.81 sequential DCT, Huffman) — arithmetic of libjpeg-turbo's defaults
// =====================================================================================================================
namespace {

const uint8_t JPEG_ZIGZAG[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct JHuff {
    bool present = false;
    uint8_t bits[17] = {0}; uint8_t vals[256] = {0};
    int mincode[17], maxcode[18], valptr[17];
    void build() {                                        // T.81 Annex C / F.2.2.3
        int code = 0, k = 0;
        for (int l = 1; l <= 16; l++) {
            valptr[l] = k; mincode[l] = code;
            code += bits[l]; k += bits[l];
            maxcode[l] = bits[l] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7FFFFFFF;
    }
};

struct JComp { int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0; int pred = 0; int plane_w = 0, plane_h = 0, real_w = 0, real_h = 0; std::vector<uint8_t> plane; };

struct JBits {
    const uint8_t* p; const uint8_t* end; uint32_t buf = 0; int cnt = 0; bool hit_marker = false;
    void fill() {
        while (cnt <= 24) {
            int b = 0;
            if (!hit_marker && p < end) {
                b = *p;
                if (b == 0xFF) {
                    if (p + 1 < end && p[1] == 0x00) p += 2;           // stuffed zero
                    else { hit_marker = true; b = 0; }                 // a marker: feed zeros until the caller deals with it
                } else p++;
            }
            buf |= (uint32_t)b << (24 - cnt); cnt += 8;
        }
    }
    int get1() { if (cnt < 1) fill(); const int b = (int)(buf >> 31); buf <<= 1; cnt--; return b; }
    int get(int n) { if (n == 0) return 0; if (cnt < n) fill(); const int v = (int)(buf >> (32 - n)); buf <<= n; cnt -= n; return v; }
    void reset() { buf = 0; cnt = 0; hit_marker = false; }
};

inline int jhuff_decode(JBits& br, const JHuff& t) {
    int code = br.get1();
    for (int l = 1; l <= 16; l++) {
        if (t.maxcode[l] >= 0 && code <= t.maxcode[l] && code >= t.mincode[l]) return t.vals[t.valptr[l] + code - t.mincode[l]];
        code = (code << 1) | br.get1();
    }
    return -1;
}
inline int jextend(int v, int t) { return v < (1 << (t - 1)) ? v - (1 << t) + 1 : v; }      // T.81 F.2.2.1

// jidctint ("islow"): Loeffler-Ligtenberg-Moschytz, 13-bit constants, 2 extra bits kept between the passes
inline int jdescale(int x, int n) { return (x + (1 << (n - 1))) >> n; }
void jpeg_idct_islow(const int* in /* dequantised, natural order */, uint8_t* out, int stride) {
    const int C_0_298 = 2446, C_0_390 = 3196, C_0_541 = 4433, C_0_765 = 6270, C_0_899 = 7373, C_1_175 = 9633, C_1_501 = 12299,
              C_1_847 = 15137, C_1_961 = 16069, C_2_053 = 16819, C_2_562 = 20995, C_3_072 = 25172;
    int ws[64];
    for (int pass = 0; pass < 2; pass++) {
        for (int i = 0; i < 8; i++) {
            const int* s = pass == 0 ? in + i : ws + i * 8;
            const int st = pass == 0 ? 8 : 1;
            int z2 = s[2 * st], z3 = s[6 * st];
            int z1 = (z2 + z3) * C_0_541;
            int tmp2 = z1 + z3 * (-C_1_847), tmp3 = z1 + z2 * C_0_765;
            z2 = s[0]; z3 = s[4 * st];
            int tmp0 = (z2 + z3) * 8192, tmp1 = (z2 - z3) * 8192;
            const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
            tmp0 = s[7 * st]; tmp1 = s[5 * st]; tmp2 = s[3 * st]; tmp3 = s[1 * st];
            z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; int z4 = tmp1 + tmp3;
            const int z5 = (z3 + z4) * C_1_175;
            tmp0 *= C_0_298; tmp1 *= C_2_053; tmp2 *= C_3_072; tmp3 *= C_1_501;
            z1 *= -C_0_899; z2 *= -C_2_562; z3 *= -C_1_961; z4 *= -C_0_390;
            z3 += z5; z4 += z5;
            tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
            const int o[8] = {tmp10 + tmp3, tmp11 + tmp2, tmp12 + tmp1, tmp13 + tmp0, tmp13 - tmp0, tmp12 - tmp1, tmp11 - tmp2, tmp10 - tmp3};
            if (pass == 0) for (int k = 0; k < 8; k++) ws[k * 8 + i] = jdescale(o[k], 13 - 2);
            else for (int k = 0; k < 8; k++) { const int v = jdescale(o[k], 13 + 2 + 3) + 128; out[i * stride + k] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }
        }
    }
}

}  // namespace

namespace {

// jdsample.c: "fancy" (triangle-filter) chroma upsampling, with the edge handling and rounding biases of libjpeg.
// in: rows x cw samples (stride = in_stride); real_rows / real_cw = the component's true downsampled size (context rows and the
// last column replicate the last REAL row / column).  out: (rows * vs) x (cw * hs).
void jpeg_upsample(const uint8_t* in, int in_stride, int real_cw, int real_rows, int hs, int vs, std::vector<uint8_t>& out, int& out_w, int& out_h) {
    out_w = real_cw * hs; out_h = real_rows * vs;
    out.assign((size_t)out_w * out_h, 0);
    auto row = [&](int r) { r = r < 0 ? 0 : (r >= real_rows ? real_rows - 1 : r); return in + (size_t)r * in_stride; };
    const bool fancy = real_cw > 2;
    for (int r = 0; r < real_rows; r++) {
        const uint8_t* p0 = row(r);
        for (int v = 0; v < vs; v++) {
            uint8_t* o = out.data() + (size_t)(r * vs + v) * out_w;
            if (hs == 1 && vs == 1) { std::memcpy(o, p0, (size_t)real_cw); continue; }
            if (hs == 1 && vs == 2) {                                   // h1v2_fancy_upsample (libjpeg-turbo)
                const uint8_t* p1 = v == 0 ? row(r - 1) : row(r + 1);
                const int bias = v == 0 ? 1 : 2;
                for (int c = 0; c < real_cw; c++) o[c] = (uint8_t)((p0[c] * 3 + p1[c] + bias) >> 2);
                continue;
            }
            if (hs == 2 && vs == 1) {
                if (!fancy) { for (int c = 0; c < real_cw; c++) o[2 * c] = o[2 * c + 1] = p0[c]; continue; }
                o[0] = p0[0]; o[1] = (uint8_t)((p0[0] * 3 + p0[1] + 2) >> 2);
                for (int c = 1; c < real_cw - 1; c++) {
                    const int iv = p0[c] * 3;
                    o[2 * c] = (uint8_t)((iv + p0[c - 1] + 1) >> 2); o[2 * c + 1] = (uint8_t)((iv + p0[c + 1] + 2) >> 2);
                }
                const int c = real_cw - 1;
                o[2 * c] = (uint8_t)((p0[c] * 3 + p0[c - 1] + 1) >> 2); o[2 * c + 1] = p0[c];
                continue;
            }
            if (hs == 2 && vs == 2) {
                if (!fancy) { for (int c = 0; c < real_cw; c++) o[2 * c] = o[2 * c + 1] = p0[c]; continue; }
                const uint8_t* p1 = v == 0 ? row(r - 1) : row(r + 1);
                int thiscol = p0[0] * 3 + p1[0], nextcol = p0[1] * 3 + p1[1], lastcol;
                o[0] = (uint8_t)((thiscol * 4 + 8) >> 4); o[1] = (uint8_t)((thiscol * 3 + nextcol + 7) >> 4);
                lastcol = thiscol; thiscol = nextcol;
                for (int c = 1; c < real_cw - 1; c++) {
                    nextcol = p0[c + 1] * 3 + p1[c + 1];
                    o[2 * c] = (uint8_t)((thiscol * 3 + lastcol + 8) >> 4); o[2 * c + 1] = (uint8_t)((thiscol * 3 + nextcol + 7) >> 4);
                    lastcol = thiscol; thiscol = nextcol;
                }
                const int c = real_cw - 1;
                o[2 * c] = (uint8_t)((thiscol * 3 + lastcol + 8) >> 4); o[2 * c + 1] = (uint8_t)((thiscol * 4 + 7) >> 4);
                continue;
            }
            for (int c = 0; c < out_w; c++) o[c] = p0[c / hs];            // other factors (3, 4): replicate, as libjpeg's int_upsample
        }
    }
}

inline uint8_t jclamp(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

}  // namespace

bool decode_jpeg(const uint8_t* data, size_t size, std::vector<uint8_t>& pixels, uint32_t& w, uint32_t& h, uint32_t& channels, std::string& err) {
    if (size < 4 || data[0] != 0xFF || data[1] != 0xD8) { err = "JPEG: missing SOI"; return false; }
    uint16_t qt[4][64]; bool have_qt[4] = {false, false, false, false};
    JHuff hdc[4], hac[4];
    std::vector<JComp> comps;
    int W = 0, H = 0, hmax = 1, vmax = 1, restart_interval = 0, adobe_transform = -1;
    bool have_sof = false, decoded_any = false;
    size_t pos = 2;
    auto be16 = [&](size_t p) { return (int)((data[p] << 8) | data[p + 1]); };
    while (pos + 4 <= size) {
        if (data[pos] != 0xFF) { pos++; continue; }
        const int m = data[pos + 1];
        if (m == 0xFF) { pos++; continue; }
        pos += 2;
        if (m == 0xD9) break;                                                     // EOI
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;                       // standalone markers
        if (pos + 2 > size) break;
        const int len = be16(pos);
        if (len < 2 || pos + (size_t)len > size) { err = "JPEG: truncated segment"; return false; }
        const uint8_t* seg = data + pos + 2; const int n = len - 2;
        if (m == 0xDB) {                                                          // DQT
            int p = 0;
            while (p < n) {
                const int pq = seg[p] >> 4, tq = seg[p] & 15; p++;
                if (tq > 3 || p + (pq ? 128 : 64) > n) { err = "JPEG: bad DQT"; return false; }
                for (int k = 0; k < 64; k++) { qt[tq][JPEG_ZIGZAG[k]] = pq ? (uint16_t)((seg[p] << 8) | seg[p + 1]) : seg[p]; p += pq ? 2 : 1; }
                have_qt[tq] = true;
            }
        } else if (m == 0xC4) {                                                   // DHT
            int p = 0;
            while (p < n) {
                if (p + 17 > n) { err = "JPEG: bad DHT"; return false; }
                const int tc = seg[p] >> 4, th = seg[p] & 15; p++;
                if (tc > 1 || th > 3) { err = "JPEG: bad DHT"; return false; }
                JHuff& t = tc ? hac[th] : hdc[th];
                int total = 0;
                for (int l = 1; l <= 16; l++) { t.bits[l] = seg[p++]; total += t.bits[l]; }
                if (total > 256 || p + total > n) { err = "JPEG: bad DHT"; return false; }
                std::memcpy(t.vals, seg + p, (size_t)total); p += total;
                t.build(); t.present = true;
            }
        } else if (m == 0xC0 || m == 0xC1) {                                      // SOF0 / SOF1
            if (n < 6 || seg[0] != 8) { err = "JPEG: only 8-bit samples are supported"; return false; }
            H = be16(pos + 3); W = be16(pos + 5);
            const int nc = seg[5];
            if (W <= 0 || H <= 0 || (nc != 1 && nc != 3) || n < 6 + 3 * nc) { err = "JPEG: unsupported frame (need 1 or 3 components; CMYK is not supported)"; return false; }
            comps.assign((size_t)nc, JComp());
            for (int c = 0; c < nc; c++) {
                comps[(size_t)c].id = seg[6 + 3 * c]; comps[(size_t)c].h = seg[7 + 3 * c] >> 4; comps[(size_t)c].v = seg[7 + 3 * c] & 15; comps[(size_t)c].tq = seg[8 + 3 * c];
                if (comps[(size_t)c].h < 1 || comps[(size_t)c].h > 4 || comps[(size_t)c].v < 1 || comps[(size_t)c].v > 4 || comps[(size_t)c].tq > 3) { err = "JPEG: bad sampling factors"; return false; }
                hmax = std::max(hmax, comps[(size_t)c].h); vmax = std::max(vmax, comps[(size_t)c].v);
            }
            const int mcus_x = (W + 8 * hmax - 1) / (8 * hmax), mcus_y = (H + 8 * vmax - 1) / (8 * vmax);
            for (auto& c : comps) {
                c.plane_w = mcus_x * c.h * 8; c.plane_h = mcus_y * c.v * 8;
                c.real_w = (W * c.h + hmax - 1) / hmax; c.real_h = (H * c.v + vmax - 1) / vmax;
                c.plane.assign((size_t)c.plane_w * c.plane_h, 0);
            }
            have_sof = true;
        } else if (m == 0xC2) { err = "JPEG: progressive files (SOF2) are not supported by this reader (baseline only)"; return false; }
        else if (m == 0xC3 || (m >= 0xC5 && m <= 0xCF && m != 0xC8 && m != 0xCC)) { err = "JPEG: lossless / hierarchical / arithmetic-coded files are not supported"; return false; }
        else if (m == 0xDD) { if (n >= 2) restart_interval = be16(pos + 2); }
        else if (m == 0xEE) { if (n >= 12 && std::memcmp(seg, "Adobe", 5) == 0) adobe_transform = seg[11]; }
        else if (m == 0xDA) {                                                     // SOS + entropy-coded data
            if (!have_sof) { err = "JPEG: SOS before SOF"; return false; }
            const int ns = seg[0];
            if (ns < 1 || ns > (int)comps.size() || n < 1 + 2 * ns + 3) { err = "JPEG: bad SOS"; return false; }
            std::vector<JComp*> sc;
            for (int k = 0; k < ns; k++) {
                JComp* found = nullptr;
                for (auto& c : comps) if (c.id == seg[1 + 2 * k]) found = &c;
                if (!found) { err = "JPEG: SOS names an unknown component"; return false; }
                found->td = seg[2 + 2 * k] >> 4; found->ta = seg[2 + 2 * k] & 15;
                if (found->td > 3 || found->ta > 3 || !hdc[found->td].present || !hac[found->ta].present || !have_qt[found->tq]) { err = "JPEG: scan uses a missing table"; return false; }
                found->pred = 0;
                sc.push_back(found);
            }
            if (seg[1 + 2 * ns] != 0 || seg[2 + 2 * ns] != 63) { err = "JPEG: spectral selection is not baseline"; return false; }
            JBits br{data + pos + (size_t)len, data + size};
            // MCU geometry: interleaved scan = hmax x vmax blocks of 8x8 per MCU; a single-component scan walks that component's blocks
            const bool interleaved = ns > 1;
            const int mx = interleaved ? (W + 8 * hmax - 1) / (8 * hmax) : (sc[0]->real_w + 7) / 8;
            const int my = interleaved ? (H + 8 * vmax - 1) / (8 * vmax) : (sc[0]->real_h + 7) / 8;
            int until_restart = restart_interval, next_rst = 0;
            int coef[64];
            for (int myi = 0; myi < my; myi++) for (int mxi = 0; mxi < mx; mxi++) {
                if (restart_interval && until_restart == 0) {
                    // byte-align, expect RSTn
                    br.reset();
                    while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) br.p++;
                    if (br.p + 1 < br.end) { next_rst = (br.p[1] - 0xD0 + 1) & 7; br.p += 2; }
                    (void)next_rst;
                    for (JComp* c : sc) c->pred = 0;
                    until_restart = restart_interval;
                }
                for (JComp* c : sc) {
                    const int bh = interleaved ? c->h : 1, bv = interleaved ? c->v : 1;
                    for (int by = 0; by < bv; by++) for (int bx = 0; bx < bh; bx++) {
                        std::memset(coef, 0, sizeof coef);
                        const int t = jhuff_decode(br, hdc[c->td]);
                        if (t < 0 || t > 11) { err = "JPEG: bad DC code"; return false; }
                        c->pred += t ? jextend(br.get(t), t) : 0;
                        coef[0] = c->pred * qt[c->tq][0];
                        for (int k = 1; k < 64;) {
                            const int rs = jhuff_decode(br, hac[c->ta]);
                            if (rs < 0) { err = "JPEG: bad AC code"; return false; }
                            const int r = rs >> 4, sbits = rs & 15;
                            if (sbits == 0) { if (r == 15) { k += 16; continue; } break; }
                            k += r;
                            if (k > 63) { err = "JPEG: AC run past the block"; return false; }
                            const int z = JPEG_ZIGZAG[k];
                            coef[z] = jextend(br.get(sbits), sbits) * qt[c->tq][z];
                            k++;
                        }
                        const int px = (mxi * bh + bx) * 8, py = (myi * bv + by) * 8;
                        if (px + 8 <= c->plane_w && py + 8 <= c->plane_h) jpeg_idct_islow(coef, c->plane.data() + (size_t)py * c->plane_w + px, c->plane_w);
                    }
                }
                if (restart_interval) until_restart--;
            }
            decoded_any = true;
            // continue after the entropy-coded segment: the bit reader stopped at (or before) the next marker
            size_t q = (size_t)(br.p - data);
            while (q + 1 < size && !(data[q] == 0xFF && data[q + 1] != 0x00 && !(data[q + 1] >= 0xD0 && data[q + 1] <= 0xD7))) q++;
            pos = q;
            continue;
        }
        pos += (size_t)len;
    }
    if (!have_sof || !decoded_any) { err = "JPEG: no image data"; return false; }
    w = (uint32_t)W; h = (uint32_t)H;
    if (comps.size() == 1) {
        channels = 1;
        pixels.resize((size_t)W * H);
        for (int y = 0; y < H; y++) std::memcpy(pixels.data() + (size_t)y * W, comps[0].plane.data() + (size_t)y * comps[0].plane_w, (size_t)W);
        return true;
    }
    std::vector<uint8_t> up[3]; int uw[3], uh[3];
    for (int c = 0; c < 3; c++) {
        const JComp& k = comps[(size_t)c];
        if (hmax % k.h || vmax % k.v) { err = "JPEG: fractional sampling ratios are not supported"; return false; }
        jpeg_upsample(k.plane.data(), k.plane_w, k.real_w, k.real_h, hmax / k.h, vmax / k.v, up[c], uw[c], uh[c]);
        if (uw[c] < W || uh[c] < H) { err = "JPEG: internal size mismatch"; return false; }
    }
    const bool rgb = adobe_transform == 0 || (comps[0].id == 'R' && comps[1].id == 'G' && comps[2].id == 'B');
    channels = 3;
    pixels.resize((size_t)W * H * 3);
    for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
        const int Y = up[0][(size_t)y * uw[0] + x], cb = up[1][(size_t)y * uw[1] + x] - 128, cr = up[2][(size_t)y * uw[2] + x] - 128;
        uint8_t* o = pixels.data() + ((size_t)y * W + x) * 3;
        if (rgb) { o[0] = (uint8_t)Y; o[1] = (uint8_t)(cb + 128); o[2] = (uint8_t)(cr + 128); continue; }
        // jdcolor.c: 16-bit fixed point, FIX(1.40200) = 91881, FIX(1.77200) = 116130, FIX(0.71414) = 46802, FIX(0.34414) = 22554
        o[0] = jclamp(Y + ((91881 * cr + 32768) >> 16));
        o[1] = jclamp(Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16));
        o[2] = jclamp(Y + ((116130 * cb + 32768) >> 16));
    }
    return true;
}

}  // namespace rt2ingest
