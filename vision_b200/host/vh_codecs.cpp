// Image codecs of the host layer (reference src/base/mgr/image_pool.cpp:13-35 loads JPG / PNG / HDR / EXR through
// ocarina's Image::load, src/base/mgr/pipeline.cpp:190-198,337-354 saves EXR / PNG). Written here, no image library:
//   JPEG  baseline / extended sequential Huffman (SOF0 / SOF1), 8-bit, 1 or 3 components, any sampling factors, restart intervals.
//         Arithmetic follows the IJG definitions so that decoded samples equal libjpeg's defaults bit for bit: the 13-bit
//         Loeffler-Ligtenberg-Moschytz integer IDCT ("islow"), triangle-filter ("fancy") chroma upsampling for 2:1 horizontally and
//         2x2, 16-bit fixed-point YCbCr -> RGB (tests compare with an independent decoder: tests/test_image_io.py).
//   PNG   all colour types and bit depths, non-interlaced; inflate by the system zlib; the five scanline filters.
//   EXR   scan-line files, HALF / FLOAT / UINT channels, compression NONE / ZIPS / ZIP (PIZ, PXR24, B44, DWA are refused by name);
//         writer: ZIP, FLOAT or HALF.
#include "vh_scene.h"
#include <zlib.h>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>

namespace vh {

std::vector<unsigned char> read_file(const std::string &path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::vector<unsigned char> d((size_t)f.tellg());
    f.seekg(0);
    f.read((char *)d.data(), (std::streamsize)d.size());
    return d;
}

// ================================================================================================ JPEG
namespace {

const unsigned char ZIGZAG[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                  35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct Huff {
    int mincode[17], maxcode[18], valptr[17];
    unsigned char vals[256];
    short fast[512];                         // 9-bit lookahead: (length << 8) | symbol, 0 = longer code
    bool present = false;
    void build(const unsigned char *counts, const unsigned char *symbols, int n) {
        std::memcpy(vals, symbols, (size_t)n);
        std::memset(fast, 0, sizeof(fast));
        int code = 0, k = 0;
        for (int l = 1; l <= 16; ++l) {
            valptr[l] = k; mincode[l] = code;
            for (int i = 0; i < counts[l - 1]; ++i, ++k, ++code)
                if (l <= 9) for (int pad = 0; pad < (1 << (9 - l)); ++pad) fast[(code << (9 - l)) | pad] = (short)((l << 8) | symbols[k]);
            maxcode[l] = counts[l - 1] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        present = true;
    }
};

struct JpegBits {
    const unsigned char *p, *end;
    uint32_t acc = 0; int n = 0;
    bool hit_marker = false;
    void fill() {
        while (n <= 24) {
            int b = 0;
            if (!hit_marker && p < end) {
                b = *p++;
                if (b == 0xFF) {
                    int b2 = p < end ? *p : 0;
                    if (b2 == 0) ++p;                         // stuffed zero
                    else { hit_marker = true; --p; b = 0; }   // a marker: feed zeros, leave it for the caller
                }
            }
            acc |= (uint32_t)b << (24 - n);
            n += 8;
        }
    }
    int peek(int k) { if (n < k) fill(); return (int)(acc >> (32 - k)); }
    void skip(int k) { acc <<= k; n -= k; }
    int get(int k) { if (k == 0) return 0; int v = peek(k); skip(k); return v; }
    void reset() { acc = 0; n = 0; hit_marker = false; }
};

inline int huff_decode(JpegBits &b, const Huff &h) {
    int look = b.peek(9);
    short f = h.fast[look];
    if (f) { b.skip(f >> 8); return f & 0xff; }
    int code = b.peek(16);
    for (int l = 10; l <= 16; ++l) {
        int c = code >> (16 - l);
        if (h.maxcode[l] >= 0 && c <= h.maxcode[l] && c >= h.mincode[l]) { b.skip(l); return h.vals[h.valptr[l] + c - h.mincode[l]]; }
    }
    throw std::runtime_error("jpeg: bad Huffman code");
}
inline int extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

inline int descale(long long x, int n) { return (int)((x + (1ll << (n - 1))) >> n); }
inline unsigned char clamp8(int v) { return (unsigned char)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

// IJG "islow": LL&M 1-D IDCT on columns then rows, CONST_BITS = 13, PASS1_BITS = 2
void idct_islow(const int *in /* dequantised, natural order */, unsigned char *out, int stride) {
    const int C_0_298 = 2446, C_0_390 = 3196, C_0_541 = 4433, C_0_765 = 6270, C_0_899 = 7373, C_1_175 = 9633, C_1_501 = 12299, C_1_847 = 15137,
              C_1_961 = 16069, C_2_053 = 16819, C_2_562 = 20995, C_3_072 = 25172;
    long long ws[64];
    for (int pass = 0; pass < 2; ++pass) {
        for (int i = 0; i < 8; ++i) {
            long long d[8];
            for (int k = 0; k < 8; ++k) d[k] = pass == 0 ? (long long)in[8 * k + i] : ws[8 * i + k];
            long long z2 = d[2], z3 = d[6];
            long long z1 = (z2 + z3) * C_0_541;
            long long tmp2 = z1 + z3 * (-C_1_847), tmp3 = z1 + z2 * C_0_765;
            long long tmp0 = (d[0] + d[4]) * 8192, tmp1 = (d[0] - d[4]) * 8192;
            long long tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
            tmp0 = d[7]; tmp1 = d[5]; tmp2 = d[3]; tmp3 = d[1];
            z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; long long z4 = tmp1 + tmp3;
            long long z5 = (z3 + z4) * C_1_175;
            tmp0 *= C_0_298; tmp1 *= C_2_053; tmp2 *= C_3_072; tmp3 *= C_1_501;
            z1 *= -C_0_899; z2 *= -C_2_562; z3 *= -C_1_961; z4 *= -C_0_390;
            z3 += z5; z4 += z5;
            tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
            long long o[8] = {tmp10 + tmp3, tmp11 + tmp2, tmp12 + tmp1, tmp13 + tmp0, tmp13 - tmp0, tmp12 - tmp1, tmp11 - tmp2, tmp10 - tmp3};
            if (pass == 0) for (int k = 0; k < 8; ++k) ws[8 * k + i] = descale(o[k], 13 - 2);
            else for (int k = 0; k < 8; ++k) out[i * stride + k] = clamp8(descale(o[k], 13 + 2 + 3) + 128);
        }
    }
}

struct JComp { int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0, pred = 0; int bw = 0, bh = 0; std::vector<unsigned char> plane; int dw = 0, dh = 0; };

// triangle-filter 2x2 upsampling of one output row: near = the closer input row, far = the other one
void h2v2_fancy_row(const unsigned char *near, const unsigned char *far, int dw, unsigned char *out) {
    if (dw == 1) { int s = near[0] * 3 + far[0]; out[0] = (unsigned char)((s * 4 + 8) >> 4); out[1] = (unsigned char)((s * 4 + 7) >> 4); return; }
    int thiscol = near[0] * 3 + far[0], nextcol = near[1] * 3 + far[1], lastcol;
    out[0] = (unsigned char)((thiscol * 4 + 8) >> 4);
    out[1] = (unsigned char)((thiscol * 3 + nextcol + 7) >> 4);
    lastcol = thiscol; thiscol = nextcol;
    for (int x = 1; x < dw - 1; ++x) {
        nextcol = near[x + 1] * 3 + far[x + 1];
        out[2 * x] = (unsigned char)((thiscol * 3 + lastcol + 8) >> 4);
        out[2 * x + 1] = (unsigned char)((thiscol * 3 + nextcol + 7) >> 4);
        lastcol = thiscol; thiscol = nextcol;
    }
    out[2 * (dw - 1)] = (unsigned char)((thiscol * 3 + lastcol + 8) >> 4);
    out[2 * (dw - 1) + 1] = (unsigned char)((thiscol * 4 + 7) >> 4);
}
void h2v1_fancy_row(const unsigned char *in, int dw, unsigned char *out) {
    if (dw == 1) { out[0] = out[1] = in[0]; return; }
    out[0] = in[0];
    out[1] = (unsigned char)((in[0] * 3 + in[1] + 2) >> 2);
    for (int x = 1; x < dw - 1; ++x) {
        out[2 * x] = (unsigned char)((in[x] * 3 + in[x - 1] + 1) >> 2);
        out[2 * x + 1] = (unsigned char)((in[x] * 3 + in[x + 1] + 2) >> 2);
    }
    out[2 * (dw - 1)] = (unsigned char)((in[dw - 1] * 3 + in[dw - 2] + 1) >> 2);
    out[2 * (dw - 1) + 1] = in[dw - 1];
}

}  // namespace

// decodes to 8-bit RGB (w * h * 3)
void decode_jpeg(const unsigned char *data, size_t size, uint32_t &w, uint32_t &h, std::vector<unsigned char> &rgb) {
    if (size < 4 || data[0] != 0xFF || data[1] != 0xD8) throw std::runtime_error("jpeg: no SOI");
    int qt[4][64]; bool have_qt[4] = {false, false, false, false};
    Huff dc[4], ac[4];
    std::vector<JComp> comps;
    int restart = 0, adobe_transform = -1;
    bool jfif = false;
    size_t i = 2;
    int W = 0, H = 0, hmax = 1, vmax = 1;
    bool decoded = false;
    while (i + 4 <= size && !decoded) {
        if (data[i] != 0xFF) { ++i; continue; }
        int m = data[i + 1];
        if (m == 0xFF) { ++i; continue; }
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { i += 2; continue; }
        if (m == 0xD9) break;
        size_t len = ((size_t)data[i + 2] << 8) | data[i + 3];
        if (len < 2 || i + 2 + len > size) throw std::runtime_error("jpeg: truncated segment");
        const unsigned char *s = data + i + 4; size_t n = len - 2;
        if (m == 0xDB) {
            size_t k = 0;
            while (k < n) {
                int pq = s[k] >> 4, tq = s[k] & 15; ++k;
                if (tq > 3 || k + (pq ? 128u : 64u) > n) throw std::runtime_error("jpeg: bad DQT");
                for (int z = 0; z < 64; ++z) { qt[tq][ZIGZAG[z]] = pq ? ((s[k] << 8) | s[k + 1]) : s[k]; k += pq ? 2 : 1; }
                have_qt[tq] = true;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (n < 6 || s[0] != 8) throw std::runtime_error("jpeg: only 8-bit samples");
            H = (s[1] << 8) | s[2]; W = (s[3] << 8) | s[4];
            int nc = s[5];
            if ((nc != 1 && nc != 3) || n < 6u + 3u * nc || !W || !H) throw std::runtime_error("jpeg: unsupported component count / size");
            comps.resize(nc);
            for (int c = 0; c < nc; ++c) {
                comps[c].id = s[6 + 3 * c]; comps[c].h = s[7 + 3 * c] >> 4; comps[c].v = s[7 + 3 * c] & 15; comps[c].tq = s[8 + 3 * c] & 3;
                if (comps[c].h < 1 || comps[c].h > 4 || comps[c].v < 1 || comps[c].v > 4) throw std::runtime_error("jpeg: bad sampling factor");
                hmax = std::max(hmax, comps[c].h); vmax = std::max(vmax, comps[c].v);
            }
        } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
            throw std::runtime_error("jpeg: progressive / lossless / arithmetic coding is not supported (baseline only)");
        } else if (m == 0xC4) {
            size_t k = 0;
            while (k + 17 <= n) {
                int tc = s[k] >> 4, th = s[k] & 15;
                int cnt = 0;
                for (int l = 0; l < 16; ++l) cnt += s[k + 1 + l];
                if (th > 3 || tc > 1 || cnt > 256 || k + 17 + cnt > n) throw std::runtime_error("jpeg: bad DHT");
                (tc ? ac[th] : dc[th]).build(s + k + 1, s + k + 17, cnt);
                k += 17 + (size_t)cnt;
            }
        } else if (m == 0xDD) {
            restart = (s[0] << 8) | s[1];
        } else if (m == 0xE0) {
            jfif = n >= 5 && !std::memcmp(s, "JFIF", 5);
        } else if (m == 0xEE) {
            if (n >= 12 && !std::memcmp(s, "Adobe", 5)) adobe_transform = s[11];
        } else if (m == 0xDA) {
            if (comps.empty()) throw std::runtime_error("jpeg: SOS before SOF");
            int ns = s[0];
            if (ns != (int)comps.size()) throw std::runtime_error("jpeg: non-interleaved multi-scan files are not supported");
            for (int c = 0; c < ns; ++c) {
                int id = s[1 + 2 * c];
                JComp *cp = nullptr;
                for (auto &cc : comps) if (cc.id == id) cp = &cc;
                if (!cp) throw std::runtime_error("jpeg: scan names an unknown component");
                cp->td = s[2 + 2 * c] >> 4; cp->ta = s[2 + 2 * c] & 15;
                if (cp->td > 3 || cp->ta > 3 || !dc[cp->td].present || !ac[cp->ta].present || !have_qt[cp->tq]) throw std::runtime_error("jpeg: missing table");
            }
            const int mcu_w = 8 * hmax, mcu_h = 8 * vmax;
            const int mx = (W + mcu_w - 1) / mcu_w, my = (H + mcu_h - 1) / mcu_h;
            const bool single = comps.size() == 1;       // a one-component scan is not interleaved: blocks in raster order
            for (auto &c : comps) {
                c.bw = single ? (W + 7) / 8 : mx * c.h; c.bh = single ? (H + 7) / 8 : my * c.v;
                c.plane.assign((size_t)c.bw * 8 * c.bh * 8, 0);
                c.dw = (W * c.h + hmax - 1) / hmax; c.dh = (H * c.v + vmax - 1) / vmax; c.pred = 0;
            }
            JpegBits bits{data + i + 2 + len, data + size};
            int coef[64], units = single ? comps[0].bw * comps[0].bh : mx * my, next_rst = 0;
            auto block = [&](JComp &c, int bx, int by) {
                std::memset(coef, 0, sizeof(coef));
                int t = huff_decode(bits, dc[c.td]);
                if (t > 15) throw std::runtime_error("jpeg: bad DC category");
                c.pred += t ? extend(bits.get(t), t) : 0;
                coef[0] = c.pred * qt[c.tq][0];
                for (int k = 1; k < 64;) {
                    int rs = huff_decode(bits, ac[c.ta]), r = rs >> 4, sz = rs & 15;
                    if (sz == 0) { if (r != 15) break; k += 16; continue; }
                    k += r;
                    if (k > 63) throw std::runtime_error("jpeg: AC index overflow");
                    coef[ZIGZAG[k]] = extend(bits.get(sz), sz) * qt[c.tq][ZIGZAG[k]];
                    ++k;
                }
                idct_islow(coef, &c.plane[((size_t)by * 8) * ((size_t)c.bw * 8) + (size_t)bx * 8], c.bw * 8);
            };
            for (int u = 0; u < units; ++u) {
                if (restart && u && u % restart == 0) {
                    bits.reset();
                    const unsigned char *q = bits.p;
                    while (q + 1 < bits.end && !(q[0] == 0xFF && q[1] >= 0xD0 && q[1] <= 0xD7)) ++q;
                    if (q + 1 >= bits.end) throw std::runtime_error("jpeg: missing restart marker");
                    (void)next_rst; bits.p = q + 2;
                    for (auto &c : comps) c.pred = 0;
                }
                if (single) block(comps[0], u % comps[0].bw, u / comps[0].bw);
                else {
                    int ux = u % mx, uy = u / mx;
                    for (auto &c : comps) for (int v = 0; v < c.v; ++v) for (int hh = 0; hh < c.h; ++hh) block(c, ux * c.h + hh, uy * c.v + v);
                }
            }
            decoded = true;
        }
        i += 2 + len;
    }
    if (!decoded) throw std::runtime_error("jpeg: no scan found");
    // upsample each component to W x H
    w = (uint32_t)W; h = (uint32_t)H;
    std::vector<std::vector<unsigned char>> full(comps.size());
    for (size_t ci = 0; ci < comps.size(); ++ci) {
        JComp &c = comps[ci];
        const int stride = c.bw * 8;
        std::vector<unsigned char> &o = full[ci];
        o.resize((size_t)W * H);
        const int fx = hmax / c.h, fy = vmax / c.v;
        if (hmax % c.h || vmax % c.v) throw std::runtime_error("jpeg: fractional sampling ratios are not supported");
        if (fx == 1 && fy == 1) {
            for (int y = 0; y < H; ++y) std::memcpy(&o[(size_t)y * W], &c.plane[(size_t)y * stride], (size_t)W);
        } else if (fx == 2 && fy == 2) {
            std::vector<unsigned char> row((size_t)2 * c.dw + 2);
            for (int y = 0; y < H; ++y) {
                int sy = y >> 1, oy = (y & 1) ? sy + 1 : sy - 1;
                oy = std::min(std::max(oy, 0), c.dh - 1);            // edge rows are replicated (the decoder's context rows)
                h2v2_fancy_row(&c.plane[(size_t)sy * stride], &c.plane[(size_t)oy * stride], c.dw, row.data());
                std::memcpy(&o[(size_t)y * W], row.data(), (size_t)W);
            }
        } else if (fx == 2 && fy == 1) {
            std::vector<unsigned char> row((size_t)2 * c.dw + 2);
            for (int y = 0; y < H; ++y) { h2v1_fancy_row(&c.plane[(size_t)y * stride], c.dw, row.data()); std::memcpy(&o[(size_t)y * W], row.data(), (size_t)W); }
        } else {                                                     // other integral ratios: replication
            for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) o[(size_t)y * W + x] = c.plane[(size_t)(y / fy) * stride + x / fx];
        }
    }
    rgb.resize((size_t)W * H * 3);
    if (comps.size() == 1) {
        for (size_t p = 0; p < (size_t)W * H; ++p) rgb[3 * p] = rgb[3 * p + 1] = rgb[3 * p + 2] = full[0][p];
        return;
    }
    bool ycc = adobe_transform >= 0 ? adobe_transform != 0 : (jfif || !(comps[0].id == 'R' && comps[1].id == 'G' && comps[2].id == 'B'));
    if (!ycc) { for (size_t p = 0; p < (size_t)W * H; ++p) for (int c = 0; c < 3; ++c) rgb[3 * p + c] = full[c][p]; return; }
    int cr_r[256], cb_b[256], cr_g[256], cb_g[256];
    for (int k = 0; k < 256; ++k) {
        int x = k - 128;
        cr_r[k] = (int)((91881ll * x + 32768) >> 16);      // FIX(1.40200)
        cb_b[k] = (int)((116130ll * x + 32768) >> 16);     // FIX(1.77200)
        cr_g[k] = -46802 * x;                               // FIX(0.71414)
        cb_g[k] = -22554 * x + 32768;                       // FIX(0.34414), rounding folded in
    }
    for (size_t p = 0; p < (size_t)W * H; ++p) {
        int y = full[0][p], cb = full[1][p], cr = full[2][p];
        rgb[3 * p] = clamp8(y + cr_r[cr]);
        rgb[3 * p + 1] = clamp8(y + ((cb_g[cb] + cr_g[cr]) >> 16));
        rgb[3 * p + 2] = clamp8(y + cb_b[cb]);
    }
}

// ================================================================================================ PNG
namespace {
std::vector<unsigned char> zlib_inflate(const unsigned char *src, size_t n, size_t expect) {
    std::vector<unsigned char> out(expect);
    z_stream z{};
    if (inflateInit(&z) != Z_OK) throw std::runtime_error("zlib: inflateInit failed");
    z.next_in = const_cast<unsigned char *>(src); z.avail_in = (uInt)n;
    z.next_out = out.data(); z.avail_out = (uInt)expect;
    int rc = inflate(&z, Z_FINISH);
    size_t got = z.total_out;
    inflateEnd(&z);
    if (rc != Z_STREAM_END && !(rc == Z_OK || rc == Z_BUF_ERROR)) throw std::runtime_error("zlib: corrupt stream");
    if (got != expect) throw std::runtime_error("zlib: unexpected decompressed size");
    return out;
}
std::vector<unsigned char> zlib_deflate(const unsigned char *src, size_t n) {
    uLongf cap = compressBound((uLong)n);
    std::vector<unsigned char> out(cap);
    if (compress2(out.data(), &cap, src, (uLong)n, 6) != Z_OK) throw std::runtime_error("zlib: deflate failed");
    out.resize(cap);
    return out;
}
inline uint32_t be32(const unsigned char *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
}  // namespace

// decodes to float RGBA in [0, 1] (w * h * 4), no colour-space conversion
void decode_png(const unsigned char *d, size_t size, uint32_t &w, uint32_t &h, std::vector<float> &rgba) {
    static const unsigned char SIG[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (size < 33 || std::memcmp(d, SIG, 8)) throw std::runtime_error("png: bad signature");
    size_t i = 8;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<unsigned char> idat, plte, trns;
    w = h = 0;
    while (i + 12 <= size) {
        uint32_t len = be32(d + i);
        if (i + 12 + len > size) throw std::runtime_error("png: truncated chunk");
        const unsigned char *t = d + i + 4, *p = d + i + 8;
        if (!std::memcmp(t, "IHDR", 4)) { w = be32(p); h = be32(p + 4); depth = p[8]; ctype = p[9]; interlace = p[12]; }
        else if (!std::memcmp(t, "PLTE", 4)) plte.assign(p, p + len);
        else if (!std::memcmp(t, "tRNS", 4)) trns.assign(p, p + len);
        else if (!std::memcmp(t, "IDAT", 4)) idat.insert(idat.end(), p, p + len);
        else if (!std::memcmp(t, "IEND", 4)) break;
        i += 12 + (size_t)len;
    }
    if (!w || !h || idat.empty()) throw std::runtime_error("png: missing IHDR / IDAT");
    if (interlace) throw std::runtime_error("png: Adam7-interlaced files are not supported");
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!channels || !(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) throw std::runtime_error("png: bad colour type / depth");
    const size_t bpp = std::max<size_t>(1, (size_t)channels * depth / 8), row_bytes = ((size_t)w * channels * depth + 7) / 8;
    std::vector<unsigned char> raw = zlib_inflate(idat.data(), idat.size(), (row_bytes + 1) * h);
    std::vector<unsigned char> prev(row_bytes, 0), cur(row_bytes);
    rgba.resize((size_t)w * h * 4);
    const float maxv = (float)((1u << depth) - 1u);
    for (uint32_t y = 0; y < h; ++y) {
        const unsigned char *src = &raw[(size_t)y * (row_bytes + 1)];
        int ft = src[0]; ++src;
        for (size_t x = 0; x < row_bytes; ++x) {
            int a = x >= bpp ? cur[x - bpp] : 0, b = prev[x], c = x >= bpp ? prev[x - bpp] : 0, v = src[x];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: { int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c); v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
                default: throw std::runtime_error("png: bad filter type");
            }
            cur[x] = (unsigned char)v;
        }
        auto sample = [&](size_t idx) -> uint32_t {                 // idx-th sample of the row
            if (depth == 8) return cur[idx];
            if (depth == 16) return ((uint32_t)cur[2 * idx] << 8) | cur[2 * idx + 1];
            size_t bit = idx * (size_t)depth;
            return (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1u);
        };
        for (uint32_t x = 0; x < w; ++x) {
            float *o = &rgba[((size_t)y * w + x) * 4];
            if (ctype == 3) {
                uint32_t k = sample(x);
                if (3 * (size_t)k + 2 >= plte.size()) throw std::runtime_error("png: palette index out of range");
                o[0] = plte[3 * k] / 255.f; o[1] = plte[3 * k + 1] / 255.f; o[2] = plte[3 * k + 2] / 255.f; o[3] = k < trns.size() ? trns[k] / 255.f : 1.f;
            } else if (ctype == 0 || ctype == 4) {
                float g = sample((size_t)x * channels) / maxv;
                o[0] = o[1] = o[2] = g; o[3] = ctype == 4 ? sample((size_t)x * 2 + 1) / maxv : 1.f;
            } else {
                for (int c = 0; c < 3; ++c) o[c] = sample((size_t)x * channels + c) / maxv;
                o[3] = ctype == 6 ? sample((size_t)x * 4 + 3) / maxv : 1.f;
            }
        }
        prev.swap(cur);
    }
}

// ================================================================================================ OpenEXR (scan lines)
namespace {
float half_to_float(uint16_t hbits) {
    uint32_t s = (uint32_t)(hbits >> 15) << 31, e = (hbits >> 10) & 31, m = hbits & 1023, bits;
    if (e == 0) {
        if (m == 0) bits = s;
        else { int sh = 0; while (!(m & 1024)) { m <<= 1; ++sh; } bits = s | ((uint32_t)(113 - sh) << 23) | ((m & 1023) << 13); }
    } else if (e == 31) bits = s | 0x7f800000u | (m << 13);
    else bits = s | ((e + 112) << 23) | (m << 13);
    float f; std::memcpy(&f, &bits, 4); return f;
}
uint16_t float_to_half(float f) {                                    // round to nearest even
    uint32_t x; std::memcpy(&x, &f, 4);
    uint32_t s = (x >> 16) & 0x8000u; int e = (int)((x >> 23) & 255) - 112; uint32_t m = x & 0x7fffffu;
    if (((x >> 23) & 255) == 255) return (uint16_t)(s | 0x7c00u | (m ? 0x200u : 0));
    if (e >= 31) return (uint16_t)(s | 0x7c00u);
    if (e <= 0) {
        if (e < -10) return (uint16_t)s;
        m |= 0x800000u;
        int sh = 14 - e;
        uint32_t r = m >> sh, rem = m & ((1u << sh) - 1), halfway = 1u << (sh - 1);
        if (rem > halfway || (rem == halfway && (r & 1))) ++r;
        return (uint16_t)(s | r);
    }
    uint32_t r = ((uint32_t)e << 10) | (m >> 13), rem = m & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (r & 1))) ++r;
    return (uint16_t)(s | r);
}
inline uint32_t le32(const unsigned char *p) { return p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
inline uint64_t le64(const unsigned char *p) { return (uint64_t)le32(p) | ((uint64_t)le32(p + 4) << 32); }
struct ExrChan { std::string name; int type; };
}  // namespace

void decode_exr(const unsigned char *d, size_t size, uint32_t &w, uint32_t &h, std::vector<float> &rgba) {
    if (size < 12 || le32(d) != 20000630u) throw std::runtime_error("exr: bad magic");
    uint32_t version = le32(d + 4);
    if (version & 0x1a00u) throw std::runtime_error("exr: tiled / deep / multi-part files are not supported");
    size_t i = 8;
    std::vector<ExrChan> chans;
    int compression = -1, line_order = 0;
    int dw[4] = {0, 0, -1, -1};
    auto cstr = [&]() { size_t e = i; while (e < size && d[e]) ++e; if (e >= size) throw std::runtime_error("exr: truncated header"); std::string s((const char *)d + i, e - i); i = e + 1; return s; };
    while (i < size && d[i]) {
        std::string name = cstr(), type = cstr();
        if (i + 4 > size) throw std::runtime_error("exr: truncated header");
        uint32_t n = le32(d + i); i += 4;
        if (i + n > size) throw std::runtime_error("exr: truncated attribute");
        const unsigned char *v = d + i;
        if (name == "channels") {
            size_t k = 0;
            while (k < n && v[k]) {
                size_t e = k; while (v[e]) ++e;
                ExrChan c{std::string((const char *)v + k, e - k), (int)le32(v + e + 1)};
                if (le32(v + e + 9) != 1 || le32(v + e + 13) != 1) throw std::runtime_error("exr: sub-sampled channels are not supported");
                chans.push_back(c);
                k = e + 17;
            }
        } else if (name == "compression") compression = v[0];
        else if (name == "dataWindow") for (int k = 0; k < 4; ++k) dw[k] = (int)le32(v + 4 * k);
        else if (name == "lineOrder") line_order = v[0];
        i += n;
    }
    ++i;
    static const char *CNAME[] = {"NONE", "RLE", "ZIPS", "ZIP", "PIZ", "PXR24", "B44", "B44A", "DWAA", "DWAB"};
    if (compression != 0 && compression != 2 && compression != 3)
        throw std::runtime_error(std::string("exr: compression ") + (compression >= 0 && compression < 10 ? CNAME[compression] : "?") + " is not supported (NONE, ZIPS, ZIP are)");
    if (chans.empty() || dw[2] < dw[0] || dw[3] < dw[1]) throw std::runtime_error("exr: missing channels / dataWindow");
    (void)line_order;
    w = (uint32_t)(dw[2] - dw[0] + 1); h = (uint32_t)(dw[3] - dw[1] + 1);
    const int block_lines = compression == 3 ? 16 : 1;
    const size_t n_blocks = (h + block_lines - 1) / block_lines;
    if (i + 8 * n_blocks > size) throw std::runtime_error("exr: truncated offset table");
    size_t line_bytes = 0;
    for (auto &c : chans) line_bytes += (size_t)w * (c.type == 1 ? 2 : 4);
    rgba.assign((size_t)w * h * 4, 0.f);
    for (size_t p = 0; p < (size_t)w * h; ++p) rgba[4 * p + 3] = 1.f;
    bool has_rgb = false;
    for (auto &c : chans) if (c.name == "R" || c.name == "G" || c.name == "B") has_rgb = true;
    std::vector<unsigned char> tmp;
    for (size_t b = 0; b < n_blocks; ++b) {
        uint64_t off = le64(d + i + 8 * b);
        if (off + 8 > size) throw std::runtime_error("exr: bad block offset");
        int y0 = (int)le32(d + off) - dw[1];
        uint32_t packed = le32(d + off + 4);
        if (off + 8 + packed > size || y0 < 0 || y0 >= (int)h) throw std::runtime_error("exr: bad block");
        int lines = std::min<int>(block_lines, (int)h - y0);
        size_t raw_bytes = line_bytes * (size_t)lines;
        const unsigned char *src = d + off + 8;
        std::vector<unsigned char> raw;
        if (compression == 0 || packed == raw_bytes) raw.assign(src, src + raw_bytes);
        else {
            tmp = zlib_inflate(src, packed, raw_bytes);
            for (size_t k = 1; k < raw_bytes; ++k) tmp[k] = (unsigned char)(tmp[k - 1] + tmp[k] - 128);      // predictor
            raw.resize(raw_bytes);
            size_t half = (raw_bytes + 1) / 2;
            for (size_t k = 0; k < raw_bytes; ++k) raw[k] = (k & 1) ? tmp[half + k / 2] : tmp[k / 2];       // de-interleave
        }
        const unsigned char *p = raw.data();
        for (int l = 0; l < lines; ++l)
            for (auto &c : chans) {
                int dst = c.name == "R" ? 0 : c.name == "G" ? 1 : c.name == "B" ? 2 : c.name == "A" ? 3 : (c.name == "Y" && !has_rgb ? 4 : -1);
                for (uint32_t x = 0; x < w; ++x) {
                    float v;
                    if (c.type == 1) { v = half_to_float((uint16_t)(p[0] | (p[1] << 8))); p += 2; }
                    else if (c.type == 2) { uint32_t u = le32(p); std::memcpy(&v, &u, 4); p += 4; }
                    else { v = (float)le32(p); p += 4; }
                    float *o = &rgba[((size_t)(y0 + l) * w + x) * 4];
                    if (dst == 4) o[0] = o[1] = o[2] = v; else if (dst >= 0) o[dst] = v;
                }
            }
    }
}

// scan-line EXR, channels A? B G R (alphabetical, as the format requires), ZIP (16-line blocks)
void save_exr(const std::string &path, const float *rgba, uint32_t w, uint32_t h, bool half) {
    std::vector<unsigned char> f;
    auto put32 = [&](uint32_t v) { for (int k = 0; k < 4; ++k) f.push_back((unsigned char)(v >> (8 * k))); };
    auto puts0 = [&](const char *s) { f.insert(f.end(), s, s + std::strlen(s) + 1); };
    auto attr = [&](const char *name, const char *type, const std::vector<unsigned char> &v) { puts0(name); puts0(type); put32((uint32_t)v.size()); f.insert(f.end(), v.begin(), v.end()); };
    auto bytes32 = [](std::initializer_list<uint32_t> vals) { std::vector<unsigned char> o; for (uint32_t v : vals) for (int k = 0; k < 4; ++k) o.push_back((unsigned char)(v >> (8 * k))); return o; };
    put32(20000630u); put32(2u);
    std::vector<unsigned char> ch;
    for (const char *n : {"B", "G", "R"}) { ch.push_back((unsigned char)n[0]); ch.push_back(0); for (uint32_t v : {half ? 1u : 2u, 0u, 1u, 1u}) for (int k = 0; k < 4; ++k) ch.push_back((unsigned char)(v >> (8 * k))); }
    ch.push_back(0);
    attr("channels", "chlist", ch);
    attr("compression", "compression", {3});
    attr("dataWindow", "box2i", bytes32({0, 0, w - 1, h - 1}));
    attr("displayWindow", "box2i", bytes32({0, 0, w - 1, h - 1}));
    attr("lineOrder", "lineOrder", {0});
    uint32_t one; { float o = 1.f; std::memcpy(&one, &o, 4); }
    attr("pixelAspectRatio", "float", bytes32({one}));
    attr("screenWindowCenter", "v2f", bytes32({0, 0}));
    attr("screenWindowWidth", "float", bytes32({one}));
    f.push_back(0);
    const size_t n_blocks = (h + 15) / 16, table = f.size();
    f.resize(f.size() + 8 * n_blocks);
    const size_t bps = half ? 2 : 4, line_bytes = 3 * (size_t)w * bps;
    std::vector<unsigned char> raw, inter;
    for (size_t b = 0; b < n_blocks; ++b) {
        uint32_t y0 = (uint32_t)b * 16, lines = std::min<uint32_t>(16, h - y0);
        raw.resize(line_bytes * lines);
        unsigned char *p = raw.data();
        for (uint32_t l = 0; l < lines; ++l)
            for (int c : {2, 1, 0})                                           // B, G, R
                for (uint32_t x = 0; x < w; ++x) {
                    float v = rgba[((size_t)(y0 + l) * w + x) * 4 + c];
                    if (half) { uint16_t hb = float_to_half(v); p[0] = (unsigned char)hb; p[1] = (unsigned char)(hb >> 8); }
                    else std::memcpy(p, &v, 4);
                    p += bps;
                }
        inter.resize(raw.size());
        size_t halfway = (raw.size() + 1) / 2;
        for (size_t k = 0; k < raw.size(); ++k) inter[(k & 1) ? halfway + k / 2 : k / 2] = raw[k];
        for (size_t k = inter.size() - 1; k >= 1; --k) inter[k] = (unsigned char)(inter[k] - inter[k - 1] + 128);
        std::vector<unsigned char> z = zlib_deflate(inter.data(), inter.size());
        const std::vector<unsigned char> &payload = z.size() < raw.size() ? z : raw;
        uint64_t off = f.size();
        for (int k = 0; k < 8; ++k) f[table + 8 * b + k] = (unsigned char)(off >> (8 * k));
        put32(y0); put32((uint32_t)payload.size());
        f.insert(f.end(), payload.begin(), payload.end());
    }
    FILE *fp = std::fopen(path.c_str(), "wb");
    if (!fp) throw std::runtime_error("cannot write " + path);
    std::fwrite(f.data(), 1, f.size(), fp);
    std::fclose(fp);
}

// Image::load by extension (image_pool.cpp:13-35): 8-bit sources are scaled by 1/255, nothing else
Image decode_image_file(const std::string &path) {
    auto ends = [&](const char *e) { size_t n = std::strlen(e); return path.size() >= n && strcasecmp(path.c_str() + path.size() - n, e) == 0; };
    std::vector<unsigned char> d = read_file(path);
    Image im;
    if (ends(".jpg") || ends(".jpeg")) {
        std::vector<unsigned char> rgb;
        decode_jpeg(d.data(), d.size(), im.w, im.h, rgb);
        im.rgba.resize((size_t)im.w * im.h * 4);
        for (size_t p = 0; p < (size_t)im.w * im.h; ++p) { for (int c = 0; c < 3; ++c) im.rgba[4 * p + c] = rgb[3 * p + c] / 255.f; im.rgba[4 * p + 3] = 1.f; }
    } else if (ends(".png")) decode_png(d.data(), d.size(), im.w, im.h, im.rgba);
    else if (ends(".exr")) decode_exr(d.data(), d.size(), im.w, im.h, im.rgba);
    else throw std::runtime_error("unsupported image format " + path);
    return im;
}

}  // namespace vh
