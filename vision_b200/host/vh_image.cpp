// Image file I/O of the host layer: what sits either side of the path - environment maps in, pictures out.
//   Radiance .hdr (RGBE; flat and new-style run-length scanlines) read + write, PNG write (8-bit RGB, zlib "stored" blocks: no
//   compressor needed), dispatch by extension with the reference's rule "gamma unless exr / hdr"
//   (reference src/base/mgr/pipeline.cpp:337-354 final_picture; src/base/mgr/image_pool.cpp:18-24 .exr / .hdr are linear on load).
// JPG / PNG / EXR decoding and EXR writing: vh_codecs.cpp.
#include "vh_scene.h"
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <strings.h>

namespace vh {

static bool ends_with_ci(const std::string &p, const char *e) {
    size_t n = std::strlen(e);
    return p.size() >= n && strcasecmp(p.c_str() + p.size() - n, e) == 0;
}
static float oetf_srgb(float x) {                                   // linear_to_srgb
    x = std::min(std::max(x, 0.f), 1.f);
    return x <= 0.0031308f ? 12.92f * x : 1.055f * std::pow(x, 1.f / 2.4f) - 0.055f;
}

// ---- Radiance RGBE ---------------------------------------------------------------------------------------------------------
// float -> RGBE (G. Ward, "Real Pixels", Graphics Gems II): shared exponent of the largest component, 8-bit mantissas
void rgbe_encode(const float *rgb, unsigned char *out) {
    float v = std::max(rgb[0], std::max(rgb[1], rgb[2]));
    if (!(v > 1e-32f)) { out[0] = out[1] = out[2] = out[3] = 0; return; }
    int e;
    float m = std::frexp(v, &e) * 256.0f / v;
    for (int c = 0; c < 3; ++c) out[c] = (unsigned char)std::max(0.f, rgb[c] * m);
    out[3] = (unsigned char)(e + 128);
}
void rgbe_decode(const unsigned char *in, float *rgb) {
    if (in[3] == 0) { rgb[0] = rgb[1] = rgb[2] = 0.f; return; }
    float f = std::ldexp(1.0f, (int)in[3] - (128 + 8));
    for (int c = 0; c < 3; ++c) rgb[c] = (float)in[c] * f;
}

void save_hdr(const std::string &path, const float *rgba, uint32_t w, uint32_t h) {
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    std::fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %u +X %u\n", h, w);
    std::vector<unsigned char> row((size_t)w * 4);
    for (uint32_t y = 0; y < h; ++y) {                              // flat (uncompressed) scanlines, top row first
        for (uint32_t x = 0; x < w; ++x) rgbe_encode(&rgba[((size_t)y * w + x) * 4], &row[(size_t)x * 4]);
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
}

Image load_hdr(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open image " + path);
    std::string line;
    std::getline(f, line);
    if (line.rfind("#?", 0) != 0) throw std::runtime_error(path + ": not a Radiance .hdr file");
    bool rgbe = false;
    while (std::getline(f, line) && !line.empty() && line != "\r") if (line.rfind("FORMAT=32-bit_rle_rgbe", 0) == 0) rgbe = true;
    if (!rgbe) throw std::runtime_error(path + ": only FORMAT=32-bit_rle_rgbe is supported");
    std::getline(f, line);
    int hh = 0, ww = 0;
    if (std::sscanf(line.c_str(), "-Y %d +X %d", &hh, &ww) != 2 || hh <= 0 || ww <= 0) throw std::runtime_error(path + ": unsupported resolution line '" + line + "'");
    Image im; im.w = (uint32_t)ww; im.h = (uint32_t)hh; im.rgba.resize((size_t)ww * hh * 4);
    std::vector<unsigned char> row((size_t)ww * 4);
    for (int y = 0; y < hh; ++y) {
        unsigned char head[4];
        f.read((char *)head, 4);
        if (!f) throw std::runtime_error(path + ": truncated");
        bool rle = ww >= 8 && ww < 32768 && head[0] == 2 && head[1] == 2 && ((head[2] << 8) | head[3]) == ww;
        if (!rle) {                                                 // flat scanline: the 4 bytes read are the first pixel
            std::memcpy(row.data(), head, 4);
            f.read((char *)row.data() + 4, (std::streamsize)row.size() - 4);
        } else {                                                    // new-style RLE: the four channels separately
            for (int c = 0; c < 4; ++c) {
                int x = 0;
                while (x < ww) {
                    int n = f.get();
                    if (n == EOF) throw std::runtime_error(path + ": truncated");
                    if (n > 128) { n -= 128; int v = f.get(); if (x + n > ww) throw std::runtime_error(path + ": bad run"); for (int k = 0; k < n; ++k) row[(size_t)(x++) * 4 + c] = (unsigned char)v; }
                    else { if (n == 0 || x + n > ww) throw std::runtime_error(path + ": bad run"); for (int k = 0; k < n; ++k) row[(size_t)(x++) * 4 + c] = (unsigned char)f.get(); }
                }
            }
        }
        if (!f) throw std::runtime_error(path + ": truncated");
        for (int x = 0; x < ww; ++x) { float *d = &im.rgba[((size_t)y * ww + x) * 4]; rgbe_decode(&row[(size_t)x * 4], d); d[3] = 1.f; }
    }
    return im;
}

// ---- PNG (write only) ------------------------------------------------------------------------------------------------------
static uint32_t crc32_update(uint32_t crc, const unsigned char *p, size_t n) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) { for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1; table[i] = c; } init = true; }
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return crc;
}
static void put_be32(std::vector<unsigned char> &v, uint32_t x) { for (int s = 24; s >= 0; s -= 8) v.push_back((unsigned char)(x >> s)); }
static void write_chunk(FILE *f, const char *tag, const std::vector<unsigned char> &data) {
    std::vector<unsigned char> head; put_be32(head, (uint32_t)data.size());
    std::fwrite(head.data(), 1, 4, f);
    std::vector<unsigned char> body(tag, tag + 4); body.insert(body.end(), data.begin(), data.end());
    std::fwrite(body.data(), 1, body.size(), f);
    std::vector<unsigned char> crc; put_be32(crc, crc32_update(0xffffffffu, body.data(), body.size()) ^ 0xffffffffu);
    std::fwrite(crc.data(), 1, 4, f);
}
// rgba01: display-referred values in [0, 1]; gamma = apply the sRGB transfer function before quantisation
void save_png(const std::string &path, const float *rgba01, uint32_t w, uint32_t h, bool gamma) {
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::fwrite(sig, 1, 8, f);
    std::vector<unsigned char> ihdr; put_be32(ihdr, w); put_be32(ihdr, h);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);     // 8-bit, colour type 2 (RGB)
    write_chunk(f, "IHDR", ihdr);
    std::vector<unsigned char> raw; raw.reserve((size_t)h * (1 + 3 * (size_t)w));
    for (uint32_t y = 0; y < h; ++y) {
        raw.push_back(0);                                           // filter type 0
        for (uint32_t x = 0; x < w; ++x)
            for (int c = 0; c < 3; ++c) {
                float v = rgba01[((size_t)y * w + x) * 4 + c];
                v = gamma ? oetf_srgb(v) : std::min(std::max(v, 0.f), 1.f);
                raw.push_back((unsigned char)(v * 255.f + 0.5f));
            }
    }
    std::vector<unsigned char> z{0x78, 0x01};                       // zlib header, then deflate "stored" blocks of <= 65535 bytes
    uint32_t a = 1, b = 0;
    for (unsigned char c : raw) { a = (a + c) % 65521u; b = (b + a) % 65521u; }
    for (size_t off = 0; off < raw.size() || off == 0; off += 65535) {
        size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n >= raw.size() ? 1 : 0);
        z.push_back((unsigned char)(n & 0xff)); z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 0xff)); z.push_back((unsigned char)((~n >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + (std::ptrdiff_t)off, raw.begin() + (std::ptrdiff_t)(off + n));
        if (raw.empty()) break;
    }
    put_be32(z, (b << 16) | a);
    write_chunk(f, "IDAT", z);
    write_chunk(f, "IEND", {});
    std::fclose(f);
}

// Pipeline::save_result / final_picture (pipeline.cpp:190-198,337-354): the tone-mapped picture gets the display gamma unless the
// file is .exr / .hdr (here: .hdr / .pfm, the float formats this build writes)
void save_image(const std::string &path, const float *rgba, uint32_t w, uint32_t h, bool display_referred) {
    if (display_referred && !ends_with_ci(path, ".png")) throw std::runtime_error("save_image: a gamma-corrected picture (render graph output) is written as .png only");
    if (ends_with_ci(path, ".hdr")) save_hdr(path, rgba, w, h);
    else if (ends_with_ci(path, ".pfm")) save_pfm(path, rgba, w, h);
    else if (ends_with_ci(path, ".png")) save_png(path, rgba, w, h, !display_referred);
    else if (ends_with_ci(path, ".ppm")) save_ppm(path, rgba, w, h);
    else if (ends_with_ci(path, ".exr")) save_exr(path, rgba, w, h, false);      // linear, like .hdr (pipeline.cpp:337-354)
    else throw std::runtime_error("save_image: unsupported output format '" + path + "' (.png, .ppm, .exr, .hdr, .pfm)");
}

}  // namespace vh
