// PNG in and out for the headless application layer.
//   gsb_png_write  = stbi_flip_vertically_on_write(true) + stbi_write_png(path, w, h, 4, fb, w * 4), the screenshot path of
//                    UiPanelToolsView::onButtonRenderSplats / onButtonRenderRtx (src/ui/tools/UiPanelToolsView.cpp:227-259)
//   gsb_png_read   = stbi_load(path, &w, &h, nullptr, STBI_rgb_alpha) for PNG textures (src/rtx/RtxHost.cpp:14-21)
// stb_image / stb_image_write are dependencies of the reference (lib/stb); the PNG container is an open format, so it is
// restated here over zlib's deflate / inflate.  Pixel parity (decode(write(x)) == x, decode == stb's decode) is the bar,
// not byte-identical files: the compressed stream is an encoder's choice.
#include "app_common.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <zlib.h>

namespace {

void put32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }

void chunk(std::vector<uint8_t>& out, const char type[4], const uint8_t* data, size_t n) {
    put32(out, (uint32_t)n);
    const size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    if (n) out.insert(out.end(), data, data + n);
    put32(out, (uint32_t)crc32(0L, out.data() + at, (uInt)(n + 4)));
}

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// filter one row with `type`; prev may be null (first row)
void filter_row(int type, const uint8_t* cur, const uint8_t* prev, size_t n, int bpp, uint8_t* out) {
    for (size_t i = 0; i < n; i++) {
        const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = prev ? prev[i] : 0, c = (prev && i >= (size_t)bpp) ? prev[i - bpp] : 0;
        int pred = 0;
        switch (type) {
            case 1: pred = a; break;
            case 2: pred = b; break;
            case 3: pred = (a + b) >> 1; break;
            case 4: pred = paeth(a, b, c); break;
            default: pred = 0;
        }
        out[i] = (uint8_t)(cur[i] - pred);
    }
}

}  // namespace

extern "C" {

void gsb_free(void* p) { free(p); }

int gsb_png_write(const char* path, const uint32_t* rgba, int width, int height, int flip_vertically) {
    if (!path || !rgba || width <= 0 || height <= 0) return gsbapp::fail("gsb_png_write: invalid argument");
    const size_t stride = (size_t)width * 4;
    std::vector<uint8_t> raw((stride + 1) * (size_t)height), trial(stride), best(stride);
    const uint8_t* src = reinterpret_cast<const uint8_t*>(rgba);        // little-endian uint32 RGBA8 = bytes R,G,B,A (make_rgba)
    for (int y = 0; y < height; y++) {
        const int sy = flip_vertically ? height - 1 - y : y;
        const int py = flip_vertically ? sy + 1 : sy - 1;                // the row written just before this one
        const uint8_t* cur = src + (size_t)sy * stride;
        const uint8_t* prev = y > 0 ? src + (size_t)py * stride : nullptr;
        int best_type = 0;
        long best_cost = -1;
        for (int type = 0; type < 5; type++) {                           // minimum sum of absolute differences, as stb does
            filter_row(type, cur, prev, stride, 4, trial.data());
            long cost = 0;
            for (size_t i = 0; i < stride; i++) cost += abs((int)(int8_t)trial[i]);
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_type = type; best.swap(trial); }
        }
        uint8_t* dst = raw.data() + (size_t)y * (stride + 1);
        dst[0] = (uint8_t)best_type;
        memcpy(dst + 1, best.data(), stride);
    }
    uLongf zlen = compressBound((uLong)raw.size());
    std::vector<uint8_t> z(zlen);
    if (compress2(z.data(), &zlen, raw.data(), (uLong)raw.size(), 6) != Z_OK) return gsbapp::fail("gsb_png_write: deflate failed");
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    std::vector<uint8_t> ihdr;
    put32(ihdr, (uint32_t)width); put32(ihdr, (uint32_t)height);
    ihdr.insert(ihdr.end(), {8, 6, 0, 0, 0});                           // 8 bit, RGBA, deflate, adaptive filtering, no interlace
    chunk(out, "IHDR", ihdr.data(), ihdr.size());
    chunk(out, "IDAT", z.data(), zlen);
    chunk(out, "IEND", nullptr, 0);
    FILE* f = fopen(path, "wb");
    if (!f) return gsbapp::fail(std::string("gsb_png_write: cannot open \"") + path + "\"");
    const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
    fclose(f);
    return ok ? GSB_OK : gsbapp::fail("gsb_png_write: short write");
}

int gsb_png_read(const char* path, uint8_t** rgba_out, int* width, int* height) {
    if (!path || !rgba_out || !width || !height) return gsbapp::fail("gsb_png_read: invalid argument");
    FILE* f = fopen(path, "rb");
    if (!f) return gsbapp::fail(std::string("gsb_png_read: cannot open \"") + path + "\"");
    std::vector<uint8_t> file;
    uint8_t buf[65536];
    size_t n;
    while ((n = fread(buf, 1, sizeof(buf), f)) > 0) file.insert(file.end(), buf, buf + n);
    fclose(f);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    if (file.size() < 8 || memcmp(file.data(), sig, 8) != 0) return gsbapp::fail("gsb_png_read: not a PNG file (only PNG textures are supported)");
    auto be32 = [&](size_t at) { return ((uint32_t)file[at] << 24) | ((uint32_t)file[at + 1] << 16) | ((uint32_t)file[at + 2] << 8) | file[at + 3]; };
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    for (size_t at = 8; at + 12 <= file.size();) {
        const uint32_t len = be32(at);
        if (at + 12 + (size_t)len > file.size()) return gsbapp::fail("gsb_png_read: truncated chunk");
        const char* type = reinterpret_cast<const char*>(&file[at + 4]);
        const uint8_t* data = &file[at + 8];
        if (!memcmp(type, "IHDR", 4) && len >= 13) { w = be32(at + 8); h = be32(at + 12); depth = data[8]; ctype = data[9]; interlace = data[12]; }
        else if (!memcmp(type, "PLTE", 4)) plte.assign(data, data + len);
        else if (!memcmp(type, "tRNS", 4)) trns.assign(data, data + len);
        else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
        else if (!memcmp(type, "IEND", 4)) break;
        at += 12 + (size_t)len;
    }
    if (w == 0 || h == 0 || w > (1u << 24) || h > (1u << 24)) return gsbapp::fail("gsb_png_read: bad image size");
    if (interlace) return gsbapp::fail("gsb_png_read: interlaced PNG files are not supported");
    int channels;
    switch (ctype) { case 0: channels = 1; break; case 2: channels = 3; break; case 3: channels = 1; break; case 4: channels = 2; break; case 6: channels = 4; break;
                     default: return gsbapp::fail("gsb_png_read: bad colour type"); }
    if (!(depth == 8 || depth == 16 || ((ctype == 0 || ctype == 3) && (depth == 1 || depth == 2 || depth == 4))))
        return gsbapp::fail("gsb_png_read: unsupported bit depth");
    if (ctype == 3 && depth == 16) return gsbapp::fail("gsb_png_read: bad palette depth");
    const size_t bits = (size_t)channels * depth, stride = (w * bits + 7) / 8, bpp = bits >= 8 ? bits / 8 : 1;
    std::vector<uint8_t> raw((stride + 1) * (size_t)h);
    uLongf rawlen = (uLongf)raw.size();
    const int zr = uncompress(raw.data(), &rawlen, idat.data(), (uLong)idat.size());
    if (zr != Z_OK || rawlen != raw.size()) return gsbapp::fail("gsb_png_read: inflate failed");
    // undo the row filters in place
    for (size_t y = 0; y < h; y++) {
        uint8_t* cur = raw.data() + y * (stride + 1) + 1;
        const uint8_t* prev = y ? cur - (stride + 1) : nullptr;
        const int type = cur[-1];
        if (type > 4) return gsbapp::fail("gsb_png_read: bad row filter");
        for (size_t i = 0; i < stride; i++) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev ? prev[i] : 0, c = (prev && i >= bpp) ? prev[i - bpp] : 0;
            int pred = 0;
            switch (type) { case 1: pred = a; break; case 2: pred = b; break; case 3: pred = (a + b) >> 1; break; case 4: pred = paeth(a, b, c); break; }
            cur[i] = (uint8_t)(cur[i] + pred);
        }
    }
    uint8_t* out = static_cast<uint8_t*>(malloc((size_t)w * h * 4));
    if (!out) return gsbapp::fail("gsb_png_read: out of memory");
    // sample s of pixel x in a row, as the raw value (0 .. 2^depth - 1; 16-bit keeps both bytes)
    auto sample = [&](const uint8_t* row, size_t x, int s) -> uint32_t {
        if (depth == 8) return row[x * channels + s];
        if (depth == 16) { const size_t k = (x * channels + s) * 2; return ((uint32_t)row[k] << 8) | row[k + 1]; }
        const size_t bit = x * depth;
        return (row[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1);
    };
    auto to8 = [&](uint32_t v) -> uint8_t {          // stb: 16 -> 8 keeps the high byte; 1/2/4-bit grey scales by 255 / (2^d - 1)
        if (depth == 16) return (uint8_t)(v >> 8);
        if (depth == 8) return (uint8_t)v;
        return (uint8_t)(v * (255u / ((1u << depth) - 1)));
    };
    for (size_t y = 0; y < h; y++) {
        const uint8_t* row = raw.data() + y * (stride + 1) + 1;
        for (size_t x = 0; x < w; x++) {
            uint8_t* px = out + (y * w + x) * 4;
            if (ctype == 3) {
                const uint32_t idx = sample(row, x, 0);
                if ((size_t)idx * 3 + 2 >= plte.size()) { free(out); return gsbapp::fail("gsb_png_read: palette index out of range"); }
                px[0] = plte[idx * 3]; px[1] = plte[idx * 3 + 1]; px[2] = plte[idx * 3 + 2];
                px[3] = idx < trns.size() ? trns[idx] : 255;
            } else if (ctype == 0 || ctype == 4) {
                const uint32_t g = sample(row, x, 0);
                px[0] = px[1] = px[2] = to8(g);
                px[3] = ctype == 4 ? to8(sample(row, x, 1)) : 255;
                if (ctype == 0 && trns.size() >= 2 && g == (((uint32_t)trns[0] << 8) | trns[1])) px[3] = 0;      // colour key
            } else {
                const uint32_t r = sample(row, x, 0), g = sample(row, x, 1), b = sample(row, x, 2);
                px[0] = to8(r); px[1] = to8(g); px[2] = to8(b);
                px[3] = ctype == 6 ? to8(sample(row, x, 3)) : 255;
                if (ctype == 2 && trns.size() >= 6 && r == (((uint32_t)trns[0] << 8) | trns[1]) && g == (((uint32_t)trns[2] << 8) | trns[3]) &&
                    b == (((uint32_t)trns[4] << 8) | trns[5])) px[3] = 0;
            }
        }
    }
    *rgba_out = out; *width = (int)w; *height = (int)h;
    return GSB_OK;
}

}  // extern "C"
