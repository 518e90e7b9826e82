/*
 * scene_io.cpp — scene files in, image files out (host only, no CUDA, no third-party code).
 *
 * Replaces, behind the C ABI of include/prt_b200.h (prt_model_*, prt_image_*, prt_write_*):
 *   Model::Model / ProcessObjFile / InitializeLightsRadiance / CreateMaterial
 *                                        Source/Model.cpp:53-394   (via tinyobjloader + tinyxml2)
 *   Camera::SetViewParametersByXmlFile   Source/Camera.cpp:339-389 (via tinyxml2)
 *   ImageTexture::ImageTexture           Source/Texture.cpp:10-21  (via stb_image)
 *   stbi_write_png / stbi_write_hdr      Source/Camera.cpp:302,327 (via stb_image_write)
 *
 * The reference's three parsers are third-party dependencies; tinyobjloader (pinned 3bb554cf) is
 * not even vendored in /root/reference. What matters for parity is restated here from their
 * published behaviour:
 *   - tinyobjloader's real-number parser (tryParseDouble: digits accumulated into a double,
 *     decimals through a power-of-ten table, exponent through pow(5, e) and ldexp) is NOT
 *     correctly rounded; vertex coordinates of the reference carry its bits, so it is the default
 *     here (PRT_LOAD_STRTOD selects strtod);
 *   - shapes: a `g` / `o` line closes the current shape if it has faces; `usemtl` does not;
 *   - faces: negative (relative) indices, v, v/vt, v//vn, v/vt/vn; triangles pass through,
 *     quads are split along the shorter diagonal, larger polygons as a fan;
 *   - materials: tinyobjloader's defaults (Kd = Ks = 0, Ns = 1), map_Kd / map_Ks with options
 *     skipped; tinyxml2's QueryDoubleAttribute / QueryIntAttribute are sscanf("%lf") / ("%d").
 * PNG / BMP / PNM decoding is lossless, so any correct decoder returns stb_image's bytes; JPEG is not, and
 * jpeg_decode.cpp restates stb_image's integer IDCT, chroma upsampling and colour conversion so that a .jpg map gives
 * the reference's texels (pinned against the stb_image.h the reference vendors). A texture that cannot be decoded
 * (CMYK / arithmetic-coded JPEG, other formats) is reported in prt_model_warnings and behaves like the reference's failed
 * load (cyan, Source/Texture.cpp:24).
 */
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <map>
#include <sstream>
#include <string>
#include <unordered_map>
#include <unordered_set>

#include "jpeg_decode.h"
#include <vector>

#include "../../include/prt_b200.h"
#include "scene_host.h"

namespace prt {
int set_error(int code, const std::string& msg); /* capi.cu: prt_last_error plumbing */
}

namespace {

using prt::set_error;

bool read_file(const std::string& path, std::vector<uint8_t>& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    size_t got = out.empty() ? 0 : std::fread(out.data(), 1, out.size(), f);
    std::fclose(f);
    return got == out.size();
}

/* ------------------------------------------------------------------------------------------------
 * tinyobjloader's tryParseDouble, restated (see the header comment). Returns false when [s, e) does
 * not start a number; tinyobjloader then keeps the caller's default (0.0).
 *
 * ATTRIBUTION: this one function follows tinyobjloader's digit-accumulating parser step for step (same state
 * machine, same order of floating-point operations), because the reference's vertex coordinates are the bits
 * THAT parser produces — strtod rounds correctly and differs in the last place on about one literal in ten.
 * tinyobjloader is Copyright (c) 2012-Present Syoyo Fujita and many contributors, released under the MIT License
 * (https://github.com/tinyobjloader/tinyobjloader, LICENSE: "Permission is hereby granted, free of charge, to any
 * person obtaining a copy of this software and associated documentation files (the "Software"), to deal in the
 * Software without restriction ... THE SOFTWARE IS PROVIDED "AS IS", WITHOUT WARRANTY OF ANY KIND ..."). The
 * reference pins it as an un-vendored submodule (absent from /root/reference), so its behaviour here is restated
 * from the published algorithm and cannot be pinned against the reference tree.
 * ---------------------------------------------------------------------------------------------- */
inline bool is_digit(char c) { return c >= '0' && c <= '9'; }

bool tinyobj_parse_double(const char* s, const char* s_end, double* result) {
    if (s >= s_end) return false;
    double mantissa = 0.0;
    int exponent = 0;
    char sign = '+', exp_sign = '+';
    const char* curr = s;
    int read = 0;
    bool end_not_reached = false;
    bool leading_decimal_dots = false;
    if (*curr == '+' || *curr == '-') {
        sign = *curr;
        curr++;
        if (curr != s_end && *curr == '.') leading_decimal_dots = true;
    } else if (is_digit(*curr)) {
    } else if (*curr == '.') {
        leading_decimal_dots = true;
    } else {
        return false;
    }
    end_not_reached = (curr != s_end);
    if (!leading_decimal_dots) {
        while (end_not_reached && is_digit(*curr)) {
            mantissa *= 10;
            mantissa += static_cast<int>(*curr - 0x30);
            curr++;
            read++;
            end_not_reached = (curr != s_end);
        }
        if (read == 0) return false;
    }
    if (!end_not_reached) goto assemble;
    if (*curr == '.') {
        curr++;
        read = 1;
        end_not_reached = (curr != s_end);
        while (end_not_reached && is_digit(*curr)) {
            static const double pow_lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
            const int lut_entries = sizeof pow_lut / sizeof pow_lut[0];
            mantissa += static_cast<int>(*curr - 0x30) * (read < lut_entries ? pow_lut[read] : std::pow(10.0, -read));
            read++;
            curr++;
            end_not_reached = (curr != s_end);
        }
    } else if (*curr == 'e' || *curr == 'E') {
    } else {
        goto assemble;
    }
    if (!end_not_reached) goto assemble;
    if (*curr == 'e' || *curr == 'E') {
        curr++;
        end_not_reached = (curr != s_end);
        if (end_not_reached && (*curr == '+' || *curr == '-')) {
            exp_sign = *curr;
            curr++;
        } else if (end_not_reached && is_digit(*curr)) {
        } else {
            return false;
        }
        read = 0;
        end_not_reached = (curr != s_end);
        while (end_not_reached && is_digit(*curr)) {
            if (exponent > std::numeric_limits<int>::max() / 10) return false;
            exponent *= 10;
            exponent += static_cast<int>(*curr - 0x30);
            curr++;
            read++;
            end_not_reached = (curr != s_end);
        }
        exponent *= (exp_sign == '+' ? 1 : -1);
        if (read == 0) return false;
    }
assemble:
    *result = (sign == '+' ? 1 : -1) * (exponent ? std::ldexp(mantissa * std::pow(5.0, exponent), exponent) : mantissa);
    return true;
}

/* tokenizer over one line */
struct Cursor {
    const char* p;
    const char* end;
    void skip_space() {
        while (p < end && (*p == ' ' || *p == '\t')) ++p;
    }
    bool at_end() {
        skip_space();
        return p >= end || *p == '\r' || *p == '\n';
    }
    /* next whitespace-delimited token */
    std::string token() {
        skip_space();
        const char* b = p;
        while (p < end && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') ++p;
        return std::string(b, p);
    }
    /* rest of the line, trailing blanks removed (names may contain spaces) */
    std::string rest() {
        skip_space();
        const char* b = p;
        const char* e = end;
        while (e > b && (e[-1] == ' ' || e[-1] == '\t' || e[-1] == '\r' || e[-1] == '\n')) --e;
        p = end;
        return std::string(b, e);
    }
    double real(bool use_strtod, double dflt = 0.0) {
        skip_space();
        const char* b = p;
        while (p < end && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') ++p;
        double v = dflt;
        if (use_strtod) {
            std::string t(b, p);
            char* q = nullptr;
            double x = std::strtod(t.c_str(), &q);
            if (q != t.c_str()) v = x;
        } else {
            double x;
            if (tinyobj_parse_double(b, p, &x)) v = x;
        }
        return v;
    }
};

/* ------------------------------------------------------------------------------------------------
 * inflate (RFC 1951) + zlib wrapper (RFC 1950), enough for PNG IDAT streams
 * ---------------------------------------------------------------------------------------------- */
struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint32_t buf = 0;
    int cnt = 0;
    bool fail = false;
    uint32_t bits(int n) {
        while (cnt < n) {
            if (p >= end) { fail = true; return 0; }
            buf |= (uint32_t)(*p++) << cnt;
            cnt += 8;
        }
        uint32_t v = n ? (buf & ((1u << n) - 1u)) : 0;
        buf >>= n;
        cnt -= n;
        return v;
    }
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
        for (int i = 1; i < 15; ++i) offs[i + 1] = offs[i] + count[i];
        for (int i = 0; i < n; ++i)
            if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
    }
    int decode(BitReader& br) const {
        int code = 0, first = 0, index = 0;
        for (int len = 1; len <= 15; ++len) {
            code |= (int)br.bits(1);
            if (br.fail) return -1;
            int c = count[len];
            if (code - c < first) return symbol[index + (code - first)];
            index += c;
            first += c;
            first <<= 1;
            code <<= 1;
        }
        return -1;
    }
};
bool inflate_raw(BitReader& br, std::vector<uint8_t>& out) {
    static const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const uint16_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    static const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static const uint16_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    for (;;) {
        const uint32_t last = br.bits(1), type = br.bits(2);
        if (br.fail) return false;
        if (type == 0) {
            br.buf = 0; br.cnt = 0; /* byte align */
            if (br.end - br.p < 4) return false;
            const uint32_t len = br.p[0] | (br.p[1] << 8), nlen = br.p[2] | (br.p[3] << 8);
            br.p += 4;
            if ((len ^ 0xffffu) != nlen || (size_t)(br.end - br.p) < len) return false;
            out.insert(out.end(), br.p, br.p + len);
            br.p += len;
        } else if (type == 1 || type == 2) {
            Huffman hl, hd;
            uint8_t lengths[320];
            if (type == 1) {
                int i = 0;
                for (; i < 144; ++i) lengths[i] = 8;
                for (; i < 256; ++i) lengths[i] = 9;
                for (; i < 280; ++i) lengths[i] = 7;
                for (; i < 288; ++i) lengths[i] = 8;
                hl.build(lengths, 288);
                for (i = 0; i < 30; ++i) lengths[i] = 5;
                hd.build(lengths, 30);
            } else {
                static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
                const int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
                if (br.fail || nlen > 286 || ndist > 30) return false;
                uint8_t cl[19];
                std::memset(cl, 0, sizeof(cl));
                for (int i = 0; i < ncode; ++i) cl[order[i]] = (uint8_t)br.bits(3);
                Huffman hc;
                hc.build(cl, 19);
                int idx = 0;
                while (idx < nlen + ndist) {
                    int sym = hc.decode(br);
                    if (sym < 0) return false;
                    if (sym < 16) {
                        lengths[idx++] = (uint8_t)sym;
                    } else {
                        uint8_t prev = 0;
                        int rep;
                        if (sym == 16) {
                            if (idx == 0) return false;
                            prev = lengths[idx - 1];
                            rep = 3 + (int)br.bits(2);
                        } else if (sym == 17) {
                            rep = 3 + (int)br.bits(3);
                        } else {
                            rep = 11 + (int)br.bits(7);
                        }
                        if (idx + rep > nlen + ndist) return false;
                        while (rep--) lengths[idx++] = prev;
                    }
                }
                hl.build(lengths, nlen);
                hd.build(lengths + nlen, ndist);
            }
            for (;;) {
                int sym = hl.decode(br);
                if (sym < 0) return false;
                if (sym < 256) {
                    out.push_back((uint8_t)sym);
                } else if (sym == 256) {
                    break;
                } else {
                    sym -= 257;
                    if (sym >= 29) return false;
                    const int len = lbase[sym] + (int)br.bits(lext[sym]);
                    const int ds = hd.decode(br);
                    if (ds < 0 || ds >= 30) return false;
                    const size_t dist = dbase[ds] + br.bits(dext[ds]);
                    if (br.fail || dist > out.size()) return false;
                    const size_t from = out.size() - dist;
                    for (int i = 0; i < len; ++i) out.push_back(out[from + i]);
                }
            }
        } else {
            return false;
        }
        if (last) break;
    }
    return !br.fail;
}
bool zlib_inflate(const uint8_t* data, size_t n, std::vector<uint8_t>& out) {
    if (n < 6) return false;
    if ((data[0] & 0x0f) != 8 || ((data[0] << 8) | data[1]) % 31 != 0 || (data[1] & 0x20)) return false;
    BitReader br{data + 2, data + n};
    return inflate_raw(br, out);
}

uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

/* ------------------------------------------------------------------------------------------------
 * image decoding with stbi_load(path, &w, &h, &channels, 0) semantics (Source/Texture.cpp:12)
 * ---------------------------------------------------------------------------------------------- */
struct Image {
    int w = 0, h = 0, c = 0;
    std::vector<uint8_t> px;
};

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    if (pa <= pb && pa <= pc) return a;
    if (pb <= pc) return b;
    return c;
}

bool decode_png(const std::vector<uint8_t>& f, Image& im, std::string& why) {
    static const uint8_t sig[8] = {137, 80, 78, 71, 13, 10, 26, 10};
    if (f.size() < 8 || std::memcmp(f.data(), sig, 8) != 0) { why = "not a PNG"; return false; }
    size_t pos = 8;
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    bool seen_ihdr = false;
    while (pos + 8 <= f.size()) {
        const uint32_t len = be32(&f[pos]);
        const char* ty = (const char*)&f[pos + 4];
        if (pos + 12 + (size_t)len > f.size()) { why = "truncated PNG chunk"; return false; }
        const uint8_t* d = &f[pos + 8];
        if (!std::memcmp(ty, "IHDR", 4)) {
            if (len < 13) { why = "bad IHDR"; return false; }
            im.w = (int)be32(d); im.h = (int)be32(d + 4);
            depth = d[8]; ctype = d[9]; interlace = d[12];
            seen_ihdr = true;
        } else if (!std::memcmp(ty, "PLTE", 4)) {
            plte.assign(d, d + len);
        } else if (!std::memcmp(ty, "tRNS", 4)) {
            trns.assign(d, d + len);
        } else if (!std::memcmp(ty, "IDAT", 4)) {
            idat.insert(idat.end(), d, d + len);
        } else if (!std::memcmp(ty, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)len;
    }
    if (!seen_ihdr || im.w <= 0 || im.h <= 0) { why = "PNG without IHDR"; return false; }
    if (interlace) { why = "interlaced PNG is not supported"; return false; }
    int src_c;
    switch (ctype) {
        case 0: src_c = 1; break;
        case 2: src_c = 3; break;
        case 3: src_c = 1; break;
        case 4: src_c = 2; break;
        case 6: src_c = 4; break;
        default: why = "bad PNG colour type"; return false;
    }
    if (!(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) { why = "bad PNG bit depth"; return false; }
    std::vector<uint8_t> raw;
    if (!zlib_inflate(idat.data(), idat.size(), raw)) { why = "corrupt PNG data (inflate)"; return false; }
    const size_t bpp_bits = (size_t)src_c * depth;
    const size_t stride = ((size_t)im.w * bpp_bits + 7) / 8;
    const size_t fbpp = std::max<size_t>(1, bpp_bits / 8);
    if (raw.size() < (stride + 1) * (size_t)im.h) { why = "PNG data too short"; return false; }
    std::vector<uint8_t> cur(stride), prev(stride, 0), rows(stride * (size_t)im.h);
    for (int y = 0; y < im.h; ++y) {
        const uint8_t* src = &raw[(stride + 1) * (size_t)y];
        const int ft = src[0];
        ++src;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= fbpp ? cur[i - fbpp] : 0, b = prev[i], c = i >= fbpp ? prev[i - fbpp] : 0;
            int v;
            switch (ft) {
                case 0: v = src[i]; break;
                case 1: v = src[i] + a; break;
                case 2: v = src[i] + b; break;
                case 3: v = src[i] + ((a + b) >> 1); break;
                case 4: v = src[i] + paeth(a, b, c); break;
                default: why = "bad PNG filter"; return false;
            }
            cur[i] = (uint8_t)v;
        }
        std::memcpy(&rows[stride * (size_t)y], cur.data(), stride);
        prev = cur;
    }
    /* expand to 8-bit samples (16-bit keeps the high byte, like stbi__convert_16_to_8) */
    const size_t npx = (size_t)im.w * im.h;
    std::vector<uint8_t> s8(npx * src_c);
    static const int depth_scale[9] = {0, 0xff, 0x55, 0, 0x11, 0, 0, 0, 0x01};
    for (int y = 0; y < im.h; ++y) {
        const uint8_t* r = &rows[stride * (size_t)y];
        uint8_t* o = &s8[(size_t)y * im.w * src_c];
        const size_t n = (size_t)im.w * src_c;
        if (depth == 8) {
            std::memcpy(o, r, n);
        } else if (depth == 16) {
            for (size_t i = 0; i < n; ++i) o[i] = r[2 * i];
        } else {
            const int scale = ctype == 3 ? 1 : depth_scale[depth];
            const int mask = (1 << depth) - 1;
            for (size_t i = 0; i < n; ++i) {
                const size_t bit = i * depth;
                const int v = (r[bit >> 3] >> (8 - depth - (int)(bit & 7))) & mask;
                o[i] = (uint8_t)(v * scale);
            }
        }
    }
    if (ctype == 3) {
        const int oc = trns.empty() ? 3 : 4;
        im.c = oc;
        im.px.resize(npx * oc);
        for (size_t i = 0; i < npx; ++i) {
            const size_t k = s8[i];
            for (int ch = 0; ch < 3; ++ch) im.px[i * oc + ch] = 3 * k + ch < plte.size() ? plte[3 * k + ch] : 0;
            if (oc == 4) im.px[i * 4 + 3] = k < trns.size() ? trns[k] : 255;
        }
    } else if (!trns.empty() && (ctype == 0 || ctype == 2)) {
        /* colour-key transparency adds an alpha channel */
        const int oc = src_c + 1;
        im.c = oc;
        im.px.resize(npx * oc);
        uint8_t key[3] = {0, 0, 0};
        for (int ch = 0; ch < src_c && (size_t)(2 * ch + 1) < trns.size(); ++ch) {
            const int v16 = (trns[2 * ch] << 8) | trns[2 * ch + 1];
            key[ch] = depth == 16 ? (uint8_t)(v16 >> 8) : (uint8_t)((v16 & 255) * (depth < 8 ? depth_scale[depth] : 1));
        }
        for (size_t i = 0; i < npx; ++i) {
            bool match = true;
            for (int ch = 0; ch < src_c; ++ch) {
                im.px[i * oc + ch] = s8[i * src_c + ch];
                match = match && s8[i * src_c + ch] == key[ch];
            }
            im.px[i * oc + src_c] = match ? 0 : 255;
        }
    } else {
        im.c = src_c;
        im.px.swap(s8);
    }
    return true;
}

bool decode_bmp(const std::vector<uint8_t>& f, Image& im, std::string& why) {
    if (f.size() < 54 || f[0] != 'B' || f[1] != 'M') { why = "not a BMP"; return false; }
    auto le32 = [&](size_t o) { return (int32_t)(f[o] | (f[o + 1] << 8) | (f[o + 2] << 16) | ((uint32_t)f[o + 3] << 24)); };
    const int32_t off = le32(10), hsz = le32(14), w = le32(18), h0 = le32(22);
    const int bpp = f[28] | (f[29] << 8), comp = le32(30);
    if (hsz < 40 || (bpp != 24 && bpp != 32) || (comp != 0 && comp != 3) || w <= 0 || h0 == 0) { why = "unsupported BMP variant"; return false; }
    const bool flip = h0 > 0;
    const int h = std::abs(h0);
    const size_t stride = ((size_t)w * (bpp / 8) + 3) & ~(size_t)3;
    if ((size_t)off + stride * h > f.size()) { why = "truncated BMP"; return false; }
    im.w = w; im.h = h; im.c = bpp == 32 ? 4 : 3;
    im.px.resize((size_t)w * h * im.c);
    for (int y = 0; y < h; ++y) {
        const uint8_t* r = &f[(size_t)off + stride * (size_t)(flip ? h - 1 - y : y)];
        for (int x = 0; x < w; ++x) {
            uint8_t* o = &im.px[((size_t)y * w + x) * im.c];
            o[0] = r[x * (bpp / 8) + 2]; o[1] = r[x * (bpp / 8) + 1]; o[2] = r[x * (bpp / 8)];
            if (im.c == 4) o[3] = r[x * 4 + 3];
        }
    }
    return true;
}

bool decode_pnm(const std::vector<uint8_t>& f, Image& im, std::string& why) {
    if (f.size() < 7 || f[0] != 'P' || (f[1] != '5' && f[1] != '6')) { why = "not a binary PNM"; return false; }
    size_t pos = 2;
    int vals[3], got = 0;
    while (got < 3 && pos < f.size()) {
        while (pos < f.size() && (f[pos] == ' ' || f[pos] == '\n' || f[pos] == '\r' || f[pos] == '\t')) ++pos;
        if (pos < f.size() && f[pos] == '#') { while (pos < f.size() && f[pos] != '\n') ++pos; continue; }
        int v = 0, nd = 0;
        while (pos < f.size() && is_digit((char)f[pos])) { v = v * 10 + (f[pos] - '0'); ++pos; ++nd; }
        if (!nd) { why = "bad PNM header"; return false; }
        vals[got++] = v;
    }
    ++pos; /* single whitespace after maxval */
    if (got < 3 || vals[2] != 255) { why = "only 8-bit PNM is supported"; return false; }
    im.w = vals[0]; im.h = vals[1]; im.c = f[1] == '6' ? 3 : 1;
    const size_t n = (size_t)im.w * im.h * im.c;
    if (pos + n > f.size()) { why = "truncated PNM"; return false; }
    im.px.assign(f.begin() + pos, f.begin() + pos + n);
    return true;
}

bool load_image(const std::string& path, Image& im, std::string& why) {
    std::vector<uint8_t> f;
    if (!read_file(path, f)) { why = "cannot open " + path; return false; }
    if (f.size() >= 8 && f[0] == 137 && f[1] == 'P') return decode_png(f, im, why);
    if (f.size() >= 2 && f[0] == 'B' && f[1] == 'M') return decode_bmp(f, im, why);
    if (f.size() >= 2 && f[0] == 'P' && (f[1] == '5' || f[1] == '6')) return decode_pnm(f, im, why);
    if (f.size() >= 3 && f[0] == 0xff && f[1] == 0xd8) {
        if (prt::decode_jpeg(f.data(), f.size(), im.w, im.h, im.c, im.px, why)) return true;
        why += " (" + path + ")";
        return false;
    }
    why = "unknown image format (" + path + ")";
    return false;
}

/* ------------------------------------------------------------------------------------------------
 * XML: the subset the scene files use (a forest of elements with attributes; comments, the
 * declaration and character data are skipped). tinyxml2 semantics kept: FirstChildElement(name)
 * is the first element of that name in document order, attributes are strings, QueryXAttribute is
 * sscanf.
 * ---------------------------------------------------------------------------------------------- */
struct XmlElement {
    std::string name;
    std::vector<std::pair<std::string, std::string>> attrs;
    std::vector<XmlElement> children;
    const char* attr(const char* key) const {
        for (auto& a : attrs)
            if (a.first == key) return a.second.c_str();
        return nullptr;
    }
    const XmlElement* child(const char* key) const {
        for (auto& c : children)
            if (c.name == key) return &c;
        return nullptr;
    }
    bool query_double(const char* key, double* out) const { /* QueryDoubleAttribute: sscanf("%lf") */
        const char* v = attr(key);
        return v && std::sscanf(v, "%lf", out) == 1;
    }
    bool query_int(const char* key, int* out) const { /* QueryIntAttribute: sscanf("%d") */
        const char* v = attr(key);
        return v && std::sscanf(v, "%d", out) == 1;
    }
};

std::string xml_unescape(const std::string& s) {
    std::string o;
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] == '&') {
            static const char* ent[5][2] = {{"&lt;", "<"}, {"&gt;", ">"}, {"&amp;", "&"}, {"&quot;", "\""}, {"&apos;", "'"}};
            bool done = false;
            for (auto& e : ent) {
                const size_t n = std::strlen(e[0]);
                if (s.compare(i, n, e[0]) == 0) { o += e[1]; i += n - 1; done = true; break; }
            }
            if (done) continue;
        }
        o += s[i];
    }
    return o;
}

struct XmlParser {
    const std::string& s;
    size_t i = 0;
    explicit XmlParser(const std::string& text) : s(text) {}
    void skip_ws() { while (i < s.size() && std::isspace((unsigned char)s[i])) ++i; }
    bool starts(const char* t) const { return s.compare(i, std::strlen(t), t) == 0; }
    /* skips text, comments, declarations, doctype; stops at '<' of an element or a close tag */
    void skip_misc() {
        for (;;) {
            while (i < s.size() && s[i] != '<') ++i;
            if (i >= s.size()) return;
            if (starts("<!--")) { size_t e = s.find("-->", i + 4); i = e == std::string::npos ? s.size() : e + 3; continue; }
            if (starts("<?")) { size_t e = s.find("?>", i + 2); i = e == std::string::npos ? s.size() : e + 2; continue; }
            if (starts("<![CDATA[")) { size_t e = s.find("]]>", i); i = e == std::string::npos ? s.size() : e + 3; continue; }
            if (starts("<!")) { size_t e = s.find('>', i); i = e == std::string::npos ? s.size() : e + 1; continue; }
            return;
        }
    }
    std::string name() {
        size_t b = i;
        while (i < s.size() && !std::isspace((unsigned char)s[i]) && s[i] != '>' && s[i] != '/' && s[i] != '=') ++i;
        return s.substr(b, i - b);
    }
    bool element(XmlElement& el) { /* at '<' of an open tag */
        ++i;
        el.name = name();
        for (;;) {
            skip_ws();
            if (i >= s.size()) return false;
            if (s[i] == '/') { i = std::min(s.size(), i + 2); return true; } /* <a ... /> */
            if (s[i] == '>') { ++i; break; }
            std::string key = name();
            skip_ws();
            if (i < s.size() && s[i] == '=') {
                ++i;
                skip_ws();
                if (i < s.size() && (s[i] == '"' || s[i] == '\'')) {
                    const char q = s[i++];
                    size_t e = s.find(q, i);
                    if (e == std::string::npos) return false;
                    el.attrs.emplace_back(key, xml_unescape(s.substr(i, e - i)));
                    i = e + 1;
                }
            } else if (key.empty()) {
                ++i; /* stray character */
            }
        }
        for (;;) { /* children until the matching close tag */
            skip_misc();
            if (i >= s.size()) return true;
            if (starts("</")) { size_t e = s.find('>', i); i = e == std::string::npos ? s.size() : e + 1; return true; }
            XmlElement c;
            if (!element(c)) return false;
            el.children.push_back(std::move(c));
        }
    }
    std::vector<XmlElement> forest() {
        std::vector<XmlElement> top;
        for (;;) {
            skip_misc();
            if (i >= s.size()) break;
            if (starts("</")) { size_t e = s.find('>', i); i = e == std::string::npos ? s.size() : e + 1; continue; }
            XmlElement el;
            if (!element(el)) break;
            top.push_back(std::move(el));
        }
        return top;
    }
};

/* ------------------------------------------------------------------------------------------------
 * the reference's material name table, Source/Model.cpp:16-51
 * ---------------------------------------------------------------------------------------------- */
int material_type_of(const std::string& name) {
    static const std::unordered_map<std::string, int> table = {
        {"material0", PRT_MAT_PHONG}, {"material1", PRT_MAT_PHONG}, {"material2", PRT_MAT_PHONG}, {"material3", PRT_MAT_PHONG},
        {"material4", PRT_MAT_PHONG}, {"light1", PRT_MAT_DIFFUSE_LIGHT}, {"light2", PRT_MAT_DIFFUSE_LIGHT},
        {"light3", PRT_MAT_DIFFUSE_LIGHT}, {"light4", PRT_MAT_DIFFUSE_LIGHT},
        {"DiffuseWhite", PRT_MAT_LAMBERTIAN}, {"DiffuseBall", PRT_MAT_LAMBERTIAN}, {"DiffuseYellow", PRT_MAT_LAMBERTIAN},
        {"LeftWall", PRT_MAT_LAMBERTIAN}, {"RightWall", PRT_MAT_LAMBERTIAN}, {"Light", PRT_MAT_DIFFUSE_LIGHT},
        {"Wall", PRT_MAT_LAMBERTIAN}, {"quad1", PRT_MAT_EMPTY}, {"Mirror", PRT_MAT_MIRROR}, {"StainlessRough", PRT_MAT_LAMBERTIAN},
        {"Towel", PRT_MAT_LAMBERTIAN}, {"BlackWoodLacquer", PRT_MAT_LAMBERTIAN}, {"Wood", PRT_MAT_LAMBERTIAN},
        {"WoodFloor", PRT_MAT_LAMBERTIAN}, {"Label", PRT_MAT_LAMBERTIAN}, {"RoughGlass", PRT_MAT_LAMBERTIAN},
        {"Plastic", PRT_MAT_LAMBERTIAN}, {"DarkPlastic", PRT_MAT_LAMBERTIAN}, {"Bin", PRT_MAT_MIRROR},
        {"WallRight", PRT_MAT_LAMBERTIAN}, {"DarkBorder", PRT_MAT_LAMBERTIAN}, {"Trims", PRT_MAT_LAMBERTIAN},
        {"Ceramic", PRT_MAT_LAMBERTIAN}};
    auto it = table.find(name);
    return it == table.end() ? PRT_MAT_LAMBERTIAN : it->second; /* Model.cpp:274-277 */
}

/* ------------------------------------------------------------------------------------------------
 * MTL (tinyobjloader's LoadMtl, the fields the reference reads: name, Kd, Ks, Ns, map_Kd, map_Ks)
 * ---------------------------------------------------------------------------------------------- */
struct RawMaterial {
    std::string name, diffuse_texname, specular_texname;
    double diffuse[3] = {0, 0, 0}, specular[3] = {0, 0, 0};
    double shininess = 1.0;
};

/* texture name of a map_* statement: options (-blendu on, -s u v w, ...) are skipped */
std::string parse_texname(Cursor c) {
    static const std::map<std::string, int> nargs = {{"-blendu", 1}, {"-blendv", 1}, {"-clamp", 1}, {"-boost", 1}, {"-bm", 1},
                                                     {"-o", 3}, {"-s", 3}, {"-t", 3}, {"-type", 1}, {"-imfchan", 1}, {"-mm", 2},
                                                     {"-colorspace", 1}, {"-texres", 1}, {"-cc", 1}};
    for (;;) {
        Cursor save = c;
        std::string t = c.token();
        if (t.empty()) return "";
        auto it = nargs.find(t);
        if (it == nargs.end()) return save.rest(); /* the file name is the rest of the line */
        for (int k = 0; k < it->second; ++k) c.token();
    }
}

void parse_mtl(const std::string& text, bool use_strtod, std::vector<RawMaterial>& out) {
    RawMaterial cur;
    bool have = false;
    size_t pos = 0;
    while (pos < text.size()) {
        size_t e = text.find('\n', pos);
        if (e == std::string::npos) e = text.size();
        Cursor c{text.data() + pos, text.data() + e};
        pos = e + 1;
        if (c.at_end() || *c.p == '#') continue;
        std::string key = c.token();
        if (key == "newmtl") {
            if (have) out.push_back(cur);
            cur = RawMaterial();
            cur.name = c.rest();
            have = true;
        } else if (key == "Kd") {
            for (int k = 0; k < 3; ++k) cur.diffuse[k] = c.real(use_strtod);
        } else if (key == "Ks") {
            for (int k = 0; k < 3; ++k) cur.specular[k] = c.real(use_strtod);
        } else if (key == "Ns") {
            cur.shininess = c.real(use_strtod);
        } else if (key == "map_Kd") {
            cur.diffuse_texname = parse_texname(c);
        } else if (key == "map_Ks") {
            cur.specular_texname = parse_texname(c);
        }
    }
    if (have) out.push_back(cur);
}

/* ------------------------------------------------------------------------------------------------
 * OBJ (tinyobjloader's LoadObj with its default ObjReaderConfig: triangulate = true)
 * ---------------------------------------------------------------------------------------------- */
struct ObjIndex { int v, vt, vn; };
struct ObjShape {
    std::string name;
    std::vector<ObjIndex> indices; /* 3 per triangle */
    std::vector<int> material_ids; /* per triangle */
};
struct ObjData {
    std::vector<double> v, vn, vt;
    std::vector<ObjShape> shapes;
    std::vector<RawMaterial> materials;
};

/* fixIndex: 1-based -> 0-based, negative = relative to the current count, 0 is invalid */
bool fix_index(int idx, int n, int* out) {
    if (idx > 0) { *out = idx - 1; return true; }
    if (idx == 0) return false;
    *out = n + idx;
    return true;
}
bool parse_triple(Cursor& c, int nv, int nvt, int nvn, ObjIndex* out) {
    std::string t = c.token();
    if (t.empty()) return false;
    out->v = out->vt = out->vn = -1;
    const char* p = t.c_str();
    char* q;
    long a = std::strtol(p, &q, 10);
    if (q == p || !fix_index((int)a, nv, &out->v)) return false;
    if (*q != '/') return true;
    p = q + 1;
    if (*p == '/') { /* v//vn */
        ++p;
        long n = std::strtol(p, &q, 10);
        if (q == p || !fix_index((int)n, nvn, &out->vn)) return false;
        return true;
    }
    long b = std::strtol(p, &q, 10);
    if (q == p || !fix_index((int)b, nvt, &out->vt)) return false;
    if (*q != '/') return true;
    p = q + 1;
    long n = std::strtol(p, &q, 10);
    if (q == p || !fix_index((int)n, nvn, &out->vn)) return false;
    return true;
}

struct ObjParser {
    ObjData& d;
    bool use_strtod;
    std::vector<std::string>& warnings;
    const std::string& dir;
    std::unordered_map<std::string, int> material_map;
    ObjShape shape;
    std::string name;
    int material = -1;
    std::unordered_set<std::string> group_names;
    int anonymous = -1;

    void add_face(const std::vector<ObjIndex>& f) {
        auto tri = [&](int a, int b, int c) {
            shape.indices.push_back(f[a]); shape.indices.push_back(f[b]); shape.indices.push_back(f[c]);
            shape.material_ids.push_back(material);
        };
        if (f.size() == 3) {
            tri(0, 1, 2);
        } else if (f.size() == 4) {
            /* split along the shorter diagonal */
            auto P = [&](int k, int ax) { const size_t i = (size_t)f[k].v; return 3 * i + ax < d.v.size() ? d.v[3 * i + ax] : 0.0; };
            double s02 = 0, s13 = 0;
            for (int ax = 0; ax < 3; ++ax) {
                const double e02 = P(2, ax) - P(0, ax), e13 = P(3, ax) - P(1, ax);
                s02 += e02 * e02;
                s13 += e13 * e13;
            }
            if (s02 < s13) { tri(0, 1, 2); tri(0, 2, 3); }
            else { tri(0, 1, 3); tri(1, 2, 3); }
        } else {
            for (size_t k = 1; k + 1 < f.size(); ++k) tri(0, (int)k, (int)k + 1);
        }
    }
    void flush_shape() {
        if (!shape.indices.empty()) {
            shape.name = name;
            d.shapes.push_back(std::move(shape));
        }
        shape = ObjShape();
    }
    std::string fresh_group_name() { /* Model.cpp:222-227 */
        std::string n;
        do { n = "Group_" + std::to_string(++anonymous); } while (group_names.count(n));
        group_names.insert(n);
        return n;
    }
    void parse(const std::string& text) {
        /* pass 1 of the reference (ProcessObjFile) collects every non-empty group name before it
         * generates names, but generates while scanning: a later explicit "Group_0" does not stop an
         * earlier anonymous group from taking that name. Same order here. */
        size_t pos = 0;
        while (pos < text.size()) {
            size_t e = text.find('\n', pos);
            if (e == std::string::npos) e = text.size();
            const char* lb = text.data() + pos;
            const char* le = text.data() + e;
            pos = e + 1;
            while (le > lb && (le[-1] == '\r')) --le;
            Cursor c{lb, le};
            if (c.at_end() || *c.p == '#') continue;
            std::string key = c.token();
            if (key == "v") {
                for (int k = 0; k < 3; ++k) d.v.push_back(c.real(use_strtod));
            } else if (key == "vn") {
                for (int k = 0; k < 3; ++k) d.vn.push_back(c.real(use_strtod));
            } else if (key == "vt") {
                for (int k = 0; k < 2; ++k) d.vt.push_back(c.real(use_strtod));
            } else if (key == "f") {
                std::vector<ObjIndex> f;
                bool ok = true;
                while (!c.at_end()) {
                    ObjIndex ix;
                    if (!parse_triple(c, (int)(d.v.size() / 3), (int)(d.vt.size() / 2), (int)(d.vn.size() / 3), &ix)) { ok = false; break; }
                    f.push_back(ix);
                }
                if (!ok || f.size() < 3) { warnings.push_back("face with an invalid vertex index or fewer than 3 vertices skipped"); continue; }
                add_face(f);
            } else if (key == "usemtl") {
                std::string m = c.token();
                auto it = material_map.find(m);
                int id = -1;
                if (it != material_map.end()) id = it->second;
                else warnings.push_back("material [ '" + m + "' ] not found in .mtl");
                material = id; /* a material change does not start a new shape */
            } else if (key == "mtllib") {
                while (!c.at_end()) {
                    std::string fn = c.token();
                    std::vector<uint8_t> raw;
                    if (!read_file(dir + "/" + fn, raw)) { warnings.push_back("material file [ " + fn + " ] not found"); continue; }
                    std::string mt((const char*)raw.data(), raw.size());
                    const size_t before = d.materials.size();
                    parse_mtl(mt, use_strtod, d.materials);
                    for (size_t i = before; i < d.materials.size(); ++i) material_map.insert({d.materials[i].name, (int)i});
                    break; /* the first file that loads wins */
                }
            } else if (key == "g") {
                flush_shape();
                std::string joined;
                while (!c.at_end()) {
                    std::string t = c.token();
                    joined += (joined.empty() ? "" : " ") + t;
                }
                if (joined.empty()) joined = fresh_group_name(); /* Model.cpp:213-232, in memory */
                else group_names.insert(joined);
                name = joined;
            } else if (key == "o") {
                flush_shape();
                name = c.rest();
            }
        }
        flush_shape();
    }
};

}  // namespace

/* ------------------------------------------------------------------------------------------------
 * prt_model: Model::Model (Source/Model.cpp:53-193) as flat arrays
 * ---------------------------------------------------------------------------------------------- */
struct prt_model {
    std::vector<double> vertices, normals, texcoords;
    std::vector<int64_t> mesh_first;
    std::vector<int32_t> mesh_material;
    std::vector<std::string> mesh_names, material_names;
    std::vector<prt_material> materials;
    std::vector<Image> images;
    std::vector<prt_texture> textures;
    prt_scene_desc desc;
    prt_camera camera;
    bool has_camera = false;
    std::string warnings;
};

extern "C" {

int prt_model_load(const char* directory, const char* name, uint32_t flags, prt_model** out) {
    if (!directory || !name || !out) return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_model_load: null argument");
    *out = nullptr;
    const std::string dir = directory, base = name;
    const bool use_strtod = (flags & PRT_LOAD_STRTOD) != 0;
    std::vector<std::string> warn;
    std::vector<uint8_t> raw;

    /* ---- <name>.xml: light radiance (Model.cpp:320-347) and camera (Camera.cpp:339-389) ---- */
    std::unordered_map<std::string, std::array<double, 3>> radiance;
    prt_model* m = new prt_model();
    std::memset(&m->camera, 0, sizeof(m->camera));
    if (read_file(dir + "/" + base + ".xml", raw)) {
        std::string text((const char*)raw.data(), raw.size());
        XmlParser xp(text);
        std::vector<XmlElement> top = xp.forest();
        for (const XmlElement& el : top) {
            if (el.name != "light") continue;
            const char* mtl = el.attr("mtlname");
            const char* rad = el.attr("radiance");
            if (mtl && rad) {
                /* ParseRadianceStr: `ss >> x >> comma >> y >> comma >> z` */
                double x = 0., y = 0., z = 0.;
                std::stringstream ss(rad);
                char comma;
                ss >> x >> comma >> y >> comma >> z;
                radiance[mtl] = {x, y, z};
            } else {
                warn.push_back("<light> without mtlname / radiance ignored");
            }
        }
        for (const XmlElement& el : top) {
            if (el.name != "camera") continue; /* FirstChildElement("camera") */
            int w = 0, h = 0;
            double fovy = 0.0;
            el.query_int("width", &w);
            el.query_int("height", &h);
            el.query_double("fovy", &fovy);
            m->camera.image_width = w;
            m->camera.image_height = h;
            m->camera.fovy = fovy;
            /* Camera.h defaults for absent children */
            const double deye[3] = {0, 0, 0}, dlook[3] = {0, 0, -1}, dup[3] = {0, 1, 0};
            for (int k = 0; k < 3; ++k) { m->camera.eye[k] = deye[k]; m->camera.look_at[k] = dlook[k]; m->camera.up[k] = dup[k]; }
            const char* axes[3] = {"x", "y", "z"};
            if (const XmlElement* e = el.child("eye")) for (int k = 0; k < 3; ++k) e->query_double(axes[k], &m->camera.eye[k]);
            if (const XmlElement* e = el.child("lookat")) for (int k = 0; k < 3; ++k) e->query_double(axes[k], &m->camera.look_at[k]);
            if (const XmlElement* e = el.child("up")) for (int k = 0; k < 3; ++k) e->query_double(axes[k], &m->camera.up[k]);
            m->has_camera = true;
            break;
        }
    } else {
        warn.push_back("Failed to load XML file: " + dir + "/" + base + ".xml");
    }

    /* ---- <name>.obj + .mtl ---- */
    if (!read_file(dir + "/" + base + ".obj", raw)) {
        delete m;
        return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_model_load: cannot open " + dir + "/" + base + ".obj");
    }
    ObjData od;
    {
        std::string text((const char*)raw.data(), raw.size());
        ObjParser op{od, use_strtod, warn, dir};
        op.parse(text);
    }

    /* ---- textures, keyed by the diffuse texture name in first-appearance order (Model.cpp:79-97) ---- */
    std::unordered_map<std::string, int> tex_index;
    for (const RawMaterial& rm : od.materials) {
        if (rm.diffuse_texname.empty() || tex_index.count(rm.diffuse_texname)) continue;
        Image im;
        std::string why;
        std::string fn = rm.diffuse_texname;
        std::replace(fn.begin(), fn.end(), '\\', '/');
        if (!load_image(dir + "/" + fn, im, why)) {
            warn.push_back("Loading Texture: " + dir + "/" + fn + " Failed!! (" + why + ")");
            im = Image(); /* texels == NULL: cyan (Texture.cpp:24) */
        }
        tex_index[rm.diffuse_texname] = (int)m->images.size();
        m->images.push_back(std::move(im));
    }

    /* ---- material instances: the first material of a name wins (Model.cpp:100-112) ---- */
    std::unordered_map<std::string, int> first_of_name;
    for (size_t i = 0; i < od.materials.size(); ++i) {
        const RawMaterial& rm = od.materials[i];
        if (!first_of_name.count(rm.name)) first_of_name[rm.name] = (int)i;
        else warn.push_back("Some Materials Have the Same Name:" + rm.name + " !");
        prt_material pm;
        std::memset(&pm, 0, sizeof(pm));
        pm.type = material_type_of(rm.name);
        pm.kd_tex = rm.diffuse_texname.empty() ? -1 : tex_index[rm.diffuse_texname];
        for (int k = 0; k < 3; ++k) { pm.kd[k] = rm.diffuse[k]; pm.ks[k] = rm.specular[k]; }
        pm.ns = rm.shininess;
        pm.alpha_x = pm.alpha_y = 0.3; /* Model.cpp:308-309 (Au) */
        const double eta[3] = {0.1, 0.5, 1.5}, kk[3] = {4.0, 0.02, 0.3};
        for (int k = 0; k < 3; ++k) { pm.eta[k] = eta[k]; pm.k[k] = kk[k]; }
        if (pm.type == PRT_MAT_DIFFUSE_LIGHT) {
            auto it = radiance.find(rm.name);
            if (it == radiance.end()) { /* lightRadianceMap.at() throws in the reference */
                delete m;
                return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_model_load: light material '" + rm.name + "' has no <light mtlname radiance> in " + base + ".xml");
            }
            for (int k = 0; k < 3; ++k) pm.emission[k] = it->second[(size_t)k];
        }
        if (pm.type != PRT_MAT_LAMBERTIAN && pm.type != PRT_MAT_PHONG) pm.kd_tex = -1; /* only these two take the texture */
        m->materials.push_back(pm);
        m->material_names.push_back(rm.name);
    }

    /* ---- meshes: one per shape, material of face 0 (Model.cpp:114-191) ---- */
    m->mesh_first.push_back(0);
    for (const ObjShape& sh : od.shapes) {
        const int mid0 = sh.material_ids.empty() ? -1 : sh.material_ids[0];
        if (mid0 < 0 || (size_t)mid0 >= od.materials.size()) {
            delete m;
            return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_model_load: shape '" + sh.name + "' has no material (the reference indexes materials[-1])");
        }
        const int mid = first_of_name[od.materials[(size_t)mid0].name];
        const size_t ntri = sh.indices.size() / 3;
        for (size_t f = 0; f < ntri; ++f) {
            double tc[3][2];
            for (int v = 0; v < 3; ++v) {
                const ObjIndex ix = sh.indices[3 * f + v];
                for (int ax = 0; ax < 3; ++ax) {
                    const size_t vi = 3 * (size_t)ix.v + ax;
                    m->vertices.push_back(ix.v >= 0 && vi < od.v.size() ? od.v[vi] : 0.0);
                    const size_t ni = 3 * (size_t)ix.vn + ax;
                    m->normals.push_back(ix.vn >= 0 && ni < od.vn.size() ? od.vn[ni] : 0.0);
                }
                for (int ax = 0; ax < 2; ++ax) {
                    const size_t ti = 2 * (size_t)ix.vt + ax;
                    tc[v][ax] = ix.vt >= 0 && ti < od.vt.size() ? od.vt[ti] : 0.0;
                }
            }
            auto same = [&](int a, int b) { return tc[a][0] == tc[b][0] && tc[a][1] == tc[b][1]; };
            if (same(0, 1) || same(1, 2) || same(0, 2)) { /* Model.cpp:170-175 */
                tc[0][0] = 0; tc[0][1] = 0; tc[1][0] = 1; tc[1][1] = 0; tc[2][0] = 1; tc[2][1] = 1;
            }
            for (int v = 0; v < 3; ++v) { m->texcoords.push_back(tc[v][0]); m->texcoords.push_back(tc[v][1]); }
        }
        m->mesh_first.push_back(m->mesh_first.back() + (int64_t)ntri);
        m->mesh_material.push_back(mid);
        m->mesh_names.push_back(sh.name);
    }

    for (Image& im : m->images) {
        prt_texture t;
        t.width = im.w; t.height = im.h; t.channels = im.c; t.reserved = 0;
        t.texels = im.px.empty() ? nullptr : im.px.data();
        m->textures.push_back(t);
    }
    prt_scene_desc& d = m->desc;
    std::memset(&d, 0, sizeof(d));
    d.n_triangles = m->mesh_first.back();
    d.vertices = m->vertices.data();
    d.normals = m->normals.data();
    d.texcoords = m->texcoords.data();
    d.n_meshes = (int32_t)m->mesh_material.size();
    d.mesh_first = m->mesh_first.data();
    d.mesh_material = m->mesh_material.data();
    d.n_materials = (int32_t)m->materials.size();
    d.materials = m->materials.data();
    d.n_textures = (int32_t)m->textures.size();
    d.textures = m->textures.data();
    for (auto& w : warn) m->warnings += w + "\n";
    *out = m;
    return PRT_OK;
}

void prt_model_destroy(prt_model* model) { delete model; }
const prt_scene_desc* prt_model_desc(const prt_model* model) { return model ? &model->desc : nullptr; }
int prt_model_camera(const prt_model* model, prt_camera* out) {
    if (!model || !out) return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_model_camera: null argument");
    if (!model->has_camera) return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_model_camera: the XML file has no <camera> element");
    *out = model->camera;
    return PRT_OK;
}
const char* prt_model_mesh_name(const prt_model* model, int32_t mesh) {
    if (!model || mesh < 0 || (size_t)mesh >= model->mesh_names.size()) return "";
    return model->mesh_names[(size_t)mesh].c_str();
}
const char* prt_model_material_name(const prt_model* model, int32_t material) {
    if (!model || material < 0 || (size_t)material >= model->material_names.size()) return "";
    return model->material_names[(size_t)material].c_str();
}
const char* prt_model_warnings(const prt_model* model) { return model ? model->warnings.c_str() : ""; }

int prt_image_load(const char* path, int32_t* width, int32_t* height, int32_t* channels, uint8_t** texels) {
    if (!path || !width || !height || !channels || !texels) return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_image_load: null argument");
    Image im;
    std::string why;
    if (!load_image(path, im, why)) return set_error(PRT_ERR_UNSUPPORTED, "prt_image_load: " + why);
    *width = im.w; *height = im.h; *channels = im.c;
    *texels = (uint8_t*)std::malloc(im.px.size() ? im.px.size() : 1);
    if (!*texels) return set_error(PRT_ERR_OUT_OF_MEMORY, "prt_image_load: out of memory");
    std::memcpy(*texels, im.px.data(), im.px.size());
    return PRT_OK;
}
void prt_image_free(uint8_t* texels) { std::free(texels); }

}  // extern "C"

/* ------------------------------------------------------------------------------------------------
 * image writers (Camera::WriteColorAttachment's two files, Source/Camera.cpp:302,327)
 * ---------------------------------------------------------------------------------------------- */
namespace {

uint32_t crc32_of(const uint8_t* p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        init = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}
void put_be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}
void put_chunk(std::vector<uint8_t>& out, const char* type, const std::vector<uint8_t>& data) {
    put_be32(out, (uint32_t)data.size());
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put_be32(out, crc32_of(&out[start], out.size() - start));
}

/* stbiw__linear_to_rgbe */
void linear_to_rgbe(uint8_t* rgbe, const float* linear) {
    const float maxcomp = std::max(linear[0], std::max(linear[1], linear[2]));
    if (maxcomp < 1e-32f) {
        rgbe[0] = rgbe[1] = rgbe[2] = rgbe[3] = 0;
    } else {
        int exponent;
        const float normalize = (float)std::frexp(maxcomp, &exponent) * 256.0f / maxcomp;
        rgbe[0] = (uint8_t)(linear[0] * normalize);
        rgbe[1] = (uint8_t)(linear[1] * normalize);
        rgbe[2] = (uint8_t)(linear[2] * normalize);
        rgbe[3] = (uint8_t)(exponent + 128);
    }
}

bool write_file(const char* path, const std::vector<uint8_t>& bytes) {
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    const size_t n = std::fwrite(bytes.data(), 1, bytes.size(), f);
    return std::fclose(f) == 0 && n == bytes.size();
}

}  // namespace

extern "C" {

int prt_write_png(const char* path, int32_t width, int32_t height, const uint8_t* rgb8) {
    if (!path || !rgb8 || width <= 0 || height <= 0) return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_write_png: bad argument");
    std::vector<uint8_t> out = {137, 80, 78, 71, 13, 10, 26, 10};
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, (uint32_t)width);
    put_be32(ihdr, (uint32_t)height);
    const uint8_t tail[5] = {8, 2, 0, 0, 0}; /* 8-bit RGB, deflate, adaptive filtering, no interlace */
    ihdr.insert(ihdr.end(), tail, tail + 5);
    put_chunk(out, "IHDR", ihdr);
    /* scanlines with filter type 0, wrapped in a zlib stream of stored blocks */
    const size_t stride = (size_t)width * 3;
    std::vector<uint8_t> raw;
    raw.reserve((stride + 1) * (size_t)height);
    for (int32_t y = 0; y < height; ++y) {
        raw.push_back(0);
        raw.insert(raw.end(), rgb8 + stride * (size_t)y, rgb8 + stride * (size_t)(y + 1));
    }
    std::vector<uint8_t> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;
    for (size_t pos = 0; pos < raw.size() || pos == 0;) {
        const size_t n = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + n >= raw.size() ? 1 : 0);
        z.push_back((uint8_t)(n & 255)); z.push_back((uint8_t)(n >> 8));
        z.push_back((uint8_t)(~n & 255)); z.push_back((uint8_t)((~n >> 8) & 255));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        for (size_t i = pos; i < pos + n; ++i) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
        pos += n;
        if (n == 0) break;
    }
    put_be32(z, (b << 16) | a);
    put_chunk(out, "IDAT", z);
    put_chunk(out, "IEND", {});
    if (!write_file(path, out)) return set_error(PRT_ERR_INVALID_ARGUMENT, std::string("prt_write_png: cannot write ") + path);
    return PRT_OK;
}

int prt_write_hdr(const char* path, int32_t width, int32_t height, const float* rgb32f) {
    if (!path || !rgb32f || width <= 0 || height <= 0) return set_error(PRT_ERR_INVALID_ARGUMENT, "prt_write_hdr: bad argument");
    char header[128];
    const int n = std::snprintf(header, sizeof(header), "#?RADIANCE\n# Written by pooraytracer_b200\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=          1.0000000000000\n\n-Y %d +X %d\n", height, width);
    std::vector<uint8_t> out(header, header + n);
    out.resize((size_t)n + (size_t)width * height * 4);
    uint8_t* p = out.data() + n;
    for (size_t i = 0; i < (size_t)width * height; ++i) linear_to_rgbe(p + 4 * i, rgb32f + 3 * i); /* flat (non run-length) scanlines */
    if (!write_file(path, out)) return set_error(PRT_ERR_INVALID_ARGUMENT, std::string("prt_write_hdr: cannot write ") + path);
    return PRT_OK;
}

}  // extern "C"
