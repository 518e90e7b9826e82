/*
 * jpeg_decode.cpp — baseline and progressive JPEG (ITU-T T.81, 8-bit, Huffman) with the texel values stb_image
 * produces, because the reference loads every texture through stbi_load(path, &w, &h, &channels, 0)
 * (Source/Texture.cpp:10-21) and a JPEG decoder's output is NOT fixed by the standard: the inverse DCT, the chroma
 * upsampling filter and the YCbCr -> RGB conversion are implementation choices. The three choices of stb_image
 * (v2.2x, default build: no STBI_JPEG_OLD, SIMD paths documented as bit-identical to the scalar ones) are restated
 * here as integer arithmetic:
 *
 *   inverse DCT   the "slow-but-accurate" integer scheme (Loeffler / Ligtenberg / Moschytz factorisation) with 12-bit
 *                 constants c = trunc(float(x) * 4096 + 0.5); the column pass rounds at 2^9 and keeps 2 extra bits
 *                 (>> 10), a column whose AC terms are all zero short-cuts to dc * 4; the row pass adds the level
 *                 shift (128 << 17) and the rounding 2^16, shifts by 17 and clamps to 0..255;
 *   upsampling    2x horizontal: (3 near + far + 2) >> 2 with the edge samples copied; 2x vertical: (3 near + far + 2)
 *                 >> 2; 2x2: the vertical 3:1 blend kept at full precision, then the horizontal one, (… + 8) >> 4, edges
 *                 (t + 2) >> 2; any other ratio: sample replication. "near" is the chroma row the output row falls in,
 *                 "far" the neighbouring row on the side the output row leans to;
 *   colour        20-bit fixed point, y * 2^20 + 2^19, coefficients trunc(x * 4096 + 0.5) << 8 for 1.402, 0.71414,
 *                 0.34414, 1.772; the Cb term of green is masked to its upper 16 bits before the sum; >> 20, clamp.
 *   An Adobe APP14 segment with transform 0, or component ids 'R','G','B' without a JFIF header, means the three
 *   planes are RGB already; one component decodes to one channel (stbi_load with req_comp = 0).
 *
 * Pinned by the test suite against the stb_image.h the reference vendors (compiled, as part of the reference's own
 * Texture.cpp, into the reference checker the tests load; this library never does) on baseline / progressive / 4:4:4 / 4:2:2 / 4:2:0 / 4:4:0 / 4:1:1 / grey files with odd
 * sizes and restart intervals (tests/test_scene_io.py, tests/golden/jpeg_*.npz). Four-component (CMYK / YCCK) and
 * arithmetic-coded files are refused, as are 12-bit ones (stb refuses those too).
 */
#include "jpeg_decode.h"

#include <cstring>

namespace prt {
namespace {

const uint8_t kZigzag[64 + 15] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13,
                                  6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31,
                                  39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
                                  /* a run that overshoots the block lands here */
                                  63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

struct Huffman {
    /* canonical code: for length l, codes [first[l], first[l] + count[l]) map to values[offset[l] ..] */
    int count[17] = {0};
    int first[17] = {0};
    int offset[17] = {0};
    uint8_t values[256] = {0};
    bool present = false;
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0;
    int hd = 0, ha = 0;
    int dc_pred = 0;
    int x = 0, y = 0, w2 = 0, h2 = 0;
    std::vector<uint8_t> data;
    std::vector<int16_t> coeff; /* progressive */
    int coeff_w = 0, coeff_h = 0;
};

struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint32_t acc = 0;
    int bits = 0;
    int marker = -1; /* a marker met inside the entropy-coded segment: the stream then reads as zeros */
    void reset() { acc = 0; bits = 0; marker = -1; }
    void fill() {
        while (bits <= 24) {
            int b = 0;
            if (marker < 0 && p < end) {
                b = *p++;
                if (b == 0xff) {
                    int c = p < end ? *p++ : 0xd9;
                    while (c == 0xff && p < end) c = *p++; /* fill bytes */
                    if (c != 0) {
                        marker = c;
                        b = 0;
                    }
                }
            }
            acc |= (uint32_t)b << (24 - bits);
            bits += 8;
        }
    }
    int get(int n) { /* n in 0..16 */
        if (n == 0) return 0;
        if (bits < n) fill();
        const int v = (int)(acc >> (32 - n));
        acc <<= n;
        bits -= n;
        return v;
    }
    int bit() { return get(1); }
};

int decode_symbol(BitReader& br, const Huffman& h) {
    int code = 0;
    for (int l = 1; l <= 16; ++l) {
        code = (code << 1) | br.bit();
        if (h.count[l] && code - h.first[l] < h.count[l] && code >= h.first[l]) return h.values[h.offset[l] + code - h.first[l]];
    }
    return -1;
}

/* T.81 F.2.2.1: the s-bit two's-complement-like magnitude code */
inline int extend_receive(BitReader& br, int s) {
    if (s == 0) return 0;
    const int v = br.get(s);
    return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v;
}

inline uint8_t clamp8(int x) { return (unsigned)x > 255u ? (x < 0 ? 0 : 255) : (uint8_t)x; }

#define PRT_F2F(x) ((int)(((x) * 4096 + 0.5)))
#define PRT_FSH(x) ((x) * 4096)
/* one 8-point pass of the integer inverse DCT: even part into x0..x3, odd part into t0..t3 */
#define PRT_IDCT_1D(s0, s1, s2, s3, s4, s5, s6, s7)       \
    int t0, t1, t2, t3, p1, p2, p3, p4, p5, x0, x1, x2, x3; \
    p2 = s2;                                              \
    p3 = s6;                                              \
    p1 = (p2 + p3) * PRT_F2F(0.5411961f);                 \
    t2 = p1 + p3 * PRT_F2F(-1.847759065f);                \
    t3 = p1 + p2 * PRT_F2F(0.765366865f);                 \
    p2 = s0;                                              \
    p3 = s4;                                              \
    t0 = PRT_FSH(p2 + p3);                                \
    t1 = PRT_FSH(p2 - p3);                                \
    x0 = t0 + t3;                                         \
    x3 = t0 - t3;                                         \
    x1 = t1 + t2;                                         \
    x2 = t1 - t2;                                         \
    t0 = s7;                                              \
    t1 = s5;                                              \
    t2 = s3;                                              \
    t3 = s1;                                              \
    p3 = t0 + t2;                                         \
    p4 = t1 + t3;                                         \
    p1 = t0 + t3;                                         \
    p2 = t1 + t2;                                         \
    p5 = (p3 + p4) * PRT_F2F(1.175875602f);               \
    t0 = t0 * PRT_F2F(0.298631336f);                      \
    t1 = t1 * PRT_F2F(2.053119869f);                      \
    t2 = t2 * PRT_F2F(3.072711026f);                      \
    t3 = t3 * PRT_F2F(1.501321110f);                      \
    p1 = p5 + p1 * PRT_F2F(-0.899976223f);                \
    p2 = p5 + p2 * PRT_F2F(-2.562915447f);                \
    p3 = p3 * PRT_F2F(-1.961570560f);                     \
    p4 = p4 * PRT_F2F(-0.390180644f);                     \
    t3 += p1 + p4;                                        \
    t2 += p2 + p3;                                        \
    t1 += p2 + p4;                                        \
    t0 += p1 + p3;

void idct_block(uint8_t* out, int out_stride, const int16_t data[64]) {
    int val[64];
    for (int i = 0; i < 8; ++i) { /* columns */
        const int16_t* d = data + i;
        int* v = val + i;
        if (d[8] == 0 && d[16] == 0 && d[24] == 0 && d[32] == 0 && d[40] == 0 && d[48] == 0 && d[56] == 0) {
            const int dcterm = d[0] * 4;
            v[0] = v[8] = v[16] = v[24] = v[32] = v[40] = v[48] = v[56] = dcterm;
        } else {
            PRT_IDCT_1D(d[0], d[8], d[16], d[24], d[32], d[40], d[48], d[56])
            x0 += 512; x1 += 512; x2 += 512; x3 += 512;
            v[0] = (x0 + t3) >> 10;
            v[56] = (x0 - t3) >> 10;
            v[8] = (x1 + t2) >> 10;
            v[48] = (x1 - t2) >> 10;
            v[16] = (x2 + t1) >> 10;
            v[40] = (x2 - t1) >> 10;
            v[24] = (x3 + t0) >> 10;
            v[32] = (x3 - t0) >> 10;
        }
    }
    for (int i = 0; i < 8; ++i) { /* rows */
        const int* v = val + 8 * i;
        uint8_t* o = out + (size_t)i * out_stride;
        PRT_IDCT_1D(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7])
        x0 += 65536 + (128 << 17);
        x1 += 65536 + (128 << 17);
        x2 += 65536 + (128 << 17);
        x3 += 65536 + (128 << 17);
        o[0] = clamp8((x0 + t3) >> 17);
        o[7] = clamp8((x0 - t3) >> 17);
        o[1] = clamp8((x1 + t2) >> 17);
        o[6] = clamp8((x1 - t2) >> 17);
        o[2] = clamp8((x2 + t1) >> 17);
        o[5] = clamp8((x2 - t1) >> 17);
        o[3] = clamp8((x3 + t0) >> 17);
        o[4] = clamp8((x3 - t0) >> 17);
    }
}

struct Decoder {
    const uint8_t* p;
    const uint8_t* end;
    std::string& why;
    int width = 0, height = 0, ncomp = 0;
    bool progressive = false, jfif = false;
    int app14_transform = -1;
    int restart_interval = 0;
    uint16_t dequant[4][64];
    bool have_q[4] = {false, false, false, false};
    Huffman hdc[4], hac[4];
    Component comp[4];
    int h_max = 1, v_max = 1, mcu_x = 0, mcu_y = 0;
    /* scan state */
    int scan_n = 0, order[4] = {0, 0, 0, 0};
    int spec_start = 0, spec_end = 63, succ_high = 0, succ_low = 0;
    int eob_run = 0;
    BitReader br;

    Decoder(const uint8_t* d, size_t n, std::string& w) : p(d), end(d + n), why(w) { std::memset(dequant, 0, sizeof(dequant)); }

    bool fail(const char* m) { why = std::string("JPEG: ") + m; return false; }
    int u8() { return p < end ? *p++ : 0; }
    int u16() { const int a = u8(); return (a << 8) | u8(); }

    int next_marker() {
        int c = u8();
        if (c != 0xff) return -1;
        while (c == 0xff && p < end) c = u8();
        return c;
    }

    bool read_dqt(int len) {
        while (len > 0) {
            const int q = u8(), prec = q >> 4, t = q & 15;
            if (prec > 1 || t > 3) return fail("bad quantisation table");
            for (int i = 0; i < 64; ++i) dequant[t][kZigzag[i]] = (uint16_t)(prec ? u16() : u8());
            have_q[t] = true;
            len -= prec ? 129 : 65;
        }
        return len == 0 || fail("bad DQT length");
    }

    bool read_dht(int len) {
        while (len > 0) {
            const int q = u8(), tc = q >> 4, th = q & 15;
            if (tc > 1 || th > 3) return fail("bad Huffman table");
            Huffman& h = tc ? hac[th] : hdc[th];
            int total = 0, code = 0;
            for (int l = 1; l <= 16; ++l) {
                h.count[l] = u8();
                h.offset[l] = total;
                h.first[l] = code;
                total += h.count[l];
                code = (code + h.count[l]) << 1;
            }
            if (total > 256) return fail("bad Huffman table");
            for (int i = 0; i < total; ++i) h.values[i] = (uint8_t)u8();
            h.present = true;
            len -= 17 + total;
        }
        return len == 0 || fail("bad DHT length");
    }

    bool read_sof(int len) {
        if (u8() != 8) return fail("only 8-bit samples");
        height = u16();
        width = u16();
        ncomp = u8();
        if (height <= 0 || width <= 0) return fail("empty image (DNL-defined heights are not supported)");
        if (ncomp != 1 && ncomp != 3) return fail("only 1- and 3-component images (CMYK / YCCK files are refused)");
        if (len != 8 + 3 * ncomp) return fail("bad SOF length");
        for (int i = 0; i < ncomp; ++i) {
            Component& c = comp[i];
            c.id = u8();
            const int q = u8();
            c.h = q >> 4;
            c.v = q & 15;
            c.tq = u8();
            if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) return fail("bad component");
            if (c.h > h_max) h_max = c.h;
            if (c.v > v_max) v_max = c.v;
        }
        for (int i = 0; i < ncomp; ++i)
            if (h_max % comp[i].h != 0 || v_max % comp[i].v != 0) return fail("fractional sampling ratios are not supported");
        mcu_x = (width + h_max * 8 - 1) / (h_max * 8);
        mcu_y = (height + v_max * 8 - 1) / (v_max * 8);
        for (int i = 0; i < ncomp; ++i) {
            Component& c = comp[i];
            c.x = (width * c.h + h_max - 1) / h_max;
            c.y = (height * c.v + v_max - 1) / v_max;
            c.w2 = mcu_x * c.h * 8;
            c.h2 = mcu_y * c.v * 8;
            c.data.assign((size_t)c.w2 * c.h2, 0);
            if (progressive) {
                c.coeff_w = c.w2 / 8;
                c.coeff_h = c.h2 / 8;
                c.coeff.assign((size_t)c.w2 * c.h2, 0);
            }
        }
        return true;
    }

    bool read_sos() {
        const int len = u16();
        scan_n = u8();
        if (scan_n < 1 || scan_n > ncomp || len != 6 + 2 * scan_n) return fail("bad SOS");
        for (int i = 0; i < scan_n; ++i) {
            const int id = u8(), q = u8();
            int which = -1;
            for (int k = 0; k < ncomp; ++k)
                if (comp[k].id == id) which = k;
            if (which < 0) return fail("SOS names an unknown component");
            comp[which].hd = q >> 4;
            comp[which].ha = q & 15;
            if (comp[which].hd > 3 || comp[which].ha > 3) return fail("bad Huffman selector");
            order[i] = which;
        }
        spec_start = u8();
        spec_end = u8();
        const int a = u8();
        succ_high = a >> 4;
        succ_low = a & 15;
        if (progressive) {
            if (spec_start > 63 || spec_end > 63 || spec_start > spec_end || succ_high > 13 || succ_low > 13) return fail("bad progressive scan");
        } else {
            if (spec_start != 0 || succ_high != 0 || succ_low != 0) return fail("bad baseline scan");
            spec_end = 63;
        }
        return true;
    }

    /* ---- blocks ---- */
    bool block_baseline(int16_t data[64], Component& c) {
        const Huffman &dc = hdc[c.hd], &ac = hac[c.ha];
        const uint16_t* dq = dequant[c.tq];
        const int t = decode_symbol(br, dc);
        if (t < 0 || t > 15) return fail("bad Huffman code");
        std::memset(data, 0, 64 * sizeof(int16_t));
        const int diff = t ? extend_receive(br, t) : 0;
        c.dc_pred += diff;
        data[0] = (int16_t)(c.dc_pred * dq[0]);
        int k = 1;
        do {
            const int rs = decode_symbol(br, ac);
            if (rs < 0) return fail("bad Huffman code");
            const int s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (rs != 0xf0) break;
                k += 16;
            } else {
                k += r;
                const int zig = kZigzag[k++];
                data[zig] = (int16_t)(extend_receive(br, s) * dq[zig]);
            }
        } while (k < 64);
        return true;
    }

    bool block_prog_dc(int16_t data[64], Component& c) {
        if (spec_end != 0) return fail("DC and AC coefficients in one progressive scan");
        if (succ_high == 0) {
            std::memset(data, 0, 64 * sizeof(int16_t));
            const int t = decode_symbol(br, hdc[c.hd]);
            if (t < 0 || t > 15) return fail("bad Huffman code");
            const int diff = t ? extend_receive(br, t) : 0;
            c.dc_pred += diff;
            data[0] = (int16_t)(c.dc_pred * (1 << succ_low));
        } else if (br.bit()) {
            data[0] = (int16_t)(data[0] + (int16_t)(1 << succ_low));
        }
        return true;
    }

    bool block_prog_ac(int16_t data[64], const Huffman& ac) {
        if (spec_start == 0) return fail("DC and AC coefficients in one progressive scan");
        if (succ_high == 0) {
            const int shift = succ_low;
            if (eob_run) { --eob_run; return true; }
            int k = spec_start;
            do {
                const int rs = decode_symbol(br, ac);
                if (rs < 0) return fail("bad Huffman code");
                const int s = rs & 15, r = rs >> 4;
                if (s == 0) {
                    if (r < 15) {
                        eob_run = 1 << r;
                        if (r) eob_run += br.get(r);
                        --eob_run;
                        break;
                    }
                    k += 16;
                } else {
                    k += r;
                    const int zig = kZigzag[k++];
                    data[zig] = (int16_t)(extend_receive(br, s) * (1 << shift));
                }
            } while (k <= spec_end);
        } else {
            const int16_t bit = (int16_t)(1 << succ_low);
            auto refine = [&](int16_t* q) {
                if (br.bit() && (*q & bit) == 0) *q = (int16_t)(*q > 0 ? *q + bit : *q - bit);
            };
            if (eob_run) {
                --eob_run;
                for (int k = spec_start; k <= spec_end; ++k) {
                    int16_t* q = &data[kZigzag[k]];
                    if (*q != 0) refine(q);
                }
            } else {
                int k = spec_start;
                do {
                    const int rs = decode_symbol(br, ac);
                    if (rs < 0) return fail("bad Huffman code");
                    int s = rs & 15, r = rs >> 4;
                    if (s == 0) {
                        if (r < 15) {
                            eob_run = (1 << r) - 1;
                            if (r) eob_run += br.get(r);
                            r = 64; /* run to the end of the band */
                        }
                        /* r == 15: sixteen zero coefficients */
                    } else {
                        if (s != 1) return fail("bad Huffman code");
                        s = br.bit() ? bit : -bit;
                    }
                    while (k <= spec_end) {
                        int16_t* q = &data[kZigzag[k++]];
                        if (*q != 0) {
                            refine(q);
                        } else {
                            if (r == 0) {
                                *q = (int16_t)s;
                                break;
                            }
                            --r;
                        }
                    }
                } while (k <= spec_end);
            }
        }
        return true;
    }

    void restart() {
        br.reset();
        for (int i = 0; i < 4; ++i) comp[i].dc_pred = 0;
        eob_run = 0;
    }

    /* after every restart_interval MCUs the stream holds an RSTn marker: realign on it */
    bool maybe_restart(int& todo) {
        if (--todo > 0) return true;
        if (br.bits < 24) br.fill();
        if (br.marker < 0xd0 || br.marker > 0xd7) return false; /* not a restart marker: the scan ends here */
        p = br.p;
        restart();
        br.p = p;
        todo = restart_interval;
        return true;
    }

    bool decode_scan() {
        restart();
        br.p = p;
        br.end = end;
        int todo = restart_interval ? restart_interval : 0x7fffffff;
        int16_t block[64];
        if (scan_n == 1) {
            Component& c = comp[order[0]];
            const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
            for (int j = 0; j < h; ++j)
                for (int i = 0; i < w; ++i) {
                    if (progressive) {
                        int16_t* d = c.coeff.data() + 64 * ((size_t)i + (size_t)j * c.coeff_w);
                        if (spec_start == 0 ? !block_prog_dc(d, c) : !block_prog_ac(d, hac[c.ha])) return false;
                    } else {
                        if (!block_baseline(block, c)) return false;
                        idct_block(c.data.data() + (size_t)c.w2 * j * 8 + (size_t)i * 8, c.w2, block);
                    }
                    if (!maybe_restart(todo)) goto done;
                }
        } else {
            for (int j = 0; j < mcu_y; ++j)
                for (int i = 0; i < mcu_x; ++i) {
                    for (int k = 0; k < scan_n; ++k) {
                        Component& c = comp[order[k]];
                        for (int y = 0; y < c.v; ++y)
                            for (int x = 0; x < c.h; ++x) {
                                const int x2 = (i * c.h + x) * 8, y2 = (j * c.v + y) * 8;
                                if (progressive) {
                                    int16_t* d = c.coeff.data() + 64 * ((size_t)(x2 >> 3) + (size_t)(y2 >> 3) * c.coeff_w);
                                    if (!block_prog_dc(d, c)) return false;
                                } else {
                                    if (!block_baseline(block, c)) return false;
                                    idct_block(c.data.data() + (size_t)c.w2 * y2 + x2, c.w2, block);
                                }
                            }
                    }
                    if (!maybe_restart(todo)) goto done;
                }
        }
    done:
        /* continue after the entropy-coded segment: at the marker the bit reader ran into, or wherever it stopped */
        if (br.marker >= 0) {
            p = br.p;
            pending_marker = br.marker;
        } else {
            p = br.p;
            pending_marker = -1;
        }
        return true;
    }
    int pending_marker = -1;

    void finish_progressive() {
        for (int n = 0; n < ncomp; ++n) {
            Component& c = comp[n];
            const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
            const uint16_t* dq = dequant[c.tq];
            for (int j = 0; j < h; ++j)
                for (int i = 0; i < w; ++i) {
                    int16_t* d = c.coeff.data() + 64 * ((size_t)i + (size_t)j * c.coeff_w);
                    for (int k = 0; k < 64; ++k) d[k] = (int16_t)(d[k] * dq[k]);
                    idct_block(c.data.data() + (size_t)c.w2 * j * 8 + (size_t)i * 8, c.w2, d);
                }
        }
    }

    bool parse() {
        if (u8() != 0xff || u8() != 0xd8) return fail("no SOI");
        bool have_frame = false;
        for (;;) {
            int m = pending_marker >= 0 ? pending_marker : next_marker();
            pending_marker = -1;
            if (m < 0) {
                /* stray bytes between segments: look for the next marker, give up at the end of the file */
                while (p < end && *p != 0xff) ++p;
                if (p >= end) return have_frame ? true : fail("no image data");
                continue;
            }
            if (m == 0xd9) break; /* EOI */
            if (m == 0xda) {
                if (!have_frame) return fail("SOS before SOF");
                if (!read_sos()) return false;
                if (!decode_scan()) return false;
                continue;
            }
            if (m >= 0xd0 && m <= 0xd7) continue; /* a restart marker outside a scan */
            if (m == 0x01 || m == 0xff) continue;
            const int len = u16() - 2;
            if (len < 0 || p + len > end) return fail("truncated segment");
            const uint8_t* seg_end = p + len;
            if (m == 0xc0 || m == 0xc1 || m == 0xc2) {
                if (have_frame) return fail("two frames");
                progressive = m == 0xc2;
                if (!read_sof(len + 2)) return false;
                have_frame = true;
            } else if (m == 0xc4) {
                if (!read_dht(len)) return false;
            } else if (m == 0xdb) {
                if (!read_dqt(len)) return false;
            } else if (m == 0xdd) {
                if (len != 2) return fail("bad DRI");
                restart_interval = u16();
            } else if (m == 0xe0) {
                if (len >= 5 && std::memcmp(p, "JFIF\0", 5) == 0) jfif = true;
            } else if (m == 0xee) {
                if (len >= 12 && std::memcmp(p, "Adobe\0", 6) == 0) app14_transform = p[11];
            } else if (m == 0xc3 || (m >= 0xc5 && m <= 0xcf && m != 0xc8)) {
                return fail("lossless, hierarchical and arithmetic-coded JPEG are not supported");
            }
            p = seg_end;
        }
        if (!have_frame) return fail("no frame");
        if (progressive) finish_progressive();
        return true;
    }
};

/* ---- chroma upsampling (one output row) ---- */
const uint8_t* upsample_row(uint8_t* out, const uint8_t* near_, const uint8_t* far_, int w, int hs, int vs) {
    if (hs == 1 && vs == 1) return near_;
    if (hs == 1 && vs == 2) {
        for (int i = 0; i < w; ++i) out[i] = (uint8_t)((3 * near_[i] + far_[i] + 2) >> 2);
        return out;
    }
    if (hs == 2 && vs == 1) {
        if (w == 1) {
            out[0] = out[1] = near_[0];
            return out;
        }
        out[0] = near_[0];
        out[1] = (uint8_t)((near_[0] * 3 + near_[1] + 2) >> 2);
        int i;
        for (i = 1; i < w - 1; ++i) {
            const int n = 3 * near_[i] + 2;
            out[i * 2 + 0] = (uint8_t)((n + near_[i - 1]) >> 2);
            out[i * 2 + 1] = (uint8_t)((n + near_[i + 1]) >> 2);
        }
        out[i * 2 + 0] = (uint8_t)((near_[w - 2] * 3 + near_[w - 1] + 2) >> 2);
        out[i * 2 + 1] = near_[w - 1];
        return out;
    }
    if (hs == 2 && vs == 2) {
        if (w == 1) {
            out[0] = out[1] = (uint8_t)((3 * near_[0] + far_[0] + 2) >> 2);
            return out;
        }
        int t1 = 3 * near_[0] + far_[0];
        out[0] = (uint8_t)((t1 + 2) >> 2);
        for (int i = 1; i < w; ++i) {
            const int t0 = t1;
            t1 = 3 * near_[i] + far_[i];
            out[i * 2 - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
            out[i * 2] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
        }
        out[w * 2 - 1] = (uint8_t)((t1 + 2) >> 2);
        return out;
    }
    for (int i = 0; i < w; ++i) /* any other ratio: replicate */
        for (int j = 0; j < hs; ++j) out[i * hs + j] = near_[i];
    return out;
}

#define PRT_FLOAT2FIXED(x) (((int)((x) * 4096.0f + 0.5f)) << 8)
void ycbcr_row(uint8_t* out, const uint8_t* y, const uint8_t* pcb, const uint8_t* pcr, int count) {
    for (int i = 0; i < count; ++i) {
        const int y_fixed = (y[i] << 20) + (1 << 19);
        const int cr = pcr[i] - 128, cb = pcb[i] - 128;
        int r = y_fixed + cr * PRT_FLOAT2FIXED(1.40200f);
        int g = y_fixed + (cr * -PRT_FLOAT2FIXED(0.71414f)) + (int)(((unsigned)(cb * -PRT_FLOAT2FIXED(0.34414f))) & 0xffff0000u);
        int b = y_fixed + cb * PRT_FLOAT2FIXED(1.77200f);
        r >>= 20;
        g >>= 20;
        b >>= 20;
        out[3 * i + 0] = clamp8(r);
        out[3 * i + 1] = clamp8(g);
        out[3 * i + 2] = clamp8(b);
    }
}

}  // namespace

bool decode_jpeg(const uint8_t* bytes, size_t n, int& width, int& height, int& channels, std::vector<uint8_t>& pixels, std::string& why) {
    Decoder d(bytes, n, why);
    if (!d.parse()) return false;
    width = d.width;
    height = d.height;
    channels = d.ncomp; /* stbi_load(..., req_comp = 0): one channel for a grey file, three otherwise */
    pixels.assign((size_t)width * height * channels, 0);
    if (d.ncomp == 1) {
        for (int j = 0; j < height; ++j) std::memcpy(pixels.data() + (size_t)j * width, d.comp[0].data.data() + (size_t)j * d.comp[0].w2, (size_t)width);
        return true;
    }
    int rgb_ids = 0;
    static const char ids[3] = {'R', 'G', 'B'};
    for (int k = 0; k < 3; ++k)
        if (d.comp[k].id == ids[k]) ++rgb_ids;
    const bool is_rgb = rgb_ids == 3 && (d.app14_transform == 0 || (d.app14_transform == -1 && !d.jfif));
    struct Up { int hs, vs, ystep, ypos, w_lores; const uint8_t *line0, *line1; std::vector<uint8_t> buf; } up[3];
    for (int k = 0; k < 3; ++k) {
        const Component& c = d.comp[k];
        up[k].hs = d.h_max / c.h;
        up[k].vs = d.v_max / c.v;
        up[k].ystep = up[k].vs >> 1;
        up[k].ypos = 0;
        up[k].w_lores = (width + up[k].hs - 1) / up[k].hs;
        up[k].line0 = up[k].line1 = c.data.data();
        up[k].buf.assign((size_t)width + 3 * 8 + 8, 0); /* a 2x row of w_lores samples may overshoot the image by hs - 1 */
        up[k].buf.resize((size_t)up[k].w_lores * up[k].hs + 8);
    }
    for (int j = 0; j < height; ++j) {
        const uint8_t* row[3];
        for (int k = 0; k < 3; ++k) {
            Up& u = up[k];
            const bool y_bot = u.ystep >= (u.vs >> 1);
            row[k] = upsample_row(u.buf.data(), y_bot ? u.line1 : u.line0, y_bot ? u.line0 : u.line1, u.w_lores, u.hs, u.vs);
            if (++u.ystep >= u.vs) {
                u.ystep = 0;
                u.line0 = u.line1;
                if (++u.ypos < d.comp[k].y) u.line1 += d.comp[k].w2;
            }
        }
        uint8_t* out = pixels.data() + (size_t)j * width * 3;
        if (is_rgb) {
            for (int i = 0; i < width; ++i) {
                out[3 * i] = row[0][i];
                out[3 * i + 1] = row[1][i];
                out[3 * i + 2] = row[2][i];
            }
        } else {
            ycbcr_row(out, row[0], row[1], row[2], width);
        }
    }
    return true;
}

}  // namespace prt
