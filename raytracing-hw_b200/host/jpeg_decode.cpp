// Baseline JPEG -> RGBA8 for glTF textures, giving the bytes stb_image gives.
//
// The reference decodes textures through tinygltf -> stb_image (thirdparty/tinygltf/tiny_gltf.h:2609,2681-2683: always
// expanded to 4 channels); the 65 JPEGs of examples/sponza are all baseline (SOF0), 8 bit, three components sampled 1x1,
// JFIF, no restart markers.  Texture parity needs the same BYTES, and a JPEG decoder's bytes depend on its arithmetic,
// so this file restates the published integer algorithms of stb_image v2.x (public domain / MIT, Sean Barrett et al.):
//   * inverse DCT: the 12-bit fixed-point LL&M-style butterfly with a DC-only shortcut per column (stbi__idct_block),
//   * YCbCr -> RGB: 20-bit fixed point with the 16-bit truncated green term (stbi__YCbCr_to_RGB_row),
//   * chroma upsampling: the 3:1 "triangle" filters for 2x1, 1x2, 2x2 and replication otherwise (stbi__resample_row_*).
// Supported: SOF0 / SOF1 Huffman, 8-bit samples, 1 or 3 components, interleaved and non-interleaved scans, restart
// intervals, 8/16-bit quantisation tables.  Progressive (SOF2) and arithmetic-coded files are refused with a message.
// Verified byte for byte against the compiled reference on the 65 Sponza JPEGs and on generated files of every
// sampling layout (tests/test_formats.py).
#include "host_scene.h"

#include <cstring>

namespace rtb {
namespace {

const uint8_t kDezigzag[64 + 15] = {
    0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
    63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};   // (a corrupt run may step past 63)

struct Huffman {
    // canonical code: for a length L, codes [first_code[L], first_code[L] + count[L]) map to symbols from first_sym[L]
    int count[17] = {0};
    int first_code[18] = {0}, first_sym[17] = {0};
    uint8_t symbols[256] = {0};
    bool present = false;
    bool build(const uint8_t counts[16], const uint8_t *syms, int n) {
        int code = 0, k = 0;
        for (int L = 1; L <= 16; ++L) {
            count[L] = counts[L - 1];
            first_code[L] = code;
            first_sym[L] = k;
            code += count[L];
            k += count[L];
            if (code > (1 << L)) return false;
            code <<= 1;
        }
        if (k != n) return false;
        std::memcpy(symbols, syms, (size_t) n);
        present = true;
        return true;
    }
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0;
    int dc_pred = 0;
    int x = 0, y = 0;      // size in samples
    int w2 = 0, h2 = 0;    // allocated (whole MCUs)
    std::vector<uint8_t> data;
};

struct Decoder {
    const uint8_t *p, *end;
    uint32_t code_buffer = 0;
    int code_bits = 0;
    int marker = -1;       // a marker met inside the entropy-coded data
    bool nomore = false;
    Huffman hdc[4], hac[4];
    uint16_t dequant[4][64];
    Component comp[4];
    int n_comp = 0, width = 0, height = 0;
    int h_max = 1, v_max = 1, mcu_w = 0, mcu_h = 0, mcu_x = 0, mcu_y = 0;
    int restart_interval = 0, todo = 0;
    bool jfif = false;
    int app14_transform = -1, rgb_ids = 0;
    int scan_n = 0, order[4] = {0, 0, 0, 0};
    std::string err;

    int get8() { return p < end ? *p++ : 0; }
    int get16() { const int a = get8(); return (a << 8) | get8(); }

    void grow() {
        do {
            unsigned b = nomore ? 0u : (unsigned) get8();
            if (b == 0xff) {
                int c = get8();
                while (c == 0xff) c = get8();
                if (c != 0) {
                    marker = c;
                    nomore = true;
                    return;
                }
            }
            code_buffer |= b << (24 - code_bits);
            code_bits += 8;
        } while (code_bits <= 24);
    }
    int get_bits(int n) {
        if (n == 0) return 0;
        if (code_bits < n) grow();
        // a marker inside the entropy data stops grow() early: out of bits on a truncated / corrupt file.  Answer zeros
        // and leave the counter alone (it would go negative and the next grow() would shift by 32 or more)
        if (code_bits < n) return 0;
        const unsigned v = code_buffer >> (32 - n);
        code_buffer <<= n;
        code_bits -= n;
        return (int) v;
    }
    int decode(const Huffman &h) {
        if (code_bits < 16) grow();
        for (int L = 1; L <= 16; ++L) {
            const int c = (int) (code_buffer >> (32 - L));
            if (c - h.first_code[L] < h.count[L] && c >= h.first_code[L]) {
                if (L > code_bits) return -1;
                code_buffer <<= L;
                code_bits -= L;
                return h.symbols[h.first_sym[L] + c - h.first_code[L]];
            }
        }
        return -1;
    }
    // ITU T.81 F.2.2.1 EXTEND: n bits, sign-extended
    int extend_receive(int n) {
        if (n == 0) return 0;
        const int v = get_bits(n);
        return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
    }
    void reset() {
        code_bits = 0;
        code_buffer = 0;
        nomore = false;
        for (auto &c : comp) c.dc_pred = 0;
        marker = -1;
        todo = restart_interval ? restart_interval : 0x7fffffff;
    }

    bool decode_block(short data[64], Component &c) {
        const Huffman &hd = hdc[c.hd], &ha = hac[c.ha];
        const uint16_t *dq = dequant[c.tq];
        const int t = decode(hd);
        if (t < 0 || t > 15) { err = "bad huffman code"; return false; }
        std::memset(data, 0, 64 * sizeof(short));
        const int diff = t ? extend_receive(t) : 0;
        const int dc = c.dc_pred + diff;
        c.dc_pred = dc;
        data[0] = (short) (dc * dq[0]);
        int k = 1;
        do {
            const int rs = decode(ha);
            if (rs < 0) { err = "bad huffman code"; return false; }
            const int s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (rs != 0xf0) break;
                k += 16;
            } else {
                k += r;
                const unsigned zig = kDezigzag[k++];
                data[zig] = (short) (extend_receive(s) * dq[zig]);
            }
        } while (k < 64);
        return true;
    }
};

inline uint8_t clamp8(int x) {
    if ((unsigned) x > 255u) return x < 0 ? 0 : 255;
    return (uint8_t) x;
}

// fixed-point constants: x * 4096 + 0.5 truncated
constexpr int f2f(double x) { return (int) (x * 4096 + 0.5); }
inline int fsh(int x) { return x * 4096; }

#define RTB_IDCT_1D(s0, s1, s2, s3, s4, s5, s6, s7)                 \
    int t0, t1, t2, t3, p1, p2, p3, p4, p5, x0, x1, x2, x3;         \
    p2 = s2;                                                        \
    p3 = s6;                                                        \
    p1 = (p2 + p3) * f2f(0.5411961f);                               \
    t2 = p1 + p3 * f2f(-1.847759065f);                              \
    t3 = p1 + p2 * f2f(0.765366865f);                               \
    p2 = s0;                                                        \
    p3 = s4;                                                        \
    t0 = fsh(p2 + p3);                                              \
    t1 = fsh(p2 - p3);                                              \
    x0 = t0 + t3;                                                   \
    x3 = t0 - t3;                                                   \
    x1 = t1 + t2;                                                   \
    x2 = t1 - t2;                                                   \
    t0 = s7;                                                        \
    t1 = s5;                                                        \
    t2 = s3;                                                        \
    t3 = s1;                                                        \
    p3 = t0 + t2;                                                   \
    p4 = t1 + t3;                                                   \
    p1 = t0 + t3;                                                   \
    p2 = t1 + t2;                                                   \
    p5 = (p3 + p4) * f2f(1.175875602f);                             \
    t0 = t0 * f2f(0.298631336f);                                    \
    t1 = t1 * f2f(2.053119869f);                                    \
    t2 = t2 * f2f(3.072711026f);                                    \
    t3 = t3 * f2f(1.501321110f);                                    \
    p1 = p5 + p1 * f2f(-0.899976223f);                              \
    p2 = p5 + p2 * f2f(-2.562915447f);                              \
    p3 = p3 * f2f(-1.961570560f);                                   \
    p4 = p4 * f2f(-0.390180644f);                                   \
    t3 += p1 + p4;                                                  \
    t2 += p2 + p3;                                                  \
    t1 += p2 + p4;                                                  \
    t0 += p1 + p3;

void idct_block(uint8_t *out, int out_stride, const short data[64]) {
    int val[64], *v = val;
    const short *d = data;
    for (int i = 0; i < 8; ++i, ++d, ++v) {
        if (d[8] == 0 && d[16] == 0 && d[24] == 0 && d[32] == 0 && d[40] == 0 && d[48] == 0 && d[56] == 0) {
            const int dcterm = d[0] * 4;   // (the columns keep two extra fraction bits)
            v[0] = v[8] = v[16] = v[24] = v[32] = v[40] = v[48] = v[56] = dcterm;
        } else {
            RTB_IDCT_1D(d[0], d[8], d[16], d[24], d[32], d[40], d[48], d[56])
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
    v = val;
    uint8_t *o = out;
    for (int i = 0; i < 8; ++i, v += 8, o += out_stride) {
        RTB_IDCT_1D(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7])
        // rounding, and the +128 level shift, folded into the constant
        x0 += 65536 + (128 << 17); x1 += 65536 + (128 << 17); x2 += 65536 + (128 << 17); x3 += 65536 + (128 << 17);
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
#undef RTB_IDCT_1D

// ---- chroma upsampling: one output row of `w` low-resolution samples stretched by hs (near / far = the two source rows)
void resample_row_1(uint8_t *out, const uint8_t *near_, const uint8_t *, int w, int) { std::memcpy(out, near_, (size_t) w); }
void resample_row_v_2(uint8_t *out, const uint8_t *near_, const uint8_t *far_, int w, int) {
    for (int i = 0; i < w; ++i) out[i] = (uint8_t) ((3 * near_[i] + far_[i] + 2) >> 2);
}
void resample_row_h_2(uint8_t *out, const uint8_t *in, const uint8_t *, int w, int) {
    if (w == 1) { out[0] = out[1] = in[0]; return; }
    out[0] = in[0];
    out[1] = (uint8_t) ((in[0] * 3 + in[1] + 2) >> 2);
    int i;
    for (i = 1; i < w - 1; ++i) {
        const int n = 3 * in[i] + 2;
        out[i * 2 + 0] = (uint8_t) ((n + in[i - 1]) >> 2);
        out[i * 2 + 1] = (uint8_t) ((n + in[i + 1]) >> 2);
    }
    out[i * 2 + 0] = (uint8_t) ((in[w - 2] * 3 + in[w - 1] + 2) >> 2);
    out[i * 2 + 1] = in[w - 1];
}
void resample_row_hv_2(uint8_t *out, const uint8_t *near_, const uint8_t *far_, int w, int) {
    if (w == 1) { out[0] = out[1] = (uint8_t) ((3 * near_[0] + far_[0] + 2) >> 2); return; }
    int t1 = 3 * near_[0] + far_[0];
    out[0] = (uint8_t) ((t1 + 2) >> 2);
    for (int i = 1; i < w; ++i) {
        const int t0 = t1;
        t1 = 3 * near_[i] + far_[i];
        out[i * 2 - 1] = (uint8_t) ((3 * t0 + t1 + 8) >> 4);
        out[i * 2] = (uint8_t) ((3 * t1 + t0 + 8) >> 4);
    }
    out[w * 2 - 1] = (uint8_t) ((t1 + 2) >> 2);
}
void resample_row_generic(uint8_t *out, const uint8_t *near_, const uint8_t *, int w, int hs) {
    for (int i = 0; i < w; ++i)
        for (int j = 0; j < hs; ++j) out[i * hs + j] = near_[i];
}

constexpr int float2fixed(double x) { return ((int) (x * 4096.0 + 0.5)) << 8; }

bool parse_and_decode(Decoder &z) {
    if (z.get8() != 0xff || z.get8() != 0xd8) { z.err = "not a JPEG"; return false; }
    bool have_frame = false;
    for (;;) {
        // next marker
        int m;
        if (z.marker >= 0) { m = z.marker; z.marker = -1; }
        else {
            int x = z.get8();
            while (x != 0xff) {
                if (z.p >= z.end) { z.err = "JPEG ends without EOI"; return false; }
                x = z.get8();
            }
            while (x == 0xff) x = z.get8();
            m = x;
        }
        if (m == 0xd9) break;   // EOI
        if (m == 0xc2) { z.err = "progressive JPEG is not supported (baseline only)"; return false; }
        if ((m >= 0xc3 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc)) { z.err = "unsupported JPEG coding process"; return false; }
        if (m == 0x01 || (m >= 0xd0 && m <= 0xd7)) continue;
        int L = z.get16();
        if (L < 2 || z.p + (L - 2) > z.end) { z.err = "bad JPEG segment length"; return false; }
        const uint8_t *seg_end = z.p + (L - 2);
        if (m == 0xdb) {   // DQT
            while (z.p < seg_end) {
                const int q = z.get8(), prec = q >> 4, t = q & 15;
                if ((prec != 0 && prec != 1) || t > 3) { z.err = "bad DQT"; return false; }
                for (int i = 0; i < 64; ++i) z.dequant[t][kDezigzag[i]] = (uint16_t) (prec ? z.get16() : z.get8());
            }
        } else if (m == 0xc4) {   // DHT
            while (z.p < seg_end) {
                const int q = z.get8(), tc = q >> 4, th = q & 15;
                if (tc > 1 || th > 3) { z.err = "bad DHT"; return false; }
                uint8_t counts[16], syms[256];
                int n = 0;
                for (int i = 0; i < 16; ++i) { counts[i] = (uint8_t) z.get8(); n += counts[i]; }
                if (n > 256) { z.err = "bad DHT"; return false; }
                for (int i = 0; i < n; ++i) syms[i] = (uint8_t) z.get8();
                if (!(tc == 0 ? z.hdc[th] : z.hac[th]).build(counts, syms, n)) { z.err = "bad DHT code lengths"; return false; }
            }
        } else if (m == 0xdd) {   // DRI
            z.restart_interval = z.get16();
        } else if (m == 0xe0) {   // APP0: JFIF?
            if (L >= 7) {
                static const char tag[5] = {'J', 'F', 'I', 'F', '\0'};
                if (std::memcmp(z.p, tag, 5) == 0) z.jfif = true;
            }
        } else if (m == 0xee) {   // APP14: Adobe colour transform flag
            if (L >= 14) {
                static const char tag[6] = {'A', 'd', 'o', 'b', 'e', '\0'};
                if (std::memcmp(z.p, tag, 6) == 0) z.app14_transform = z.p[11];
            }
        } else if (m == 0xc0 || m == 0xc1) {   // SOF0 / SOF1
            if (have_frame) { z.err = "two frames in one JPEG"; return false; }
            const int prec = z.get8();
            z.height = z.get16();
            z.width = z.get16();
            z.n_comp = z.get8();
            if (prec != 8) { z.err = "JPEG: only 8-bit samples"; return false; }
            if (z.width <= 0 || z.height <= 0) { z.err = "JPEG with zero size"; return false; }
            // (a header of a few hundred bytes must not be able to ask for gigabytes of planes: 2^28 pixels at most)
            if ((size_t) z.width > 32768 || (size_t) z.height > 32768 || (size_t) z.width * (size_t) z.height > ((size_t) 1 << 28)) { z.err = "JPEG too large"; return false; }
            if (z.n_comp != 1 && z.n_comp != 3) { z.err = "JPEG: 1 or 3 components only"; return false; }
            if (L != 8 + 3 * z.n_comp) { z.err = "bad SOF length"; return false; }
            static const char rgb[3] = {'R', 'G', 'B'};
            for (int i = 0; i < z.n_comp; ++i) {
                Component &c = z.comp[i];
                c.id = z.get8();
                if (z.n_comp == 3 && c.id == rgb[i]) ++z.rgb_ids;
                const int q = z.get8();
                c.h = q >> 4; c.v = q & 15;
                c.tq = z.get8();
                if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) { z.err = "bad SOF component"; return false; }
                z.h_max = std::max(z.h_max, c.h);
                z.v_max = std::max(z.v_max, c.v);
            }
            for (int i = 0; i < z.n_comp; ++i)
                if (z.h_max % z.comp[i].h != 0 || z.v_max % z.comp[i].v != 0) { z.err = "JPEG: fractional sampling ratios are not supported"; return false; }
            z.mcu_w = z.h_max * 8; z.mcu_h = z.v_max * 8;
            z.mcu_x = (z.width + z.mcu_w - 1) / z.mcu_w;
            z.mcu_y = (z.height + z.mcu_h - 1) / z.mcu_h;
            for (int i = 0; i < z.n_comp; ++i) {
                Component &c = z.comp[i];
                c.x = (z.width * c.h + z.h_max - 1) / z.h_max;
                c.y = (z.height * c.v + z.v_max - 1) / z.v_max;
                c.w2 = z.mcu_x * c.h * 8;
                c.h2 = z.mcu_y * c.v * 8;
                c.data.assign((size_t) c.w2 * c.h2, 0);
            }
            have_frame = true;
        } else if (m == 0xda) {   // SOS + entropy-coded data
            if (!have_frame) { z.err = "SOS before SOF"; return false; }
            z.scan_n = z.get8();
            if (z.scan_n < 1 || z.scan_n > z.n_comp || L != 6 + 2 * z.scan_n) { z.err = "bad SOS"; return false; }
            for (int i = 0; i < z.scan_n; ++i) {
                const int id = z.get8(), q = z.get8();
                int which = 0;
                for (; which < z.n_comp; ++which) if (z.comp[which].id == id) break;
                if (which == z.n_comp) { z.err = "bad SOS component"; return false; }
                z.comp[which].hd = q >> 4;
                z.comp[which].ha = q & 15;
                if (z.comp[which].hd > 3 || z.comp[which].ha > 3 || !z.hdc[z.comp[which].hd].present || !z.hac[z.comp[which].ha].present) { z.err = "SOS names a missing Huffman table"; return false; }
                z.order[i] = which;
            }
            const int ss = z.get8();
            z.get8();
            const int a = z.get8();
            if (ss != 0 || a != 0) { z.err = "bad SOS (spectral selection in a sequential file)"; return false; }
            z.reset();
            short data[64];
            if (z.scan_n == 1) {
                // non-interleaved: the component's own blocks, row by row
                Component &c = z.comp[z.order[0]];
                const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
                for (int j = 0; j < h; ++j)
                    for (int i = 0; i < w; ++i) {
                        if (!z.decode_block(data, c)) return false;
                        idct_block(&c.data[(size_t) c.w2 * j * 8 + (size_t) i * 8], c.w2, data);
                        if (--z.todo <= 0) {
                            if (z.code_bits < 24) z.grow();
                            if (!(z.marker >= 0xd0 && z.marker <= 0xd7)) goto scan_done;
                            z.reset();
                        }
                    }
            } else {
                for (int j = 0; j < z.mcu_y; ++j)
                    for (int i = 0; i < z.mcu_x; ++i) {
                        for (int k = 0; k < z.scan_n; ++k) {
                            Component &c = z.comp[z.order[k]];
                            for (int y = 0; y < c.v; ++y)
                                for (int x = 0; x < c.h; ++x) {
                                    const int x2 = (i * c.h + x) * 8, y2 = (j * c.v + y) * 8;
                                    if (!z.decode_block(data, c)) return false;
                                    idct_block(&c.data[(size_t) c.w2 * y2 + x2], c.w2, data);
                                }
                        }
                        if (--z.todo <= 0) {
                            if (z.code_bits < 24) z.grow();
                            if (!(z.marker >= 0xd0 && z.marker <= 0xd7)) goto scan_done;
                            z.reset();
                        }
                    }
            }
        scan_done:
            if (z.marker >= 0xd0 && z.marker <= 0xd7) z.marker = -1;
            continue;   // (the scan header was consumed above; its segment end is not skipped to)
        }
        z.p = seg_end;
    }
    if (!have_frame) { z.err = "JPEG without a frame"; return false; }
    return true;
}

}  // namespace

bool decode_jpeg(const std::vector<uint8_t> &bytes, HostTexture &out, std::string &err) {
    Decoder z;
    z.p = bytes.data();
    z.end = bytes.data() + bytes.size();
    std::memset(z.dequant, 0, sizeof z.dequant);
    if (!parse_and_decode(z)) { err = z.err.empty() ? "JPEG decode failed" : z.err; return false; }
    const int W = z.width, H = z.height;
    out.width = W;
    out.height = H;
    out.rgba.assign((size_t) W * H * 4, 255);
    // per component: the row resampler and its two source rows, stepped as the output rows go by
    struct Resample {
        void (*fn)(uint8_t *, const uint8_t *, const uint8_t *, int, int);
        const uint8_t *line0, *line1;
        int hs, vs, w_lores, ystep, ypos;
    } rs[3];
    std::vector<uint8_t> linebuf[3];
    for (int k = 0; k < z.n_comp; ++k) {
        Component &c = z.comp[k];
        Resample &r = rs[k];
        r.hs = z.h_max / c.h;
        r.vs = z.v_max / c.v;
        r.ystep = r.vs >> 1;
        r.w_lores = (W + r.hs - 1) / r.hs;
        r.ypos = 0;
        r.line0 = r.line1 = c.data.data();
        if (r.hs == 1 && r.vs == 1) r.fn = resample_row_1;
        else if (r.hs == 1 && r.vs == 2) r.fn = resample_row_v_2;
        else if (r.hs == 2 && r.vs == 1) r.fn = resample_row_h_2;
        else if (r.hs == 2 && r.vs == 2) r.fn = resample_row_hv_2;
        else r.fn = resample_row_generic;
        linebuf[k].assign((size_t) W + 3 + 8, 0);
    }
    const bool is_rgb = z.n_comp == 3 && (z.rgb_ids == 3 || (z.app14_transform == 0 && !z.jfif));
    for (int j = 0; j < H; ++j) {
        const uint8_t *row[3] = {nullptr, nullptr, nullptr};
        for (int k = 0; k < z.n_comp; ++k) {
            Component &c = z.comp[k];
            Resample &r = rs[k];
            const bool y_bot = r.ystep >= (r.vs >> 1);
            const uint8_t *near_ = y_bot ? r.line1 : r.line0, *far_ = y_bot ? r.line0 : r.line1;
            if (r.fn == resample_row_1) row[k] = near_;
            else {
                r.fn(linebuf[k].data(), near_, far_, r.w_lores, r.hs);
                row[k] = linebuf[k].data();
            }
            if (++r.ystep >= r.vs) {
                r.ystep = 0;
                r.line0 = r.line1;
                if (++r.ypos < c.y) r.line1 += c.w2;
            }
        }
        uint8_t *o = &out.rgba[(size_t) j * W * 4];
        if (z.n_comp == 1) {
            for (int i = 0; i < W; ++i) { o[i * 4] = o[i * 4 + 1] = o[i * 4 + 2] = row[0][i]; o[i * 4 + 3] = 255; }
        } else if (is_rgb) {
            for (int i = 0; i < W; ++i) { o[i * 4] = row[0][i]; o[i * 4 + 1] = row[1][i]; o[i * 4 + 2] = row[2][i]; o[i * 4 + 3] = 255; }
        } else {
            for (int i = 0; i < W; ++i) {
                const int y_fixed = (row[0][i] << 20) + (1 << 19);   // rounding
                const int cb = row[1][i] - 128, cr = row[2][i] - 128;
                int r = y_fixed + cr * float2fixed(1.40200f);
                int g = y_fixed + (cr * -float2fixed(0.71414f)) + (int) ((unsigned) (cb * -float2fixed(0.34414f)) & 0xffff0000u);
                int b = y_fixed + cb * float2fixed(1.77200f);
                r >>= 20; g >>= 20; b >>= 20;
                o[i * 4] = clamp8(r); o[i * 4 + 1] = clamp8(g); o[i * 4 + 2] = clamp8(b); o[i * 4 + 3] = 255;
            }
        }
    }
    return true;
}

}  // namespace rtb
