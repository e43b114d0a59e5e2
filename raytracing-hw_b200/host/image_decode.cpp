// Image decode for glTF textures -> RGBA8 (what tinygltf/stb_image hand the reference: always 4 channels, 8 bit,
// thirdparty/tinygltf/tiny_gltf.h:2609,2681-2683; scene_parser.cpp:131-142 rejects anything but 8 bits).
//   * PNG: own decoder on top of zlib's inflate (1-8 bit, grey / RGB / palette / grey+alpha / RGBA, tRNS, interlace 0).
//     16-bit samples are not "8 bit" for the reference (it throws); we throw the same message.
//   * JPEG: own baseline decoder (jpeg_decode.cpp) restating stb_image's integer arithmetic, so the bytes equal the
//     reference's (tests/test_formats.py compares both decoders with the compiled reference).
// ONE decode path per format, the same in every build: nothing here depends on files outside this repository.
#include "host_scene.h"

#include <algorithm>
#include <cstring>
#include <string>
#include <zlib.h>

namespace rtb {
namespace {

constexpr uint32_t kMaxTextureSide = 32768;   // sanity limit on either side of an image

uint32_t be32(const uint8_t *p) { return ((uint32_t) p[0] << 24) | ((uint32_t) p[1] << 16) | ((uint32_t) p[2] << 8) | p[3]; }

int paeth(int a, int b, int c) {
    const int p = a + b - c;
    const int pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    if (pa <= pb && pa <= pc) return a;
    return pb <= pc ? b : c;
}

bool decode_png(const std::vector<uint8_t> &bytes, HostTexture &out, std::string &err) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    if (bytes.size() < 8 || std::memcmp(bytes.data(), sig, 8) != 0) { err = "not a PNG"; return false; }
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, palette, trns;
    size_t p = 8;
    bool seen_ihdr = false, seen_iend = false;
    while (p + 12 <= bytes.size() && !seen_iend) {
        const uint32_t len = be32(&bytes[p]);
        const char *tag = (const char *) &bytes[p + 4];
        if (p + 12 + (size_t) len > bytes.size()) { err = "truncated PNG chunk"; return false; }
        const uint8_t *data = &bytes[p + 8];
        if (!std::memcmp(tag, "IHDR", 4)) {
            if (len < 13) { err = "bad IHDR"; return false; }
            w = be32(data); h = be32(data + 4);
            depth = data[8]; ctype = data[9]; interlace = data[12];
            seen_ihdr = true;
        } else if (!std::memcmp(tag, "PLTE", 4)) {
            palette.assign(data, data + len);
        } else if (!std::memcmp(tag, "tRNS", 4)) {
            trns.assign(data, data + len);
        } else if (!std::memcmp(tag, "IDAT", 4)) {
            idat.insert(idat.end(), data, data + len);
        } else if (!std::memcmp(tag, "IEND", 4)) {
            seen_iend = true;
        }
        p += 12 + (size_t) len;
    }
    if (!seen_ihdr || w == 0 || h == 0) { err = "PNG without IHDR"; return false; }
    if (w > kMaxTextureSide || h > kMaxTextureSide) { err = "PNG larger than " + std::to_string(kMaxTextureSide) + " pixels a side"; return false; }
    if (interlace != 0) { err = "interlaced PNG not supported"; return false; }
    if (depth == 16) { err = "Texture format not supported: only 8 bit channels are supported"; return false; }
    if (depth != 8 && !(ctype == 0 || ctype == 3)) { err = "unsupported PNG bit depth"; return false; }
    int channels = 0;
    switch (ctype) {
        case 0: channels = 1; break;
        case 2: channels = 3; break;
        case 3: channels = 1; break;
        case 4: channels = 2; break;
        case 6: channels = 4; break;
        default: err = "bad PNG colour type"; return false;
    }
    const size_t bits_pp = (size_t) channels * (size_t) depth;
    const size_t stride = ((size_t) w * bits_pp + 7) / 8;
    const size_t bpp = bits_pp >= 8 ? bits_pp / 8 : 1;
    std::vector<uint8_t> raw((stride + 1) * (size_t) h);
    {
        z_stream zs;
        std::memset(&zs, 0, sizeof(zs));
        if (inflateInit(&zs) != Z_OK) { err = "zlib init failed"; return false; }
        // zlib's counters are 32 bit: feed and drain in pieces of at most 1 GB
        const size_t kPiece = (size_t) 1 << 30;
        size_t in_off = 0, out_off = 0;
        int rc = Z_OK;
        while (rc == Z_OK) {
            if (zs.avail_in == 0 && in_off < idat.size()) {
                const size_t m = std::min(kPiece, idat.size() - in_off);
                zs.next_in = idat.data() + in_off;
                zs.avail_in = (uInt) m;
                in_off += m;
            }
            if (zs.avail_out == 0 && out_off < raw.size()) {
                const size_t m = std::min(kPiece, raw.size() - out_off);
                zs.next_out = raw.data() + out_off;
                zs.avail_out = (uInt) m;
                out_off += m;
            }
            const bool last = in_off >= idat.size();
            rc = inflate(&zs, last ? Z_FINISH : Z_NO_FLUSH);
            if (rc == Z_BUF_ERROR && zs.avail_out == 0 && out_off >= raw.size()) break;   // all rows are there
            if (rc == Z_BUF_ERROR && (zs.avail_in == 0 && !last)) rc = Z_OK;
        }
        const size_t got = out_off - zs.avail_out;
        inflateEnd(&zs);
        if ((rc != Z_STREAM_END && rc != Z_OK && rc != Z_BUF_ERROR) || got != raw.size()) { err = "PNG inflate failed"; return false; }
    }
    // un-filter in place
    std::vector<uint8_t> prev(stride, 0);
    for (uint32_t y = 0; y < h; ++y) {
        uint8_t *row = &raw[(stride + 1) * (size_t) y];
        const int ft = row[0];
        uint8_t *cur = row + 1;
        for (size_t x = 0; x < stride; ++x) {
            const int a = x >= bpp ? cur[x - bpp] : 0;
            const int b = prev[x];
            const int c = x >= bpp ? prev[x - bpp] : 0;
            int v = cur[x];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: err = "bad PNG filter"; return false;
            }
            cur[x] = (uint8_t) v;
        }
        std::memcpy(prev.data(), cur, stride);
    }
    out.width = (int) w;
    out.height = (int) h;
    out.rgba.resize((size_t) w * h * 4);
    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t *cur = &raw[(stride + 1) * (size_t) y + 1];
        uint8_t *dst = &out.rgba[(size_t) y * w * 4];
        for (uint32_t x = 0; x < w; ++x) {
            uint8_t r, g, b, a = 255;
            if (ctype == 2) {
                r = cur[x * 3]; g = cur[x * 3 + 1]; b = cur[x * 3 + 2];
                // tRNS colour key (16-bit big-endian per channel; 8-bit images use the low byte)
                if (trns.size() >= 6 && r == trns[1] && g == trns[3] && b == trns[5]) a = 0;
            }
            else if (ctype == 6) { r = cur[x * 4]; g = cur[x * 4 + 1]; b = cur[x * 4 + 2]; a = cur[x * 4 + 3]; }
            else if (ctype == 4) { r = g = b = cur[x * 2]; a = cur[x * 2 + 1]; }
            else {
                unsigned s;
                if (depth == 8) s = cur[x];
                else {
                    const size_t bit = (size_t) x * (size_t) depth;
                    s = (cur[bit >> 3] >> (8 - depth - (int) (bit & 7))) & ((1u << depth) - 1u);
                }
                if (ctype == 3) {
                    if ((size_t) s * 3 + 2 >= palette.size()) { err = "PNG palette index out of range"; return false; }
                    r = palette[s * 3]; g = palette[s * 3 + 1]; b = palette[s * 3 + 2];
                    if (s < trns.size()) a = trns[s];
                } else {
                    const unsigned maxv = (1u << depth) - 1u;
                    r = g = b = (uint8_t) (s * 255u / maxv);
                    if (trns.size() >= 2 && s == ((((unsigned) trns[0] << 8) | trns[1]) & maxv)) a = 0;   // grey colour key
                }
            }
            dst[x * 4] = r; dst[x * 4 + 1] = g; dst[x * 4 + 2] = b; dst[x * 4 + 3] = a;
        }
    }
    return true;
}

}  // namespace

bool decode_image(const std::vector<uint8_t> &bytes, HostTexture &out, std::string &err) {
    if (bytes.size() >= 8 && bytes[0] == 0x89 && bytes[1] == 'P') return decode_png(bytes, out, err);
    if (bytes.size() >= 4 && bytes[0] == 0xff && bytes[1] == 0xd8) return decode_jpeg(bytes, out, err);
    err = "unknown image format (PNG and baseline JPEG are supported)";
    return false;
}

}  // namespace rtb
