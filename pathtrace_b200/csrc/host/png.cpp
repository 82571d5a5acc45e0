// png.cpp -- the PNG reader behind "png" textures (scene_parser.h:39-55 calls lodepng::decode, which returns RGBA8).
// Chunk parsing and scanline unfiltering are written here; inflate comes from the system zlib.
#include "flat_scene.h"

#include <zlib.h>

#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace b200pt {

namespace {

uint32_t be32(const unsigned char *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

int paeth(int a, int b, int c)
{
    int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

} // namespace

void decode_png_rgba8(const std::string &path, std::vector<unsigned char> &rgba, unsigned &width, unsigned &height)
{
    std::ifstream f(path, std::ios::binary);
    if (!f.good())
        throw std::runtime_error("png: cannot open '" + path + "'");
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string file = ss.str();
    const unsigned char *d = (const unsigned char *)file.data();
    const size_t n = file.size();
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (n < 8 || memcmp(d, sig, 8) != 0)
        throw std::runtime_error("png: '" + path + "' is not a PNG file");
    unsigned depth = 0, ctype = 0, interlace = 0;
    std::vector<unsigned char> idat, palette, trns;
    bool have_ihdr = false;
    for (size_t pos = 8; pos + 12 <= n;) {
        const uint32_t len = be32(d + pos);
        const char *type = (const char *)d + pos + 4;
        const unsigned char *body = d + pos + 8;
        if (pos + 12 + (size_t)len > n)
            throw std::runtime_error("png: truncated chunk in '" + path + "'");
        if (!memcmp(type, "IHDR", 4) && len >= 13) {
            width = be32(body);
            height = be32(body + 4);
            depth = body[8];
            ctype = body[9];
            interlace = body[12];
            have_ihdr = true;
        } else if (!memcmp(type, "PLTE", 4)) {
            palette.assign(body, body + len);
        } else if (!memcmp(type, "tRNS", 4)) {
            trns.assign(body, body + len);
        } else if (!memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!memcmp(type, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)len;
    }
    if (!have_ihdr || width == 0 || height == 0 || width > 32768 || height > 32768)
        throw std::runtime_error("png: bad header in '" + path + "'");
    if (interlace)
        throw std::runtime_error("png: interlaced files are not supported ('" + path + "')");
    unsigned channels;
    switch (ctype) {
    case 0: channels = 1; break;
    case 2: channels = 3; break;
    case 3: channels = 1; break;
    case 4: channels = 2; break;
    case 6: channels = 4; break;
    default: throw std::runtime_error("png: unknown colour type in '" + path + "'");
    }
    if (!(depth == 8 || depth == 16 || (depth < 8 && (ctype == 0 || ctype == 3))))
        throw std::runtime_error("png: unsupported bit depth in '" + path + "'");
    const size_t bpp_bits = (size_t)channels * depth;
    const size_t stride = (width * bpp_bits + 7) / 8, bpp = (bpp_bits + 7) / 8;
    std::vector<unsigned char> raw((stride + 1) * (size_t)height);
    uLongf raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size())
        throw std::runtime_error("png: inflate failed for '" + path + "'");
    // undo the scanline filters in place (PNG spec 9.2); `prev` is the row above, already reconstructed
    std::vector<unsigned char> zero(stride, 0);
    for (unsigned y = 0; y < height; y++) {
        unsigned char *row = raw.data() + (stride + 1) * (size_t)y;
        const unsigned filter = row[0];
        unsigned char *cur = row + 1;
        const unsigned char *prev = y ? raw.data() + (stride + 1) * (size_t)(y - 1) + 1 : zero.data();
        for (size_t i = 0; i < stride; i++) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int x = cur[i];
            switch (filter) {
            case 0: break;
            case 1: x += a; break;
            case 2: x += b; break;
            case 3: x += (a + b) >> 1; break;
            case 4: x += paeth(a, b, c); break;
            default: throw std::runtime_error("png: bad filter type in '" + path + "'");
            }
            cur[i] = (unsigned char)x;
        }
    }
    rgba.assign((size_t)width * height * 4, 255);
    for (unsigned y = 0; y < height; y++) {
        const unsigned char *cur = raw.data() + (stride + 1) * (size_t)y + 1;
        for (unsigned x = 0; x < width; x++) {
            unsigned v[4] = {0, 0, 0, 255};
            auto sample = [&](unsigned k) -> unsigned { // k-th sample of pixel x, reduced to 8 bits
                if (depth == 8)
                    return cur[(size_t)x * channels + k];
                if (depth == 16)
                    return cur[((size_t)x * channels + k) * 2]; // most significant byte, as lodepng's RGBA8 conversion
                const size_t bit = (size_t)x * depth;
                const unsigned raw_v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1u);
                return ctype == 3 ? raw_v : raw_v * 255u / ((1u << depth) - 1u);
            };
            if (ctype == 3) {
                const unsigned idx = sample(0);
                if ((size_t)idx * 3 + 2 < palette.size()) {
                    v[0] = palette[idx * 3];
                    v[1] = palette[idx * 3 + 1];
                    v[2] = palette[idx * 3 + 2];
                } else {
                    v[0] = v[1] = v[2] = 0;
                }
                v[3] = idx < trns.size() ? trns[idx] : 255;
            } else if (ctype == 0 || ctype == 4) {
                v[0] = v[1] = v[2] = sample(0);
                if (ctype == 4)
                    v[3] = sample(1);
                else if (trns.size() >= 2 && depth == 8 && v[0] == trns[1])
                    v[3] = 0;
            } else {
                v[0] = sample(0);
                v[1] = sample(1);
                v[2] = sample(2);
                if (ctype == 6)
                    v[3] = sample(3);
                else if (trns.size() >= 6 && depth == 8 && v[0] == trns[1] && v[1] == trns[3] && v[2] == trns[5])
                    v[3] = 0;
            }
            unsigned char *o = rgba.data() + ((size_t)y * width + x) * 4;
            o[0] = (unsigned char)v[0], o[1] = (unsigned char)v[1], o[2] = (unsigned char)v[2], o[3] = (unsigned char)v[3];
        }
    }
}

} // namespace b200pt
