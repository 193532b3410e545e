// image.cpp — image files for bitmap reflectance textures: PFM, binary PPM, PNG and baseline JPEG, decoded here (no third-party code:
// the reference reads textures through stb_image, engine/Loader.cpp:66-86, in a path it never enables).
// PNG: 8-bit grey / grey+alpha / RGB / RGBA / palette, non-interlaced; zlib stream inflated by the decoder below
// (stored, fixed and dynamic Huffman blocks, RFC 1950 / 1951), the five scanline filters of the PNG specification.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "path_tracer.h"
#include "scene.h"

namespace gsbhost {
namespace {

std::vector<uint8_t> readAll(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    std::vector<uint8_t> data;
    uint8_t buf[65536];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) data.insert(data.end(), buf, buf + n);
    std::fclose(f);
    return data;
}

// ---- inflate (RFC 1951) ------------------------------------------------------------------------------------------------
struct BitReader {
    const uint8_t* p;
    size_t n, pos = 0;
    uint32_t acc = 0;
    int bits = 0;
    uint32_t get(int count) {
        while (bits < count) {
            if (pos >= n) throw std::runtime_error("png: truncated zlib stream");
            acc |= (uint32_t)p[pos++] << bits;
            bits += 8;
        }
        const uint32_t v = acc & ((count == 32) ? 0xffffffffu : ((1u << count) - 1u));
        acc >>= count;
        bits -= count;
        return v;
    }
    void alignByte() { acc = 0; bits = 0; }
};

struct Huffman {
    uint16_t count[16] = {0};
    std::vector<uint16_t> symbol;
    void build(const uint8_t* lengths, int n) {
        std::memset(count, 0, sizeof(count));
        for (int i = 0; i < n; i++) count[lengths[i]]++;
        count[0] = 0;
        uint16_t offs[16];
        offs[1] = 0;
        for (int l = 1; l < 15; l++) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
        symbol.assign((size_t)n, 0);
        for (int i = 0; i < n; i++)
            if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
    }
    int decode(BitReader& br) const {
        int code = 0, first = 0, index = 0;
        for (int len = 1; len <= 15; len++) {
            code |= (int)br.get(1);
            const int c = count[len];
            if (code - c < first) return symbol[(size_t)(index + (code - first))];
            index += c;
            first += c;
            first <<= 1;
            code <<= 1;
        }
        throw std::runtime_error("png: bad Huffman code");
    }
};

std::vector<uint8_t> inflateZlib(const std::vector<uint8_t>& z) {
    if (z.size() < 6 || (z[0] & 0x0f) != 8) throw std::runtime_error("png: not a zlib deflate stream");
    BitReader br{z.data() + 2, z.size() - 2};
    std::vector<uint8_t> out;
    static const uint16_t lenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const uint16_t lenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    static const uint16_t distBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static const uint16_t distExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    for (bool last = false; !last;) {
        last = br.get(1) != 0;
        const uint32_t type = br.get(2);
        if (type == 0) {
            br.alignByte();
            if (br.pos + 4 > br.n) throw std::runtime_error("png: truncated stored block");
            const uint32_t len = br.p[br.pos] | (br.p[br.pos + 1] << 8);
            br.pos += 4;
            if (br.pos + len > br.n) throw std::runtime_error("png: truncated stored block");
            out.insert(out.end(), br.p + br.pos, br.p + br.pos + len);
            br.pos += len;
            continue;
        }
        if (type == 3) throw std::runtime_error("png: bad deflate block type");
        Huffman lit, dist;
        if (type == 1) {
            uint8_t l[288];
            for (int i = 0; i < 144; i++) l[i] = 8;
            for (int i = 144; i < 256; i++) l[i] = 9;
            for (int i = 256; i < 280; i++) l[i] = 7;
            for (int i = 280; i < 288; i++) l[i] = 8;
            lit.build(l, 288);
            uint8_t d[30];
            for (int i = 0; i < 30; i++) d[i] = 5;
            dist.build(d, 30);
        } else {
            const int nlen = (int)br.get(5) + 257, ndist = (int)br.get(5) + 1, ncode = (int)br.get(4) + 4;
            static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            uint8_t cl[19] = {0};
            for (int i = 0; i < ncode; i++) cl[order[i]] = (uint8_t)br.get(3);
            Huffman clh;
            clh.build(cl, 19);
            uint8_t lengths[320] = {0};
            for (int i = 0; i < nlen + ndist;) {
                const int sym = clh.decode(br);
                if (sym < 16) lengths[i++] = (uint8_t)sym;
                else {
                    int rep, val = 0;
                    if (sym == 16) { if (i == 0) throw std::runtime_error("png: bad code lengths"); val = lengths[i - 1]; rep = 3 + (int)br.get(2); }
                    else if (sym == 17) rep = 3 + (int)br.get(3);
                    else rep = 11 + (int)br.get(7);
                    if (i + rep > nlen + ndist) throw std::runtime_error("png: bad code lengths");
                    while (rep--) lengths[i++] = (uint8_t)val;
                }
            }
            lit.build(lengths, nlen);
            dist.build(lengths + nlen, ndist);
        }
        for (;;) {
            const int sym = lit.decode(br);
            if (sym < 256) out.push_back((uint8_t)sym);
            else if (sym == 256) break;
            else {
                if (sym > 285) throw std::runtime_error("png: bad length symbol");
                const int len = lenBase[sym - 257] + (int)br.get(lenExtra[sym - 257]);
                const int ds = dist.decode(br);
                if (ds > 29) throw std::runtime_error("png: bad distance symbol");
                const size_t d = distBase[ds] + br.get(distExtra[ds]);
                if (d > out.size()) throw std::runtime_error("png: distance too far back");
                for (int k = 0; k < len; k++) out.push_back(out[out.size() - d]);
            }
        }
    }
    return out;
}

float srgbToLinear(uint8_t v) {
    const float c = (float)v / 255.0f;
    return c <= 0.04045f ? c / 12.92f : std::pow((c + 0.055f) / 1.055f, 2.4f);
}

void decodePng(const std::vector<uint8_t>& file, std::vector<float>& rgb, uint32_t& W, uint32_t& H) {
    auto be32 = [&](size_t o) { return ((uint32_t)file[o] << 24) | ((uint32_t)file[o + 1] << 16) | ((uint32_t)file[o + 2] << 8) | file[o + 3]; };
    size_t pos = 8;
    uint32_t depth = 0, colour = 0, interlace = 0;
    std::vector<uint8_t> idat, palette;
    W = H = 0;
    while (pos + 8 <= file.size()) {
        const uint32_t len = be32(pos);
        const std::string type((const char*)&file[pos + 4], 4);
        const size_t body = pos + 8;
        if (body + len + 4 > file.size()) throw std::runtime_error("png: truncated chunk");
        if (type == "IHDR") { W = be32(body); H = be32(body + 4); depth = file[body + 8]; colour = file[body + 9]; interlace = file[body + 12]; }
        else if (type == "PLTE") palette.assign(file.begin() + (long)body, file.begin() + (long)(body + len));
        else if (type == "IDAT") idat.insert(idat.end(), file.begin() + (long)body, file.begin() + (long)(body + len));
        else if (type == "IEND") break;
        pos = body + len + 4;
    }
    if (!W || !H || depth != 8 || interlace != 0) throw std::runtime_error("png: only 8-bit non-interlaced images are supported");
    const int ch = colour == 0 ? 1 : colour == 2 ? 3 : colour == 3 ? 1 : colour == 4 ? 2 : colour == 6 ? 4 : 0;
    if (!ch) throw std::runtime_error("png: unknown colour type");
    const std::vector<uint8_t> raw = inflateZlib(idat);
    const size_t stride = (size_t)W * ch;
    if (raw.size() < (stride + 1) * H) throw std::runtime_error("png: image data too short");
    std::vector<uint8_t> img(stride * H);
    for (uint32_t y = 0; y < H; y++) {
        const uint8_t f = raw[(stride + 1) * y];
        const uint8_t* src = &raw[(stride + 1) * y + 1];
        uint8_t* dst = &img[stride * y];
        const uint8_t* up = y ? dst - stride : nullptr;
        for (size_t x = 0; x < stride; x++) {
            const int a = x >= (size_t)ch ? dst[x - ch] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)ch) ? up[x - ch] : 0;
            int pred = 0;
            if (f == 1) pred = a;
            else if (f == 2) pred = b;
            else if (f == 3) pred = (a + b) / 2;
            else if (f == 4) {
                const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
                pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
            } else if (f != 0) throw std::runtime_error("png: bad scanline filter");
            dst[x] = (uint8_t)(src[x] + pred);
        }
    }
    rgb.resize((size_t)W * H * 3);
    for (size_t i = 0; i < (size_t)W * H; i++) {
        uint8_t r, g, b;
        if (colour == 3) {
            const size_t k = 3 * (size_t)img[i];
            if (k + 2 >= palette.size()) throw std::runtime_error("png: palette index out of range");
            r = palette[k]; g = palette[k + 1]; b = palette[k + 2];
        } else if (ch <= 2) { r = g = b = img[i * ch]; }
        else { r = img[i * ch]; g = img[i * ch + 1]; b = img[i * ch + 2]; }
        rgb[3 * i] = srgbToLinear(r); rgb[3 * i + 1] = srgbToLinear(g); rgb[3 * i + 2] = srgbToLinear(b);
    }
}

// ---- baseline JPEG (ITU T.81 sequential DCT, Huffman, 8 bit; grey or YCbCr with any sampling factors, restart intervals) ------
// What the reference's own textured scene needs (staircase2/textures/*.jpg are baseline files).  Progressive files are refused.
struct JpegHuff {
    uint8_t bits[17] = {0};
    uint8_t vals[256] = {0};
    int mincode[17], maxcode[18], valptr[17];
    bool present = false;
    void build() {
        int code = 0, k = 0;
        for (int l = 1; l <= 16; l++) {
            valptr[l] = k;
            mincode[l] = code;
            code += bits[l];
            k += bits[l];
            maxcode[l] = bits[l] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        present = true;
    }
};

struct JpegBits {
    const uint8_t* p;
    size_t n, pos;
    uint32_t acc = 0;
    int cnt = 0;
    int bit() {
        if (cnt == 0) {
            uint8_t b = 0;
            if (pos < n) {
                b = p[pos++];
                if (b == 0xff) {
                    if (pos < n && p[pos] == 0x00) pos++;          // stuffed zero
                    else { pos--; b = 0; }                          // a marker: feed zeros, the caller handles it
                }
            }
            acc = b;
            cnt = 8;
        }
        cnt--;
        return (int)((acc >> cnt) & 1u);
    }
    int receive(int s) {
        int v = 0;
        for (int i = 0; i < s; i++) v = (v << 1) | bit();
        return v;
    }
    int decode(const JpegHuff& h) {
        int code = 0;
        for (int l = 1; l <= 16; l++) {
            code = (code << 1) | bit();
            if (h.maxcode[l] >= 0 && code <= h.maxcode[l] && code >= h.mincode[l]) return h.vals[h.valptr[l] + code - h.mincode[l]];
        }
        throw std::runtime_error("jpeg: bad Huffman code");
    }
    void restart() {
        cnt = 0;
        // skip to just behind the RSTn marker
        while (pos + 1 < n && !(p[pos] == 0xff && p[pos + 1] >= 0xd0 && p[pos + 1] <= 0xd7)) pos++;
        if (pos + 1 < n) pos += 2;
    }
};

void decodeJpeg(const std::vector<uint8_t>& f, std::vector<float>& rgb, uint32_t& W, uint32_t& H) {
    static const uint8_t zigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                       35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    uint16_t quant[4][64] = {{0}};
    JpegHuff dc[4], ac[4];
    struct Comp { int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0, pred = 0, stride = 0; std::vector<uint8_t> plane; } comp[3];
    int ncomp = 0, restartInterval = 0;
    W = H = 0;
    size_t pos = 2;
    auto be16 = [&](size_t o) { return (int)((f[o] << 8) | f[o + 1]); };
    size_t scanStart = 0;
    while (pos + 4 <= f.size()) {
        if (f[pos] != 0xff) { pos++; continue; }
        const uint8_t m = f[pos + 1];
        if (m == 0xd8 || m == 0x01 || (m >= 0xd0 && m <= 0xd7) || m == 0xff) { pos += (m == 0xff) ? 1 : 2; continue; }
        const size_t len = (size_t)be16(pos + 2), body = pos + 4, end = pos + 2 + len;
        if (end > f.size()) throw std::runtime_error("jpeg: truncated segment");
        if (m == 0xdb) {
            for (size_t q = body; q < end;) {
                const int pq = f[q] >> 4, tq = f[q] & 15;
                q++;
                if (tq > 3) throw std::runtime_error("jpeg: bad quantisation table");
                for (int i = 0; i < 64; i++) { quant[tq][zigzag[i]] = (uint16_t)(pq ? be16(q) : f[q]); q += pq ? 2 : 1; }
            }
        } else if (m == 0xc4) {
            for (size_t q = body; q < end;) {
                const int tc = f[q] >> 4, th = f[q] & 15;
                q++;
                if (th > 3) throw std::runtime_error("jpeg: bad Huffman table");
                JpegHuff& h = tc ? ac[th] : dc[th];
                int total = 0;
                for (int l = 1; l <= 16; l++) { h.bits[l] = f[q++]; total += h.bits[l]; }
                if (total > 256 || q + total > end) throw std::runtime_error("jpeg: bad Huffman table");
                for (int i = 0; i < total; i++) h.vals[i] = f[q++];
                h.build();
            }
        } else if (m == 0xc0 || m == 0xc1) {
            if (f[body] != 8) throw std::runtime_error("jpeg: only 8-bit samples are supported");
            H = (uint32_t)be16(body + 1);
            W = (uint32_t)be16(body + 3);
            ncomp = f[body + 5];
            if (ncomp != 1 && ncomp != 3) throw std::runtime_error("jpeg: only grey and three-component images are supported");
            for (int c = 0; c < ncomp; c++) {
                comp[c].id = f[body + 6 + 3 * c];
                comp[c].h = f[body + 7 + 3 * c] >> 4;
                comp[c].v = f[body + 7 + 3 * c] & 15;
                comp[c].tq = f[body + 8 + 3 * c] & 3;
                if (comp[c].h < 1 || comp[c].h > 4 || comp[c].v < 1 || comp[c].v > 4) throw std::runtime_error("jpeg: bad sampling factors");
            }
        } else if (m == 0xc2 || (m >= 0xc3 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc)) {
            throw std::runtime_error("jpeg: only baseline (sequential Huffman) files are supported");
        } else if (m == 0xdd) {
            restartInterval = be16(body);
        } else if (m == 0xda) {
            const int ns = f[body];
            if (ns != ncomp) throw std::runtime_error("jpeg: non-interleaved scans are not supported");
            for (int k = 0; k < ns; k++) {
                const int id = f[body + 1 + 2 * k];
                for (int c = 0; c < ncomp; c++)
                    if (comp[c].id == id) { comp[c].td = f[body + 2 + 2 * k] >> 4; comp[c].ta = f[body + 2 + 2 * k] & 15; }
            }
            scanStart = end;
            break;
        }
        pos = end;
    }
    if (!W || !H || !scanStart) throw std::runtime_error("jpeg: no image data");
    int hmax = 1, vmax = 1;
    for (int c = 0; c < ncomp; c++) { hmax = std::max(hmax, comp[c].h); vmax = std::max(vmax, comp[c].v); }
    const int mcusX = (int)((W + 8 * hmax - 1) / (8 * hmax)), mcusY = (int)((H + 8 * vmax - 1) / (8 * vmax));
    for (int c = 0; c < ncomp; c++) {
        comp[c].stride = mcusX * comp[c].h * 8;
        comp[c].plane.assign((size_t)comp[c].stride * mcusY * comp[c].v * 8, 0);
        if (!dc[comp[c].td & 3].present || !ac[comp[c].ta & 3].present) throw std::runtime_error("jpeg: missing Huffman table");
    }
    float cosT[8][8];
    for (int x = 0; x < 8; x++)
        for (int u = 0; u < 8; u++) cosT[x][u] = (u == 0 ? 0.70710678f : 1.0f) * std::cos((2 * x + 1) * u * 3.14159265358979f / 16.0f) * 0.5f;
    JpegBits br{f.data(), f.size(), scanStart};
    int sinceRestart = 0;
    for (int my = 0; my < mcusY; my++)
        for (int mx = 0; mx < mcusX; mx++) {
            if (restartInterval && sinceRestart == restartInterval) {
                br.restart();
                for (int c = 0; c < ncomp; c++) comp[c].pred = 0;
                sinceRestart = 0;
            }
            sinceRestart++;
            for (int c = 0; c < ncomp; c++)
                for (int by = 0; by < comp[c].v; by++)
                    for (int bx = 0; bx < comp[c].h; bx++) {
                        float F[64] = {0.0f};
                        const int t = br.decode(dc[comp[c].td & 3]);
                        int diff = t ? br.receive(t) : 0;
                        if (t && diff < (1 << (t - 1))) diff -= (1 << t) - 1;
                        comp[c].pred += diff;
                        F[0] = (float)(comp[c].pred * quant[comp[c].tq][0]);
                        for (int k = 1; k < 64;) {
                            const int rs = br.decode(ac[comp[c].ta & 3]), r = rs >> 4, sbits = rs & 15;
                            if (sbits == 0) {
                                if (r == 15) { k += 16; continue; }
                                break;
                            }
                            k += r;
                            if (k > 63) throw std::runtime_error("jpeg: bad coefficient run");
                            int v = br.receive(sbits);
                            if (v < (1 << (sbits - 1))) v -= (1 << sbits) - 1;
                            F[zigzag[k]] = (float)(v * quant[comp[c].tq][zigzag[k]]);
                            k++;
                        }
                        float tmp[64];
                        for (int v = 0; v < 8; v++)
                            for (int x = 0; x < 8; x++) {
                                float a = 0.0f;
                                for (int u = 0; u < 8; u++) a += F[8 * v + u] * cosT[x][u];
                                tmp[8 * v + x] = a;
                            }
                        uint8_t* dst = &comp[c].plane[(size_t)((my * comp[c].v + by) * 8) * comp[c].stride + (size_t)(mx * comp[c].h + bx) * 8];
                        for (int y = 0; y < 8; y++)
                            for (int x = 0; x < 8; x++) {
                                float a = 0.0f;
                                for (int v = 0; v < 8; v++) a += tmp[8 * v + x] * cosT[y][v];
                                const int q = (int)std::floor(a + 128.5f);
                                dst[(size_t)y * comp[c].stride + x] = (uint8_t)std::min(std::max(q, 0), 255);
                            }
                    }
        }
    rgb.resize((size_t)W * H * 3);
    for (uint32_t y = 0; y < H; y++)
        for (uint32_t x = 0; x < W; x++) {
            float s[3] = {0, 128, 128};
            for (int c = 0; c < ncomp; c++)   // chroma planes: nearest sample (no filtering)
                s[c] = comp[c].plane[(size_t)(y * comp[c].v / vmax) * comp[c].stride + (size_t)(x * comp[c].h / hmax)];
            float r = s[0], g = s[0], b = s[0];
            if (ncomp == 3) {
                r = s[0] + 1.402f * (s[2] - 128.0f);
                g = s[0] - 0.344136f * (s[1] - 128.0f) - 0.714136f * (s[2] - 128.0f);
                b = s[0] + 1.772f * (s[1] - 128.0f);
            }
            float* o = &rgb[3 * ((size_t)y * W + x)];
            o[0] = srgbToLinear((uint8_t)std::min(std::max((int)std::floor(r + 0.5f), 0), 255));
            o[1] = srgbToLinear((uint8_t)std::min(std::max((int)std::floor(g + 0.5f), 0), 255));
            o[2] = srgbToLinear((uint8_t)std::min(std::max((int)std::floor(b + 0.5f), 0), 255));
        }
}

void decodePpm(const std::vector<uint8_t>& file, std::vector<float>& rgb, uint32_t& W, uint32_t& H) {
    size_t pos = 2;
    auto number = [&]() {
        for (;;) {
            while (pos < file.size() && std::isspace(file[pos])) pos++;
            if (pos < file.size() && file[pos] == '#') { while (pos < file.size() && file[pos] != '\n') pos++; continue; }
            break;
        }
        uint32_t v = 0;
        bool any = false;
        while (pos < file.size() && file[pos] >= '0' && file[pos] <= '9') { v = v * 10 + (file[pos++] - '0'); any = true; }
        if (!any) throw std::runtime_error("ppm: bad header");
        return v;
    };
    W = number(); H = number();
    const uint32_t maxv = number();
    pos++;   // the single whitespace byte after maxval
    if (!W || !H || maxv != 255 || pos + (size_t)W * H * 3 > file.size()) throw std::runtime_error("ppm: only 8-bit binary P6 files are supported");
    rgb.resize((size_t)W * H * 3);
    for (size_t i = 0; i < rgb.size(); i++) rgb[i] = srgbToLinear(file[pos + i]);
}

}  // namespace

void loadImageLinearRGB(const std::string& path, std::vector<float>& rgb, uint32_t& width, uint32_t& height) {
    const std::vector<uint8_t> file = readAll(path);
    static const uint8_t pngMagic[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (file.size() > 8 && !std::memcmp(file.data(), pngMagic, 8)) return decodePng(file, rgb, width, height);
    if (file.size() > 2 && file[0] == 'P' && file[1] == '6') return decodePpm(file, rgb, width, height);
    if (file.size() > 3 && file[0] == 0xff && file[1] == 0xd8 && file[2] == 0xff) return decodeJpeg(file, rgb, width, height);
    if (file.size() > 2 && file[0] == 'P' && (file[1] == 'F' || file[1] == 'f')) {
        std::vector<float> rgba;
        readPFM(path, rgba, width, height);   // linear already
        rgb.resize((size_t)width * height * 3);
        for (size_t i = 0; i < (size_t)width * height; i++)
            for (int c = 0; c < 3; c++) rgb[3 * i + c] = rgba[4 * i + c];
        return;
    }
    throw std::runtime_error(path + ": unsupported image format (PNG, baseline JPEG, binary PPM and PFM are read)");
}

}  // namespace gsbhost
