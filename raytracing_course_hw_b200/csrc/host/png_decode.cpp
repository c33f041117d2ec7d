// Self-contained PNG decoder (8/16-bit, all colour types, non-interlaced + Adam7 not supported
// => falls through to stb when available).  Output is forced to 3 channels RGB8 with the same
// conversions stbi_load(..., 3) applies: grey -> replicate, alpha dropped, 16-bit -> high byte,
// palette -> RGB.  PNG is lossless, so the texels equal what the reference's stb_image yields
// (hw8/src/sceneio.cpp:374-379).
#include <cstdint>
#include <cstring>
#include <vector>

namespace b200rt_host {
namespace {

struct BitReader {
    const uint8_t *p, *end;
    uint32_t buf = 0; int cnt = 0;
    bool ok = true;
    uint32_t bits(int n) {
        while (cnt < n) {
            if (p >= end) { ok = false; return 0; }
            buf |= (uint32_t)(*p++) << cnt; cnt += 8;
        }
        uint32_t v = buf & ((1u << n) - 1);
        buf >>= n; cnt -= n;
        return n ? v : 0;
    }
    void align() { buf = 0; cnt = 0; }
};

struct Huff {
    uint16_t count[16] = {0}, symbol[320] = {0};
    void build(const uint8_t *len, int n) {
        memset(count, 0, sizeof(count));
        for (int i = 0; i < n; i++) count[len[i]]++;
        count[0] = 0;
        uint16_t offs[16]; offs[1] = 0;
        for (int i = 1; i < 15; i++) offs[i + 1] = offs[i] + count[i];
        for (int i = 0; i < n; i++) if (len[i]) symbol[offs[len[i]]++] = (uint16_t)i;
    }
    int decode(BitReader &br) const {
        int code = 0, first = 0, index = 0;
        for (int l = 1; l <= 15; l++) {
            code |= (int)br.bits(1);
            if (!br.ok) return -1;
            int c = count[l];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
};

bool inflate_zlib(const uint8_t *src, size_t n, std::vector<uint8_t> &out) {
    if (n < 6) return false;
    BitReader br{src + 2, src + n};
    static const uint16_t lbase[] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
    static const uint16_t lext[] = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
    static const uint16_t dbase[] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
    static const uint16_t dext[] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
    int last;
    do {
        last = (int)br.bits(1);
        int type = (int)br.bits(2);
        if (!br.ok) return false;
        if (type == 0) {
            br.align();
            if (br.end - br.p < 4) return false;
            uint32_t len = br.p[0] | (br.p[1] << 8);
            br.p += 4;
            if ((size_t)(br.end - br.p) < len) return false;
            out.insert(out.end(), br.p, br.p + len);
            br.p += len;
            continue;
        }
        Huff hl, hd;
        uint8_t lens[320];
        if (type == 1) {
            int i = 0;
            for (; i < 144; i++) lens[i] = 8;
            for (; i < 256; i++) lens[i] = 9;
            for (; i < 280; i++) lens[i] = 7;
            for (; i < 288; i++) lens[i] = 8;
            hl.build(lens, 288);
            for (i = 0; i < 30; i++) lens[i] = 5;
            hd.build(lens, 30);
        } else if (type == 2) {
            int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
            static const uint8_t order[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};
            uint8_t cl[19] = {0};
            for (int i = 0; i < ncode; i++) cl[order[i]] = (uint8_t)br.bits(3);
            Huff hc; hc.build(cl, 19);
            int i = 0;
            while (i < nlen + ndist) {
                int sym = hc.decode(br);
                if (sym < 0) return false;
                if (sym < 16) lens[i++] = (uint8_t)sym;
                else {
                    int rep, val = 0;
                    if (sym == 16) { if (i == 0) return false; val = lens[i - 1]; rep = 3 + (int)br.bits(2); }
                    else if (sym == 17) rep = 3 + (int)br.bits(3);
                    else rep = 11 + (int)br.bits(7);
                    if (i + rep > nlen + ndist) return false;
                    while (rep--) lens[i++] = (uint8_t)val;
                }
            }
            hl.build(lens, nlen);
            hd.build(lens + nlen, ndist);
        } else return false;
        while (true) {
            int sym = hl.decode(br);
            if (sym < 0 || !br.ok) return false;
            if (sym < 256) out.push_back((uint8_t)sym);
            else if (sym == 256) break;
            else {
                sym -= 257;
                if (sym >= 29) return false;
                int len = lbase[sym] + (int)br.bits(lext[sym]);
                int ds = hd.decode(br);
                if (ds < 0 || ds >= 30) return false;
                size_t dist = dbase[ds] + br.bits(dext[ds]);
                if (dist > out.size()) return false;
                size_t from = out.size() - dist;
                for (int k = 0; k < len; k++) out.push_back(out[from + k]);
            }
        }
    } while (!last);
    return br.ok;
}

inline uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }
inline int paeth(int a, int b, int c) {
    int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

}  // namespace

bool decode_png_rgb8(const std::vector<uint8_t> &f, int32_t &w, int32_t &h, std::vector<uint8_t> &rgb) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (f.size() < 33 || memcmp(f.data(), sig, 8) != 0) return false;
    size_t pos = 8;
    uint32_t W = 0, H = 0; int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, palette;
    while (pos + 12 <= f.size()) {
        uint32_t len = be32(&f[pos]);
        const uint8_t *type = &f[pos + 4], *data = &f[pos + 8];
        if (pos + 12 + (size_t)len > f.size()) return false;
        if (!memcmp(type, "IHDR", 4) && len >= 13) {
            W = be32(data); H = be32(data + 4); depth = data[8]; ctype = data[9]; interlace = data[12];
        } else if (!memcmp(type, "PLTE", 4)) palette.assign(data, data + len);
        else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
        else if (!memcmp(type, "IEND", 4)) break;
        pos += 12 + (size_t)len;
    }
    if (!W || !H || interlace || W > (1u << 15) || H > (1u << 15)) return false;
    int chans = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!chans || (depth != 8 && depth != 16 && !(depth < 8 && (ctype == 0 || ctype == 3)))) return false;
    size_t bpp_bits = (size_t)chans * depth, stride = (W * bpp_bits + 7) / 8, bpp = (bpp_bits + 7) / 8;
    std::vector<uint8_t> raw;
    raw.reserve((stride + 1) * H);
    if (!inflate_zlib(idat.data(), idat.size(), raw) || raw.size() < (stride + 1) * H) return false;
    std::vector<uint8_t> prev(stride, 0), cur(stride);
    rgb.assign((size_t)W * H * 3, 0);
    for (uint32_t y = 0; y < H; y++) {
        const uint8_t *src = &raw[y * (stride + 1)];
        int ft = src[0];
        for (size_t i = 0; i < stride; i++) {
            int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0, x = src[1 + i];
            switch (ft) {
                case 0: break; case 1: x += a; break; case 2: x += b; break;
                case 3: x += (a + b) >> 1; break; case 4: x += paeth(a, b, c); break;
                default: return false;
            }
            cur[i] = (uint8_t)x;
        }
        uint8_t *dst = &rgb[(size_t)y * W * 3];
        for (uint32_t x = 0; x < W; x++) {
            uint8_t s[4] = {0, 0, 0, 0};
            if (depth >= 8) {
                size_t step = depth / 8;
                for (int k = 0; k < chans; k++) s[k] = cur[(x * chans + k) * step];  // 16-bit: high byte
            } else {
                size_t bit = (size_t)x * depth;
                int v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1);
                s[0] = ctype == 3 ? (uint8_t)v : (uint8_t)(v * (255 / ((1 << depth) - 1)));
            }
            if (ctype == 3) {
                size_t pi = (size_t)s[0] * 3;
                if (pi + 2 < palette.size()) { dst[0] = palette[pi]; dst[1] = palette[pi + 1]; dst[2] = palette[pi + 2]; }
            } else if (ctype == 0 || ctype == 4) { dst[0] = dst[1] = dst[2] = s[0]; }
            else { dst[0] = s[0]; dst[1] = s[1]; dst[2] = s[2]; }
            dst += 3;
        }
        prev.swap(cur);
    }
    w = (int32_t)W; h = (int32_t)H;
    return true;
}

}  // namespace b200rt_host
