// Self-contained PNG decoder and encoder (8/16-bit, all colour types, non-interlaced; Adam7 is not supported
// => falls through to stb when available).  Decoded output is forced to 3 channels RGB8 with the same
// conversions stbi_load(..., 3) applies: grey -> replicate, alpha dropped, 16-bit -> high byte,
// palette -> RGB.  PNG is lossless, so the texels equal what the reference's stb_image yields
// (hw8/src/sceneio.cpp:374-379).
//
// Inflate is table driven (11-bit primary table per Huffman code, canonical bit-by-bit walk for the rare longer
// codes, 64-bit bit buffer): the first version decoded every symbol one bit at a time and PNG decoding was 60 % of
// the 250 ms glTF parse of the benchmark scene.  Images of one scene are decoded concurrently (gltf_scene.cpp).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

namespace b200rt_host {
namespace {

struct BitReader {
    const uint8_t *p, *end;
    uint64_t buf = 0;
    int cnt = 0;      // bits in `buf`
    int pad = 0;      // of which zero bits invented past the end of the stream (always the topmost ones)
    void refill() {
        while (cnt <= 56) {
            if (p < end) buf |= (uint64_t)(*p++) << cnt;
            else pad += 8;
            cnt += 8;
        }
    }
    // true once invented bits have been consumed: the stream is truncated
    bool overrun() const { return cnt < pad; }
    uint32_t peek(int n) { if (cnt < n) refill(); return (uint32_t)(buf & ((1ull << n) - 1)); }
    void drop(int n) { buf >>= n; cnt -= n; }
    uint32_t bits(int n) {
        if (n == 0) return 0;
        const uint32_t v = peek(n);
        drop(n);
        return v;
    }
};

constexpr int FAST = 11;

struct Huff {
    uint16_t count[16] = {0}, symbol[320] = {0};
    uint16_t fast[1 << FAST];          // (symbol << 4) | length, 0 = not in the primary table
    static uint32_t reverse(uint32_t c, int n) {
        uint32_t r = 0;
        for (int i = 0; i < n; i++) { r = (r << 1) | (c & 1); c >>= 1; }
        return r;
    }
    bool build(const uint8_t *len, int n) {
        memset(count, 0, sizeof(count));
        memset(fast, 0, sizeof(fast));
        for (int i = 0; i < n; i++) count[len[i]]++;
        count[0] = 0;
        uint16_t offs[16];
        offs[1] = 0;
        for (int i = 1; i < 15; i++) offs[i + 1] = offs[i] + count[i];
        uint32_t next_code[16] = {0}, code = 0;
        for (int l = 1; l <= 15; l++) { code = (code + count[l - 1]) << 1; next_code[l] = code; }
        for (int i = 0; i < n; i++) {
            const int l = len[i];
            if (!l) continue;
            symbol[offs[l]++] = (uint16_t)i;
            const uint32_t c = next_code[l]++;
            if (l <= FAST) {
                if (c >> l) return false;          // over-subscribed code
                const uint32_t r = reverse(c, l);
                for (uint32_t k = r; k < (1u << FAST); k += 1u << l) fast[k] = (uint16_t)((i << 4) | l);
            }
        }
        return true;
    }
    int decode(BitReader &br) const {
        const uint16_t e = fast[br.peek(FAST)];
        if (e) { br.drop(e & 15); return e >> 4; }
        // canonical walk, MSB of the code first (= LSB of the stream first)
        int code = 0, first = 0, index = 0;
        for (int l = 1; l <= 15; l++) {
            code |= (int)br.bits(1);
            const int c = count[l];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
};

// `limit`: the caller knows how many bytes a well-formed stream yields; anything longer is malformed.
bool inflate_zlib(const uint8_t *src, size_t n, std::vector<uint8_t> &out, size_t limit) {
    if (n < 6) return false;
    BitReader br{src + 2, src + n};
    static const uint16_t lbase[] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
    static const uint16_t lext[] = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
    static const uint16_t dbase[] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
    static const uint16_t dext[] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
    out.resize(limit);
    uint8_t *dst = out.data();
    size_t pos = 0;
    std::vector<Huff> tables(2);
    Huff &hl = tables[0], &hd = tables[1];
    int last;
    do {
        last = (int)br.bits(1);
        const int type = (int)br.bits(2);
        if (br.overrun()) return false;
        if (type == 0) {
            // stored block: back to the byte stream
            br.drop(br.cnt & 7);
            if (br.overrun()) return false;
            const uint8_t *q = br.p - (br.cnt - br.pad) / 8;      // first byte not yet consumed
            if (q < src || br.end - q < 4) return false;
            const uint32_t len = q[0] | (q[1] << 8);
            q += 4;
            if ((size_t)(br.end - q) < len || pos + len > limit) return false;
            memcpy(dst + pos, q, len);
            pos += len;
            br.p = q + len; br.buf = 0; br.cnt = 0; br.pad = 0;
            continue;
        }
        uint8_t lens[320];
        if (type == 1) {
            int i = 0;
            for (; i < 144; i++) lens[i] = 8;
            for (; i < 256; i++) lens[i] = 9;
            for (; i < 280; i++) lens[i] = 7;
            for (; i < 288; i++) lens[i] = 8;
            if (!hl.build(lens, 288)) return false;
            for (i = 0; i < 30; i++) lens[i] = 5;
            if (!hd.build(lens, 30)) return false;
        } else if (type == 2) {
            const int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
            static const uint8_t order[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};
            uint8_t cl[19] = {0};
            for (int i = 0; i < ncode; i++) cl[order[i]] = (uint8_t)br.bits(3);
            Huff &hc = hd;         // the code-length code is done before hd is built
            if (!hc.build(cl, 19)) return false;
            int i = 0;
            while (i < nlen + ndist) {
                const int sym = hc.decode(br);
                if (sym < 0 || br.overrun()) return false;
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
            uint8_t dl[32];
            memcpy(dl, lens + nlen, (size_t)ndist);
            if (!hl.build(lens, nlen) || !hd.build(dl, ndist)) return false;
        } else return false;
        while (true) {
            int sym = hl.decode(br);
            if (sym < 0) return false;
            if (sym < 256) {
                if (pos >= limit) return false;
                dst[pos++] = (uint8_t)sym;
            } else if (sym == 256) break;
            else {
                sym -= 257;
                if (sym >= 29) return false;
                const size_t len = lbase[sym] + br.bits(lext[sym]);
                const int ds = hd.decode(br);
                if (ds < 0 || ds >= 30) return false;
                const size_t dist = dbase[ds] + br.bits(dext[ds]);
                if (dist > pos || pos + len > limit) return false;
                const uint8_t *from = dst + pos - dist;
                uint8_t *to = dst + pos;
                if (dist >= len) memcpy(to, from, len);
                else for (size_t k = 0; k < len; k++) to[k] = from[k];      // overlapping run
                pos += len;
            }
            if (br.overrun()) return false;
        }
    } while (!last);
    out.resize(pos);
    return !br.overrun();
}

inline uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }
inline int paeth(int a, int b, int c) {
    int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// Reverses the scanline filter of row `cur` (filtered bytes in place) given the previous reconstructed row.
bool unfilter_row(int ft, uint8_t *cur, const uint8_t *prev, size_t stride, size_t bpp) {
    switch (ft) {
        case 0: return true;
        case 1: for (size_t i = bpp; i < stride; i++) cur[i] = (uint8_t)(cur[i] + cur[i - bpp]); return true;
        case 2: for (size_t i = 0; i < stride; i++) cur[i] = (uint8_t)(cur[i] + prev[i]); return true;
        case 3:
            for (size_t i = 0; i < bpp && i < stride; i++) cur[i] = (uint8_t)(cur[i] + (prev[i] >> 1));
            for (size_t i = bpp; i < stride; i++) cur[i] = (uint8_t)(cur[i] + ((cur[i - bpp] + prev[i]) >> 1));
            return true;
        case 4:
            for (size_t i = 0; i < bpp && i < stride; i++) cur[i] = (uint8_t)(cur[i] + paeth(0, prev[i], 0));
            for (size_t i = bpp; i < stride; i++) cur[i] = (uint8_t)(cur[i] + paeth(cur[i - bpp], prev[i], prev[i - bpp]));
            return true;
        default: return false;
    }
}

const uint32_t *crc_table() {
    static uint32_t t[256];
    static bool done = false;
    if (!done) {
        for (uint32_t n = 0; n < 256; n++) {
            uint32_t c = n;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            t[n] = c;
        }
        done = true;
    }
    return t;
}
uint32_t crc32(uint32_t crc, const uint8_t *p, size_t n) {
    const uint32_t *t = crc_table();
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = t[(crc ^ p[i]) & 255] ^ (crc >> 8);
    return ~crc;
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
    const size_t bpp_bits = (size_t)chans * depth, stride = (W * bpp_bits + 7) / 8, bpp = (bpp_bits + 7) / 8;
    const size_t expect = (stride + 1) * H;
    // The header is untrusted: deflate expands at most 1032 : 1, so a stream this short cannot hold the image
    // it announces (no allocation is sized from the IHDR alone).
    if (expect > idat.size() * 1032 + 1024) return false;
    std::vector<uint8_t> raw;
    if (!inflate_zlib(idat.data(), idat.size(), raw, expect) || raw.size() < expect) return false;
    rgb.assign((size_t)W * H * 3, 0);
    std::vector<uint8_t> zero(stride, 0);
    for (uint32_t y = 0; y < H; y++) {
        uint8_t *cur = &raw[y * (stride + 1) + 1];
        const uint8_t *prev = y ? cur - (stride + 1) : zero.data();
        if (!unfilter_row(cur[-1], cur, prev, stride, bpp)) return false;
        uint8_t *dst = &rgb[(size_t)y * W * 3];
        if (depth == 8 && ctype == 2) { memcpy(dst, cur, (size_t)W * 3); continue; }
        if (depth == 8 && ctype == 6) {
            for (uint32_t x = 0; x < W; x++) { dst[3 * x] = cur[4 * x]; dst[3 * x + 1] = cur[4 * x + 1]; dst[3 * x + 2] = cur[4 * x + 2]; }
            continue;
        }
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
    }
    w = (int32_t)W; h = (int32_t)H;
    return true;
}

// Minimal PNG writer (RGB8, filter 0, stored deflate blocks): the counterpart of the reference's optional
// stb_image_write path.  Lossless by construction; size = payload + 0.01 %.
void encode_png_rgb8(int32_t w, int32_t h, const uint8_t *rgb, std::vector<uint8_t> &out) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    out.assign(sig, sig + 8);
    auto put32 = [](std::vector<uint8_t> &v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); };
    auto chunk = [&](const char *type, const std::vector<uint8_t> &data) {
        put32(out, (uint32_t)data.size());
        const size_t at = out.size();
        out.insert(out.end(), type, type + 4);
        out.insert(out.end(), data.begin(), data.end());
        put32(out, crc32(0, &out[at], out.size() - at));
    };
    std::vector<uint8_t> ihdr;
    put32(ihdr, (uint32_t)w); put32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk("IHDR", ihdr);
    const size_t stride = (size_t)w * 3 + 1;
    std::vector<uint8_t> raw(stride * h);
    for (int32_t y = 0; y < h; y++) {
        raw[y * stride] = 0;
        memcpy(&raw[y * stride + 1], rgb + (size_t)y * w * 3, (size_t)w * 3);
    }
    std::vector<uint8_t> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t s1 = 1, s2 = 0;
    for (size_t off = 0; off < raw.size() || off == 0;) {
        const size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n >= raw.size() ? 1 : 0);
        z.push_back(n & 255); z.push_back(n >> 8); z.push_back(~n & 255); z.push_back((~n >> 8) & 255);
        z.insert(z.end(), raw.begin() + (long)off, raw.begin() + (long)(off + n));
        off += n;
        if (n == 0) break;
    }
    for (size_t i = 0; i < raw.size();) {            // adler32, reduced every 5552 bytes
        const size_t n = std::min<size_t>(5552, raw.size() - i);
        for (size_t k = 0; k < n; k++) { s1 += raw[i + k]; s2 += s1; }
        s1 %= 65521; s2 %= 65521;
        i += n;
    }
    put32(z, (s2 << 16) | s1);
    chunk("IDAT", z);
    chunk("IEND", {});
}

}  // namespace b200rt_host
