// vrt_images.cpp — the file inputs of the operator surface, host code (no CUDA):
//   * the skybox: VoxelScene(engine, voxPath, skyboxFilename) loads it with stbi_loadf(..., 4) into an RGBA32F texture
//     (source/voxels/resource/voxel_scene.cpp:132, source/engine/resource/texture_2d.cpp:22-38): a Radiance RGBE .hdr file;
//   * the blue-noise texture: VoxelRenderer loads resource/blue_noise_rgba.png with stbi_load(..., 4) into RGBA8
//     (source/voxels/voxel_renderer.cpp:22).
// stb_image (cci.20210910 in the reference's Conan file) is not in the checkout: its published decoding rules are restated here —
// RGBE -> float is ldexp(1, e - 136) * mantissa with alpha 1 (stbi__hdr_convert), flat and new-style run-length scanlines
// (stbi__hdr_load); PNG -> RGBA8 for 8 / 16-bit grey, grey+alpha, RGB, RGBA and palette images incl. tRNS, Adam7 not supported.
// Also here: SHA-256 and the git-LFS pointer check, so that a real asset (sha256 == the pointer's oid) replaces a stand-in and a
// pointer stub is rejected with VRT_E_LFS_POINTER instead of being decoded as pixels.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include <string>
#include <vector>
#include <zlib.h>
#include "../../include/vrt.h"

extern thread_local std::string g_vrt_err;
static int bad(int code, const std::string& msg) { g_vrt_err = msg; return code; }

static int slurp(const char* path, std::vector<uint8_t>& buf) {
    FILE* f = path ? fopen(path, "rb") : nullptr;
    if (!f) return bad(VRT_E_IO, std::string("Could not load image ") + (path ? path : "(null)"));
    fseek(f, 0, SEEK_END); const long sz = ftell(f); fseek(f, 0, SEEK_SET);
    if (sz < 0) { fclose(f); return bad(VRT_E_IO, "Could not load image (size)"); }
    buf.resize((size_t)sz);
    const size_t got = sz ? fread(buf.data(), 1, (size_t)sz, f) : 0;
    fclose(f);
    if (got != (size_t)sz) return bad(VRT_E_IO, "Could not load image (short read)");
    return VRT_OK;
}

// ---------------------------------------------------------------------------------------------------------------- SHA-256
namespace {
struct Sha256 {
    uint32_t h[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au, 0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};
    uint8_t block[64]; size_t fill = 0; uint64_t total = 0;
    static uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
    void compress(const uint8_t* p) {
        static const uint32_t K[64] = {
            0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u, 0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u,
            0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, 0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u,
            0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u, 0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
            0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u, 0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};
        uint32_t w[64];
        for (int i = 0; i < 16; i++) w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
        for (int i = 16; i < 64; i++) {
            const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3), s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
            w[i] = w[i - 16] + s0 + w[i - 7] + s1;
        }
        uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
        for (int i = 0; i < 64; i++) {
            const uint32_t t1 = hh + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) + ((e & f) ^ (~e & g)) + K[i] + w[i];
            const uint32_t t2 = (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
            hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
        }
        h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
    }
    void update(const uint8_t* p, size_t n) {
        total += n;
        while (n) {
            const size_t k = n < 64 - fill ? n : 64 - fill;
            memcpy(block + fill, p, k); fill += k; p += k; n -= k;
            if (fill == 64) { compress(block); fill = 0; }
        }
    }
    void finish(char hex[65]) {
        const uint64_t bits = total * 8;
        const uint8_t one = 0x80, zero = 0;
        update(&one, 1);
        while (fill != 56) update(&zero, 1);
        uint8_t len[8]; for (int i = 0; i < 8; i++) len[i] = (uint8_t)(bits >> (56 - 8 * i));
        update(len, 8);
        for (int i = 0; i < 8; i++) snprintf(hex + 8 * i, 9, "%08x", h[i]);
    }
};
}  // namespace

extern "C" int vrt_sha256(const void* data, size_t n, char hex[65]) {
    if ((!data && n) || !hex) return bad(VRT_E_INVALID, "null argument");
    Sha256 s; s.update((const uint8_t*)data, n); s.finish(hex);
    return VRT_OK;
}
extern "C" int vrt_file_sha256(const char* path, char hex[65]) {
    if (!hex) return bad(VRT_E_INVALID, "null argument");
    FILE* f = path ? fopen(path, "rb") : nullptr;
    if (!f) return bad(VRT_E_IO, std::string("cannot open ") + (path ? path : "(null)"));
    Sha256 s; uint8_t buf[1 << 16]; size_t k;
    while ((k = fread(buf, 1, sizeof(buf), f)) > 0) s.update(buf, k);
    fclose(f);
    s.finish(hex);
    return VRT_OK;
}

// git-LFS pointer: "version https://git-lfs.github.com/spec/v1\noid sha256:<64 hex>\nsize <n>\n". VRT_OK and the fields if `path` is one.
static bool parse_lfs(const std::vector<uint8_t>& b, char oid[65], uint64_t* size) {
    static const char head[] = "version https://git-lfs";
    if (b.size() < sizeof(head) - 1 || b.size() > 1024 || memcmp(b.data(), head, sizeof(head) - 1) != 0) return false;
    const std::string s((const char*)b.data(), b.size());
    const size_t o = s.find("oid sha256:"), z = s.find("size ");
    if (o == std::string::npos || z == std::string::npos || o + 11 + 64 > s.size()) return false;
    if (oid) { memcpy(oid, s.data() + o + 11, 64); oid[64] = 0; }
    if (size) *size = strtoull(s.c_str() + z + 5, nullptr, 10);
    return true;
}
extern "C" int vrt_lfs_pointer_info(const char* path, char oid_hex[65], uint64_t* size) {
    std::vector<uint8_t> b;
    FILE* f = path ? fopen(path, "rb") : nullptr;
    if (!f) return bad(VRT_E_IO, std::string("cannot open ") + (path ? path : "(null)"));
    b.resize(1025); b.resize(fread(b.data(), 1, 1025, f)); fclose(f);
    if (!parse_lfs(b, oid_hex, size)) return bad(VRT_E_INVALID, "not a git-LFS pointer file");
    return VRT_OK;
}
// The asset a config names: `candidate` if it is real data whose sha256 and size are those of the LFS pointer at `pointer_path`
// (the reference checkout's stub), i.e. the real file has appeared. VRT_OK: use it and drop the "stand-in" label.
extern "C" int vrt_asset_matches_pointer(const char* candidate, const char* pointer_path) {
    char oid[65], hex[65]; uint64_t size = 0;
    int rc = vrt_lfs_pointer_info(pointer_path, oid, &size);
    if (rc) return rc;
    std::vector<uint8_t> head;
    FILE* f = candidate ? fopen(candidate, "rb") : nullptr;
    if (!f) return bad(VRT_E_IO, std::string("cannot open ") + (candidate ? candidate : "(null)"));
    head.resize(1025); head.resize(fread(head.data(), 1, 1025, f));
    fseek(f, 0, SEEK_END); const long sz = ftell(f); fclose(f);
    if (parse_lfs(head, nullptr, nullptr)) return bad(VRT_E_LFS_POINTER, "the candidate is itself a git-LFS pointer stub");
    if ((uint64_t)sz != size) return bad(VRT_E_INVALID, "size differs from the LFS pointer's");
    rc = vrt_file_sha256(candidate, hex);
    if (rc) return rc;
    if (strcmp(hex, oid) != 0) return bad(VRT_E_INVALID, std::string("sha256 ") + hex + " differs from the LFS pointer's oid " + oid);
    return VRT_OK;
}

// ---------------------------------------------------------------------------------------------------------------- Radiance RGBE
static bool hdr_line(const std::vector<uint8_t>& b, size_t& at, std::string& line) {
    line.clear();
    if (at >= b.size()) return false;
    while (at < b.size() && b[at] != '\n') { if (line.size() < 1023) line.push_back((char)b[at]); at++; }
    if (at < b.size()) at++;
    return true;
}
static void rgbe_to_float(const uint8_t* p, float* out) {   // stbi__hdr_convert, req_comp 4
    if (p[3] != 0) {
        const float f1 = ldexpf(1.0f, (int)p[3] - (128 + 8));
        out[0] = (float)p[0] * f1; out[1] = (float)p[1] * f1; out[2] = (float)p[2] * f1;
    } else out[0] = out[1] = out[2] = 0.0f;
    out[3] = 1.0f;
}
extern "C" int vrt_image_decode_hdr(const uint8_t* bytes, size_t n, float** rgba32f_out, uint32_t* w_out, uint32_t* h_out) {
    if (!bytes || !rgba32f_out || !w_out || !h_out) return bad(VRT_E_INVALID, "null argument");
    try {
        const std::vector<uint8_t> b(bytes, bytes + n);
        if (parse_lfs(b, nullptr, nullptr)) return bad(VRT_E_LFS_POINTER, "file is a git-LFS pointer stub, not image data");
        size_t at = 0; std::string line;
        if (!hdr_line(b, at, line) || (line != "#?RADIANCE" && line != "#?RGBE")) return bad(VRT_E_PARSE, "Could not load image (not a Radiance HDR file)");
        bool fmt = false;
        for (;;) {
            if (!hdr_line(b, at, line)) return bad(VRT_E_PARSE, "Could not load image (HDR header)");
            if (line.empty()) break;
            if (line == "FORMAT=32-bit_rle_rgbe") fmt = true;
        }
        if (!fmt) return bad(VRT_E_PARSE, "Could not load image (unsupported HDR format)");
        if (!hdr_line(b, at, line) || line.compare(0, 3, "-Y ") != 0) return bad(VRT_E_PARSE, "Could not load image (unsupported HDR data layout)");
        char* end = nullptr;
        const long H = strtol(line.c_str() + 3, &end, 10);
        while (*end == ' ') end++;
        if (strncmp(end, "+X ", 3) != 0) return bad(VRT_E_PARSE, "Could not load image (unsupported HDR data layout)");
        const long W = strtol(end + 3, nullptr, 10);
        if (W <= 0 || H <= 0 || W > (1 << 24) || H > (1 << 24) || (uint64_t)W * H > ((uint64_t)1 << 28)) return bad(VRT_E_PARSE, "Could not load image (HDR size)");
        float* out = (float*)malloc((size_t)W * H * 16);
        if (!out) return bad(VRT_E_OOM, "host allocation failed");
        bool flat = W < 8 || W >= 32768;
        std::vector<uint8_t> scan((size_t)W * 4);
        for (long y = 0; y < H; y++) {
            if (!flat) {
                if (at + 4 > b.size()) { free(out); return bad(VRT_E_PARSE, "Could not load image (HDR data truncated)"); }
                const uint8_t c1 = b[at], c2 = b[at + 1], len = b[at + 2];
                if (c1 != 2 || c2 != 2 || (len & 0x80)) {
                    if (y != 0) { free(out); return bad(VRT_E_PARSE, "Could not load image (HDR scanline)"); }
                    flat = true;                                            // not run-length encoded: the whole image is flat RGBE
                } else {
                    if ((((long)len << 8) | b[at + 3]) != W) { free(out); return bad(VRT_E_PARSE, "Could not load image (invalid decoded scanline length)"); }
                    at += 4;
                    for (int k = 0; k < 4; k++) {                           // each of R, G, B, E run-length coded on its own
                        long i = 0;
                        while (i < W) {
                            if (at >= b.size()) { free(out); return bad(VRT_E_PARSE, "Could not load image (HDR data truncated)"); }
                            uint8_t count = b[at++];
                            if (count > 128) {
                                count -= 128;
                                if (count == 0 || count > W - i || at >= b.size()) { free(out); return bad(VRT_E_PARSE, "Could not load image (corrupt HDR run)"); }
                                const uint8_t v = b[at++];
                                for (int z = 0; z < count; z++) scan[(size_t)(i++) * 4 + k] = v;
                            } else {
                                if (count == 0 || count > W - i || at + count > b.size()) { free(out); return bad(VRT_E_PARSE, "Could not load image (corrupt HDR run)"); }
                                for (int z = 0; z < count; z++) scan[(size_t)(i++) * 4 + k] = b[at++];
                            }
                        }
                    }
                    for (long x = 0; x < W; x++) rgbe_to_float(&scan[(size_t)x * 4], out + ((size_t)y * W + x) * 4);
                    continue;
                }
            }
            if (at + (size_t)W * 4 > b.size()) { free(out); return bad(VRT_E_PARSE, "Could not load image (HDR data truncated)"); }
            for (long x = 0; x < W; x++) rgbe_to_float(&b[at + (size_t)x * 4], out + ((size_t)y * W + x) * 4);
            at += (size_t)W * 4;
        }
        *rgba32f_out = out; *w_out = (uint32_t)W; *h_out = (uint32_t)H;
        return VRT_OK;
    } catch (const std::bad_alloc&) { return bad(VRT_E_OOM, "out of host memory"); }
}

// ---------------------------------------------------------------------------------------------------------------- PNG
static uint32_t rd32(const uint8_t* p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }
static int paeth(int a, int b, int c) { const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c); return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); }

extern "C" int vrt_image_decode_png(const uint8_t* bytes, size_t n, uint8_t** rgba8_out, uint32_t* w_out, uint32_t* h_out) {
    if (!bytes || !rgba8_out || !w_out || !h_out) return bad(VRT_E_INVALID, "null argument");
    try {
        const std::vector<uint8_t> b(bytes, bytes + n);
        if (parse_lfs(b, nullptr, nullptr)) return bad(VRT_E_LFS_POINTER, "file is a git-LFS pointer stub, not image data");
        static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
        if (n < 8 || memcmp(bytes, sig, 8) != 0) return bad(VRT_E_PARSE, "Could not load image (not a PNG file)");
        uint32_t W = 0, H = 0; int depth = 0, ctype = -1, interlace = 0;
        std::vector<uint8_t> idat, plte; uint8_t trnsPal[256]; memset(trnsPal, 255, sizeof(trnsPal));
        bool haveTrns = false; uint16_t trnsKey[3] = {0, 0, 0};
        size_t at = 8; bool end = false;
        while (!end && at + 12 <= n) {
            const uint32_t len = rd32(&b[at]); const uint8_t* type = &b[at + 4];
            if ((uint64_t)at + 12 + len > n) return bad(VRT_E_PARSE, "Could not load image (PNG chunk truncated)");
            const uint8_t* d = &b[at + 8];
            if (!memcmp(type, "IHDR", 4)) {
                if (len != 13) return bad(VRT_E_PARSE, "Could not load image (bad IHDR)");
                W = rd32(d); H = rd32(d + 4); depth = d[8]; ctype = d[9]; interlace = d[12];
                if (d[10] != 0 || d[11] != 0) return bad(VRT_E_PARSE, "Could not load image (bad PNG compression / filter method)");
            } else if (!memcmp(type, "PLTE", 4)) plte.assign(d, d + len);
            else if (!memcmp(type, "tRNS", 4)) {
                haveTrns = true;
                if (ctype == 3) memcpy(trnsPal, d, len < 256 ? len : 256);
                else if (ctype == 0 && len >= 2) trnsKey[0] = (uint16_t)(d[0] << 8 | d[1]);
                else if (ctype == 2 && len >= 6) for (int k = 0; k < 3; k++) trnsKey[k] = (uint16_t)(d[2 * k] << 8 | d[2 * k + 1]);
            } else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), d, d + len);
            else if (!memcmp(type, "IEND", 4)) end = true;
            at += 12 + (size_t)len;
        }
        static const int chans[7] = {1, 0, 3, 1, 2, 0, 4};
        if (ctype < 0 || ctype > 6 || !chans[ctype] || W == 0 || H == 0 || (uint64_t)W * H > ((uint64_t)1 << 28)) return bad(VRT_E_PARSE, "Could not load image (bad PNG header)");
        if (!(depth == 8 || depth == 16 || ((ctype == 0 || ctype == 3) && (depth == 1 || depth == 2 || depth == 4)))) return bad(VRT_E_PARSE, "Could not load image (bad PNG bit depth)");
        if (interlace) return bad(VRT_E_UNSUPPORTED, "interlaced (Adam7) PNG files are not supported");
        const int ch = chans[ctype];
        const size_t bpp = (size_t)(ch * depth + 7) / 8, stride = ((size_t)W * ch * depth + 7) / 8;
        std::vector<uint8_t> raw((stride + 1) * H);
        uLongf rawLen = (uLongf)raw.size();
        if (uncompress(raw.data(), &rawLen, idat.data(), (uLong)idat.size()) != Z_OK || rawLen != raw.size()) return bad(VRT_E_PARSE, "Could not load image (PNG data does not inflate)");
        // undo the scanline filters in place
        for (uint32_t y = 0; y < H; y++) {
            uint8_t* cur = &raw[(stride + 1) * y + 1]; const uint8_t* up = y ? cur - (stride + 1) : nullptr;
            const int filter = raw[(stride + 1) * y];
            for (size_t i = 0; i < stride; i++) {
                const int a = i >= bpp ? cur[i - bpp] : 0, bb = up ? up[i] : 0, c = (up && i >= bpp) ? up[i - bpp] : 0;
                int v = cur[i];
                switch (filter) { case 0: break; case 1: v += a; break; case 2: v += bb; break; case 3: v += (a + bb) >> 1; break; case 4: v += paeth(a, bb, c); break;
                                  default: return bad(VRT_E_PARSE, "Could not load image (bad PNG filter)"); }
                cur[i] = (uint8_t)v;
            }
        }
        uint8_t* out = (uint8_t*)malloc((size_t)W * H * 4);
        if (!out) return bad(VRT_E_OOM, "host allocation failed");
        const int maxv = (1 << (depth < 8 ? depth : 8)) - 1;
        for (uint32_t y = 0; y < H; y++) {
            const uint8_t* row = &raw[(stride + 1) * y + 1];
            for (uint32_t x = 0; x < W; x++) {
                uint16_t s[4] = {0, 0, 0, 0};                              // samples at file depth
                for (int k = 0; k < ch; k++) {
                    if (depth == 16) s[k] = (uint16_t)(row[((size_t)x * ch + k) * 2] << 8 | row[((size_t)x * ch + k) * 2 + 1]);
                    else if (depth == 8) s[k] = row[(size_t)x * ch + k];
                    else { const size_t bit = (size_t)x * depth; s[k] = (row[bit >> 3] >> (8 - depth - (bit & 7))) & maxv; }
                }
                auto to8 = [&](uint16_t v) -> uint8_t { return depth == 16 ? (uint8_t)(v >> 8) : (depth == 8 ? (uint8_t)v : (uint8_t)(v * (255 / maxv))); };
                uint8_t* o = out + ((size_t)y * W + x) * 4;
                if (ctype == 3) {
                    const size_t pi = s[0];
                    if (pi * 3 + 2 >= plte.size()) { o[0] = o[1] = o[2] = 0; o[3] = 255; }
                    else { o[0] = plte[pi * 3]; o[1] = plte[pi * 3 + 1]; o[2] = plte[pi * 3 + 2]; o[3] = trnsPal[pi & 255]; }
                } else if (ctype == 0) { o[0] = o[1] = o[2] = to8(s[0]); o[3] = (haveTrns && s[0] == trnsKey[0]) ? 0 : 255; }
                else if (ctype == 4) { o[0] = o[1] = o[2] = to8(s[0]); o[3] = to8(s[1]); }
                else if (ctype == 2) { o[0] = to8(s[0]); o[1] = to8(s[1]); o[2] = to8(s[2]); o[3] = (haveTrns && s[0] == trnsKey[0] && s[1] == trnsKey[1] && s[2] == trnsKey[2]) ? 0 : 255; }
                else { o[0] = to8(s[0]); o[1] = to8(s[1]); o[2] = to8(s[2]); o[3] = to8(s[3]); }
            }
        }
        *rgba8_out = out; *w_out = W; *h_out = H;
        return VRT_OK;
    } catch (const std::bad_alloc&) { return bad(VRT_E_OOM, "out of host memory"); }
}

extern "C" int vrt_image_read_hdr(const char* path, float** rgba32f_out, uint32_t* w, uint32_t* h) {
    std::vector<uint8_t> b;
    const int rc = slurp(path, b);
    return rc ? rc : vrt_image_decode_hdr(b.data(), b.size(), rgba32f_out, w, h);
}
extern "C" int vrt_image_read_png(const char* path, uint8_t** rgba8_out, uint32_t* w, uint32_t* h) {
    std::vector<uint8_t> b;
    const int rc = slurp(path, b);
    return rc ? rc : vrt_image_decode_png(b.data(), b.size(), rgba8_out, w, h);
}

// VoxelScene's skybox (voxel_scene.cpp:132) and VoxelRenderer's noise texture (voxel_renderer.cpp:22) from their files
extern "C" int vrt_scene_load_skybox(vrt_scene* s, const char* path) {
    if (!s) return bad(VRT_E_INVALID, "null scene");
    float* px = nullptr; uint32_t w = 0, h = 0;
    int rc = vrt_image_read_hdr(path, &px, &w, &h);
    if (rc) return rc;
    rc = vrt_scene_set_skybox(s, px, w, h);
    free(px);
    return rc;
}
extern "C" int vrt_scene_load_noise(vrt_scene* s, const char* path) {
    if (!s) return bad(VRT_E_INVALID, "null scene");
    uint8_t* px = nullptr; uint32_t w = 0, h = 0;
    int rc = vrt_image_read_png(path, &px, &w, &h);
    if (rc) return rc;
    rc = vrt_scene_set_noise(s, px, w, h);                                 // rejects anything but 512 x 512 (NOISE_SIZE, voxel_volume.frag:70)
    free(px);
    return rc;
}
