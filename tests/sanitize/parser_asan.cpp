// parser_asan — the product's host-side parsers (csrc/vrt_vox.cpp: MagicaVoxel .vox; csrc/vrt_images.cpp: Radiance .hdr, PNG, sha256,
// LFS pointers; csrc/vrt_hostside.cpp: image writers, jitter / frame counter) under AddressSanitizer + UndefinedBehaviorSanitizer.
// These read files a user hands in (a crafted .vox once crashed the loader), so they are fed the golden files and several
// thousand damaged copies: truncations, random byte flips, 32-bit fields overwritten with extreme values. Built by
// tests/test_sanitize.py from the product's own sources with g++ -fsanitize=address,undefined -fno-sanitize-recover=all: any report
// ends the process with a non-zero code. The parsers must answer every input with VRT_OK or an error code — never a crash, a leak, an
// out-of-bounds access or an allocation beyond their documented caps.
//   usage: parser_asan <golden.vox> [more.vox ...]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <zlib.h>
#include "../../voxel-raytracing_b200/csrc/vrt_vox.h"

thread_local std::string g_vrt_err;                                         // defined in vrt_api.cu in the library
extern "C" int vrt_scene_set_skybox(vrt_scene*, const float*, uint32_t, uint32_t) { return VRT_OK; }   // (device uploads: not part of this check)
extern "C" int vrt_scene_set_noise(vrt_scene*, const uint8_t*, uint32_t, uint32_t) { return VRT_OK; }

static uint32_t g_seed = 2024u;
static uint32_t rnd() { g_seed = g_seed * 1664525u + 1013904223u; return g_seed >> 8; }

static std::vector<uint8_t> slurp(const char* path) {
    std::vector<uint8_t> b;
    FILE* f = fopen(path, "rb");
    if (!f) return b;
    fseek(f, 0, SEEK_END); const long n = ftell(f); fseek(f, 0, SEEK_SET);
    b.resize((size_t)n);
    if (n > 0 && fread(b.data(), 1, (size_t)n, f) != (size_t)n) b.clear();
    fclose(f);
    return b;
}
static std::vector<uint8_t> damaged(const std::vector<uint8_t>& good, int k) {
    std::vector<uint8_t> b = good;
    if (k == 0 || b.size() < 8) return b;
    switch (rnd() % 4u) {
        case 0: b.resize(rnd() % b.size()); break;
        case 1: for (int j = 0, n = 1 + (int)(rnd() % 8u); j < n; j++) b[rnd() % b.size()] = (uint8_t)rnd(); break;
        case 2: { const size_t at = rnd() % (b.size() - 3); const uint32_t ext[] = {0xffffffffu, 0x80000000u, 0x7fffffffu, 0u, 0x00010000u, rnd()}; memcpy(&b[at], &ext[rnd() % 6u], 4); break; }
        default: { const size_t at = rnd() % b.size(); b.insert(b.begin() + (long)at, (size_t)(rnd() % 64u), (uint8_t)rnd()); break; }
    }
    return b;
}

static void put32be(std::vector<uint8_t>& v, uint32_t x) { for (int s = 24; s >= 0; s -= 8) v.push_back((uint8_t)(x >> s)); }
static void png_chunk(std::vector<uint8_t>& out, const char* type, const std::vector<uint8_t>& body) {
    put32be(out, (uint32_t)body.size());
    std::vector<uint8_t> t(type, type + 4); t.insert(t.end(), body.begin(), body.end());
    out.insert(out.end(), t.begin(), t.end());
    put32be(out, (uint32_t)crc32(0L, t.data(), (uInt)t.size()));
}
// a valid PNG of the given colour type (0 grey, 2 RGB, 3 palette, 4 grey+alpha, 6 RGBA), 8 bits, filter types cycling per row
static std::vector<uint8_t> make_png(int w, int h, int ctype) {
    const int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;
    std::vector<uint8_t> raw;
    for (int y = 0; y < h; y++) { raw.push_back((uint8_t)(y % 5)); for (int i = 0; i < w * ch; i++) raw.push_back((uint8_t)rnd()); }
    std::vector<uint8_t> z(compressBound((uLong)raw.size())); uLongf zn = (uLongf)z.size();
    compress2(z.data(), &zn, raw.data(), (uLong)raw.size(), 6); z.resize(zn);
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a}, ihdr;
    put32be(ihdr, (uint32_t)w); put32be(ihdr, (uint32_t)h); ihdr.push_back(8); ihdr.push_back((uint8_t)ctype); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    png_chunk(out, "IHDR", ihdr);
    if (ctype == 3) { std::vector<uint8_t> pal(256 * 3); for (auto& b : pal) b = (uint8_t)rnd(); png_chunk(out, "PLTE", pal); }
    const size_t half = z.size() / 2;                                      // two IDAT chunks
    png_chunk(out, "IDAT", std::vector<uint8_t>(z.begin(), z.begin() + (long)half));
    png_chunk(out, "IDAT", std::vector<uint8_t>(z.begin() + (long)half, z.end()));
    png_chunk(out, "IEND", {});
    return out;
}
// a valid Radiance picture: flat pixels, or new-style run-length scanlines (width 8..32767)
static std::vector<uint8_t> make_hdr(int w, int h, bool rle) {
    char head[128]; const int hn = snprintf(head, sizeof(head), "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", h, w);
    std::vector<uint8_t> out(head, head + hn);
    for (int y = 0; y < h; y++) {
        if (!rle) { for (int i = 0; i < w * 4; i++) out.push_back((uint8_t)(1 + rnd() % 254u)); continue; }
        out.push_back(2); out.push_back(2); out.push_back((uint8_t)(w >> 8)); out.push_back((uint8_t)(w & 255));
        for (int c = 0; c < 4; c++)
            for (int x = 0; x < w;) {
                const int left = w - x;
                if (rnd() & 1u) { const int n = 1 + (int)(rnd() % (uint32_t)(left < 127 ? left : 127)); out.push_back((uint8_t)(128 + n)); out.push_back((uint8_t)rnd()); x += n; }
                else { const int n = 1 + (int)(rnd() % (uint32_t)(left < 128 ? left : 128)); out.push_back((uint8_t)n); for (int i = 0; i < n; i++) out.push_back((uint8_t)rnd()); x += n; }
            }
    }
    return out;
}

int main(int argc, char** argv) {
    long accepted = 0, rejected = 0;
    // ---- .vox: the golden files and damaged copies, through both routes (layout for the device flatten; dense volume on the host)
    for (int a = 1; a < argc; a++) {
        const std::vector<uint8_t> good = slurp(argv[a]);
        if (good.empty()) { fprintf(stderr, "cannot read %s\n", argv[a]); return 6; }
        for (int k = 0; k < 1000; k++) {
            const std::vector<uint8_t> b = damaged(good, k);
            std::string err;
            vrt::VoxLayout lay;
            const int rc = vrt::parse_vox_bytes(b.data(), b.size(), lay, err);
            if (k == 0 && rc != VRT_OK) { fprintf(stderr, "%s: the golden file was rejected: %s\n", argv[a], err.c_str()); return 7; }
            if (rc == VRT_OK) {
                accepted++;
                if ((unsigned long long)lay.w * lay.h * lay.d <= (64ull << 20)) {      // (a damaged translation can make a sparse giant scene: parse only)
                    std::vector<uint8_t> vol((size_t)lay.w * lay.h * lay.d);
                    vrt::flatten_host(lay, vol.data());
                    for (size_t i = 0; i < lay.instances.size(); i++) (void)vrt::flatten_instance_params(lay, i);
                }
            } else rejected++;
            if ((k & 7) == 0) { vrt::DenseScene ds; std::string e2; (void)vrt::load_vox_bytes(b.data(), b.size(), ds, e2); }   // what vrt_vox_decode runs
        }
    }
    // ---- .png / .hdr: generated valid pictures of every layout the readers accept, and damaged copies
    for (int ctype : {0, 2, 3, 4, 6}) {
        const std::vector<uint8_t> good = make_png(13 + ctype, 9, ctype);
        for (int k = 0; k < 800; k++) {
            const std::vector<uint8_t> b = damaged(good, k);
            uint8_t* px = nullptr; uint32_t w = 0, h = 0;
            const int rc = vrt_image_decode_png(b.data(), b.size(), &px, &w, &h);
            if (k == 0 && rc != VRT_OK) { fprintf(stderr, "a valid PNG of colour type %d was rejected: %s\n", ctype, g_vrt_err.c_str()); return 8; }
            if (rc == VRT_OK) { accepted++; volatile uint8_t t = px[(size_t)w * h * 4 - 1]; (void)t; free(px); } else rejected++;
        }
    }
    for (int rle = 0; rle < 2; rle++) {
        const std::vector<uint8_t> good = make_hdr(rle ? 40 : 7, 6, rle != 0);
        for (int k = 0; k < 1500; k++) {
            const std::vector<uint8_t> b = damaged(good, k);
            float* px = nullptr; uint32_t w = 0, h = 0;
            const int rc = vrt_image_decode_hdr(b.data(), b.size(), &px, &w, &h);
            if (k == 0 && rc != VRT_OK) { fprintf(stderr, "a valid %s .hdr was rejected: %s\n", rle ? "run-length" : "flat", g_vrt_err.c_str()); return 9; }
            if (rc == VRT_OK) { accepted++; volatile float t = px[(size_t)w * h * 4 - 1]; (void)t; free(px); } else rejected++;
        }
    }
    // ---- sha256 / LFS pointers / writers / frame driver
    char hex[65];
    if (vrt_sha256("abc", 3, hex) != VRT_OK || strcmp(hex, "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad") != 0) { fprintf(stderr, "sha256(abc) is wrong\n"); return 10; }
    for (size_t n : {0u, 55u, 56u, 63u, 64u, 65u, 1000u}) { std::vector<uint8_t> m(n, 0x61); vrt_sha256(m.data(), n, hex); }
    for (int rw = 1; rw <= 3840; rw = rw * 3 + 1) {
        const int pc = vrt_jitter_phase_count(rw, 3840);
        float x, y; vrt_jitter_offset(rw, pc, &x, &y);
        int fc = 0; for (int k = 0; k < 5; k++) vrt_upscaler_update(rw, 3840, &fc, &x, &y);
        const uint32_t tgt[2] = {3840, 2160}; uint32_t rr[2];
        for (uint32_t scal = 0; scal < 20; scal += 5) vrt_render_resolution(1, scal, tgt, rr);
    }
    printf("parser sanitize ok: %ld inputs accepted, %ld rejected\n", accepted, rejected);
    return 0;
}
