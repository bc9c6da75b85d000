// vrt_hostside.cpp — host code either side of the render pass, no CUDA:
//   * the frame driver's jitter sequence and frame counter: UpscalerStage::update (source/voxels/stages/upscaler_stage.cpp:59-70) on
//     ffxFsr2GetJitterPhaseCount / ffxFsr2GetJitterOffset (FSR2 v2.1.0, declared thirdparty/fsr2/include/ffx_fsr2.h:356,395; only headers
//     and LFS stubs of the Windows .lib are in the checkout, so the library's published Halton(2,3) algorithm is implemented here);
//   * VoxelRenderSettings::renderResolution (source/voxels/voxel_render_settings.cpp:3-13);
//   * image files for the headless path (the reference presents to a swap chain instead): binary PPM and PNG (stored deflate blocks).
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/vrt.h"

extern "C" int vrt_jitter_phase_count(int renderWidth, int displayWidth) {
    if (renderWidth <= 0 || displayWidth <= 0) return 0;
    const float ratio = (float)displayWidth / (float)renderWidth;
    return (int32_t)(8.0f * powf(ratio, 2.0f));
}

// radical inverse in `base`, accumulated digit by digit in binary32 exactly as the library does (the jitter is a shader input: its
// bits reach cameraJitter and from there the primary ray)
static float radical_inverse(int32_t index, int32_t base) {
    float digitWeight = 1.0f, value = 0.0f;
    while (index > 0) {
        digitWeight /= (float)base;
        value = value + digitWeight * (float)(index % base);
        index = (int32_t)(uint32_t)floorf((float)index / (float)base);
    }
    return value;
}

extern "C" int vrt_jitter_offset(int index, int phaseCount, float* outX, float* outY) {
    if (phaseCount <= 0 || index < 0 || !outX || !outY) return VRT_E_INVALID;
    const int32_t k = (index % phaseCount) + 1;
    *outX = radical_inverse(k, 2) - 0.5f;
    *outY = radical_inverse(k, 3) - 0.5f;
    return VRT_OK;
}

extern "C" int vrt_upscaler_update(int renderWidth, int targetWidth, int* frameCount, float* jitterX, float* jitterY) {
    if (!frameCount || *frameCount < 0) return VRT_E_INVALID;
    const int phases = vrt_jitter_phase_count(renderWidth, targetWidth);
    if (phases <= 0) return VRT_E_INVALID;
    const int rc = vrt_jitter_offset(*frameCount % phases, phases, jitterX, jitterY);
    if (rc) return rc;
    *frameCount += 1;
    if (*frameCount > phases) *frameCount = 0;       // '>' as in the reference: the wrap happens one frame late
    return VRT_OK;
}

extern "C" void vrt_render_resolution(int fsrEnable, uint32_t scaling, const uint32_t target[2], uint32_t out[2]) {
    for (int k = 0; k < 2; k++) out[k] = (fsrEnable && scaling) ? (uint32_t)((10.0f / (float)scaling) * (float)target[k]) : target[k];
}

// ---- image files ----
static uint32_t crc32_of(const uint8_t* p, size_t n, uint32_t crc) {
    static uint32_t table[256]; static bool ready = false;
    if (!ready) { for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; } ready = true; }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFF] ^ (crc >> 8);
    return ~crc;
}
static void be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
static void png_chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
    be32(out, (uint32_t)data.size());
    const size_t at = out.size();
    out.insert(out.end(), type, type + 4); out.insert(out.end(), data.begin(), data.end());
    be32(out, crc32_of(out.data() + at, out.size() - at, 0));
}

extern "C" int vrt_image_write(const char* path, const void* rgba8, int w, int h, int format) {
    if (!path || !rgba8 || w <= 0 || h <= 0) return VRT_E_INVALID;
    const uint8_t* px = (const uint8_t*)rgba8;
    std::vector<uint8_t> out;
    if (format == VRT_IMAGE_PPM) {
        char head[64]; const int n = snprintf(head, sizeof(head), "P6\n%d %d\n255\n", w, h);
        out.assign(head, head + n);
        out.reserve(out.size() + (size_t)w * h * 3);
        for (size_t i = 0; i < (size_t)w * h; i++) { out.push_back(px[4 * i]); out.push_back(px[4 * i + 1]); out.push_back(px[4 * i + 2]); }
    } else if (format == VRT_IMAGE_PNG_RGB || format == VRT_IMAGE_PNG_RGBA) {
        const int ch = format == VRT_IMAGE_PNG_RGBA ? 4 : 3;
        static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
        out.assign(sig, sig + 8);
        std::vector<uint8_t> ihdr; be32(ihdr, (uint32_t)w); be32(ihdr, (uint32_t)h);
        ihdr.push_back(8); ihdr.push_back(ch == 4 ? 6 : 2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
        png_chunk(out, "IHDR", ihdr);
        if (format == VRT_IMAGE_PNG_RGB) { std::vector<uint8_t> srgb(1, 0); png_chunk(out, "sRGB", srgb); }
        // scanlines (filter type 0) in a zlib stream of stored blocks
        std::vector<uint8_t> raw; raw.reserve((size_t)h * ((size_t)w * ch + 1));
        for (int y = 0; y < h; y++) { raw.push_back(0); for (int x = 0; x < w; x++) for (int c = 0; c < ch; c++) raw.push_back(px[4 * ((size_t)y * w + x) + c]); }
        std::vector<uint8_t> z; z.push_back(0x78); z.push_back(0x01);
        uint32_t a = 1, b = 0;
        size_t at = 0;
        do {
            const size_t n = raw.size() - at < 65535 ? raw.size() - at : 65535;
            z.push_back(at + n == raw.size() ? 1 : 0);
            z.push_back(n & 0xFF); z.push_back(n >> 8); z.push_back(~n & 0xFF); z.push_back((~n >> 8) & 0xFF);
            z.insert(z.end(), raw.begin() + at, raw.begin() + at + n);
            for (size_t i = at; i < at + n; i++) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
            at += n;
        } while (at < raw.size());
        be32(z, (b << 16) | a);
        png_chunk(out, "IDAT", z);
        png_chunk(out, "IEND", std::vector<uint8_t>());
    } else return VRT_E_INVALID;
    FILE* f = fopen(path, "wb");
    if (!f) return VRT_E_IO;
    const size_t n = fwrite(out.data(), 1, out.size(), f);
    const int rc = fclose(f);
    return (n == out.size() && rc == 0) ? VRT_OK : VRT_E_IO;
}
