// oracle_asan — SURVEY.md 8(c) self-check (v): the CPU restatement runs clean under AddressSanitizer + UndefinedBehaviorSanitizer.
// Built by tests/test_sanitize.py from the oracle's own sources with -fsanitize=address,undefined -fno-sanitize-recover=all, so any
// report ends the process with a non-zero code. Exercises every exported entry: generators, the full pass with every pass subset on
// ragged sizes, both denoiser modes with integral and fractional step widths, the loader on the golden files and on truncated copies of
// them, blit, camera, jitter sequence.
//   usage: oracle_asan <golden.vox> [more.vox ...]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../oracle/vrt_oracle.h"

static uint32_t g_seed = 12345u;
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

int main(int argc, char** argv) {
    // ---- a small generated scene, every pass subset, ragged frame sizes
    const uint32_t W = 37, H = 29, D = 41;
    std::vector<uint8_t> vol((size_t)W * H * D);
    for (int gen = 1; gen <= 2; gen++) if (orc_generate(gen, 42u + gen, vol.data(), W, H, D) != 0) { fprintf(stderr, "orc_generate failed\n"); return 2; }
    for (size_t i = 0; i < vol.size(); i++) if ((rnd() & 63u) == 0) vol[i] = (uint8_t)(1 + rnd() % 255);   // some clutter for short rays
    std::vector<float> palette(256 * 8, 0.0f);
    for (int m = 0; m < 256; m++) { for (int k = 0; k < 3; k++) palette[m * 8 + k] = (float)((m * 37 + k * 91) % 256) / 255.0f; palette[m * 8 + 3] = 1.0f; palette[m * 8 + 4] = (m % 3 == 0) ? 0.5f : 0.0f; }
    std::vector<uint8_t> noise(512 * 512 * 4);
    for (auto& b : noise) b = (uint8_t)rnd();
    const uint32_t skyW = 16, skyH = 8;
    std::vector<float> sky(skyW * skyH * 4);
    for (auto& v : sky) v = (float)(rnd() % 1000) / 500.0f;
    orc_scene sc{vol.data(), W, H, D, palette.data(), noise.data(), sky.data(), skyW, skyH};
    orc_settings st{};
    st.aoSamples = 4; st.ambientIntensity = 0.1f; st.lightDir[0] = st.lightDir[1] = st.lightDir[2] = 0.57735026f; st.lightIntensity = 1.0f;
    st.lightColor[0] = st.lightColor[1] = st.lightColor[2] = st.lightColor[3] = 1.0f;
    const int sizes[][2] = {{33, 17}, {1, 1}, {64, 5}};
    const uint32_t masks[] = {ORC_PASS_ALL, ORC_PASS_PRIMARY, ORC_PASS_PRIMARY | ORC_PASS_SHADE | ORC_PASS_REFLECT, ORC_PASS_PRIMARY | ORC_PASS_SHADE | ORC_PASS_AO | ORC_PASS_SHADOW};
    for (const auto& sz : sizes) for (uint32_t mask : masks) for (int pose = 0; pose < 3; pose++) {
        const int w = sz[0], h = sz[1];
        orc_camera cam{};
        float d[3], r[3], u[3];
        orc_camera_vectors(pose == 0 ? 37.0f : pose == 1 ? -90.0f : 200.5f, pose == 2 ? 89.0f : -25.0f, 1.920982f, d, r, u);
        const float pos[3][3] = {{-20.3f, 40.2f, -11.1f}, {18.5f, 14.5f, 20.5f} /* inside the volume */, {18.0f, -50.0f, 20.0f}};
        for (int k = 0; k < 3; k++) { cam.camPos[k] = pos[pose][k]; cam.camDir[k] = d[k]; cam.camRight[k] = r[k]; cam.camUp[k] = u[k]; }
        cam.camPos[3] = 1.0f; cam.volumeBounds[0] = W; cam.volumeBounds[1] = H; cam.volumeBounds[2] = D; cam.frame = 5u + (uint32_t)pose;
        cam.screenSize[0] = w; cam.screenSize[1] = h; cam.cameraJitter[0] = 0.25f * (float)pose; cam.cameraJitter[1] = -0.125f;
        const size_t px = (size_t)w * h;
        std::vector<uint8_t> color(px * 4), maskp(px), hp(px), dn(px * 4);
        std::vector<float> depth(px), motion(px * 2), position(px * 4), cf(px * 3);
        std::vector<int8_t> normal(px * 4);
        std::vector<int16_t> hit(px * 4);
        std::vector<uint32_t> rays(px), fetches(px);
        orc_planes pl{color.data(), depth.data(), motion.data(), maskp.data(), position.data(), normal.data(), hit.data(), hp.data(), cf.data(), rays.data(), fetches.data()};
        st.pass_mask = mask;
        if (orc_render(&sc, &cam, &st, &pl, 0, h, 1, 2) != 0) { fprintf(stderr, "orc_render failed\n"); return 3; }
        orc_planes sparse{};                                       // NULL planes are allowed
        sparse.color = color.data();
        if (orc_render(&sc, &cam, &st, &sparse, 0, h, 2, 1) != 0) { fprintf(stderr, "orc_render (colour only, every 2nd row) failed\n"); return 3; }
        const float sws[] = {2.0f, 1.5f, 0.0f, 4.75f};
        for (int mode = 0; mode < 2; mode++) for (float sw : sws) for (int it = 1; it <= 3; it++) {
            orc_denoise_settings ds{it, 20.4f, 1e-2f, 1e-1f, sw, mode};
            if (orc_denoise(color.data(), normal.data(), position.data(), w, h, &ds, dn.data(), 2) != 0) { fprintf(stderr, "orc_denoise failed\n"); return 4; }
        }
        std::vector<uint8_t> win((size_t)50 * 23 * 4);
        for (int flags = 0; flags < 4; flags++) if (orc_blit(color.data(), w, h, win.data(), 50, 23, flags, 2) != 0) { fprintf(stderr, "orc_blit failed\n"); return 5; }
    }
    // single rays, including degenerate directions
    const float dirs[][3] = {{1, 0, 0}, {0, -1, 0}, {0.6f, 0.8f, 0}, {0.57735026f, 0.57735026f, 0.57735026f}, {0, 0, 0}, {NAN, 1, 0}, {1e-30f, 1, 1e-30f}};
    for (const auto& dd : dirs) {
        const float start[3] = {3.5f, 40.5f, 7.25f};
        int32_t iout[6]; float fout[6];
        orc_trace_one(&sc, start, dd, 512, iout, fout);
    }
    // frame driver arithmetic
    for (int rw = 1; rw <= 3840; rw = rw * 3 + 1) {
        const int pc = orc_jitter_phase_count(rw, 3840);
        float x, y; orc_jitter_offset(&x, &y, rw % (pc > 0 ? pc : 1), pc);
        int fc = 0; for (int k = 0; k < 5; k++) orc_upscaler_update(rw, 3840, &fc, &x, &y);
        const uint32_t tgt[2] = {3840, 2160}; uint32_t rr[2];
        for (uint32_t scal = 0; scal < 20; scal += 5) orc_render_resolution(1, scal, tgt, rr);
    }
    // ---- the loader: the golden files and truncated copies of them
    int loaded = 0, rejected = 0;
    for (int a = 1; a < argc; a++) {
        const std::vector<uint8_t> good = slurp(argv[a]);
        if (good.empty()) { fprintf(stderr, "cannot read %s\n", argv[a]); return 6; }
        for (int k = 0; k < 400; k++) {
            std::vector<uint8_t> b = good;
            if (k > 0) b.resize(rnd() % b.size());            // (the oracle's loader trusts size fields like the reference's parser does: damaged
                                                              // size fields are the PRODUCT parser's business, parser_asan.cpp)
            uint8_t* v = nullptr; uint32_t whd[3] = {0, 0, 0}; std::vector<float> pal(256 * 8);
            const int rc = orc_load_vox(b.data(), b.size(), &v, whd, pal.data());
            if (rc == 0) { loaded++; volatile uint8_t touch = v[(size_t)whd[0] * whd[1] * whd[2] - 1]; (void)touch; orc_free(v); } else rejected++;
        }
    }
    printf("oracle sanitize ok: %d loads accepted, %d rejected\n", loaded, rejected);
    return 0;
}
