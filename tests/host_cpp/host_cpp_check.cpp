// host_cpp_check — drives libvrt through the C++ host mirror (include/vrt_host.hpp) the way the reference's
// VoxelRenderer drives GeometryStage / DenoiserStage, and dumps the resulting planes so tests/test_gpu.py can compare
// them byte for byte with the oracle. usage: host_cpp_check <scene.vox> <W> <H> <out_prefix>
//   --link-only : construct nothing, just prove the binary links against every symbol it uses (CPU tier)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "vrt_host.hpp"

static void dump(const std::string& path, const void* p, size_t n) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f || fwrite(p, 1, n, f) != n) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(2); }
    fclose(f);
}

static void check_image_file(const std::string& path, const vrt::RenderImage& win, vrt::VoxelRenderer& renderer, int W, int H) {
    std::vector<uint8_t> host(win.bytes);
    renderer.blitStage().download(host.data());
    vrt::check(vrt_image_write(path.c_str(), host.data(), W, H, VRT_IMAGE_PNG_RGB), "vrt_image_write");
}

int main(int argc, char** argv) {
    if (argc == 2 && !strcmp(argv[1], "--link-only")) { printf("%s\n", vrt_version()); return 0; }
    if (argc == 2 && !strcmp(argv[1], "--no-gpu")) {   // error behaviour without a device: an exception, never a fallback
        try { vrt::Engine e(0); } catch (const std::runtime_error& ex) { printf("threw: %s\n", ex.what()); return 0; }
        return 3;
    }
    if (argc == 2 && !strcmp(argv[1], "--camera")) {   // CameraController input handling: host arithmetic only, no device
        vrt::CameraController cam;
        using K = vrt::CameraController;
        const struct { int kind; float delta; unsigned keys; double x, y; int button; } script[] = {
            {0, 0.016f, K::KEY_W, 0, 0, 0}, {0, 0.016f, K::KEY_W | K::KEY_D, 0, 0, 0}, {1, 0, 0, 400.0, 300.0, 1}, {1, 0, 0, 417.5, 288.25, 1},
            {0, 0.033f, K::KEY_S | K::KEY_A, 0, 0, 0}, {1, 0, 0, 10.0, 500.0, 1}, {1, 0, 0, 0, 0, 0}, {1, 0, 0, 90.0, 90.0, 1}, {1, 0, 0, 95.0, 700.0, 1},
            {0, 0.5f, K::KEY_A, 0, 0, 0}, {0, 0.25f, K::KEY_W, 0, 0, 0}};
        for (const auto& st : script) {
            if (st.kind == 0) cam.update(st.delta, st.keys); else cam.mouseCallback(st.x, st.y, st.button != 0);
            printf("%a %a %a %a %a %a %a %a %a %a %a %a %a %a\n", cam.position[0], cam.position[1], cam.position[2], cam.yaw, cam.pitch, cam.direction[0], cam.direction[1],
                   cam.direction[2], cam.right[0], cam.right[1], cam.right[2], cam.up[0], cam.up[1], cam.up[2]);
        }
        return 0;
    }
    if (argc != 5) { fprintf(stderr, "usage: %s scene.vox W H out_prefix\n", argv[0]); return 1; }
    try {
        auto engine = std::make_shared<vrt::Engine>(0);
        auto settings = std::make_shared<vrt::VoxelRenderSettings>();
        settings->voxPath = argv[1];
        settings->targetResolution[0] = (uint32_t)atoi(argv[2]); settings->targetResolution[1] = (uint32_t)atoi(argv[3]);
        const int W = atoi(argv[2]), H = atoi(argv[3]);
        // stand-in noise (counter hash) and sky (gradient) generated here so the test needs no side files
        std::vector<uint8_t> noise(512 * 512 * 4);
        for (size_t i = 0; i < noise.size(); i++) { uint32_t x = (uint32_t)i * 2654435761u; x ^= x >> 15; x *= 2246822519u; x ^= x >> 13; noise[i] = (uint8_t)(x >> 24); }
        std::vector<float> sky(64 * 32 * 4);
        for (int y = 0; y < 32; y++) for (int x = 0; x < 64; x++) { float* p = &sky[(y * 64 + x) * 4]; p[0] = x / 64.0f; p[1] = y / 32.0f; p[2] = 0.75f; p[3] = 1.0f; }
        dump(std::string(argv[4]) + "_noise.bin", noise.data(), noise.size());
        dump(std::string(argv[4]) + "_sky.bin", sky.data(), sky.size() * 4);

        auto scene = std::make_shared<vrt::VoxelScene>(engine, settings->voxPath);
        scene->setNoise(noise.data(), 512, 512);
        scene->setSkybox(sky.data(), 64, 32);
        vrt::VoxelRenderer renderer(engine, settings, scene);
        renderer.camera.position[0] = -20.5f; renderer.camera.position[1] = scene->height * 0.7f; renderer.camera.position[2] = -30.25f;
        renderer.camera.yaw = 55.0f; renderer.camera.pitch = -12.0f; renderer.camera.updateDirectionVectors();
        const vrt::RenderImage final = renderer.recordCommands(/*frameCount*/ 3);
        engine->waitIdle();

        std::vector<uint8_t> buf;
        auto& gs = renderer.geometryStage();
        for (int plane : {VRT_PLANE_COLOR_RGBA8, VRT_PLANE_DEPTH_R32F, VRT_PLANE_MASK_R8, VRT_PLANE_POSITION_RGBA32F, VRT_PLANE_NORMAL_RGBA8S}) {
            const vrt::RenderImage img = gs.image(plane);
            buf.resize(img.bytes); gs.download(img, buf.data());
            dump(std::string(argv[4]) + "_plane" + std::to_string(plane) + ".bin", buf.data(), buf.size());
        }
        buf.resize(final.bytes); renderer.denoiserStage().download(buf.data());
        dump(std::string(argv[4]) + "_denoised.bin", buf.data(), buf.size());
        // the whole frame as the reference records it: upscaler-stage jitter + frame counter, FSR render resolution, blit into a window
        settings->fsrSetttings.enable = true; settings->fsrSetttings.scaling = vrt::FsrScaling::QUALITY;
        renderer.update(0.016f); renderer.update(0.016f);
        const uint32_t window[2] = {(uint32_t)W, (uint32_t)H};
        const vrt::RenderImage& win = renderer.recordCommands(window);
        engine->waitIdle();
        buf.resize(win.bytes); renderer.blitStage().download(buf.data());
        dump(std::string(argv[4]) + "_window.bin", buf.data(), buf.size());
        const uint32_t* rr = settings->renderResolution();
        const vrt_camera push2 = renderer.camera.push(*scene, rr, (uint32_t)renderer.upscalerStage().frameCount, renderer.upscalerStage().jitterX, renderer.upscalerStage().jitterY);
        dump(std::string(argv[4]) + "_camera2.bin", &push2, sizeof(push2));
        buf.resize((size_t)rr[0] * rr[1] * 4); renderer.denoiserStage().download(buf.data());   // the image the blit sampled
        dump(std::string(argv[4]) + "_final2.bin", buf.data(), buf.size());
        for (int plane : {VRT_PLANE_COLOR_RGBA8, VRT_PLANE_POSITION_RGBA32F, VRT_PLANE_NORMAL_RGBA8S}) {
            const vrt::RenderImage img = gs.image(plane);
            buf.resize(img.bytes); gs.download(img, buf.data());
            dump(std::string(argv[4]) + "_fsr_plane" + std::to_string(plane) + ".bin", buf.data(), buf.size());
        }
        check_image_file(std::string(argv[4]) + "_window.png", win, renderer, W, H);
        settings->fsrSetttings.enable = false;
        const vrt_frame_stats st = gs.stats();
        const vrt_camera push = renderer.camera.push(*scene, settings->renderResolution(), 3);
        dump(std::string(argv[4]) + "_camera.bin", &push, sizeof(push));
        printf("ok %dx%d bounds %ux%ux%u rays %llu launches %u\n", W, H, scene->width, scene->height, scene->depth, (unsigned long long)st.rays, st.kernel_launches);
    } catch (const std::exception& ex) {
        fprintf(stderr, "error: %s\n", ex.what());
        return 1;
    }
    return 0;
}
