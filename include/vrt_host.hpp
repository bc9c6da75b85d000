// vrt_host.hpp — C++ host side above the C ABI (include/vrt.h), header-only. It mirrors the reference's own operator
// surface for this path so that a caller of the reference's classes finds the same names, argument meaning and error
// behaviour (exceptions: the reference throws std::runtime_error from VoxelScene and lets vk::resultCheck throw;
// /root/reference/source/voxels/resource/voxel_scene.cpp:42,46,50):
//
//   vrt::VoxelRenderSettings  == VoxelRenderSettings           source/voxels/voxel_render_settings.hpp:15-59
//   vrt::VoxelScene           == VoxelScene                    source/voxels/resource/voxel_scene.hpp:18-34
//   vrt::CameraController     == CameraController              source/voxels/resource/camera_controller.cpp:15-28
//   vrt::GeometryStage        == GeometryStage                 source/voxels/stages/geometry_stage.hpp:30-56
//   vrt::GeometryBuffer       == GeometryBuffer                source/voxels/stages/geometry_stage.hpp:19-27
//   vrt::DenoiserStage        == DenoiserStage                 source/voxels/stages/denoiser_stage.hpp:21-47
//   vrt::UpscalerStage        == UpscalerStage::update         source/voxels/stages/upscaler_stage.cpp:59-70 (jitter + frame counter)
//   vrt::BlitStage            == BlitStage::record             source/voxels/stages/blit_stage.cpp:25-57 (+ shader/blit.frag)
//   vrt::VoxelRenderer        == VoxelRenderer::update / recordCommands  source/voxels/voxel_renderer.cpp:33-105
//
// What differs, deliberately: there is no vk::CommandBuffer / flightFrame (the context's CUDA stream plays the role of
// the command buffer; record() is asynchronous exactly like the reference's, results are valid after sync()), render
// targets are device pointers instead of RenderImage, the "window" of the blit is a device buffer, and FSR2's upscaling pass is out
// of scope (closed source) — with fsrSetttings.enable the pass renders at the reduced resolution with FSR2's jitter
// sequence and the blit's LINEAR sampler does the enlarging.
#pragma once
#include <cmath>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "vrt.h"

namespace vrt {

inline void check(int rc, const char* what) {
    if (rc != VRT_OK) throw std::runtime_error(std::string(what) + ": " + vrt_last_error() + " (vrt_status " + std::to_string(rc) + ")");
}

// The reference's Engine owns the device and queue; here: one CUDA device + stream (+ the tile shard of this process).
class Engine {
public:
    explicit Engine(int device = 0, int rank = 0, int world = 1) {
        check(vrt_context_create(device, &_ctx), "vrt_context_create");   // throws when there is no GPU: no CPU fallback
        if (world > 1) check(vrt_context_set_shard(_ctx, rank, world), "vrt_context_set_shard");
    }
    ~Engine() { vrt_context_destroy(_ctx); }
    Engine(const Engine&) = delete;
    Engine& operator=(const Engine&) = delete;
    vrt_context* handle() const { return _ctx; }
    void waitIdle() const { check(vrt_context_sync(_ctx), "vrt_context_sync"); }   // == device.waitIdle()
    void* stream() const { return vrt_context_stream(_ctx); }
    // MAX_FRAMES_IN_FLIGHT of the reference (source/engine/engine.hpp:19): GeometryStage objects created after this call record
    // into frames that run on streams of their own, so that consecutive frames overlap on the GPU; joinFrames() makes the
    // engine's stream wait for all of them (a queue-level fence), waitIdle() waits on the host.
    void setFramesInFlight(bool concurrent) const { check(vrt_context_set_concurrent_frames(_ctx, concurrent ? 1 : 0), "vrt_context_set_concurrent_frames"); }
    void joinFrames() const { check(vrt_context_join(_ctx), "vrt_context_join"); }
private:
    vrt_context* _ctx = nullptr;
};

// voxel_render_settings.hpp:6-59. FsrSettings::enable defaults to false here (true in the reference): see the note above.
enum class FsrScaling : uint32_t { NONE = 10, QUALITY = 15, BALANCED = 17, PERFORMANCE = 20, ULTRA_PERFORMANCE = 30 };
struct FsrSettings { bool enable = false; FsrScaling scaling = FsrScaling::BALANCED; };
struct DenoiserSettings { bool enable = true; int iterations = 2; float phiColor0 = 20.4f, phiNormal0 = 1E-2f, phiPos0 = 1E-1f, stepWidth = 2.0f; };
struct AmbientOcclusionSettings { int numSamples = 4; float intensity = 1.0f; };
struct LightSettings {
    float direction[3] = {0.57735026f, 0.57735026f, 0.57735026f};   // overwritten by the constructor below (glm::normalize(vec3(1)))
    float color[4] = {1.0f, 1.0f, 1.0f, 1.0f};
    float intensity = 1.0f;
};
class VoxelRenderSettings {
public:
    uint32_t targetResolution[2] = {1920, 1080};
    FsrSettings fsrSetttings = {};                       // (sic) the reference's spelling, voxel_render_settings.hpp:49
    DenoiserSettings denoiserSettings = {};
    AmbientOcclusionSettings occlusionSettings = {};
    LightSettings lightSettings = {};
    std::string voxPath = "../resource/treehouse.vox";
    std::string skyboxPath = "../resource/rustig_koppie.hdr";
    // additions of this build (the reference behaves as VRT_PASS_ALL / G-buffer on / AS_EXECUTED)
    uint32_t passMask = VRT_PASS_ALL, outMask = VRT_OUT_GBUFFER, traversal = VRT_TRAVERSAL_DEFAULT, pipeline = VRT_PIPELINE_DEFAULT;
    int denoiseMode = VRT_DENOISE_AS_EXECUTED;

    VoxelRenderSettings() {
        vrt_render_settings d; vrt_default_settings(&d);
        std::memcpy(lightSettings.direction, d.light.direction, sizeof(lightSettings.direction));
    }
    // voxel_render_settings.cpp:8-13; the pointer stays valid until the next call
    const uint32_t* renderResolution() const {
        vrt_render_resolution(fsrSetttings.enable ? 1 : 0, (uint32_t)fsrSetttings.scaling, targetResolution, _render);
        return _render;
    }
    vrt_render_settings lower(bool withDenoiser) const {
        vrt_render_settings s; vrt_default_settings(&s);
        s.pass_mask = passMask; s.out_mask = outMask; s.traversal = traversal; s.pipeline = pipeline;
        s.vol.aoSamples = (uint32_t)occlusionSettings.numSamples; s.vol.ambientIntensity = occlusionSettings.intensity;   // geometry_stage.cpp:135-137
        std::memcpy(s.light.direction, lightSettings.direction, sizeof(s.light.direction));                               // geometry_stage.cpp:138-142
        s.light.intensity = lightSettings.intensity; std::memcpy(s.light.color, lightSettings.color, sizeof(s.light.color));
        s.denoise.enable = withDenoiser && denoiserSettings.enable; s.denoise.iterations = denoiserSettings.iterations;
        s.denoise.phiColor0 = denoiserSettings.phiColor0; s.denoise.phiNormal0 = denoiserSettings.phiNormal0;
        s.denoise.phiPos0 = denoiserSettings.phiPos0; s.denoise.stepWidth = denoiserSettings.stepWidth; s.denoise.mode = denoiseMode;
        return s;
    }
private:
    mutable uint32_t _render[2] = {0, 0};
};

// A complete voxel scene: dense volume + palette (+ noise + skybox) resident on the GPU. voxel_scene.hpp:18-34.
class VoxelScene {
public:
    uint32_t width = 0, height = 0, depth = 0;
    // VoxelScene(engine, filename, skyboxFilename), voxel_scene.cpp:33: the .vox file and the Radiance .hdr skybox (stbi_loadf in the
    // reference, voxel_scene.cpp:132 -> texture_2d.cpp:31-38). A git-LFS pointer stub in either place throws (VRT_E_LFS_POINTER).
    VoxelScene(const std::shared_ptr<Engine>& engine, const std::string& filename, const std::string& skyboxFilename) : _engine(engine) {
        check(vrt_scene_load_vox(engine->handle(), filename.c_str(), &_scene), "VoxelScene");   // messages of voxel_scene.cpp:42,46,50
        bounds();
        const int rc = vrt_scene_load_skybox(_scene, skyboxFilename.c_str());
        if (rc != VRT_OK) { destroy(); check(rc, "VoxelScene skybox"); }
    }
    // without a skybox file: pixels are passed with setSkybox (stand-ins, tests)
    VoxelScene(const std::shared_ptr<Engine>& engine, const std::string& filename) : _engine(engine) {
        check(vrt_scene_load_vox(engine->handle(), filename.c_str(), &_scene), "VoxelScene");
        bounds();
    }
    VoxelScene(const std::shared_ptr<Engine>& engine, const uint8_t* dense, uint32_t w, uint32_t h, uint32_t d, const vrt_material palette[256])
        : _engine(engine) {
        check(vrt_scene_from_dense(engine->handle(), dense, w, h, d, palette, &_scene), "VoxelScene(dense)");
        bounds();
    }
    VoxelScene(const std::shared_ptr<Engine>& engine, int generator, uint64_t seed, uint32_t w, uint32_t h, uint32_t d) : _engine(engine) {
        check(vrt_scene_generate(engine->handle(), generator, seed, w, h, d, &_scene), "VoxelScene(generate)");
        bounds();
    }
    ~VoxelScene() { destroy(); }
    VoxelScene(const VoxelScene&) = delete;
    VoxelScene& operator=(const VoxelScene&) = delete;
    void destroy() { if (_scene) { vrt_scene_destroy(_scene); _scene = nullptr; } }   // == AResource::destroy()
    void setNoise(const uint8_t* rgba8, uint32_t w, uint32_t h) { check(vrt_scene_set_noise(_scene, rgba8, w, h), "setNoise"); }
    void loadNoise(const std::string& pngPath) { check(vrt_scene_load_noise(_scene, pngPath.c_str()), "noise texture"); }   // voxel_renderer.cpp:22
    void setSkybox(const float* rgba32f, uint32_t w, uint32_t h) { check(vrt_scene_set_skybox(_scene, rgba32f, w, h), "setSkybox"); }
    vrt_scene* handle() const { return _scene; }
    const std::shared_ptr<Engine>& engine() const { return _engine; }
private:
    void bounds() { uint32_t b[3]; check(vrt_scene_bounds(_scene, b), "vrt_scene_bounds"); width = b[0]; height = b[1]; depth = b[2]; }
    std::shared_ptr<Engine> _engine;
    vrt_scene* _scene = nullptr;
};

// camera_controller.cpp:15-71 + the push-constant fill of voxel_renderer.cpp:72-83. The reference polls GLFW; headless, the caller passes
// the key state and cursor events in (same arithmetic, same order of operations).
class CameraController {
public:
    enum Key : unsigned { KEY_W = 1, KEY_S = 2, KEY_A = 4, KEY_D = 8 };
    float position[3] = {8.0f, 8.0f, -50.0f};
    float yaw = 90.0f, pitch = 0.0f, focalLength = 1.920982f;    // voxel_renderer.cpp:20 (fov 55 degrees)
    float direction[3], right[3], up[3], normalDir[3];
    CameraController() { updateDirectionVectors(); }
    void updateDirectionVectors() {
        float r2[3], u2[3];
        vrt_camera_vectors(yaw, pitch, 1.0f, normalDir, r2, u2);          // normalDir * 1.0f == normalDir
        vrt_camera_vectors(yaw, pitch, focalLength, direction, right, up);
    }
    // CameraController::update, camera_controller.cpp:30-44: cameraSpeed 50; W/S along normalDir, A/D along right
    void update(float delta, unsigned keys) {
        const float cameraSpeed = 50.0f;
        for (int k = 0; k < 3; k++) {
            if (keys & KEY_W) position[k] += (cameraSpeed * normalDir[k]) * delta;
            if (keys & KEY_S) position[k] -= (cameraSpeed * normalDir[k]) * delta;
            if (keys & KEY_A) position[k] -= (right[k] * cameraSpeed) * delta;
            if (keys & KEY_D) position[k] += (right[k] * cameraSpeed) * delta;
        }
        updateDirectionVectors();
    }
    // CameraController::mouseCallback, camera_controller.cpp:46-71: drag with the right button held; one degree per pixel
    void mouseCallback(double cursorX, double cursorY, bool rightButton) {
        if (rightButton) {
            if (_lastCursorX > 0 && _lastCursorY > 0) {
                const float offsetX = static_cast<float>(cursorX - _lastCursorX), offsetY = static_cast<float>(cursorY - _lastCursorY);
                yaw -= offsetX;
                const float p = pitch - offsetY;
                pitch = p < -90.0f ? -90.0f : (p > 90.0f ? 90.0f : p);
                updateDirectionVectors();
            }
            _lastCursorX = cursorX; _lastCursorY = cursorY;
        } else { _lastCursorX = -1; _lastCursorY = -1; }
    }
    vrt_camera push(const VoxelScene& scene, const uint32_t res[2], uint32_t frame, float jx = 0.0f, float jy = 0.0f) const {
        vrt_camera c; std::memset(&c, 0, sizeof(c));
        for (int i = 0; i < 3; i++) { c.camPos[i] = position[i]; c.camDir[i] = direction[i]; c.camRight[i] = right[i]; c.camUp[i] = up[i]; }
        c.camPos[3] = 1.0f;
        c.volumeBounds[0] = scene.width; c.volumeBounds[1] = scene.height; c.volumeBounds[2] = scene.depth;
        c.frame = frame; c.screenSize[0] = (int32_t)res[0]; c.screenSize[1] = (int32_t)res[1];
        c.cameraJitter[0] = jx; c.cameraJitter[1] = jy;
        return c;
    }
private:
    double _lastCursorX = -1, _lastCursorY = -1;
};

// One render target on the device (RenderImage stand-in).
struct RenderImage { void* devicePtr = nullptr; size_t bytes = 0; int plane = -1; };

struct GeometryBuffer { RenderImage color, depth, motion, mask, normal, position; };   // geometry_stage.hpp:19-27

class GeometryStage {
public:
    GeometryStage(const std::shared_ptr<Engine>& engine, const std::shared_ptr<VoxelRenderSettings>& settings, const std::shared_ptr<VoxelScene>& scene)
        : _engine(engine), _settings(settings), _scene(scene) {}
    ~GeometryStage() { vrt_frame_release(_frame); }
    GeometryStage(const GeometryStage&) = delete;
    GeometryStage& operator=(const GeometryStage&) = delete;
    // Records the pass on the engine's stream with the LIVE settings (geometry_stage.cpp:135-141) and the pushed camera block
    // (the reference's caller pushes ScreenQuadPush before record: voxel_renderer.cpp:72-83). Asynchronous.
    GeometryBuffer record(const vrt_camera& push) {
        const vrt_render_settings s = _settings->lower(false);
        if (_frame && (push.screenSize[0] != _w || push.screenSize[1] != _h)) { vrt_frame_release(_frame); _frame = nullptr; }   // RENDER_RESIZE
        _w = push.screenSize[0]; _h = push.screenSize[1];
        check(vrt_render(_engine->handle(), _scene->handle(), &push, &s, &_frame), "GeometryStage::record");
        GeometryBuffer g;
        g.color = image(VRT_PLANE_COLOR_RGBA8);
        if (s.out_mask & VRT_OUT_GBUFFER) {
            g.depth = image(VRT_PLANE_DEPTH_R32F); g.motion = image(VRT_PLANE_MOTION_RG32F); g.mask = image(VRT_PLANE_MASK_R8);
            g.normal = image(VRT_PLANE_NORMAL_RGBA8S); g.position = image(VRT_PLANE_POSITION_RGBA32F);
        }
        return g;
    }
    RenderImage image(int plane) const {
        RenderImage r; r.plane = plane;
        check(vrt_frame_plane(_frame, plane, &r.devicePtr, &r.bytes), "vrt_frame_plane");
        return r;
    }
    void download(const RenderImage& img, void* host) const { check(vrt_frame_download(_frame, img.plane, host, img.bytes), "vrt_frame_download"); }
    vrt_frame_stats stats() const { vrt_frame_stats s; check(vrt_frame_stats_get(_frame, &s), "vrt_frame_stats_get"); return s; }
    vrt_frame* frame() const { return _frame; }
private:
    std::shared_ptr<Engine> _engine;
    std::shared_ptr<VoxelRenderSettings> _settings;
    std::shared_ptr<VoxelScene> _scene;
    vrt_frame* _frame = nullptr;
    int32_t _w = 0, _h = 0;
};

class DenoiserStage {
public:
    DenoiserStage(const std::shared_ptr<Engine>& engine, const std::shared_ptr<VoxelRenderSettings>& settings) : _engine(engine), _settings(settings) {}
    ~DenoiserStage() { vrt_buffer_free(_engine->handle(), _out.devicePtr); }
    DenoiserStage(const DenoiserStage&) = delete;
    DenoiserStage& operator=(const DenoiserStage&) = delete;
    // denoiser_stage.hpp:45-46. The output target is owned by the stage (re-created when the resolution changes, like the
    // reference's RENDER_RESIZE handling); ping-pong scratch lives in the library. Runs with the LIVE settings.
    const RenderImage& record(const RenderImage& colorInput, const RenderImage& normalInput, const RenderImage& posInput, int w, int h) {
        const size_t bytes = (size_t)w * (size_t)h * 4u;
        if (_out.bytes != bytes) {
            vrt_buffer_free(_engine->handle(), _out.devicePtr); _out = RenderImage();
            check(vrt_buffer_alloc(_engine->handle(), bytes, &_out.devicePtr), "DenoiserStage target");
            _out.bytes = bytes; _out.plane = VRT_PLANE_DENOISED_RGBA8;
        }
        const vrt_render_settings s = _settings->lower(true);
        check(vrt_denoise(_engine->handle(), colorInput.devicePtr, normalInput.devicePtr, posInput.devicePtr, w, h, &s.denoise, _out.devicePtr),
              "DenoiserStage::record");
        return _out;
    }
    void download(void* host) const { check(vrt_buffer_download(_engine->handle(), _out.devicePtr, host, _out.bytes), "vrt_buffer_download"); }
private:
    std::shared_ptr<Engine> _engine;
    std::shared_ptr<VoxelRenderSettings> _settings;
    RenderImage _out;
};

// The frame-driver half of UpscalerStage (upscaler_stage.cpp:59-70): update() once per frame computes this frame's jitter from
// frameCount and advances it; VoxelRenderer pushes both (voxel_renderer.cpp:80-82). record() — FSR2 itself — does not exist here.
class UpscalerStage {
public:
    float jitterX = 0.0f, jitterY = 0.0f;
    int frameCount = 0;
    explicit UpscalerStage(const std::shared_ptr<VoxelRenderSettings>& settings) : _settings(settings) {}
    // UpscalerStage(engine, settings), upscaler_stage.cpp:18: with an engine the stage can also record() its pass
    UpscalerStage(const std::shared_ptr<Engine>& engine, const std::shared_ptr<VoxelRenderSettings>& settings) : _settings(settings), _engine(engine) {}
    ~UpscalerStage() { if (_up) vrt_upscaler_destroy(_up); if (_engine && _target.devicePtr) vrt_buffer_free(_engine->handle(), _target.devicePtr); }
    UpscalerStage(const UpscalerStage&) = delete;
    UpscalerStage& operator=(const UpscalerStage&) = delete;
    void update(float /*delta*/) {
        check(vrt_upscaler_update((int)_settings->renderResolution()[0], (int)_settings->targetResolution[0], &frameCount, &jitterX, &jitterY),
              "UpscalerStage::update");
    }
    // UpscalerStage::record (upscaler_stage.cpp:78-172): colour, depth, motion and the reactive mask of the render resolution in, the
    // target-resolution image out. The reference dispatches FSR2 here (closed source); this is libvrt's own temporal upscaler.
    const RenderImage& record(const RenderImage& color, const RenderImage& depth, const RenderImage& motion, const RenderImage& mask) {
        if (!_engine) throw std::runtime_error("UpscalerStage::record needs the engine constructor");
        const int rw = (int)_settings->renderResolution()[0], rh = (int)_settings->renderResolution()[1];
        const int tw = (int)_settings->targetResolution[0], th = (int)_settings->targetResolution[1];
        if (!_up || _rw != rw || _rh != rh || _tw != tw || _th != th) {   // RENDER_RESIZE / TARGET_RESIZE
            if (_up) { vrt_upscaler_destroy(_up); _up = nullptr; }
            if (_target.devicePtr) { vrt_buffer_free(_engine->handle(), _target.devicePtr); _target.devicePtr = nullptr; }
            check(vrt_upscaler_create(_engine->handle(), rw, rh, tw, th, &_up), "UpscalerStage");
            _target.bytes = (size_t)tw * th * 4; _tw = tw; _th = th;
            check(vrt_buffer_alloc(_engine->handle(), _target.bytes, &_target.devicePtr), "UpscalerStage target");
            _rw = rw; _rh = rh; _first = true;
        }
        check(vrt_upscaler_dispatch(_up, color.devicePtr, depth.devicePtr, motion.devicePtr, mask.devicePtr, jitterX, jitterY, 1.0f, 1.0f, _first ? 1 : 0,
                                    _target.devicePtr, nullptr), "UpscalerStage::record");
        _first = false;
        return _target;
    }
private:
    std::shared_ptr<VoxelRenderSettings> _settings;
    std::shared_ptr<Engine> _engine;
    vrt_upscaler* _up = nullptr;
    RenderImage _target{};
    int _rw = 0, _rh = 0, _tw = 0, _th = 0;
    bool _first = true;
};

// BlitStage::record (blit_stage.cpp:25-57): the final image, sampled LINEAR / CLAMP_TO_EDGE, scaled to fill the window and centre-cropped
// (blit.frag:14-22), stored sRGB-encoded like the reference's swap-chain format. The window is a device buffer owned by the stage.
class BlitStage {
public:
    uint32_t flags = VRT_BLIT_SRGB;
    explicit BlitStage(const std::shared_ptr<Engine>& engine) : _engine(engine) {}
    ~BlitStage() { vrt_buffer_free(_engine->handle(), _window.devicePtr); }
    BlitStage(const BlitStage&) = delete;
    BlitStage& operator=(const BlitStage&) = delete;
    const RenderImage& record(const RenderImage& image, const uint32_t sourceSize[2], const uint32_t windowSize[2]) {
        const size_t bytes = (size_t)windowSize[0] * windowSize[1] * 4u;
        if (_window.bytes != bytes) {
            vrt_buffer_free(_engine->handle(), _window.devicePtr); _window = RenderImage();
            check(vrt_buffer_alloc(_engine->handle(), bytes, &_window.devicePtr), "BlitStage window");
            _window.bytes = bytes;
        }
        check(vrt_blit(_engine->handle(), image.devicePtr, (int)sourceSize[0], (int)sourceSize[1], _window.devicePtr, (int)windowSize[0], (int)windowSize[1], flags),
              "BlitStage::record");
        return _window;
    }
    void download(void* host) const { check(vrt_buffer_download(_engine->handle(), _window.devicePtr, host, _window.bytes), "vrt_buffer_download"); }
private:
    std::shared_ptr<Engine> _engine;
    RenderImage _window;
};

// VoxelRenderer::update / recordCommands (voxel_renderer.cpp:33-105) without FSR2's pass: push constants -> geometry stage -> denoiser
// stage (when enabled, voxel_renderer.cpp:87-92) -> blit into the window.
class VoxelRenderer {
public:
    CameraController camera;
    VoxelRenderer(const std::shared_ptr<Engine>& engine, const std::shared_ptr<VoxelRenderSettings>& settings, const std::shared_ptr<VoxelScene>& scene)
        : _engine(engine), _settings(settings), _scene(scene), _geometry(engine, settings, scene), _denoiser(engine, settings), _upscaler(settings),
          _blit(engine) {}
    // voxel_renderer.cpp:33-39: per-frame host update (the camera controller's input handling has no headless counterpart)
    void update(float delta, unsigned keys = 0) { camera.update(delta, keys); _upscaler.update(delta); }
    // one whole frame as the reference records it: frame counter and jitter from the upscaler stage, final image blitted into a window
    // of windowSize (engine->windowSize in the reference). Returns the window image (RGBA8, sRGB-encoded).
    const RenderImage& recordCommands(const uint32_t windowSize[2]) {
        const RenderImage img = recordCommands((uint32_t)_upscaler.frameCount, _upscaler.jitterX, _upscaler.jitterY);
        uint32_t src[2] = {_settings->renderResolution()[0], _settings->renderResolution()[1]};
        return _blit.record(img, src, windowSize);
    }
    // one frame; returns the final colour image (denoised when the denoiser is enabled)
    RenderImage recordCommands(uint32_t frameCount, float jitterX = 0.0f, float jitterY = 0.0f) {
        const uint32_t* res = _settings->renderResolution();
        const vrt_camera push = camera.push(*_scene, res, frameCount, jitterX, jitterY);
        const GeometryBuffer g = _geometry.record(push);
        if (_settings->denoiserSettings.enable && (_settings->outMask & VRT_OUT_GBUFFER))
            return _denoiser.record(g.color, g.normal, g.position, (int)res[0], (int)res[1]);
        return g.color;
    }
    GeometryStage& geometryStage() { return _geometry; }
    DenoiserStage& denoiserStage() { return _denoiser; }
    UpscalerStage& upscalerStage() { return _upscaler; }
    BlitStage& blitStage() { return _blit; }
private:
    std::shared_ptr<Engine> _engine;
    std::shared_ptr<VoxelRenderSettings> _settings;
    std::shared_ptr<VoxelScene> _scene;
    GeometryStage _geometry;
    DenoiserStage _denoiser;
    UpscalerStage _upscaler;
    BlitStage _blit;
};

}  // namespace vrt
