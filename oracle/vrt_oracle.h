/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h header). C interface of the CPU restatement of
 * /root/reference/shader/voxel_volume.frag + shader/denoiser.frag and the host code that feeds them.
 * PARITY UNPINNED: the reference holds no golden vectors / tests for this path.
 * Plain C types only so tests can bind it with ctypes. */
#ifndef VRT_ORACLE_H
#define VRT_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* == ScreenQuadPush, source/voxels/resource/screen_quad_push.hpp:5-15 (96 bytes) */
typedef struct {
    float camPos[4], camDir[4], camRight[4], camUp[4];
    uint32_t volumeBounds[3];
    uint32_t frame;
    int32_t screenSize[2];
    float cameraJitter[2];
} orc_camera;

/* pass bits: reference behaviour == all bits set */
enum { ORC_PASS_PRIMARY = 1, ORC_PASS_SHADOW = 2, ORC_PASS_AO = 4, ORC_PASS_REFLECT = 8, ORC_PASS_SHADE = 16,
       ORC_PASS_ALL = 31 };
enum { ORC_DENOISE_AS_EXECUTED = 0, ORC_DENOISE_AS_INTENDED = 1 };

typedef struct {
    uint32_t pass_mask;
    uint32_t aoSamples;        /* VolumeParameters, parameters.hpp:5-9 */
    float ambientIntensity;
    float lightDir[3];         /* Light, voxel_scene.hpp:10-15 (normalised by the host) */
    float lightIntensity;
    float lightColor[4];
} orc_settings;

typedef struct {
    const uint8_t* volume;     /* dense u8[W*H*D], x fastest, then y, then z (voxel_scene.cpp:99) */
    uint32_t w, h, d;
    const float* palette;      /* 256 x {diffuse rgba, metallic, pad x3} = 256*8 floats (material.hpp:5-12) */
    const uint8_t* noise;      /* 512x512 RGBA8 blue noise, row-major */
    const float* sky;          /* skyW x skyH RGBA32F equirect, row-major (row 0 = v 0) */
    uint32_t skyW, skyH;
} orc_scene;

/* Output planes; any pointer may be NULL. All are full-frame W*H row-major (row 0 = top). */
typedef struct {
    uint8_t* color;      /* RGBA8_UNORM, a = 0 */
    float* depth;        /* R32F */
    float* motion;       /* RG32F (always 0) */
    uint8_t* mask;       /* R8_UNORM */
    float* position;     /* RGBA32F, a = 0 */
    int8_t* normal;      /* RGBA8_SNORM, a = 0 */
    /* instrumentation (not produced by the reference shader; required by BASELINE.json) */
    int16_t* hit;        /* x,y,z of mapPos at loop exit of the PRIMARY ray, and its step count i */
    uint8_t* hit_palette;/* material at loop exit of the primary ray */
    float* color_f32;    /* RGB before the UNORM8 store (3 floats) */
    uint32_t* rays;      /* rays traced for this pixel (primary + all traceRay/traceRayHit calls) */
    uint32_t* fetches;   /* getVoxel calls over all rays of this pixel (== algorithmic bytes) */
} orc_planes;

/* Render rows [row_begin,row_end) with stride row_step (pixels of other rows untouched). nthreads<=0: all cores. */
int orc_render(const orc_scene* scene, const orc_camera* cam, const orc_settings* st, const orc_planes* out,
               int row_begin, int row_end, int row_step, int nthreads);

/* Trace one ray (traceRayInt + traceRay, voxel_volume.frag:127-196); out[0..2]=mapPos, out[3]=steps i,
 * out[4]=material, out[5]=fetches; fout[0..2]=pos, [3..5]=normal. */
int orc_trace_one(const orc_scene* scene, const float start[3], const float dir[3], uint32_t maxSteps,
                  int32_t out[6], float fout[6]);

/* Denoiser: denoiser.frag:38-73 with host parameters of denoiser_stage.cpp:37-61,143-154.
 * color_in RGBA8, normal RGBA8_SNORM, position RGBA32F; writes the final ping-pong target to color_out. */
typedef struct { int iterations; float phiColor0, phiNormal0, phiPos0, stepWidth; int mode; } orc_denoise_settings;
int orc_denoise(const uint8_t* color_in, const int8_t* normal, const float* position, int w, int h,
                const orc_denoise_settings* ds, uint8_t* color_out, int nthreads);

/* CameraController::updateDirectionVectors, camera_controller.cpp:15-28. dir = normalDir*focalLength. */
void orc_camera_vectors(float yaw_deg, float pitch_deg, float focal, float dir[3], float right[3], float up[3]);

/* .vox -> dense volume + palette, restating voxel_scene.cpp:33-117 on an independent .vox reader.
 * Returns 0, or <0 on error (-2 = git-LFS pointer stub). Caller frees with orc_free. */
int orc_load_vox(const uint8_t* bytes, size_t n, uint8_t** volume, uint32_t whd[3], float palette[256 * 8]);
void orc_free(void* p);
/* synthetic volumes of BASELINE configs 4/5: 1 = noise terrain, 2 = sparse lattice shapes */
int orc_generate(int generator, uint32_t seed, uint8_t* vol, uint32_t W, uint32_t H, uint32_t D);
int orc_num_threads(void);

/* ---- the steps either side of the pass (orc_output.cpp) ---- */
/* ffxFsr2GetJitterPhaseCount / ffxFsr2GetJitterOffset (FSR2 v2.1.0, thirdparty/fsr2/include/ffx_fsr2.h:356,395) */
int orc_jitter_phase_count(int renderWidth, int displayWidth);
void orc_jitter_offset(float* outX, float* outY, int index, int phaseCount);
/* UpscalerStage::update, upscaler_stage.cpp:59-70 (jitter of this frame; advances *frameCount) */
void orc_upscaler_update(int renderWidth, int targetWidth, int* frameCount, float* jitterX, float* jitterY);
/* VoxelRenderSettings::renderResolution, voxel_render_settings.cpp:3-13 */
void orc_render_resolution(int fsrEnable, uint32_t scaling, const uint32_t target[2], uint32_t out[2]);
/* blit.frag:14-22 through a LINEAR / CLAMP_TO_EDGE sampler; RGBA8 in and out; flags bit 0: sRGB-encode rgb on store */
int orc_blit(const uint8_t* src, int sw, int sh, uint8_t* dst, int tw, int th, int flags, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
