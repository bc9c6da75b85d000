/* vrt.h — C ABI of libvrt, the B200-native (sm_100a) drop-in for the reference's voxel_volume render pass
 * (+ denoiser). One vrt_context per process and GPU; multi-GPU = one process per GPU, each context owning an
 * interleaved shard of image tiles (see vrt_context_set_shard), frame assembled on rank 0.
 *
 * Every function returns VRT_OK (0) or a negative vrt_status; nothing throws across the ABI. vrt_last_error()
 * returns a thread-local message for the last failing call. Handles are opaque; the caller owns every host
 * buffer it passes, the library owns all device memory. Calls on one context are not re-entrant (same as the
 * reference, which is single-threaded: SURVEY.md §8b); independent contexts may be used from different threads.
 *
 * Reference interfaces replaced (paths relative to /root/reference):
 *   vrt_camera              == ScreenQuadPush           source/voxels/resource/screen_quad_push.hpp:5-15
 *   vrt_volume_params       == VolumeParameters         source/voxels/resource/parameters.hpp:5-9
 *   vrt_light               == Light                    source/voxels/resource/voxel_scene.hpp:10-15
 *   vrt_material            == Material                 source/voxels/resource/material.hpp:5-12
 *   vrt_denoiser_settings   == DenoiserSettings         source/voxels/voxel_render_settings.hpp:21-29
 *   vrt_scene_load_vox      == VoxelScene::VoxelScene   source/voxels/resource/voxel_scene.cpp:33-133
 *   vrt_render              == GeometryStage::record    source/voxels/stages/geometry_stage.cpp:106-174
 *                            + DenoiserStage::record    source/voxels/stages/denoiser_stage.cpp:156-258
 *                              as sequenced by VoxelRenderer::recordCommands  source/voxels/voxel_renderer.cpp:55-105
 *   vrt_frame_plane         == GeometryBuffer           source/voxels/stages/geometry_stage.hpp:19-27
 *   vrt_camera_vectors      == CameraController::updateDirectionVectors  source/voxels/resource/camera_controller.cpp:15-28
 *   vrt_blit                == BlitStage::record + shader/blit.frag      source/voxels/stages/blit_stage.cpp:25-57, shader/blit.frag:14-22
 *   vrt_upscaler_update     == UpscalerStage::update (jitter + frame counter only; FSR2 itself is closed source and out of scope)
 *                                                       source/voxels/stages/upscaler_stage.cpp:59-70
 *   vrt_jitter_phase_count / vrt_jitter_offset == ffxFsr2GetJitterPhaseCount / ffxFsr2GetJitterOffset  thirdparty/fsr2/include/ffx_fsr2.h:356,395
 *   vrt_render_resolution   == VoxelRenderSettings::renderResolution     source/voxels/voxel_render_settings.cpp:3-13
 */
#ifndef VRT_H
#define VRT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    VRT_OK = 0,
    VRT_E_INVALID = -1,       /* bad argument */
    VRT_E_LFS_POINTER = -2,   /* the file is a git-LFS pointer stub, not data */
    VRT_E_NO_INSTANCE = -3,   /* "Voxel scene does not contain an instance."  voxel_scene.cpp:49-50 */
    VRT_E_PARSE = -4,         /* "Could not parse voxel scene"               voxel_scene.cpp:45-46 */
    VRT_E_IO = -5,            /* "Failed to read voxel scene"                voxel_scene.cpp:40-43 */
    VRT_E_CUDA = -6,          /* CUDA runtime error (message in vrt_last_error) */
    VRT_E_NO_DEVICE = -7,     /* no CUDA device: there is NO CPU fallback */
    VRT_E_UNSUPPORTED = -8,
    VRT_E_OOM = -9,
    VRT_E_COMM = -10          /* multi-GPU exchange: rendezvous failed, a rank reported an error, or a wait timed out */
} vrt_status;

typedef struct vrt_context vrt_context;
typedef struct vrt_scene vrt_scene;
typedef struct vrt_frame vrt_frame;
typedef struct vrt_upscaler vrt_upscaler;

/* == ScreenQuadPush: 96 bytes, offsets 0,16,32,48,64,76,80,88 */
typedef struct {
    float camPos[4], camDir[4], camRight[4], camUp[4];
    uint32_t volumeBounds[3];
    uint32_t frame;
    int32_t screenSize[2];
    float cameraJitter[2];
} vrt_camera;

typedef struct { uint32_t aoSamples; float ambientIntensity; } vrt_volume_params;
typedef struct { float direction[3]; float intensity; float color[4]; } vrt_light;
typedef struct { float diffuse[4]; float metallic, pad1, pad2, pad3; } vrt_material;

enum { VRT_DENOISE_AS_EXECUTED = 0, VRT_DENOISE_AS_INTENDED = 1 };   /* SURVEY.md F6 */
typedef struct {
    int32_t enable;
    int32_t iterations;                 /* 1..10 (MAX_DENOISER_PASSES, denoiser_stage.hpp:9) */
    float phiColor0, phiNormal0, phiPos0, stepWidth;
    int32_t mode;
} vrt_denoiser_settings;

/* pass_mask bits (new; the reference behaves as VRT_PASS_ALL). SHADE off => colour = palette albedo. */
enum { VRT_PASS_PRIMARY = 1, VRT_PASS_SHADOW = 2, VRT_PASS_AO = 4, VRT_PASS_REFLECT = 8, VRT_PASS_SHADE = 16, VRT_PASS_ALL = 31 };
/* output selection bits */
enum {
    VRT_OUT_GBUFFER = 1,      /* depth, motion, mask, position, normal planes (the reference always writes them) */
    VRT_OUT_INSTRUMENT = 2,   /* hit voxel + step count, palette index, float colour, per-pixel ray / fetch counts */
};
/* traversal back-ends (all bit-identical; for measurement). 0 = library default. */
enum { VRT_TRAVERSAL_DEFAULT = 0, VRT_TRAVERSAL_DENSE_U8 = 1, VRT_TRAVERSAL_BRICK = 2, VRT_TRAVERSAL_FINE = 3, VRT_TRAVERSAL_LINEAR = 4, VRT_TRAVERSAL_OCTANT = 5, VRT_TRAVERSAL_SPARSE = 6 };

/* how the pass is scheduled on the GPU (all bit-identical; for measurement). 0 = library default (wavefront).
 *   MEGAKERNEL: one thread runs a pixel's whole ray tree (the shader's own shape).
 *   WAVEFRONT : primary / reflection-bounce / AO+shadow / resolve kernels over compact hit records. */
enum { VRT_PIPELINE_DEFAULT = 0, VRT_PIPELINE_MEGAKERNEL = 1, VRT_PIPELINE_WAVEFRONT = 2 };

typedef struct {
    uint32_t pass_mask;
    uint32_t out_mask;
    uint32_t traversal;
    uint32_t pipeline;
    vrt_volume_params vol;
    vrt_light light;
    vrt_denoiser_settings denoise;
} vrt_render_settings;

/* planes of a frame */
typedef enum {
    VRT_PLANE_COLOR_RGBA8 = 0,      /* 4 B/px, a = 0 */
    VRT_PLANE_DENOISED_RGBA8 = 1,   /* 4 B/px (valid if denoise.enable) */
    VRT_PLANE_DEPTH_R32F = 2,
    VRT_PLANE_MOTION_RG32F = 3,
    VRT_PLANE_MASK_R8 = 4,
    VRT_PLANE_POSITION_RGBA32F = 5,
    VRT_PLANE_NORMAL_RGBA8S = 6,
    VRT_PLANE_HIT_XYZS_I16X4 = 7,   /* instrumented: primary-ray mapPos x,y,z at loop exit + step count */
    VRT_PLANE_HIT_PALETTE_U8 = 8,   /* instrumented: material at loop exit */
    VRT_PLANE_COLOR_RGB32F = 9,     /* instrumented: colour before the UNORM8 store */
    VRT_PLANE_RAYS_U32 = 10,        /* instrumented: rays traced for the pixel */
    VRT_PLANE_FETCHES_U32 = 11,     /* instrumented: getVoxel calls of the reference algorithm for the pixel */
    VRT_PLANE_COUNT = 12
} vrt_plane;

typedef struct {
    uint64_t rays;            /* rays the reference algorithm traces for this frame (this rank's pixels) */
    uint64_t fetches;         /* getVoxel calls of the reference algorithm == algorithmic bytes of the traversal */
    uint64_t pixels;          /* pixels rendered by this rank */
    float ms_trace;           /* CUDA-event time of the trace+shade kernel(s) */
    float ms_denoise;
    float ms_total;           /* first launch to last kernel of the frame on the context stream */
    uint32_t kernel_launches; /* kernels launched for this frame */
    uint32_t reserved;
    /* per stage of the wavefront pipeline: 0 primary, 1 reflection bounces, 2 AO + shadow (any-hit), 3 resolve.
     * Megakernel pipeline: everything is stage 0. */
    float ms_stage[4];
    uint64_t stage_rays[3];
    uint64_t stage_fetches[3];
} vrt_frame_stats;

/* ---- context ---- */
int vrt_context_create(int device, vrt_context** out);          /* VRT_E_NO_DEVICE when there is no GPU */
void vrt_context_destroy(vrt_context*);
/* tile shard of this context: tiles t with (t % world) == rank are rendered (interleaved for load balance). */
int vrt_context_set_shard(vrt_context*, int rank, int world);
const char* vrt_last_error(void);
const char* vrt_version(void);
int vrt_device_info(vrt_context*, int* sm_count, size_t* l2_bytes, size_t* hbm_bytes);
int vrt_context_sync(vrt_context*);                             /* wait for everything queued on the context stream */
void* vrt_context_stream(vrt_context*);                         /* cudaStream_t the context launches on */
/* Frames in flight (the reference keeps two: source/engine/engine.hpp:19). With this on, every frame object created afterwards
 * renders on a stream of its own, forked from the context stream at each vrt_render, so the kernels of one frame fill the tails
 * of another's. vrt_context_join: the context stream waits for everything queued on the frames' streams so far;
 * vrt_context_sync waits on the host for all of it; vrt_frame_stream: the cudaStream_t a frame's work is queued on. */
int vrt_context_set_concurrent_frames(vrt_context*, int on);
int vrt_context_join(vrt_context*);
/* device scratch owned by the caller (outputs of vrt_denoise / vrt_assemble, gather staging) */
int vrt_buffer_alloc(vrt_context*, size_t bytes, void** d_ptr);
void vrt_buffer_free(vrt_context*, void* d_ptr);
int vrt_buffer_download(vrt_context*, const void* d_src, void* host_dst, size_t bytes);   /* on the context stream; syncs */

/* ---- scene ---- */
int vrt_scene_load_vox(vrt_context*, const char* vox_path, vrt_scene** out);
int vrt_scene_load_vox_memory(vrt_context*, const uint8_t* bytes, size_t n, vrt_scene** out);
int vrt_scene_from_dense(vrt_context*, const uint8_t* volume, uint32_t w, uint32_t h, uint32_t d,
                         const vrt_material palette[256], vrt_scene** out);
/* on-GPU procedural volumes: generator 1 = noise terrain, 2 = sparse lattice shapes (DESIGN.md "synthetic volumes") */
int vrt_scene_generate(vrt_context*, int generator, uint64_t seed, uint32_t w, uint32_t h, uint32_t d, vrt_scene** out);
int vrt_scene_set_noise(vrt_scene*, const uint8_t* rgba8, uint32_t w, uint32_t h);      /* must be 512x512 */
int vrt_scene_set_skybox(vrt_scene*, const float* rgba32f, uint32_t w, uint32_t h);
/* The reference's file inputs (VoxelScene(engine, voxPath, skyboxFilename): voxel_scene.cpp:33,132; VoxelRenderer's noise texture:
 * voxel_renderer.cpp:22; both through Texture2D = stbi_loadf / stbi_load with 4 channels, texture_2d.cpp:22-38):
 * skybox = Radiance RGBE .hdr -> RGBA32F (flat and run-length scanlines), noise = PNG -> RGBA8 (must be 512 x 512).
 * VRT_E_LFS_POINTER when the file is a git-LFS pointer stub, VRT_E_IO / VRT_E_PARSE like "Could not load image". */
int vrt_scene_load_skybox(vrt_scene*, const char* hdr_path);
int vrt_scene_load_noise(vrt_scene*, const char* png_path);
int vrt_scene_bounds(const vrt_scene*, uint32_t whd[3]);
int vrt_scene_palette(const vrt_scene*, vrt_material palette[256]);
int vrt_scene_download_volume(const vrt_scene*, uint8_t* dst, size_t bytes);
/* Only the volume and the palette are resident for every scene; the safe-step fields of a traversal back-end are built the first
 * time it is asked for. The default back-end is prepared by the loaders, so this is only needed before measuring another one
 * (0 = default). Thread-safe per scene. */
int vrt_scene_prepare(vrt_scene*, uint32_t traversal);
/* device bytes of the scene: the dense volume, the fields built so far, the peak during their build; host wall time of load + builds */
int vrt_scene_memory(const vrt_scene*, size_t* volume_bytes, size_t* field_bytes, size_t* peak_build_bytes, double* load_ms);
int vrt_scene_accel_info(const vrt_scene*, size_t* dense_bytes, size_t* brick_mask_bytes, size_t* brick_dist_bytes);
/* brick levels back to the host (tests): bmask = 8 B per 4x4x4 brick, bdist = 1 B per brick; either may be NULL */
int vrt_scene_download_accel(const vrt_scene*, uint64_t* bmask_dst, uint8_t* bdist_dst);
/* per-voxel safe-step field (1 B per voxel, 4x4x4-bricked: 64 B per brick, (x&3) + (y&3)*4 + (z&3)*16 inside) */
int vrt_scene_download_fdist(const vrt_scene*, uint8_t* dst, size_t bytes);
/* directional safe-step field of one direction octant (0..7), aproned linear: (w+2)*(h+2)*(d+2) bytes; built on first use */
int vrt_scene_download_odist(const vrt_scene*, int octant, uint8_t* dst, size_t bytes);
/* the two-level (sparse) directional field of one octant, expanded to the same aproned linear layout (tests; built on first use) */
int vrt_scene_download_sdist(const vrt_scene*, int octant, uint8_t* dst, size_t bytes);
void vrt_scene_destroy(vrt_scene*);
/* Host-only .vox decode (no GPU involved): the dense volume + palette vrt_scene_load_vox_memory would upload.
 * *volume_out is malloc'ed (x + y*w + z*w*h); free it with vrt_host_free. */
int vrt_vox_decode(const uint8_t* bytes, size_t n, uint8_t** volume_out, uint32_t whd[3], vrt_material palette[256]);
void vrt_host_free(void*);

/* ---- render ---- */
/* Asynchronous on the context stream. If *frame is NULL a frame is allocated, otherwise it is reused
 * (must match the screen size). cam->volumeBounds must equal the scene bounds. */
int vrt_render(vrt_context*, const vrt_scene*, const vrt_camera*, const vrt_render_settings*, vrt_frame** frame);
int vrt_frame_sync(vrt_frame*);
int vrt_frame_plane(vrt_frame*, int plane, void** device_ptr, size_t* bytes);
int vrt_frame_download(vrt_frame*, int plane, void* host_dst, size_t bytes);
/* Read-back that overlaps the NEXT frame (the reference keeps 2 frames in flight, source/engine/engine.hpp:19): the copy is
 * queued on the context's copy stream behind this frame's kernels and the call returns at once; host_dst should be pinned
 * memory and must stay untouched until vrt_frame_download_wait returns. Rendering into the same frame again waits for the
 * copy on the device by itself, so alternate between two frames to keep the GPU busy. */
int vrt_frame_download_async(vrt_frame*, int plane, void* host_dst, size_t bytes);
int vrt_frame_download_wait(vrt_frame*);
int vrt_frame_stats_get(vrt_frame*, vrt_frame_stats* out);       /* syncs */
void* vrt_frame_stream(vrt_frame*);
void vrt_frame_release(vrt_frame*);

/* Stand-alone denoiser on host planes already resident on device: DenoiserStage::record. */
int vrt_denoise(vrt_context*, const void* d_color_rgba8, const void* d_normal_rgba8s, const void* d_position_rgba32f,
                int w, int h, const vrt_denoiser_settings*, void* d_out_rgba8);
/* The same without synchronising: on `stream` (NULL: the context stream), with a caller-owned scratch image of w*h*4 bytes
 * (needed when iterations > 1). This is what rank 0 runs on the assembled planes of a sharded frame. */
int vrt_denoise_on(vrt_context*, const void* d_color_rgba8, const void* d_normal_rgba8s, const void* d_position_rgba32f,
                   int w, int h, const vrt_denoiser_settings*, void* d_out_rgba8, void* d_scratch_rgba8, void* stream);
/* vrt_denoise_on for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same inputs, output, stream meaning
 * and results as vrt_denoise_on, but the intermediate image the iterations ping-pong through lives in a file-scope __device__
 * array of the library (csrc/vrt_denoise.cu) instead of caller memory, and the call performs no cudaMalloc / cudaFree or any
 * other allocating routine. Limits and rules:
 *   - w * h <= VRT_DENOISE_STATIC_MAX_PIXELS (one 3840 x 2160 image: 33 177 600 B, held by every GPU that loads the library
 *     until the process exits); larger images are rejected with VRT_E_INVALID and a message naming the limit.
 *   - every call on a device shares the one array and it keeps the last call's contents. Nothing in it has to start at zero
 *     (each iteration writes every pixel of its target before the next one reads it), so nothing is cleared; but two calls
 *     that overlap on the GPU (different streams, no ordering between them) race on it: order them.
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress, once per device): make one
 *     call before starting a capture. */
#define VRT_DENOISE_STATIC_MAX_PIXELS (3840 * 2160)
int vrt_denoise_static(vrt_context*, const void* d_color_rgba8, const void* d_normal_rgba8s, const void* d_position_rgba32f,
                       int w, int h, const vrt_denoiser_settings*, void* d_out_rgba8, void* stream);

/* Multi-GPU assembly on the gathering rank: d_gathered holds `world` shard buffers back to back (rank-major), each
 * of vrt_frame_plane() size as produced under vrt_context_set_shard(rank, world); writes a row-major w x h plane. */
int vrt_assemble(vrt_context*, const void* d_gathered, int world, int w, int h, int plane, void* d_out);
/* the same on a cudaStream_t of the caller's (the stream its gather runs on), leaving the context stream free for the next frame */
int vrt_assemble_on(vrt_context*, const void* d_gathered, int world, int w, int h, int plane, void* d_out, void* stream);

/* Assembly without a collective: this rank's tiles of `plane` are stored straight into the row-major w x h frame at d_frame, which
 * may be peer memory (another GPU's buffer mapped into this process over NVLink). On `stream` (NULL: the frame's own stream); the
 * caller orders the stores against the consumer (device-side barrier / signal). With world == 1 it is a plain de-tile copy. */
int vrt_frame_push_tiles(vrt_frame*, int plane, void* d_frame, void* stream);

/* ---- multi-GPU exchange inside the library: one process per GPU, every rank traces its interleaved tile shard,
 * the frame is assembled on rank 0. No collective call on the data path: rank 0 owns the assembled planes of `slots` frames in flight
 * and exports them (CUDA IPC); every other rank maps them as peer memory (NVLink / NVSwitch) and stores its tiles straight into the
 * row-major frame, then an arrival flag; rank 0 orders its consumer behind a wait kernel on the flags. The same stores can go to a
 * frame in pinned HOST memory shared by all ranks, so a read-back uses every GPU's PCIe link at once.
 *   key: names the POSIX shared-memory rendezvous segment; the same on every rank, unique per job (rank 0 creates it).
 *   vrt_comm_create also sets the tile shard (vrt_context_set_shard(rank, world)); call it before the first frame.
 *   vrt_comm_alloc_frames is collective (same arguments on every rank); plane masks are bit (1 << vrt_plane).
 *   Per frame, every rank: vrt_render_exchange(..., slot, ...) (fused), or vrt_render then vrt_frame_exchange(frame, slot, ...) —
 *   asynchronous; the call itself may block on the host until rank 0 has released what the slot held before (use more slots than
 *   frames in flight and it never does).
 *   Rank 0 then: vrt_comm_wait(slot, stream) [device-side], its own work on that stream (vrt_denoise_on, copies, ...),
 *   vrt_comm_release_on(slot, stream); or vrt_comm_wait_host(slot) / vrt_comm_release(slot) for the host frame.
 * Every wait has a time limit (VRT_COMM_TIMEOUT_S, default 120 s) and returns VRT_E_COMM instead of hanging when a rank is gone. */
int vrt_comm_create(vrt_context*, int rank, int world, const char* key);
int vrt_comm_destroy(vrt_context*);                       /* collective; vrt_context_destroy calls it */
int vrt_comm_barrier(vrt_context*);                       /* host barrier over all ranks */
int vrt_comm_alloc_frames(vrt_context*, int w, int h, int slots, uint32_t plane_mask, uint32_t host_plane_mask);
int vrt_frame_exchange(vrt_frame*, int slot, uint32_t plane_mask, uint32_t host_plane_mask, void* stream);
/* vrt_render + vrt_frame_exchange in one call with the exchange FUSED into the pass: the colour / G-buffer planes of plane_mask are
 * stored by the pass's own kernels straight into rank 0's frame (peer stores over NVLink beside the local ones) and the arrival flag
 * follows the last kernel as a stream memory operation: no push kernel, the last kernel of every rank is the collective. */
int vrt_render_exchange(vrt_context*, const vrt_scene*, const vrt_camera*, const vrt_render_settings*, int slot,
                        uint32_t plane_mask, uint32_t host_plane_mask, vrt_frame** frame);
/* The denoiser of a sharded frame, distributed over the ranks ("strips"; SURVEY.md 8(e), in place of route (a) — all inputs to rank 0,
 * denoiser there — which stays available through vrt_render_exchange + vrt_denoise_on). Replaces VoxelRenderer::recordCommands'
 * geometry + denoiser stages (source/voxels/voxel_renderer.cpp:85-88) for world > 1.
 *   vrt_comm_alloc_strips: collective, after vrt_comm_alloc_frames with VRT_PLANE_DENOISED_RGBA8 in plane_mask (and in host_plane_mask
 *     if to_host will be used). Rank q will denoise rows [q * ceil(h / world), ...) and holds those rows plus a halo (the summed reach
 *     of the iterations' taps) of colour, normal and position for every slot. VRT_E_UNSUPPORTED if a strip is thinner than the halo.
 *   vrt_render_denoise_exchange: every rank, per frame. The pass's kernels store this rank's pixels into the strips of the ranks that
 *     own their rows (NVLink peer stores); each rank denoises its rows once all contributions are in (settings: those given to
 *     vrt_comm_alloc_strips; st->denoise is ignored) and copies them into the denoised plane of `slot` on rank 0 (to_host != 0: and of
 *     the shared host frame). Rank 0 then uses vrt_comm_wait / vrt_comm_wait_host / vrt_comm_release as for any exchange. The assembled
 *     plane is byte-identical to the single-GPU denoised frame. Wavefront pipeline only.
 *     to_host: 0 = into rank 0's device frame, 1 = there and into the shared host frame, 2 = into the shared host frame only (then only
 *     vrt_comm_wait_host sees the frame arrive).
 *   Without a denoiser (vrt_comm_alloc_strips(ctx, NULL)) the strips carry the colour plane alone and deliver it unchanged: the read-back
 *     path of a sharded frame — tiles cross NVLink into row strips, every rank's copy engine writes whole rows into the shared host frame
 *     over its own PCIe link, and no push kernel occupies an SM. The frames then need VRT_PLANE_COLOR_RGBA8 in the masks used. */
/* rows above and below a range of rows that the denoiser's output for that range depends on (the summed reach of all iterations' taps): the
 * halo of a strip. -1: unbounded / bad settings. Pure host arithmetic (no device needed). New; derived from the per-iteration stepWidth
 * of DenoiserStage (source/voxels/stages/denoiser_stage.cpp:143-154) and the 3x3 tap offsets (:37-61). */
int vrt_denoise_reach_rows(const vrt_denoiser_settings*);
int vrt_comm_alloc_strips(vrt_context*, const vrt_denoiser_settings*);
int vrt_render_denoise_exchange(vrt_context*, const vrt_scene*, const vrt_camera*, const vrt_render_settings*, int slot, int to_host, vrt_frame** frame);
int vrt_comm_wait(vrt_context*, int slot, void* stream);  /* rank 0; stream NULL: vrt_comm_stream() */
int vrt_comm_wait_host(vrt_context*, int slot);           /* rank 0 */
int vrt_comm_release(vrt_context*, int slot);             /* rank 0 */
int vrt_comm_release_on(vrt_context*, int slot, void* stream);
int vrt_comm_frame_plane(vrt_context*, int slot, int plane, void** d_ptr, void** h_ptr, size_t* bytes);
void* vrt_comm_stream(vrt_context*);                      /* rank 0's consumer stream; vrt_context_join / _sync include it */

/* ---- temporal upscaler: the stage after the denoiser, UpscalerStage::record -> ffxFsr2ContextDispatch
 * (source/voxels/stages/upscaler_stage.cpp:78-172). FSR2 is closed source in the reference checkout, so this is an OWN DEFINITION of
 * the same interface (PARITY UNPINNED): jittered temporal accumulation with neighbourhood rectification, documented in
 * csrc/vrt_upscale.cu. Inputs are row-major device planes of the render resolution (what vrt_frame_plane returns on an unsharded
 * frame, or rank 0's assembled planes): colour RGBA8, depth R32F, motion RG32F (may be NULL), reactive mask R8 (may be NULL); jitter
 * as UpscalerStage::update / vrt_upscaler_update produced it for this frame; reset: ignore the history. Output: target-resolution RGBA8.
 * Deterministic; exact pass-through at scale 1 with jitter (0, 0). Asynchronous on `stream` (NULL: the context stream). */
int vrt_upscaler_create(vrt_context*, int renderW, int renderH, int targetW, int targetH, vrt_upscaler** out);
int vrt_upscaler_dispatch(vrt_upscaler*, const void* d_color_rgba8, const void* d_depth_r32f, const void* d_motion_rg32f, const void* d_mask_r8,
                          float jitterX, float jitterY, float mvScaleX, float mvScaleY, int reset, void* d_out_rgba8, void* stream);
void vrt_upscaler_destroy(vrt_upscaler*);

/* ---- output stage (the step downstream of the pass): BlitStage::record running blit.frag, headless ----
 * d_src: sw x sh RGBA8 on the device (the colour or denoised plane of a full frame, or any RGBA8 image); d_dst: tw x th RGBA8 device
 * buffer standing in for the swap-chain image. Sampling is the reference's: LINEAR, CLAMP_TO_EDGE; the source is scaled to fill
 * the target and centre-cropped (blit.frag:17-19). VRT_BLIT_SRGB: encode rgb linear -> sRGB on store, as the reference's swap chain
 * does (vk-bootstrap default format B8G8R8A8_SRGB, source/engine/swapchain.cpp:15); VRT_BLIT_BGRA: that format's byte order.
 * vrt_blit runs on the context stream and returns without synchronising (vrt_context_sync / vrt_buffer_download order after it). */
enum { VRT_BLIT_SRGB = 1, VRT_BLIT_BGRA = 2 };
int vrt_blit(vrt_context*, const void* d_src_rgba8, int sw, int sh, void* d_dst_rgba8, int tw, int th, uint32_t flags);
int vrt_blit_on(vrt_context*, const void* d_src_rgba8, int sw, int sh, void* d_dst_rgba8, int tw, int th, uint32_t flags, void* stream);
/* image files for the headless path (host memory in, RGBA8 row-major): binary PPM (rgb), PNG rgb (tagged sRGB) or PNG rgba */
enum { VRT_IMAGE_PPM = 0, VRT_IMAGE_PNG_RGB = 1, VRT_IMAGE_PNG_RGBA = 2 };
int vrt_image_write(const char* path, const void* rgba8, int w, int h, int format);

/* image decoders behind vrt_scene_load_skybox / vrt_scene_load_noise (host only; *out is malloc'ed, free with vrt_host_free) */
int vrt_image_read_hdr(const char* path, float** rgba32f_out, uint32_t* w, uint32_t* h);
int vrt_image_decode_hdr(const uint8_t* bytes, size_t n, float** rgba32f_out, uint32_t* w, uint32_t* h);
int vrt_image_read_png(const char* path, uint8_t** rgba8_out, uint32_t* w, uint32_t* h);
int vrt_image_decode_png(const uint8_t* bytes, size_t n, uint8_t** rgba8_out, uint32_t* w, uint32_t* h);
/* Every bundled asset of the reference checkout is a git-LFS pointer stub. These decide whether a file that has been
 * supplied IS the asset a stub names: sha256 and size must equal the pointer's oid and size. VRT_OK = use it instead of the stand-in. */
int vrt_sha256(const void* data, size_t n, char hex[65]);
int vrt_file_sha256(const char* path, char hex[65]);
int vrt_lfs_pointer_info(const char* path, char oid_hex[65], uint64_t* size);   /* VRT_E_INVALID if `path` is not a pointer file */
int vrt_asset_matches_pointer(const char* candidate_path, const char* pointer_path);

/* ---- frame driver (the caller of the pass): jitter sequence and frame counter of UpscalerStage::update ----
 * Pure host functions, usable without a GPU. vrt_upscaler_update: jitter of this frame from *frameCount, then advances it the way the
 * reference does (wraps to 0 only once it exceeds the phase count); the values go to vrt_camera.cameraJitter / .frame. */
int vrt_jitter_phase_count(int renderWidth, int displayWidth);
int vrt_jitter_offset(int index, int phaseCount, float* outX, float* outY);
int vrt_upscaler_update(int renderWidth, int targetWidth, int* frameCount, float* jitterX, float* jitterY);
/* scaling = FsrScaling (10 NONE, 15 QUALITY, 17 BALANCED, 20 PERFORMANCE, 30 ULTRA_PERFORMANCE; voxel_render_settings.hpp:6-13) */
void vrt_render_resolution(int fsrEnable, uint32_t scaling, const uint32_t target[2], uint32_t out[2]);

/* ---- host helpers mirroring reference host code ---- */
void vrt_camera_vectors(float yaw_deg, float pitch_deg, float focal, float dir[3], float right[3], float up[3]);
void vrt_default_settings(vrt_render_settings* out);            /* VoxelRenderSettings{} defaults, voxel_render_settings.hpp:15-59 */

#ifdef __cplusplus
}
#endif
#endif
