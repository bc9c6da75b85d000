// vrt_internal.h — the objects behind the opaque handles of include/vrt.h, shared by the translation units that implement the ABI
// (vrt_api.cu: context / scene / frame / render; vrt_comm.cu: the multi-GPU exchange). Internal: nothing here is ABI.
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include <string>
#include <vector>
#include "vrt_launch.h"

extern thread_local std::string g_vrt_err;
inline int fail(int code, const std::string& msg) { g_vrt_err = msg; return code; }
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(VRT_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

struct vrt_comm;

using namespace vrt;

struct vrt_context {
    int device = 0, rank = 0, world = 1;
    cudaStream_t stream = nullptr;
    cudaStream_t copyStream = nullptr;   // asynchronous read-backs (vrt_frame_download_async)
    WaveSide side{};                     // low-priority stream + fork/join events (launch_wavefront)
    cudaDeviceProp prop{};
    bool concurrent = false;             // frames created from now on render on their own streams (vrt_context_set_concurrent_frames)
    std::vector<struct vrt_frame*> frames;
    vrt_comm* comm = nullptr;            // multi-GPU exchange (vrt_comm.cu), or null
};

// Only the dense volume and the palette are resident for every scene. The fields of a traversal back-end are built when that
// back-end is first asked for (vrt_scene_prepare, or the first vrt_render that names it), under the scene's mutex: the default
// back-end (OCTANT) is prepared at load time so that no frame pays for it.
struct vrt_scene {
    vrt_context* ctx = nullptr;
    uint32_t w = 0, h = 0, d = 0, Bx = 0, By = 0, Bz = 0;
    uint8_t* vol = nullptr; unsigned long long* bmask = nullptr; uint8_t* bdist = nullptr; uint8_t* fdist = nullptr; uint8_t* ldist = nullptr;
    uint8_t* odist[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    uint32_t* stop[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // two-level directional fields: top level
    uint8_t* sfine[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   //   per-voxel bricks (512 B each)
    uint32_t sfineBricks[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t TB[3] = {0, 0, 0};
    bool sparseReady = false;
    uint32_t* vtop = nullptr; uint8_t* vfine = nullptr; uint32_t vfineBricks = 0;   // the palette volume, two-level (vol == nullptr then)
    size_t volumeBytes = 0;              // device bytes of the volume as stored (dense, or top + bricks)
    vrt_material* palette = nullptr; vrt_material hostPalette[256];
    uint8_t* noise = nullptr; float* sky = nullptr; uint32_t skyW = 0, skyH = 0;
    int accelLaunches = 0;
    std::mutex fieldMutex;               // guards the lazy field builds below
    size_t fieldBytes = 0;               // safe-step fields resident now
    size_t peakBuildBytes = 0;           // largest scratch + fields footprint seen during a build
    double loadMs = 0.0;                 // host wall time of upload / generate / flatten + field builds so far
};

static const int kPlaneBpp[VRT_PLANE_COUNT] = {4, 4, 4, 8, 1, 16, 4, 8, 1, 12, 4, 4};   // bytes per pixel of each vrt_plane

struct vrt_frame {
    vrt_context* ctx = nullptr;
    int W = 0, H = 0;
    size_t pixels = 0;                 // pixels stored per plane on this rank (full frame, or localTiles*1024 when sharded)
    int tilesX = 0, tilesY = 0, localTiles = 0;
    void* planes[VRT_PLANE_COUNT] = {};
    void* ping[2] = {nullptr, nullptr};
    Counters* ctr = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr;
    cudaEvent_t es[3] = {nullptr, nullptr, nullptr};   // wavefront stage boundaries
    cudaEvent_t rendered = nullptr, copied = nullptr;  // asynchronous read-back: frame complete on the main stream / copy complete
    bool copyPending = false;
    bool staged = false;
    bool denoised = false;
    uint32_t launches = 0, out_mask = 0;
    cudaStream_t stream = nullptr;     // the stream this frame's work is queued on: the context's, or its own (concurrent frames)
    WaveSide side{};
    cudaEvent_t fork = nullptr;        // own stream: orders the frame after what the context stream held when vrt_render was called
    cudaEvent_t joined = nullptr;      // own stream: vrt_context_join
    bool own = false;
    WaveBuf wb{};                      // wavefront pipeline buffers (allocated on first use)
    uint32_t wbRays = 0;               // result planes allocated in wb.res
};


// vrt_render with the fused exchange targets (vrt_api.cu; used by vrt_render_exchange in vrt_comm.cu)
int vrt_render_impl(vrt_context* c, const vrt_scene* s, const vrt_camera* cam, const vrt_render_settings* st, vrt_frame** frame, const XPlanes* x);
// the denoiser's iteration loop (vrt_api.cu): iteration i writes ping[i % 2]; used by vrt_denoise, vrt_denoise_on and vrt_denoise_static
int run_denoise(vrt_context* c, cudaStream_t stream, const void* color, const void* normal, const void* position, int W, int H,
                const vrt_denoiser_settings* ds, void* ping[2], int* lastOut, uint32_t* launches);
