// vrt_api.cu — implementation of the C ABI in include/vrt.h: headless CUDA context, scene residency in HBM,
// frame planes as device buffers (instead of Vulkan render targets), per-frame orchestration that mirrors
// VoxelRenderer::recordCommands (voxel_renderer.cpp:55-105): geometry pass, then optional denoiser passes.
// There is NO CPU fallback: without a CUDA device context creation fails with VRT_E_NO_DEVICE.
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <chrono>
#include <mutex>
#include <new>
#include <string>
#include <vector>
#include "vrt_internal.h"
#include "vrt_vox.h"

thread_local std::string g_vrt_err;

extern "C" const char* vrt_last_error(void) { return g_vrt_err.c_str(); }
extern "C" const char* vrt_version(void) { return "vrt 0.1 (sm_100a)"; }

extern "C" void vrt_context_destroy(vrt_context* c);
extern "C" int vrt_context_create(int device, vrt_context** out) {
    if (!out) return fail(VRT_E_INVALID, "out is null");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) return fail(VRT_E_NO_DEVICE, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(VRT_E_INVALID, "device index out of range");
    CU(cudaSetDevice(device));
    vrt_context* c = new (std::nothrow) vrt_context();
    if (!c) return fail(VRT_E_OOM, "host allocation failed");
    c->device = device;
    int prLow = 0, prHigh = 0;
    e = cudaGetDeviceProperties(&c->prop, device);
    if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prLow, &prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&c->side.stream, cudaStreamNonBlocking, prLow);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->side.fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->side.join, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { vrt_context_destroy(c); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }   // frees whatever was created
    *out = c;
    return VRT_OK;
}
int vrt_comm_join_into(vrt_context* c);   // vrt_comm.cu
int vrt_comm_sync(vrt_context* c);
extern "C" int vrt_comm_destroy(vrt_context* c);
extern "C" void vrt_context_destroy(vrt_context* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    vrt_comm_destroy(c);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->copyStream) cudaStreamDestroy(c->copyStream);
    if (c->side.stream) cudaStreamDestroy(c->side.stream);
    if (c->side.fork) cudaEventDestroy(c->side.fork);
    if (c->side.join) cudaEventDestroy(c->side.join);
    delete c;
}
extern "C" int vrt_context_set_shard(vrt_context* c, int rank, int world) {
    if (!c || world < 1 || rank < 0 || rank >= world) return fail(VRT_E_INVALID, "bad shard");
    c->rank = rank; c->world = world;
    return VRT_OK;
}
extern "C" int vrt_device_info(vrt_context* c, int* sm_count, size_t* l2_bytes, size_t* hbm_bytes) {
    if (!c) return fail(VRT_E_INVALID, "ctx is null");
    if (sm_count) *sm_count = c->prop.multiProcessorCount;
    if (l2_bytes) *l2_bytes = (size_t)c->prop.l2CacheSize;
    if (hbm_bytes) *hbm_bytes = c->prop.totalGlobalMem;
    return VRT_OK;
}
extern "C" int vrt_context_sync(vrt_context* c) {
    if (!c) return fail(VRT_E_INVALID, "ctx is null");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    for (vrt_frame* f : c->frames) if (f->own) CU(cudaStreamSynchronize(f->stream));
    return vrt_comm_sync(c);
}
extern "C" void* vrt_context_stream(vrt_context* c) { return c ? (void*)c->stream : nullptr; }
// Frames in flight: with this on, every frame object created afterwards renders on a stream of its own (forked from the context
// stream at each vrt_render), so that the kernels of one frame fill the tails of another's. vrt_context_join makes the context
// stream wait for everything queued on the frames' streams so far; vrt_context_sync waits on the host for all of it.
extern "C" int vrt_context_set_concurrent_frames(vrt_context* c, int on) {
    if (!c) return fail(VRT_E_INVALID, "ctx is null");
    c->concurrent = on != 0;
    return VRT_OK;
}
extern "C" int vrt_context_join(vrt_context* c) {
    if (!c) return fail(VRT_E_INVALID, "ctx is null");
    CU(cudaSetDevice(c->device));
    for (vrt_frame* f : c->frames) if (f->own) { CU(cudaEventRecord(f->joined, f->stream)); CU(cudaStreamWaitEvent(c->stream, f->joined, 0)); }
    return vrt_comm_join_into(c);
}

// ------------------------------------------------------------------------------------------------ scene
extern "C" int vrt_buffer_alloc(vrt_context* c, size_t bytes, void** d_ptr) {
    if (!c || !d_ptr || bytes == 0) return fail(VRT_E_INVALID, "bad argument");
    CU(cudaSetDevice(c->device));
    const cudaError_t e = cudaMalloc(d_ptr, bytes);
    if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e));
    return VRT_OK;
}
extern "C" void vrt_buffer_free(vrt_context* c, void* d_ptr) {
    if (!c || !d_ptr) return;
    cudaSetDevice(c->device);
    cudaFree(d_ptr);
}
extern "C" int vrt_buffer_download(vrt_context* c, const void* d_src, void* host_dst, size_t bytes) {
    if (!c || !d_src || !host_dst) return fail(VRT_E_INVALID, "null argument");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(host_dst, d_src, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return VRT_OK;
}

extern "C" void vrt_scene_destroy(vrt_scene* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    cudaFree(s->vol); cudaFree(s->bmask); cudaFree(s->bdist); cudaFree(s->fdist); cudaFree(s->ldist); for (uint8_t* p : s->odist) cudaFree(p); for (uint32_t* p : s->stop) cudaFree(p); for (uint8_t* p : s->sfine) cudaFree(p); cudaFree(s->vtop); cudaFree(s->vfine); cudaFree(s->palette); cudaFree(s->noise); cudaFree(s->sky);
    delete s;
}

// A volume larger than this stays two-level (bricks of one value + 512-byte bricks, like the sparse fields): cfg5's 8 GiB dense
// volume is then never materialised by vrt_scene_generate. VRT_VOLUME_DENSE_MAX_GIB overrides (0: every scene two-level; tests).
static bool volume_wants_sparse(uint32_t w, uint32_t h, uint32_t d) {
    const char* e = getenv("VRT_VOLUME_DENSE_MAX_GIB");
    const double maxGiB = e ? atof(e) : 4.0;
    return (double)w * (double)h * (double)d / 1073741824.0 > maxGiB;
}
static int scene_alloc(vrt_context* c, uint32_t w, uint32_t h, uint32_t d, vrt_scene** out, bool dense = true) {
    if (!c || !out) return fail(VRT_E_INVALID, "null argument");
    if (!w || !h || !d || w > 32767 || h > 32767 || d > 32767) return fail(VRT_E_INVALID, "volume bounds must be 1..32767 per axis");
    CU(cudaSetDevice(c->device));
    vrt_scene* s = new (std::nothrow) vrt_scene();
    if (!s) return fail(VRT_E_OOM, "host allocation failed");
    s->ctx = c; s->w = w; s->h = h; s->d = d;
    s->Bx = (w + 3) / 4; s->By = (h + 3) / 4; s->Bz = (d + 3) / 4;
    sparse_top_dims(w, h, d, s->TB);
    cudaError_t e = dense ? cudaMalloc(&s->vol, (size_t)w * h * d) : cudaSuccess;
    if (dense) s->volumeBytes = (size_t)w * h * d;
    if (e == cudaSuccess) e = cudaMalloc(&s->palette, sizeof(vrt_material) * 256);
    if (e != cudaSuccess) { vrt_scene_destroy(s); return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e)); }
    *out = s;
    return VRT_OK;
}
// The palette volume in the two-level layout: from the dense volume the scene holds (which is then freed) or, with generator != 0,
// straight from the procedural function (no dense volume ever exists).
static int scene_make_sparse_volume(vrt_scene* s, int generator, uint32_t seed) {
    vrt_context* c = s->ctx;
    const size_t topWords = (size_t)s->TB[0] * s->TB[1] * s->TB[2];
    uint32_t* counter = nullptr; uint32_t n = 0;
    cudaError_t e = cudaMalloc(&s->vtop, topWords * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&counter, sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemsetAsync(counter, 0, sizeof(uint32_t), c->stream);
    if (e == cudaSuccess) e = launch_sparse_volume_classify(s->vol, generator, seed, s->w, s->h, s->d, s->vtop, counter, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&n, counter, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess && n) e = cudaMalloc(&s->vfine, (size_t)n * 512u);
    if (e == cudaSuccess && n) e = launch_sparse_volume_fill(s->vol, generator, seed, s->w, s->h, s->d, s->vtop, s->vfine, c->stream);
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);
    cudaFree(counter);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e));
    s->vfineBricks = n; s->accelLaunches += 2;
    if (s->vol) { s->peakBuildBytes = std::max(s->peakBuildBytes, (size_t)s->w * s->h * s->d); cudaFree(s->vol); s->vol = nullptr; }
    s->volumeBytes = topWords * sizeof(uint32_t) + (size_t)n * 512u;
    return VRT_OK;
}

static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static int cuda_fail(cudaError_t e) { return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e)); }

// brick occupancy + brick distance (back-end BRICK; far-field term of FINE / LINEAR). Caller holds fieldMutex.
static int dense_only(const vrt_scene* s) {
    return s->vol ? VRT_OK : fail(VRT_E_UNSUPPORTED, "this traversal back-end is kept for measurement and needs the dense volume; the scene is over the dense budget and stored two-level (use the default back-end)");
}
static int scene_ensure_bricks(vrt_scene* s) {
    if (s->bmask) return VRT_OK;
    { const int rc = dense_only(s); if (rc) return rc; }
    vrt_context* c = s->ctx;
    const size_t nb = (size_t)s->Bx * s->By * s->Bz;
    uint8_t *tmpA = nullptr, *tmpB = nullptr;
    cudaError_t e = cudaMalloc(&s->bmask, nb * 8);
    if (e == cudaSuccess) e = cudaMalloc(&s->bdist, nb);
    if (e == cudaSuccess) e = cudaMalloc(&tmpA, nb);
    if (e == cudaSuccess) e = cudaMalloc(&tmpB, nb);
    if (e == cudaSuccess) e = launch_build_accel(s->vol, s->w, s->h, s->d, s->bmask, s->bdist, tmpA, tmpB, c->stream, &s->accelLaunches);
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);
    cudaFree(tmpA); cudaFree(tmpB);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) { cudaFree(s->bmask); cudaFree(s->bdist); s->bmask = nullptr; s->bdist = nullptr; return cuda_fail(e); }
    s->peakBuildBytes = std::max(s->peakBuildBytes, s->fieldBytes + nb * 11);
    s->fieldBytes += nb * 9;
    return VRT_OK;
}
// per-voxel isotropic safe-step field: bricked (FINE) and / or aproned linear (LINEAR); three volume-sized scratch buffers
static int scene_ensure_iso(vrt_scene* s, bool wantFine, bool wantLinear) {
    wantFine = wantFine && !s->fdist; wantLinear = wantLinear && !s->ldist;
    if (!wantFine && !wantLinear) return VRT_OK;
    int rc = scene_ensure_bricks(s);
    if (rc) return rc;
    vrt_context* c = s->ctx;
    const size_t nv = (size_t)s->w * s->h * s->d, nb = (size_t)s->Bx * s->By * s->Bz, nl = (size_t)(s->w + 2) * (s->h + 2) * (s->d + 2);
    uint8_t *a = nullptr, *t1 = nullptr, *t2 = nullptr, *fd = nullptr, *ld = nullptr;
    cudaError_t e = cudaMalloc(&a, nv);
    if (e == cudaSuccess) e = cudaMalloc(&t1, nv);
    if (e == cudaSuccess) e = cudaMalloc(&t2, nv);
    if (e == cudaSuccess && wantFine) e = cudaMalloc(&fd, nb * 64);
    if (e == cudaSuccess && wantLinear) e = cudaMalloc(&ld, nl);
    if (e == cudaSuccess) e = launch_build_fdist(s->vol, s->bdist, s->w, s->h, s->d, fd, ld, a, t1, t2, c->stream, &s->accelLaunches);
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);
    cudaFree(a); cudaFree(t1); cudaFree(t2);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) { cudaFree(fd); cudaFree(ld); return cuda_fail(e); }
    const size_t add = (fd ? nb * 64 : 0) + (ld ? nl : 0);
    s->peakBuildBytes = std::max(s->peakBuildBytes, s->fieldBytes + add + 3 * nv);
    s->fieldBytes += add;
    if (fd) s->fdist = fd;
    if (ld) s->ldist = ld;
    return VRT_OK;
}
// directional safe-step fields (OCTANT): 8 fields of one byte per voxel
static int scene_ensure_octant(vrt_scene* s) {
    if (s->odist[0]) return VRT_OK;
    vrt_context* c = s->ctx;
    const size_t fieldBytes = (size_t)(s->w + 2) * (s->h + 2) * (s->d + 2), scratchBytes = octant_scratch_bytes(s->w, s->h, s->d);
    uint32_t* scratch = nullptr;
    cudaError_t e = cudaMalloc(&scratch, scratchBytes);
    for (int o = 0; o < 8 && e == cudaSuccess; o++) e = cudaMalloc(&s->odist[o], fieldBytes);
    if (e == cudaSuccess) e = s->vol ? launch_odist_init(s->vol, s->w, s->h, s->d, scratch, c->stream, &s->accelLaunches)
                                     : launch_odist_init_sparse(s->vtop, s->vfine, s->w, s->h, s->d, scratch, c->stream, &s->accelLaunches);
    for (int o = 0; o < 8 && e == cudaSuccess; o++) e = launch_odist_octant(s->w, s->h, s->d, o, s->odist[o], scratch, c->stream, &s->accelLaunches);
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);
    cudaFree(scratch);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) {
        for (uint8_t*& p : s->odist) { cudaFree(p); p = nullptr; }
        return cuda_fail(e);
    }
    s->peakBuildBytes = std::max(s->peakBuildBytes, s->fieldBytes + 8 * fieldBytes + scratchBytes);
    s->fieldBytes += 8 * fieldBytes;
    return VRT_OK;
}
// two-level (sparse) directional fields: every octant's linear field is built into ONE scratch buffer in turn and compacted
// (top level + per-voxel bricks near surfaces only); what stays resident is a few per cent of the eight linear fields
static int scene_ensure_sparse(vrt_scene* s) {
    if (s->sparseReady) return VRT_OK;
    vrt_context* c = s->ctx;
    const size_t fieldBytes = (size_t)(s->w + 2) * (s->h + 2) * (s->d + 2), scratchBytes = octant_scratch_bytes(s->w, s->h, s->d);
    sparse_top_dims(s->w, s->h, s->d, s->TB);
    const size_t topWords = (size_t)s->TB[0] * s->TB[1] * s->TB[2];
    uint32_t *scratch = nullptr, *counter = nullptr; uint8_t* lin = nullptr;
    size_t added = 0;
    cudaError_t e = cudaMalloc(&scratch, scratchBytes);
    if (e == cudaSuccess) e = cudaMalloc(&lin, fieldBytes);
    if (e == cudaSuccess) e = cudaMalloc(&counter, sizeof(uint32_t));
    if (e == cudaSuccess) e = s->vol ? launch_odist_init(s->vol, s->w, s->h, s->d, scratch, c->stream, &s->accelLaunches)
                                     : launch_odist_init_sparse(s->vtop, s->vfine, s->w, s->h, s->d, scratch, c->stream, &s->accelLaunches);
    for (int o = 0; o < 8 && e == cudaSuccess; o++) {
        e = launch_odist_octant(s->w, s->h, s->d, o, lin, scratch, c->stream, &s->accelLaunches);
        if (e == cudaSuccess) e = cudaMalloc(&s->stop[o], topWords * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMemsetAsync(counter, 0, sizeof(uint32_t), c->stream);
        if (e == cudaSuccess) e = launch_sparse_classify(lin, s->w, s->h, s->d, s->stop[o], counter, c->stream);
        uint32_t n = 0;
        if (e == cudaSuccess) e = cudaMemcpyAsync(&n, counter, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e == cudaSuccess && n) e = cudaMalloc(&s->sfine[o], (size_t)n * 512u);
        if (e == cudaSuccess && n) e = launch_sparse_fill(lin, s->w, s->h, s->d, s->stop[o], s->sfine[o], c->stream);
        s->sfineBricks[o] = n;
        s->accelLaunches += 2;
        added += topWords * sizeof(uint32_t) + (size_t)n * 512u;
    }
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);
    cudaFree(scratch); cudaFree(lin); cudaFree(counter);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) {
        for (int o = 0; o < 8; o++) { cudaFree(s->stop[o]); cudaFree(s->sfine[o]); s->stop[o] = nullptr; s->sfine[o] = nullptr; s->sfineBricks[o] = 0; }
        return cuda_fail(e);
    }
    s->peakBuildBytes = std::max(s->peakBuildBytes, s->fieldBytes + added + fieldBytes + scratchBytes);
    s->fieldBytes += added;
    s->sparseReady = true;
    return VRT_OK;
}
// the library's choice when the caller says VRT_TRAVERSAL_DEFAULT: the linear directional fields while eight of them fit a budget,
// the two-level layout beyond it (VRT_DEFAULT_TRAVERSAL / VRT_FIELD_BUDGET_GIB override, for measurement)
static uint32_t resolve_traversal(const vrt_scene* s, uint32_t traversal) {
    if (traversal != VRT_TRAVERSAL_DEFAULT) return traversal;
    static const int forced = getenv("VRT_DEFAULT_TRAVERSAL") ? atoi(getenv("VRT_DEFAULT_TRAVERSAL")) : 0;
    if (forced >= VRT_TRAVERSAL_DENSE_U8 && forced <= VRT_TRAVERSAL_SPARSE) return (uint32_t)forced;
    static const double budgetGiB = getenv("VRT_FIELD_BUDGET_GIB") ? atof(getenv("VRT_FIELD_BUDGET_GIB")) : 16.0;
    const double linearGiB = 8.0 * (double)(s->w + 2) * (double)(s->h + 2) * (double)(s->d + 2) / 1073741824.0;
    return linearGiB <= budgetGiB ? (uint32_t)VRT_TRAVERSAL_OCTANT : (uint32_t)VRT_TRAVERSAL_SPARSE;
}

static int scene_ensure(vrt_scene* s, uint32_t traversal) {
    std::lock_guard<std::mutex> lock(s->fieldMutex);
    const double t0 = now_ms();
    int rc = VRT_OK;
    switch (resolve_traversal(s, traversal)) {
        case VRT_TRAVERSAL_DENSE_U8: break;
        case VRT_TRAVERSAL_BRICK: rc = scene_ensure_bricks(s); break;
        case VRT_TRAVERSAL_FINE: rc = scene_ensure_iso(s, true, false); break;
        case VRT_TRAVERSAL_LINEAR: rc = scene_ensure_iso(s, false, true); break;
        case VRT_TRAVERSAL_OCTANT: rc = scene_ensure_octant(s); break;
        case VRT_TRAVERSAL_SPARSE: rc = scene_ensure_sparse(s); break;
        default: rc = fail(VRT_E_INVALID, "unknown traversal");
    }
    s->loadMs += now_ms() - t0;
    return rc;
}
extern "C" int vrt_scene_prepare(vrt_scene* s, uint32_t traversal) {
    if (!s) return fail(VRT_E_INVALID, "null argument");
    CU(cudaSetDevice(s->ctx->device));
    return scene_ensure(s, traversal);
}
extern "C" int vrt_scene_memory(const vrt_scene* s, size_t* volume_bytes, size_t* field_bytes, size_t* peak_build_bytes, double* load_ms) {
    if (!s) return fail(VRT_E_INVALID, "null argument");
    if (volume_bytes) *volume_bytes = s->volumeBytes;
    if (field_bytes) *field_bytes = s->fieldBytes;
    if (peak_build_bytes) *peak_build_bytes = s->peakBuildBytes + s->volumeBytes;
    if (load_ms) *load_ms = s->loadMs;
    return VRT_OK;
}

// palette upload + the default back-end's fields (so that no frame pays for them)
static int scene_finish(vrt_scene* s, const vrt_material palette[256], double t0) {
    vrt_context* c = s->ctx;
    memcpy(s->hostPalette, palette, sizeof(vrt_material) * 256);
    CU(cudaMemcpyAsync(s->palette, palette, sizeof(vrt_material) * 256, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    s->loadMs += now_ms() - t0;
    return scene_ensure(s, VRT_TRAVERSAL_DEFAULT);
}

extern "C" int vrt_scene_from_dense(vrt_context* c, const uint8_t* volume, uint32_t w, uint32_t h, uint32_t d,
                                    const vrt_material palette[256], vrt_scene** out) {
    if (!volume || !palette) return fail(VRT_E_INVALID, "null argument");
    vrt_scene* s = nullptr;
    const double t0 = now_ms();
    int rc = scene_alloc(c, w, h, d, &s);
    if (rc) return rc;
    cudaError_t e = cudaMemcpyAsync(s->vol, volume, (size_t)w * h * d, cudaMemcpyHostToDevice, c->stream);
    if (e != cudaSuccess) { vrt_scene_destroy(s); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    if (volume_wants_sparse(w, h, d)) { rc = scene_make_sparse_volume(s, 0, 0); if (rc) { vrt_scene_destroy(s); return rc; } }
    rc = scene_finish(s, palette, t0);
    if (rc) { vrt_scene_destroy(s); return rc; }
    *out = s;
    return VRT_OK;
}

// VoxelScene::VoxelScene (voxel_scene.cpp:33-133): parse on the host, flatten on the device. Every model is uploaded once and each
// instance is one kernel launch, in the reference's order. A scene with a non-permutation transform (an invalid _r yields a zero row,
// as in the reference's parser) can collide inside one instance, where the reference's sequential loop order decides: such a scene
// is flattened by the host loop and uploaded, like vrt_scene_from_dense.
static int scene_from_layout(vrt_context* c, const VoxLayout& lay, vrt_scene** out) {
    const double t0 = now_ms();
    bool device = true;
    for (const VoxLayout::Instance& in : lay.instances) device = device && in.permutation;
    if (!device) {
        std::vector<uint8_t> vol((size_t)lay.w * lay.h * lay.d);
        flatten_host(lay, vol.data());
        return vrt_scene_from_dense(c, vol.data(), lay.w, lay.h, lay.d, lay.palette, out);
    }
    vrt_scene* s = nullptr;
    int rc = scene_alloc(c, lay.w, lay.h, lay.d, &s);
    if (rc) return rc;
    std::vector<uint8_t*> dmodels(lay.models.size(), nullptr);
    cudaError_t e = cudaMemsetAsync(s->vol, 0, (size_t)lay.w * lay.h * lay.d, c->stream);
    for (size_t i = 0; i < lay.instances.size() && e == cudaSuccess; i++) {
        const uint32_t mi = lay.instances[i].model;
        const VoxLayout::Model& m = lay.models[mi];
        if (m.cells.empty()) continue;
        if (!dmodels[mi]) {
            e = cudaMalloc(&dmodels[mi], m.cells.size());
            if (e == cudaSuccess) e = cudaMemcpyAsync(dmodels[mi], m.cells.data(), m.cells.size(), cudaMemcpyHostToDevice, c->stream);
        }
        if (e == cudaSuccess) e = launch_flatten_instance(dmodels[mi], s->vol, flatten_instance_params(lay, i), c->stream);
        if (e == cudaSuccess) s->accelLaunches++;
    }
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);   // lay's host buffers are pageable: the copies must be done before they go away
    for (uint8_t* p : dmodels) cudaFree(p);
    if (e != cudaSuccess || e2 != cudaSuccess) { vrt_scene_destroy(s); return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e != cudaSuccess ? e : e2)); }
    if (volume_wants_sparse(lay.w, lay.h, lay.d)) { rc = scene_make_sparse_volume(s, 0, 0); if (rc) { vrt_scene_destroy(s); return rc; } }
    rc = scene_finish(s, lay.palette, t0);
    if (rc) { vrt_scene_destroy(s); return rc; }
    *out = s;
    return VRT_OK;
}

// Nothing throws across the ABI: host containers sized from the file can (bad_alloc, length_error).
#define VRT_NOTHROW(body) try { body } catch (const std::bad_alloc&) { return fail(VRT_E_OOM, "out of host memory"); } catch (const std::exception& ex) { return fail(VRT_E_PARSE, ex.what()); }
extern "C" int vrt_scene_load_vox_memory(vrt_context* c, const uint8_t* bytes, size_t n, vrt_scene** out) {
    if (!c || !out) return fail(VRT_E_INVALID, "null argument");
    VRT_NOTHROW(
        VoxLayout lay; std::string err;
        int rc = parse_vox_bytes(bytes, n, lay, err);
        if (rc) return fail(rc, err);
        return scene_from_layout(c, lay, out);
    )
}
extern "C" int vrt_scene_load_vox(vrt_context* c, const char* path, vrt_scene** out) {
    if (!c || !out) return fail(VRT_E_INVALID, "null argument");
    VRT_NOTHROW(
        VoxLayout lay; std::string err;
        int rc = parse_vox_file(path, lay, err);
        if (rc) return fail(rc, err);
        return scene_from_layout(c, lay, out);
    )
}

extern "C" int vrt_scene_generate(vrt_context* c, int generator, uint64_t seed, uint32_t w, uint32_t h, uint32_t d, vrt_scene** out) {
    if (generator != 1 && generator != 2) return fail(VRT_E_INVALID, "unknown generator");
    vrt_scene* s = nullptr;
    const double t0 = now_ms();
    const bool sparse = volume_wants_sparse(w, h, d);
    int rc = scene_alloc(c, w, h, d, &s, !sparse);
    if (rc) return rc;
    if (sparse) {                                                   // brick by brick from the generator: no dense volume, ever
        rc = scene_make_sparse_volume(s, generator, (uint32_t)seed);
        if (rc) { vrt_scene_destroy(s); return rc; }
    } else {
        cudaError_t e = launch_generate(generator, (uint32_t)seed, s->vol, w, h, d, c->stream);
        if (e != cudaSuccess) { vrt_scene_destroy(s); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    }
    vrt_material pal[256];
    generator_palette(generator, (uint32_t)seed, pal);
    rc = scene_finish(s, pal, t0);
    if (rc) { vrt_scene_destroy(s); return rc; }
    *out = s;
    return VRT_OK;
}

extern "C" int vrt_scene_set_noise(vrt_scene* s, const uint8_t* rgba8, uint32_t w, uint32_t h) {
    if (!s || !rgba8) return fail(VRT_E_INVALID, "null argument");
    if (w != 512 || h != 512) return fail(VRT_E_INVALID, "blue-noise texture must be 512x512 (NOISE_SIZE, voxel_volume.frag:70)");
    CU(cudaSetDevice(s->ctx->device));
    if (!s->noise) CU(cudaMalloc(&s->noise, 512 * 512 * 4));
    CU(cudaMemcpy(s->noise, rgba8, 512 * 512 * 4, cudaMemcpyHostToDevice));
    return VRT_OK;
}
extern "C" int vrt_scene_set_skybox(vrt_scene* s, const float* rgba32f, uint32_t w, uint32_t h) {
    if (!s || !rgba32f || !w || !h) return fail(VRT_E_INVALID, "null argument");
    CU(cudaSetDevice(s->ctx->device));
    if (s->sky) { cudaFree(s->sky); s->sky = nullptr; }
    CU(cudaMalloc(&s->sky, (size_t)w * h * 16));
    CU(cudaMemcpy(s->sky, rgba32f, (size_t)w * h * 16, cudaMemcpyHostToDevice));
    s->skyW = w; s->skyH = h;
    return VRT_OK;
}
extern "C" int vrt_scene_bounds(const vrt_scene* s, uint32_t whd[3]) {
    if (!s || !whd) return fail(VRT_E_INVALID, "null argument");
    whd[0] = s->w; whd[1] = s->h; whd[2] = s->d;
    return VRT_OK;
}
extern "C" int vrt_scene_palette(const vrt_scene* s, vrt_material palette[256]) {
    if (!s || !palette) return fail(VRT_E_INVALID, "null argument");
    memcpy(palette, s->hostPalette, sizeof(vrt_material) * 256);
    return VRT_OK;
}
extern "C" int vrt_scene_download_volume(const vrt_scene* s, uint8_t* dst, size_t bytes) {
    if (!s || !dst || bytes != (size_t)s->w * s->h * s->d) return fail(VRT_E_INVALID, "bad size");
    CU(cudaSetDevice(s->ctx->device));
    if (s->vol) { CU(cudaMemcpy(dst, s->vol, bytes, cudaMemcpyDeviceToHost)); return VRT_OK; }
    // two-level volume: expanded slab by slab through a bounded device buffer (tests only)
    const size_t slice = (size_t)s->w * s->h;
    const uint32_t slabD = (uint32_t)std::max<size_t>(1, std::min<size_t>(s->d, ((size_t)256 << 20) / slice));
    uint8_t* tmp = nullptr;
    CU(cudaMalloc(&tmp, slice * slabD));
    cudaError_t e = cudaSuccess;
    for (uint32_t z0 = 0; z0 < s->d && e == cudaSuccess; z0 += slabD) {
        const uint32_t nz = std::min(slabD, s->d - z0);
        e = launch_sparse_volume_expand(s->vtop, s->vfine, s->w, s->h, s->d, z0, nz, tmp, s->ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dst + slice * z0, tmp, slice * nz, cudaMemcpyDeviceToHost, s->ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->ctx->stream);
    }
    cudaFree(tmp);
    if (e != cudaSuccess) return cuda_fail(e);
    return VRT_OK;
}
extern "C" int vrt_scene_accel_info(const vrt_scene* s, size_t* dense_bytes, size_t* brick_mask_bytes, size_t* brick_dist_bytes) {
    if (!s) return fail(VRT_E_INVALID, "null argument");
    const size_t nb = (size_t)s->Bx * s->By * s->Bz;
    if (dense_bytes) *dense_bytes = (size_t)s->w * s->h * s->d;
    if (brick_mask_bytes) *brick_mask_bytes = s->bmask ? nb * 8 : 0;
    if (brick_dist_bytes) *brick_dist_bytes = (s->bdist ? nb : 0) + (s->fdist ? nb * 64 : 0) + (s->ldist ? (size_t)(s->w + 2) * (s->h + 2) * (s->d + 2) : 0);
    return VRT_OK;
}

extern "C" int vrt_scene_download_accel(const vrt_scene* s, uint64_t* bmask_dst, uint8_t* bdist_dst) {
    if (!s) return fail(VRT_E_INVALID, "null argument");
    const size_t nb = (size_t)s->Bx * s->By * s->Bz;
    CU(cudaSetDevice(s->ctx->device));
    { const int rc = scene_ensure(const_cast<vrt_scene*>(s), VRT_TRAVERSAL_BRICK); if (rc) return rc; }
    if (bmask_dst) CU(cudaMemcpy(bmask_dst, s->bmask, nb * 8, cudaMemcpyDeviceToHost));
    if (bdist_dst) CU(cudaMemcpy(bdist_dst, s->bdist, nb, cudaMemcpyDeviceToHost));
    return VRT_OK;
}

extern "C" int vrt_scene_download_fdist(const vrt_scene* s, uint8_t* dst, size_t bytes) {
    if (!s || !dst || bytes != (size_t)s->Bx * s->By * s->Bz * 64u) return fail(VRT_E_INVALID, "bad size");
    CU(cudaSetDevice(s->ctx->device));
    { const int rc = scene_ensure(const_cast<vrt_scene*>(s), VRT_TRAVERSAL_FINE); if (rc) return rc; }
    CU(cudaMemcpy(dst, s->fdist, bytes, cudaMemcpyDeviceToHost));
    return VRT_OK;
}

extern "C" int vrt_scene_download_odist(const vrt_scene* s, int octant, uint8_t* dst, size_t bytes) {
    if (!s || !dst || octant < 0 || octant > 7 || bytes != (size_t)(s->w + 2) * (s->h + 2) * (s->d + 2)) return fail(VRT_E_INVALID, "bad argument");
    CU(cudaSetDevice(s->ctx->device));
    const int rc = scene_ensure(const_cast<vrt_scene*>(s), VRT_TRAVERSAL_OCTANT);
    if (rc) return rc;
    CU(cudaMemcpy(dst, s->odist[octant], bytes, cudaMemcpyDeviceToHost));
    return VRT_OK;
}

extern "C" int vrt_scene_download_sdist(const vrt_scene* s, int octant, uint8_t* dst, size_t bytes) {
    if (!s || !dst || octant < 0 || octant > 7 || bytes != (size_t)(s->w + 2) * (s->h + 2) * (s->d + 2)) return fail(VRT_E_INVALID, "bad argument");
    CU(cudaSetDevice(s->ctx->device));
    const int rc = scene_ensure(const_cast<vrt_scene*>(s), VRT_TRAVERSAL_SPARSE);
    if (rc) return rc;
    uint8_t* tmp = nullptr;
    CU(cudaMalloc(&tmp, bytes));
    cudaError_t e = launch_sparse_expand(s->stop[octant], s->sfine[octant], s->w, s->h, s->d, tmp, s->ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dst, tmp, bytes, cudaMemcpyDeviceToHost, s->ctx->stream);
    const cudaError_t e2 = cudaStreamSynchronize(s->ctx->stream);
    cudaFree(tmp);
    if (e != cudaSuccess || e2 != cudaSuccess) return cuda_fail(e != cudaSuccess ? e : e2);
    return VRT_OK;
}

// ------------------------------------------------------------------------------------------------ frames
extern "C" void vrt_frame_release(vrt_frame* f) {
    if (!f) return;
    cudaSetDevice(f->ctx->device);
    for (void* p : f->planes) cudaFree(p);
    cudaFree(f->ping[0]); cudaFree(f->ping[1]); cudaFree(f->ctr);
    cudaFree(f->wb.rec); cudaFree(f->wb.bq[0]); cudaFree(f->wb.bq[1]); cudaFree(f->wb.sq); cudaFree(f->wb.res); cudaFree(f->wb.ctr);
    if (f->e0) cudaEventDestroy(f->e0);
    if (f->e1) cudaEventDestroy(f->e1);
    if (f->e2) cudaEventDestroy(f->e2);
    for (cudaEvent_t e : f->es) if (e) cudaEventDestroy(e);
    if (f->copyPending) cudaEventSynchronize(f->copied);
    if (f->rendered) cudaEventDestroy(f->rendered);
    if (f->copied) cudaEventDestroy(f->copied);
    if (f->own) {
        if (f->stream) { cudaStreamSynchronize(f->stream); cudaStreamDestroy(f->stream); }
        if (f->fork) cudaEventDestroy(f->fork);
        if (f->joined) cudaEventDestroy(f->joined);
    }
    if (f->ctx) { auto& v = f->ctx->frames; for (size_t i = 0; i < v.size(); i++) if (v[i] == f) { v.erase(v.begin() + i); break; } }
    delete f;
}

static int frame_ensure_plane(vrt_frame* f, int plane) {
    if (f->planes[plane]) return VRT_OK;
    CU(cudaMalloc(&f->planes[plane], f->pixels * kPlaneBpp[plane]));
    CU(cudaMemsetAsync(f->planes[plane], 0, f->pixels * kPlaneBpp[plane], f->stream));
    return VRT_OK;
}

// Wavefront buffers, sized for the worst case (every pixel a hit with a full 5-bounce chain).
static int frame_ensure_wave(vrt_frame* f, uint32_t raysPerRecord) {
    WaveBuf& wb = f->wb;
    if (!wb.ctr) CU(cudaMalloc(&wb.ctr, sizeof(WaveCtr)));
    if (raysPerRecord == 0) return VRT_OK;      // no shading: the primary kernel allocates no records
    if (f->pixels * 6 >= 0xffffffffull) return fail(VRT_E_UNSUPPORTED, "frame too large for 32-bit record indices");
    if (!wb.rec) {
        wb.pixCap = (uint32_t)f->pixels; wb.recCap = 6u * wb.pixCap;
        cudaError_t e = cudaMalloc(&wb.rec, (size_t)wb.recCap * 3 * sizeof(RecQ));
        if (e == cudaSuccess) e = cudaMalloc(&wb.bq[0], (size_t)wb.pixCap * sizeof(BounceEnt));
        if (e == cudaSuccess) e = cudaMalloc(&wb.bq[1], (size_t)wb.pixCap * sizeof(BounceEnt));
        if (e == cudaSuccess) e = cudaMalloc(&wb.sq, (size_t)wb.pixCap * 5 * sizeof(uint32_t));
        if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e));
    }
    if (raysPerRecord > f->wbRays) {
        cudaFree(wb.res); wb.res = nullptr; f->wbRays = 0;
        cudaError_t e = cudaMalloc(&wb.res, (size_t)wb.recCap * raysPerRecord);
        if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e));
        f->wbRays = raysPerRecord;
    }
    return VRT_OK;
}

static int frame_create(vrt_context* c, int W, int H, vrt_frame** out) {
    vrt_frame* f = new (std::nothrow) vrt_frame();
    if (!f) return fail(VRT_E_OOM, "host allocation failed");
    f->ctx = c; f->W = W; f->H = H;
    f->tilesX = (W + VRT_TILE - 1) / VRT_TILE; f->tilesY = (H + VRT_TILE - 1) / VRT_TILE;
    const int tiles = f->tilesX * f->tilesY;
    f->localTiles = tiles > c->rank ? (tiles - c->rank + c->world - 1) / c->world : 0;
    // sharded frames are padded to the same tile count on every rank so the gather is one equal-size collective
    const int maxLocal = (tiles + c->world - 1) / c->world;
    f->pixels = c->world == 1 ? (size_t)W * H : (size_t)maxLocal * VRT_TILE * VRT_TILE;
    cudaError_t e = cudaMalloc(&f->ctr, sizeof(Counters));
    if (e == cudaSuccess) e = cudaEventCreate(&f->e0);
    if (e == cudaSuccess) e = cudaEventCreate(&f->e1);
    if (e == cudaSuccess) e = cudaEventCreate(&f->e2);
    for (int k = 0; k < 3; k++) if (e == cudaSuccess) e = cudaEventCreate(&f->es[k]);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->rendered, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->copied, cudaEventDisableTiming);
    if (c->concurrent) {
        f->own = true;
        int prLow = 0, prHigh = 0;
        if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prLow, &prHigh);
        if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&f->stream, cudaStreamNonBlocking, prHigh);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->fork, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->joined, cudaEventDisableTiming);
    } else { f->stream = c->stream; f->side = c->side; }
    if (e != cudaSuccess) { vrt_frame_release(f); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    c->frames.push_back(f);
    *out = f;
    return VRT_OK;
}

// Host tables of DenoiserStage (denoiser_stage.cpp:37-61) as seen by the shader under std140.
static void denoise_taps(int mode, DenoiseTaps& t) {
    float w[9]; float o[18]; int i = 0;
    for (int y = -1; y <= 1; y++)
        for (int x = -1; x <= 1; x++, i++) {
            o[2 * i] = (float)x; o[2 * i + 1] = (float)y;
            const float sx = ((float)x * (float)x) / (2.0f * 2.0f * 2.0f), sy = ((float)y * (float)y) / (2.0f * 2.0f * 2.0f);
            w[i] = expf(-(sx + sy));   // glm::gauss(vec2(x,y), vec2(0), vec2(2))
        }
    t.n = 0;
    for (int k = 0; k < 25; k++) {
        float kw, ox, oy;
        if (mode == VRT_DENOISE_AS_INTENDED) { if (k >= 9) break; kw = w[k]; ox = o[2 * k]; oy = o[2 * k + 1]; }
        else { const int fi = 4 * k; kw = fi < 9 ? w[fi] : 0.0f; ox = fi < 18 ? o[fi] : 0.0f; oy = fi + 1 < 18 ? o[fi + 1] : 0.0f; }
        if (kw == 0.0f) continue;   // zero-weight taps add +0 to sum and total: elided
        t.ox[t.n] = (int)ox; t.oy[t.n] = (int)oy; t.k[t.n] = kw; t.n++;
    }
}

int run_denoise(vrt_context* c, cudaStream_t stream, const void* color, const void* normal, const void* position, int W, int H,
                       const vrt_denoiser_settings* ds, void* ping[2], int* lastOut, uint32_t* launches) {
    if (ds->iterations < 1 || ds->iterations > 10) return fail(VRT_E_INVALID, "denoiser iterations must be 1..10 (MAX_DENOISER_PASSES)");
    if (ds->mode != VRT_DENOISE_AS_EXECUTED && ds->mode != VRT_DENOISE_AS_INTENDED) return fail(VRT_E_INVALID, "bad denoise mode");
    DenoiseTaps taps; denoise_taps(ds->mode, taps);
    const void* src = color;
    int last = 0;
    for (int i = 0; i < ds->iterations; i++) {   // denoiser_stage.cpp:143-154 and :203-257
        DenoiseIter it;
        it.phiColor = 1.0f / (float)i * ds->phiColor0;
        it.phiNormal = 1.0f / (float)i * ds->phiNormal0;
        it.phiPos = 1.0f / (float)i * ds->phiPos0;
        it.stepWidth = (float)i * ds->stepWidth + 1.0f;
        const int tgt = i % 2;
        cudaError_t e = launch_denoise((const uint8_t*)src, (const int8_t*)normal, (const float*)position, (uint8_t*)ping[tgt], W, H, taps, it, stream);
        if (e != cudaSuccess) return fail(VRT_E_CUDA, cudaGetErrorString(e));
        if (launches) (*launches)++;
        src = ping[tgt]; last = tgt;
    }
    *lastOut = last;
    return VRT_OK;
}

extern "C" int vrt_render(vrt_context* c, const vrt_scene* s, const vrt_camera* cam, const vrt_render_settings* st, vrt_frame** frame) {
    return vrt_render_impl(c, s, cam, st, frame, nullptr);
}
// x: planes of the assembled frame the pass's kernels store into as well (vrt_render_exchange), or null
int vrt_render_impl(vrt_context* c, const vrt_scene* s, const vrt_camera* cam, const vrt_render_settings* st, vrt_frame** frame, const XPlanes* x) {
    if (!c || !s || !cam || !st || !frame) return fail(VRT_E_INVALID, "null argument");
    if (s->ctx != c) return fail(VRT_E_INVALID, "scene belongs to another context");
    const int W = cam->screenSize[0], H = cam->screenSize[1];
    if (W <= 0 || H <= 0 || W > 32768 || H > 32768) return fail(VRT_E_INVALID, "bad screenSize");
    if (cam->volumeBounds[0] != s->w || cam->volumeBounds[1] != s->h || cam->volumeBounds[2] != s->d)
        return fail(VRT_E_INVALID, "camera volumeBounds differ from the scene bounds");
    if (!(st->pass_mask & VRT_PASS_PRIMARY)) return fail(VRT_E_INVALID, "pass_mask must include VRT_PASS_PRIMARY");
    if (st->traversal > VRT_TRAVERSAL_SPARSE) return fail(VRT_E_INVALID, "unknown traversal");
    if (st->pipeline > VRT_PIPELINE_WAVEFRONT) return fail(VRT_E_INVALID, "unknown pipeline");
    if (st->vol.aoSamples > 255) return fail(VRT_E_INVALID, "aoSamples must be <= 255 (the reference GUI allows 0..16)");
    const bool wavefront = st->pipeline != VRT_PIPELINE_MEGAKERNEL;
    if (st->denoise.enable && c->world != 1) return fail(VRT_E_UNSUPPORTED, "denoise runs on the assembled frame: exchange colour, normal and position (vrt_frame_exchange) and call vrt_denoise_on on rank 0");
    CU(cudaSetDevice(c->device));
    vrt_frame* f = *frame;
    if (f && (f->W != W || f->H != H || f->ctx != c)) return fail(VRT_E_INVALID, "frame does not match screenSize/context");
    if (!f) { int rc = frame_create(c, W, H, &f); if (rc) return rc; }
    uint32_t out_mask = st->out_mask;
    if (st->denoise.enable) out_mask |= VRT_OUT_GBUFFER;
    int rc = frame_ensure_plane(f, VRT_PLANE_COLOR_RGBA8);
    if (!rc && (out_mask & VRT_OUT_GBUFFER))
        for (int p : {VRT_PLANE_DEPTH_R32F, VRT_PLANE_MOTION_RG32F, VRT_PLANE_MASK_R8, VRT_PLANE_POSITION_RGBA32F, VRT_PLANE_NORMAL_RGBA8S}) if (!rc) rc = frame_ensure_plane(f, p);
    if (!rc && (out_mask & VRT_OUT_INSTRUMENT))
        for (int p : {VRT_PLANE_HIT_XYZS_I16X4, VRT_PLANE_HIT_PALETTE_U8, VRT_PLANE_COLOR_RGB32F, VRT_PLANE_RAYS_U32, VRT_PLANE_FETCHES_U32}) if (!rc) rc = frame_ensure_plane(f, p);
    if (!rc && st->denoise.enable)
        for (int k = 0; k < 2 && !rc; k++) if (!f->ping[k]) { cudaError_t e = cudaMalloc(&f->ping[k], f->pixels * 4); if (e != cudaSuccess) rc = fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    if (!rc && wavefront)
        rc = frame_ensure_wave(f, (st->pass_mask & VRT_PASS_SHADE) ? ((st->pass_mask & VRT_PASS_AO) ? st->vol.aoSamples : 0u) + 1u : 0u);
    if (rc) { if (!*frame) vrt_frame_release(f); return rc; }

    const uint32_t traversal = resolve_traversal(s, st->traversal);
    rc = scene_ensure(const_cast<vrt_scene*>(s), traversal);   // a no-op once the back-end's fields exist (vrt_scene_prepare / load time)
    if (rc) { if (!*frame) vrt_frame_release(f); return rc; }
    SceneDev sc{};
    for (int o = 0; o < 8; o++) { sc.odist[o] = s->odist[o]; sc.sdist[o].top = s->stop[o]; sc.sdist[o].fine = s->sfine[o]; }
    sc.TBx = s->TB[0]; sc.TBy = s->TB[1]; sc.TBz = s->TB[2];
    sc.svol.top = s->vtop; sc.svol.fine = s->vfine;
    sc.vol = s->vol; sc.bmask = s->bmask; sc.bdist = s->bdist; sc.fdist = s->fdist; sc.ldist = s->ldist; sc.W = s->w; sc.H = s->h; sc.D = s->d; sc.Bx = s->Bx; sc.By = s->By; sc.Bz = s->Bz;
    sc.palette = s->palette; sc.noise = s->noise; sc.sky = s->sky; sc.skyW = s->skyW; sc.skyH = s->skyH;
    FrameParams fp{};
    if (x) fp.x = *x;
    fp.cam = *cam; fp.pass_mask = st->pass_mask; fp.out_mask = out_mask;
    fp.traversal = traversal;
    fp.aoSamples = st->vol.aoSamples; fp.ambientIntensity = st->vol.ambientIntensity;
    for (int k = 0; k < 3; k++) { fp.lightDir[k] = st->light.direction[k]; fp.lightColor[k] = st->light.color[k]; }
    fp.lightIntensity = st->light.intensity;
    fp.rank = c->rank; fp.world = c->world; fp.tilesX = f->tilesX; fp.tilesY = f->tilesY; fp.localTiles = f->localTiles;
    PlanesDev pl{};
    pl.color = (uint8_t*)f->planes[VRT_PLANE_COLOR_RGBA8]; pl.depth = (float*)f->planes[VRT_PLANE_DEPTH_R32F];
    pl.motion = (float*)f->planes[VRT_PLANE_MOTION_RG32F]; pl.mask = (uint8_t*)f->planes[VRT_PLANE_MASK_R8];
    pl.position = (float*)f->planes[VRT_PLANE_POSITION_RGBA32F]; pl.normal = (int8_t*)f->planes[VRT_PLANE_NORMAL_RGBA8S];
    pl.hit = (int16_t*)f->planes[VRT_PLANE_HIT_XYZS_I16X4]; pl.hit_palette = (uint8_t*)f->planes[VRT_PLANE_HIT_PALETTE_U8];
    pl.color_f32 = (float*)f->planes[VRT_PLANE_COLOR_RGB32F]; pl.rays = (uint32_t*)f->planes[VRT_PLANE_RAYS_U32];
    pl.fetches = (uint32_t*)f->planes[VRT_PLANE_FETCHES_U32];

    const cudaStream_t fs = f->stream;
    if (f->own) { CU(cudaEventRecord(f->fork, c->stream)); CU(cudaStreamWaitEvent(fs, f->fork, 0)); }
    if (f->copyPending) { CU(cudaStreamWaitEvent(fs, f->copied, 0)); f->copyPending = false; }   // a read-back of the previous content
    f->launches = 0; f->out_mask = out_mask; f->denoised = false;
    CU(cudaMemsetAsync(f->ctr, 0, sizeof(Counters), fs));
    CU(cudaEventRecord(f->e0, fs));
    cudaError_t e;
    f->staged = wavefront && f->localTiles > 0;
    // frames on their own streams fill each other's tails: the any-hit side stream (launch_wave_t) is for a frame that runs alone
    if (wavefront) e = launch_wavefront(sc, fp, pl, f->ctr, f->wb, c->prop.multiProcessorCount, fs, f->own ? nullptr : &f->side, &f->launches, f->es);
    else { e = launch_render(sc, fp, pl, f->ctr, fs); f->launches++; }
    if (e != cudaSuccess) { if (!*frame) vrt_frame_release(f); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    CU(cudaEventRecord(f->e1, fs));
    if (st->denoise.enable) {
        int last = 0;
        rc = run_denoise(c, fs, pl.color, pl.normal, pl.position, W, H, &st->denoise, f->ping, &last, &f->launches);
        if (rc) { if (!*frame) vrt_frame_release(f); return rc; }
        // expose the final ping-pong target as the DENOISED plane without a copy
        if (last == 1) { void* t = f->ping[0]; f->ping[0] = f->ping[1]; f->ping[1] = t; }
        f->denoised = true;
    }
    CU(cudaEventRecord(f->e2, fs));
    *frame = f;
    return VRT_OK;
}

extern "C" int vrt_frame_sync(vrt_frame* f) {
    if (!f) return fail(VRT_E_INVALID, "null frame");
    CU(cudaSetDevice(f->ctx->device));
    CU(cudaStreamSynchronize(f->stream));
    return VRT_OK;
}
extern "C" void* vrt_frame_stream(vrt_frame* f) { return f ? (void*)f->stream : nullptr; }
extern "C" int vrt_frame_plane(vrt_frame* f, int plane, void** device_ptr, size_t* bytes) {
    if (!f || plane < 0 || plane >= VRT_PLANE_COUNT) return fail(VRT_E_INVALID, "bad plane");
    void* p = plane == VRT_PLANE_DENOISED_RGBA8 ? (f->denoised ? f->ping[0] : nullptr) : f->planes[plane];
    if (!p) return fail(VRT_E_INVALID, "plane was not produced for this frame (check out_mask / denoise.enable)");
    if (device_ptr) *device_ptr = p;
    if (bytes) *bytes = f->pixels * kPlaneBpp[plane];
    return VRT_OK;
}
extern "C" int vrt_frame_download(vrt_frame* f, int plane, void* host_dst, size_t bytes) {
    void* p = nullptr; size_t n = 0;
    int rc = vrt_frame_plane(f, plane, &p, &n);
    if (rc) return rc;
    if (!host_dst || bytes != n) return fail(VRT_E_INVALID, "size mismatch");
    CU(cudaSetDevice(f->ctx->device));
    CU(cudaMemcpyAsync(host_dst, p, n, cudaMemcpyDeviceToHost, f->stream));
    CU(cudaStreamSynchronize(f->stream));
    return VRT_OK;
}
extern "C" int vrt_frame_download_async(vrt_frame* f, int plane, void* host_dst, size_t bytes) {
    void* p = nullptr; size_t n = 0;
    int rc = vrt_frame_plane(f, plane, &p, &n);
    if (rc) return rc;
    if (!host_dst || bytes != n) return fail(VRT_E_INVALID, "size mismatch");
    vrt_context* c = f->ctx;
    CU(cudaSetDevice(c->device));
    CU(cudaEventRecord(f->rendered, f->stream));
    CU(cudaStreamWaitEvent(c->copyStream, f->rendered, 0));
    CU(cudaMemcpyAsync(host_dst, p, n, cudaMemcpyDeviceToHost, c->copyStream));
    CU(cudaEventRecord(f->copied, c->copyStream));
    f->copyPending = true;
    return VRT_OK;
}
extern "C" int vrt_frame_download_wait(vrt_frame* f) {
    if (!f) return fail(VRT_E_INVALID, "null frame");
    if (!f->copyPending) return VRT_OK;
    CU(cudaSetDevice(f->ctx->device));
    CU(cudaEventSynchronize(f->copied));
    return VRT_OK;
}
extern "C" int vrt_frame_stats_get(vrt_frame* f, vrt_frame_stats* out) {
    if (!f || !out) return fail(VRT_E_INVALID, "null argument");
    CU(cudaSetDevice(f->ctx->device));
    CU(cudaStreamSynchronize(f->stream));
    Counters h{};
    CU(cudaMemcpy(&h, f->ctr, sizeof(h), cudaMemcpyDeviceToHost));
    memset(out, 0, sizeof(*out));
    out->rays = h.rays; out->fetches = h.fetches; out->pixels = h.pixels;
    CU(cudaEventElapsedTime(&out->ms_trace, f->e0, f->e1));
    CU(cudaEventElapsedTime(&out->ms_denoise, f->e1, f->e2));
    CU(cudaEventElapsedTime(&out->ms_total, f->e0, f->e2));
    out->kernel_launches = f->launches;
    for (int k = 0; k < 3; k++) { out->stage_rays[k] = h.stageRays[k]; out->stage_fetches[k] = h.stageFetches[k]; }
#ifdef VRT_ENGINE_STATS
    if (getenv("VRT_DEBUG_STATS")) {
        fprintf(stderr, "engine stats:");
        for (int k = 0; k < 24; k++) fprintf(stderr, " %llu", h.dbg[k]);
        fprintf(stderr, "\n");
    }
#endif
    if (f->staged) {
        CU(cudaEventElapsedTime(&out->ms_stage[0], f->e0, f->es[0]));
        CU(cudaEventElapsedTime(&out->ms_stage[1], f->es[0], f->es[1]));
        CU(cudaEventElapsedTime(&out->ms_stage[2], f->es[1], f->es[2]));
        CU(cudaEventElapsedTime(&out->ms_stage[3], f->es[2], f->e1));
    } else out->ms_stage[0] = out->ms_trace;
    return VRT_OK;
}

extern "C" int vrt_denoise(vrt_context* c, const void* d_color, const void* d_normal, const void* d_position, int w, int h,
                           const vrt_denoiser_settings* ds, void* d_out) {
    if (!c || !d_color || !d_normal || !d_position || !ds || !d_out || w <= 0 || h <= 0) return fail(VRT_E_INVALID, "null argument");
    if (d_out == d_color) return fail(VRT_E_INVALID, "d_out must not alias d_color (iteration 0 reads neighbours while it writes)");
    CU(cudaSetDevice(c->device));
    void* ping[2] = {d_out, nullptr};
    if (ds->iterations > 1) CU(cudaMalloc(&ping[1], (size_t)w * h * 4));
    // make the last iteration land in d_out: iteration i writes ping[i%2]; swap if the last index is odd
    if (((ds->iterations - 1) % 2) == 1) { void* t = ping[0]; ping[0] = ping[1]; ping[1] = t; }
    int last = 0; uint32_t n = 0;
    int rc = run_denoise(c, c->stream, d_color, d_normal, d_position, w, h, ds, ping, &last, &n);
    cudaError_t e = cudaStreamSynchronize(c->stream);
    void* scratch = ping[0] == d_out ? ping[1] : ping[0];
    if (scratch) cudaFree(scratch);
    if (rc) return rc;
    if (e != cudaSuccess) return fail(VRT_E_CUDA, cudaGetErrorString(e));
    return VRT_OK;
}

extern "C" int vrt_denoise_on(vrt_context* c, const void* d_color, const void* d_normal, const void* d_position, int w, int h,
                              const vrt_denoiser_settings* ds, void* d_out, void* d_scratch, void* stream) {
    if (!c || !d_color || !d_normal || !d_position || !ds || !d_out || w <= 0 || h <= 0) return fail(VRT_E_INVALID, "null argument");
    if (d_out == d_color || d_scratch == d_color || d_out == d_scratch) return fail(VRT_E_INVALID, "d_out / d_scratch must not alias d_color or each other");
    if (ds->iterations > 1 && !d_scratch) return fail(VRT_E_INVALID, "more than one iteration needs a scratch image");
    CU(cudaSetDevice(c->device));
    void* ping[2] = {d_out, d_scratch};
    if (((ds->iterations - 1) % 2) == 1) { ping[0] = d_scratch; ping[1] = d_out; }   // the last iteration lands in d_out
    int last = 0; uint32_t n = 0;
    return run_denoise(c, stream ? (cudaStream_t)stream : c->stream, d_color, d_normal, d_position, w, h, ds, ping, &last, &n);
}

// BlitStage::record (blit_stage.cpp:25-57): source image -> "window" buffer through blit.frag. Asynchronous on the given stream.
extern "C" int vrt_blit_on(vrt_context* c, const void* d_src, int sw, int sh, void* d_dst, int tw, int th, uint32_t flags, void* stream) {
    if (!c || !d_src || !d_dst || sw <= 0 || sh <= 0 || tw <= 0 || th <= 0 || (flags & ~(uint32_t)(VRT_BLIT_SRGB | VRT_BLIT_BGRA)))
        return fail(VRT_E_INVALID, "bad argument");
    CU(cudaSetDevice(c->device));
    cudaError_t e = launch_blit(d_src, sw, sh, d_dst, tw, th, flags, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(VRT_E_CUDA, cudaGetErrorString(e));
    return VRT_OK;
}
extern "C" int vrt_blit(vrt_context* c, const void* d_src, int sw, int sh, void* d_dst, int tw, int th, uint32_t flags) {
    return vrt_blit_on(c, d_src, sw, sh, d_dst, tw, th, flags, c ? (void*)c->stream : nullptr);
}

// Assemble a sharded plane on the gathering rank: `gathered` holds world shard buffers back to back (rank-major),
// each of vrt_frame_plane() size; writes a row-major W x H plane.
extern "C" int vrt_assemble_on(vrt_context* c, const void* d_gathered, int world, int w, int h, int plane, void* d_out, void* stream);
extern "C" int vrt_assemble(vrt_context* c, const void* d_gathered, int world, int w, int h, int plane, void* d_out) {
    return vrt_assemble_on(c, d_gathered, world, w, h, plane, d_out, c ? (void*)c->stream : nullptr);
}
// the same on a stream of the caller's (e.g. the one its gather runs on, so that the context stream can trace the next frame)
extern "C" int vrt_assemble_on(vrt_context* c, const void* d_gathered, int world, int w, int h, int plane, void* d_out, void* stream) {
    if (!c || !d_gathered || !d_out || world < 1 || plane < 0 || plane >= VRT_PLANE_COUNT) return fail(VRT_E_INVALID, "bad argument");
    CU(cudaSetDevice(c->device));
    const int tiles = ((w + VRT_TILE - 1) / VRT_TILE) * ((h + VRT_TILE - 1) / VRT_TILE);
    const size_t stride = (size_t)((tiles + world - 1) / world) * VRT_TILE * VRT_TILE;
    cudaError_t e = launch_detile(d_gathered, d_out, w, h, world, stride, kPlaneBpp[plane], (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(VRT_E_CUDA, cudaGetErrorString(e));
    return VRT_OK;
}

// Multi-GPU assembly without a collective: this rank's tiles of `plane` go straight into a row-major w x h frame at d_frame, which
// may be PEER memory (another GPU's buffer mapped into this process: CUDA IPC / symmetric memory over NVLink). Runs on `stream`
// (null: the frame's stream); the caller orders it against the consumer (a device-side barrier or signal).
extern "C" int vrt_frame_push_tiles(vrt_frame* f, int plane, void* d_frame, void* stream) {
    if (!f || !d_frame || plane < 0 || plane >= VRT_PLANE_COUNT || plane == VRT_PLANE_DENOISED_RGBA8) return fail(VRT_E_INVALID, "bad argument");
    if (!f->planes[plane]) return fail(VRT_E_INVALID, "plane was not produced for this frame");
    vrt_context* c = f->ctx;
    CU(cudaSetDevice(c->device));
    const cudaError_t e = launch_push_tiles(f->planes[plane], d_frame, f->W, f->H, c->rank, c->world, f->localTiles, kPlaneBpp[plane],
                                            stream ? (cudaStream_t)stream : f->stream);
    if (e != cudaSuccess) return fail(VRT_E_CUDA, cudaGetErrorString(e));
    return VRT_OK;
}

// ---- host helpers ----
extern "C" void vrt_camera_vectors(float yaw, float pitch, float focal, float dir[3], float right[3], float up[3]) {
    // CameraController::updateDirectionVectors (camera_controller.cpp:15-28), worldUp = (0,-1,0), glm normalize/cross
    const float k = 0.01745329251994329576923690768489f;
    float n[3] = {cosf(yaw * k) * cosf(pitch * k), sinf(pitch * k), sinf(yaw * k) * cosf(pitch * k)};
    auto norm = [](float v[3]) { const float inv = 1.0f / sqrtf((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]); v[0] *= inv; v[1] *= inv; v[2] *= inv; };
    auto cross = [](const float a[3], const float b[3], float o[3]) { o[0] = a[1] * b[2] - b[1] * a[2]; o[1] = a[2] * b[0] - b[2] * a[0]; o[2] = a[0] * b[1] - b[0] * a[1]; };
    norm(n);
    const float wu[3] = {0.0f, -1.0f, 0.0f};
    float r[3], u[3];
    cross(n, wu, r); norm(r);
    cross(r, n, u); norm(u);
    for (int i = 0; i < 3; i++) { dir[i] = n[i] * focal; right[i] = r[i]; up[i] = u[i]; }
}

extern "C" void vrt_default_settings(vrt_render_settings* s) {
    if (!s) return;
    memset(s, 0, sizeof(*s));
    s->pass_mask = VRT_PASS_ALL; s->out_mask = VRT_OUT_GBUFFER; s->traversal = VRT_TRAVERSAL_DEFAULT; s->pipeline = VRT_PIPELINE_DEFAULT;
    s->vol.aoSamples = 4; s->vol.ambientIntensity = 1.0f;                  // AmbientOcclusionSettings
    const float inv = 1.0f / sqrtf((1.0f * 1.0f + 1.0f * 1.0f) + 1.0f * 1.0f);   // glm::normalize(vec3(1))
    s->light.direction[0] = s->light.direction[1] = s->light.direction[2] = 1.0f * inv;
    s->light.intensity = 1.0f;
    for (float& v : s->light.color) v = 1.0f;
    s->denoise.enable = 1; s->denoise.iterations = 2; s->denoise.phiColor0 = 20.4f; s->denoise.phiNormal0 = 1E-2f;
    s->denoise.phiPos0 = 1E-1f; s->denoise.stepWidth = 2.0f; s->denoise.mode = VRT_DENOISE_AS_EXECUTED;
}

// Host-only .vox decode (no GPU needed): what vrt_scene_load_vox_memory uploads. Free *volume_out with vrt_host_free.
extern "C" int vrt_vox_decode(const uint8_t* bytes, size_t n, uint8_t** volume_out, uint32_t whd[3], vrt_material palette[256]) {
    if (!volume_out || !whd || !palette) return fail(VRT_E_INVALID, "null argument");
    VRT_NOTHROW(
        DenseScene ds; std::string err;
        int rc = load_vox_bytes(bytes, n, ds, err);
        if (rc) return fail(rc, err);
        uint8_t* v = (uint8_t*)malloc(ds.volume.size());
        if (!v) return fail(VRT_E_OOM, "host allocation failed");
        memcpy(v, ds.volume.data(), ds.volume.size());
        memcpy(palette, ds.palette, sizeof(vrt_material) * 256);
        *volume_out = v; whd[0] = ds.w; whd[1] = ds.h; whd[2] = ds.d;
        return VRT_OK;
    )
}
extern "C" void vrt_host_free(void* p) { free(p); }
