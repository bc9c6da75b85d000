// vrt_comm.cu — the multi-GPU exchange, inside the library: one process per GPU, every rank traces its
// interleaved tile shard, the frame is assembled on rank 0.
//
// There is no collective call on the data path. Rank 0 owns the assembled planes of `slots` frames in flight (one cudaMalloc block,
// exported with cudaIpcGetMemHandle); every other rank maps that block (cudaIpcOpenMemHandle: peer memory, NVLink / NVSwitch on an
// 8xB200 box) and its push kernel stores its tiles straight into the row-major frame — a tile row is 128 contiguous bytes in both
// layouts, so the stores are full lines — followed by a one-word arrival flag. Rank 0 orders its consumer (denoiser, read-back,
// present) behind a wait kernel that polls the `world` flags of the slot. The exchange therefore overlaps the tracing of the next
// frames tile shard by tile shard, and the de-tile pass a gather would need on rank 0 does not exist.
// The same kernel can store a rank's tiles into a frame in PINNED HOST memory shared by all ranks (POSIX shm, registered with CUDA in
// every process): the read-back of a frame then runs over every GPU's PCIe link at once instead of funnelling through rank 0's.
// Rendezvous, barrier and the slot hand-shake (a consumer releases a slot before the ranks may overwrite it) live in a small POSIX
// shared-memory header; nothing on the data path touches the host.
#include <cuda_runtime.h>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <new>
#include <string>
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include "vrt_internal.h"

namespace {

constexpr int kMaxWorld = 16, kMaxSlots = 8;
constexpr uint32_t kMagic = 0x56525443u;   // "VRTC"

struct CommShm {
    std::atomic<uint32_t> magic;
    uint32_t world;
    std::atomic<uint32_t> joined;
    std::atomic<uint32_t> barCount, barGen;
    std::atomic<int32_t> error;                       // a rank that failed sets it; every wait gives up
    // frame layout, written by rank 0 in vrt_comm_alloc_frames, valid once framesGen == the caller's generation
    std::atomic<uint32_t> framesGen;
    cudaIpcMemHandle_t devHandle;
    uint64_t devBytes, hostBytes;
    int32_t w, h, slots;
    uint32_t planeMask, hostPlaneMask;
    uint64_t planeOff[VRT_PLANE_COUNT], slotStride, arriveOff;
    uint64_t hostPlaneOff[VRT_PLANE_COUNT], hostSlotStride;
    std::atomic<uint64_t> released[kMaxSlots];        // generations of the slot the consumer on rank 0 is done with (host store / host function)
    volatile uint32_t releasedOp[kMaxSlots];          // the same, written by a stream memory operation of rank 0's consumer stream
    volatile uint32_t hostArrive[kMaxSlots][kMaxWorld];   // generation whose tiles rank r has stored into the HOST frame (written by kernels)
    // strips (vrt_comm_alloc_strips): every rank's buffer of the denoiser inputs of its rows, exported to all ranks
    cudaIpcMemHandle_t stripHandle[kMaxWorld];
};
static_assert(kMaxWorld == VRT_MAX_WORLD, "StripTable and the rendezvous header are sized for the same world");

// Stream memory operations (cuStreamWaitValue32 / cuStreamWriteValue32), fetched through the runtime so that libvrt has no link-time
// dependency on libcuda (it must load on a machine without a driver). A wait on a memory word needs no SM: the persistent kernels
// of the frames in flight own every block slot, and a one-warp wait KERNEL would queue behind them (measured: the exchange then adds a
// frame time of latency and throttles every rank through the slot hand-shake).
typedef int (*StreamValue32Fn)(cudaStream_t, unsigned long long, uint32_t, unsigned int);
StreamValue32Fn g_waitValue32 = nullptr, g_writeValue32 = nullptr;
void load_stream_memops() {
    static bool done = false;
    if (done) return;
    done = true;
    if (getenv("VRT_COMM_NO_MEMOPS")) return;
    void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess) g_waitValue32 = (StreamValue32Fn)fn;
    fn = nullptr;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess) g_writeValue32 = (StreamValue32Fn)fn;
    cudaGetLastError();
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
double timeout_s() { static const double t = getenv("VRT_COMM_TIMEOUT_S") ? atof(getenv("VRT_COMM_TIMEOUT_S")) : 120.0; return t; }

}  // namespace

struct vrt_comm {
    int rank = 0, world = 1;
    std::string key;
    CommShm* shm = nullptr;                           // host mapping of the header
    uint32_t* shmDevArrive = nullptr;                 // device view of shm->hostArrive
    uint32_t* shmDevReleased = nullptr;               // device view of shm->releasedOp
    uint32_t framesGen = 0;
    uint8_t* devBase = nullptr;                       // rank 0: its allocation; others: the peer mapping of it
    uint8_t* hostBase = nullptr; uint8_t* hostBaseDev = nullptr; size_t hostBytes = 0;
    uint64_t gen[kMaxSlots] = {};                     // exchanges issued on each slot (identical on every rank)
    cudaStream_t stream = nullptr;                    // rank 0's consumer stream (arrival wait, caller's work, release)
    cudaStream_t pushStream = nullptr;                // highest priority: the tile stores of a finished frame never queue behind the next frame's kernels
    cudaStream_t stripStream = nullptr;               // highest priority: this rank's share of the denoiser (strips)
    cudaEvent_t joined = nullptr;
    uint32_t* devError = nullptr;                     // set by a wait kernel that gave up
    bool joinedOk = false;
    // strips: the denoiser runs on every rank, on the rows it owns (vrt_render_denoise_exchange)
    uint8_t* stripBase[kMaxWorld] = {};               // [rank]: own allocation, the others' as peer mappings
    StripTable* stripTables = nullptr;                // device, one per slot
    uint8_t* dnOut = nullptr; uint8_t* dnScratch = nullptr;   // the denoiser's ping-pong images of this rank's strip (+ halo)
    int stripRows = 0, stripHalo = 0, stripBufRows = 0;
    uint64_t stripSlotBytes = 0, stripFlagsOff = 0;
    vrt_denoiser_settings stripDs{};
    bool stripPlain = false;                          // strips without a denoiser: only the colour plane is transposed (read-back path)
};

// ---------------------------------------------------------------------------------------------------------------- kernels
namespace vrt {

// arrival flags of one exchange: generation g of slot s is complete for this rank (device frame on rank 0 and / or host frame).
// A kernel of its own: it starts after the push kernels before it on the stream have completed, i.e. after their stores are performed.
__global__ void comm_signal_kernel(volatile uint32_t* devFlag, volatile uint32_t* hostFlag, uint32_t g) {
    __threadfence_system();
    if (devFlag) *devFlag = g;
    if (hostFlag) *hostFlag = g;
}

// rank 0: returns once the `world` arrival flags of the slot have reached generation g. Gives up after ~20 s (sets *err).
__global__ void comm_wait_kernel(const volatile uint32_t* flags, int world, uint32_t g, uint32_t* err) {
    const int r = threadIdx.x;
    if (r < world) {
        unsigned long long t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + r) : "memory");
            if ((int32_t)(v - g) >= 0) break;
            __nanosleep(200);
            unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > 20000000000ull) { *err = 1u; break; }
        }
    }
    __threadfence_system();
}

}  // namespace vrt

// ---------------------------------------------------------------------------------------------------------------- host
static int comm_spin(vrt_comm* m, const char* what, bool (*done)(vrt_comm*, const void*), const void* arg, double limit = -1.0) {
    const double t0 = now_s();
    if (limit < 0) limit = timeout_s();
    for (unsigned it = 0;; it++) {
        if (done(m, arg)) return VRT_OK;
        if (m->shm && m->shm->error.load(std::memory_order_relaxed)) return fail(VRT_E_COMM, std::string("another rank reported an error while this rank waited for ") + what);
        if ((it & 63u) == 63u) { sched_yield(); if (now_s() - t0 > limit) { if (m->shm) m->shm->error.store(1); return fail(VRT_E_COMM, std::string("timed out waiting for ") + what); } }
    }
}

static int comm_barrier(vrt_comm* m, double limit = -1.0) {
    CommShm* h = m->shm;
    const uint32_t gen = h->barGen.load();
    if (h->barCount.fetch_add(1) + 1 == (uint32_t)m->world) { h->barCount.store(0); h->barGen.store(gen + 1); return VRT_OK; }
    return comm_spin(m, "the barrier", [](vrt_comm* mm, const void* g) { return mm->shm->barGen.load() != *(const uint32_t*)g; }, &gen, limit);
}

static void comm_unmap_strips(vrt_comm* m) {
    for (int r = 0; r < m->world; r++)
        if (m->stripBase[r]) { if (r == m->rank) cudaFree(m->stripBase[r]); else cudaIpcCloseMemHandle(m->stripBase[r]); m->stripBase[r] = nullptr; }
    cudaFree(m->stripTables); cudaFree(m->dnOut); cudaFree(m->dnScratch);
    m->stripTables = nullptr; m->dnOut = m->dnScratch = nullptr; m->stripRows = 0;
}
static void comm_unmap_frames(vrt_comm* m) {
    comm_unmap_strips(m);
    if (m->devBase) { if (m->rank == 0) cudaFree(m->devBase); else cudaIpcCloseMemHandle(m->devBase); m->devBase = nullptr; }
    if (m->hostBase) { cudaHostUnregister(m->hostBase); munmap(m->hostBase, m->hostBytes); m->hostBase = nullptr; m->hostBaseDev = nullptr; }
    if (m->rank == 0 && m->hostBytes) shm_unlink(("/vrt_" + m->key + "_f" + std::to_string(m->framesGen)).c_str());
    m->hostBytes = 0;
}

extern "C" int vrt_comm_destroy(vrt_context* c) {
    if (!c || !c->comm) return VRT_OK;
    vrt_comm* m = c->comm;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    if (m->shm && m->joinedOk && !m->shm->error.load()) comm_barrier(m, 15.0);   // nobody unmaps memory a peer may still be storing into
    comm_unmap_frames(m);
    if (m->stream) cudaStreamDestroy(m->stream);
    if (m->pushStream) cudaStreamDestroy(m->pushStream);
    if (m->stripStream) cudaStreamDestroy(m->stripStream);
    if (m->joined) cudaEventDestroy(m->joined);
    cudaFree(m->devError);
    if (m->shm) { cudaHostUnregister(m->shm); munmap(m->shm, sizeof(CommShm)); }
    if (m->rank == 0) shm_unlink(("/vrt_" + m->key).c_str());
    delete m;
    c->comm = nullptr;
    return VRT_OK;
}

extern "C" int vrt_comm_create(vrt_context* c, int rank, int world, const char* key) {
    if (!c || !key || !*key || world < 1 || world > kMaxWorld || rank < 0 || rank >= world) return fail(VRT_E_INVALID, "bad communicator arguments (world <= 16)");
    if (c->comm) return fail(VRT_E_INVALID, "the context already has a communicator");
    if (!c->frames.empty()) return fail(VRT_E_INVALID, "create the communicator before the first frame (it sets the tile shard)");
    CU(cudaSetDevice(c->device));
    vrt_comm* m = new (std::nothrow) vrt_comm();
    if (!m) return fail(VRT_E_OOM, "host allocation failed");
    m->rank = rank; m->world = world; m->key = key;
    const std::string name = "/vrt_" + m->key;
    int fd = -1;
    if (rank == 0) {
        shm_unlink(name.c_str());                                          // a stale segment of a crashed run with the same key
        fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd >= 0 && ftruncate(fd, sizeof(CommShm)) != 0) { close(fd); fd = -1; }
    } else {
        const double t0 = now_s();
        while ((fd = shm_open(name.c_str(), O_RDWR, 0600)) < 0 && now_s() - t0 < timeout_s()) usleep(1000);
        struct stat st{};
        while (fd >= 0 && (fstat(fd, &st) != 0 || (size_t)st.st_size < sizeof(CommShm)) && now_s() - t0 < timeout_s()) usleep(1000);
    }
    if (fd < 0) { delete m; return fail(VRT_E_COMM, "cannot open the shared-memory rendezvous segment " + name); }
    void* p = mmap(nullptr, sizeof(CommShm), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) { delete m; return fail(VRT_E_COMM, "mmap of the rendezvous segment failed"); }
    m->shm = (CommShm*)p;
    c->comm = m;
    if (rank == 0) {
        memset((void*)m->shm, 0, sizeof(CommShm));
        m->shm->world = (uint32_t)world;
        m->shm->magic.store(kMagic, std::memory_order_release);
    } else {
        int rc = comm_spin(m, "rank 0 to publish the rendezvous segment", [](vrt_comm* mm, const void*) { return mm->shm->magic.load(std::memory_order_acquire) == kMagic; }, nullptr);
        if (rc == VRT_OK && m->shm->world != (uint32_t)world) { m->shm->error.store(1); rc = fail(VRT_E_COMM, "world size differs from rank 0's"); }
        if (rc) { const std::string msg = g_vrt_err; vrt_comm_destroy(c); return fail(rc, msg); }
    }
    // the arrival flags of the host frames are written by kernels: the header is mapped into the device's address space
    cudaError_t e = cudaHostRegister(m->shm, sizeof(CommShm), cudaHostRegisterMapped | cudaHostRegisterPortable);
    void* dv = nullptr;
    if (e == cudaSuccess) e = cudaHostGetDevicePointer(&dv, (void*)&m->shm->hostArrive[0][0], 0);
    int prLow = 0, prHigh = 0;
    if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prLow, &prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&m->stream, cudaStreamNonBlocking, prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&m->pushStream, cudaStreamNonBlocking, prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&m->stripStream, cudaStreamNonBlocking, prHigh);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->joined, cudaEventDisableTiming);
    load_stream_memops();
    if (e == cudaSuccess) e = cudaMalloc(&m->devError, sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(m->devError, 0, sizeof(uint32_t));
    if (e != cudaSuccess) { m->shm->error.store(1); const std::string msg = cudaGetErrorString(e); vrt_comm_destroy(c); return fail(VRT_E_CUDA, msg); }
    m->shmDevArrive = (uint32_t*)dv;
    if (cudaHostGetDevicePointer(&dv, (void*)&m->shm->releasedOp[0], 0) == cudaSuccess) m->shmDevReleased = (uint32_t*)dv;
    m->shm->joined.fetch_add(1);
    const int rc = comm_barrier(m);
    if (rc) { const std::string msg = g_vrt_err; vrt_comm_destroy(c); return fail(rc, msg); }
    m->joinedOk = true;
    c->rank = rank; c->world = world;
    return VRT_OK;
}

extern "C" int vrt_comm_barrier(vrt_context* c) {
    if (!c || !c->comm) return fail(VRT_E_INVALID, "no communicator");
    return comm_barrier(c->comm);
}
extern "C" void* vrt_comm_stream(vrt_context* c) { return c && c->comm ? (void*)c->comm->stream : nullptr; }

// Collective: every rank calls it with the same arguments. Rank 0 allocates the assembled planes of `slots` frames in flight on its
// GPU (plane_mask) and in shared pinned host memory (host_plane_mask); the others map both.
extern "C" int vrt_comm_alloc_frames(vrt_context* c, int w, int h, int slots, uint32_t plane_mask, uint32_t host_plane_mask) {
    if (!c || !c->comm) return fail(VRT_E_INVALID, "no communicator");
    if (w <= 0 || h <= 0 || slots < 1 || slots > kMaxSlots || (plane_mask | host_plane_mask) >> VRT_PLANE_COUNT)
        return fail(VRT_E_INVALID, "bad frame layout (slots <= 8)");   // (the denoised plane of a slot is filled by vrt_render_denoise_exchange only)
    vrt_comm* m = c->comm; CommShm* sh = m->shm;
    CU(cudaSetDevice(c->device));
    CU(cudaDeviceSynchronize());
    int rc = comm_barrier(m);                                              // nobody is still storing into the old frames
    if (rc) return rc;
    comm_unmap_frames(m);
    m->framesGen++;
    memset(m->gen, 0, sizeof(m->gen));
    const std::string hname = "/vrt_" + m->key + "_f" + std::to_string(m->framesGen);
    cudaError_t e = cudaSuccess;
    if (m->rank == 0) {
        uint64_t off = 0, hoff = 0;
        for (int p = 0; p < VRT_PLANE_COUNT; p++) {
            sh->planeOff[p] = off; sh->hostPlaneOff[p] = hoff;
            const uint64_t bytes = ((uint64_t)w * h * kPlaneBpp[p] + 255u) & ~(uint64_t)255u;
            if (plane_mask >> p & 1u) off += bytes;
            if (host_plane_mask >> p & 1u) hoff += bytes;
        }
        sh->slotStride = off; sh->hostSlotStride = hoff;
        sh->arriveOff = off * slots;
        sh->devBytes = sh->arriveOff + sizeof(uint32_t) * kMaxSlots * kMaxWorld;
        sh->hostBytes = hoff * slots;
        sh->w = w; sh->h = h; sh->slots = slots; sh->planeMask = plane_mask; sh->hostPlaneMask = host_plane_mask;
        for (int s = 0; s < kMaxSlots; s++) { sh->released[s].store(0); sh->releasedOp[s] = 0; for (int r = 0; r < kMaxWorld; r++) sh->hostArrive[s][r] = 0; }
        e = cudaMalloc(&m->devBase, sh->devBytes);
        if (e == cudaSuccess) e = cudaMemset(m->devBase, 0, sh->devBytes);
        if (e == cudaSuccess && m->world > 1) e = cudaIpcGetMemHandle(&sh->devHandle, m->devBase);
        if (e == cudaSuccess && sh->hostBytes) {
            shm_unlink(hname.c_str());
            const int fd = shm_open(hname.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
            if (fd < 0 || ftruncate(fd, (off_t)sh->hostBytes) != 0) { if (fd >= 0) close(fd); sh->error.store(1); return fail(VRT_E_COMM, "cannot create the shared host frame segment"); }
            close(fd);
        }
        if (e != cudaSuccess) { sh->error.store(1); return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e)); }
        sh->framesGen.store(m->framesGen, std::memory_order_release);
    } else {
        rc = comm_spin(m, "rank 0 to allocate the frames", [](vrt_comm* mm, const void*) { return mm->shm->framesGen.load(std::memory_order_acquire) == mm->framesGen; }, nullptr);
        if (rc) return rc;
        if (sh->w != w || sh->h != h || sh->slots != slots || sh->planeMask != plane_mask || sh->hostPlaneMask != host_plane_mask) { sh->error.store(1); return fail(VRT_E_COMM, "frame layout differs from rank 0's"); }
        void* base = nullptr;
        e = cudaIpcOpenMemHandle(&base, sh->devHandle, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { sh->error.store(1); return fail(VRT_E_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e)); }
        m->devBase = (uint8_t*)base;
    }
    if (sh->hostBytes) {
        const int fd = shm_open(hname.c_str(), O_RDWR, 0600);
        void* p = fd >= 0 ? mmap(nullptr, sh->hostBytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0) : MAP_FAILED;
        if (fd >= 0) close(fd);
        if (p == MAP_FAILED) { sh->error.store(1); return fail(VRT_E_COMM, "cannot map the shared host frame segment"); }
        m->hostBase = (uint8_t*)p; m->hostBytes = sh->hostBytes;
        if (m->rank == 0) memset(p, 0, sh->hostBytes);                     // touch every page once, before the timed use
        void* dv = nullptr;
        e = cudaHostRegister(p, sh->hostBytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
        if (e == cudaSuccess) e = cudaHostGetDevicePointer(&dv, p, 0);
        if (e != cudaSuccess) { sh->error.store(1); return fail(VRT_E_CUDA, std::string("cudaHostRegister of the shared host frames: ") + cudaGetErrorString(e)); }
        m->hostBaseDev = (uint8_t*)dv;
    }
    return comm_barrier(m);
}

// ---- one exchange = begin (slot hand-shake on the host) + the stores (push kernels, or the pass's own kernels) + finish (arrival flags)
static int exchange_begin(vrt_context* c, int W, int H, int slot, uint32_t plane_mask, uint32_t host_plane_mask, uint64_t* gOut) {
    if (!c || !c->comm) return fail(VRT_E_INVALID, "no communicator");
    vrt_comm* m = c->comm; CommShm* sh = m->shm;
    if (!m->devBase) return fail(VRT_E_INVALID, "call vrt_comm_alloc_frames first");
    if (slot < 0 || slot >= sh->slots || W != sh->w || H != sh->h) return fail(VRT_E_INVALID, "slot or frame size does not match vrt_comm_alloc_frames");
    if ((plane_mask & ~sh->planeMask) || (host_plane_mask & ~sh->hostPlaneMask)) return fail(VRT_E_INVALID, "plane was not allocated by vrt_comm_alloc_frames");
    CU(cudaSetDevice(c->device));
    const uint64_t g = ++m->gen[slot];
    struct Arg { int slot; uint64_t g; } a{slot, g};
    const int rc = comm_spin(m, "rank 0 to release the slot", [](vrt_comm* mm, const void* q) { const Arg* x = (const Arg*)q;
        const uint64_t r = mm->shm->released[x->slot].load(std::memory_order_acquire), r2 = mm->shm->releasedOp[x->slot];
        return (r > r2 ? r : r2) + 1 >= x->g; }, &a);
    *gOut = g;
    return rc;
}
// push kernels for the planes in plane_mask / host_plane_mask (none: the pass's kernels stored them already), then the arrival flags
static int exchange_finish(vrt_frame* f, int slot, uint64_t g, uint32_t plane_mask, uint32_t host_plane_mask, bool devFlagWanted, bool hostFlagWanted, void* stream) {
    vrt_context* c = f->ctx; vrt_comm* m = c->comm; CommShm* sh = m->shm;
    for (int p = 0; p < VRT_PLANE_COUNT; p++) if (((plane_mask | host_plane_mask) >> p & 1u) && !f->planes[p]) return fail(VRT_E_INVALID, "plane was not produced for this frame");
    // Push kernels run on the communicator's high-priority stream behind the frame's kernels; the frame's next render waits for
    // them (f->copied), as it does for an asynchronous read-back. Without push kernels the flags follow the pass on its own stream.
    static const bool ownStream = !getenv("VRT_COMM_PUSH_ON_FRAME_STREAM");
    cudaStream_t s = stream ? (cudaStream_t)stream : f->stream;
    cudaError_t e = cudaSuccess;
    const bool hop = !stream && ownStream && (plane_mask | host_plane_mask) != 0 && f->localTiles > 0;
    if (hop) {
        e = cudaEventRecord(f->rendered, f->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(m->pushStream, f->rendered, 0);
        s = m->pushStream;
    }
    for (int p = 0; p < VRT_PLANE_COUNT && e == cudaSuccess; p++) {
        if (plane_mask >> p & 1u)
            e = launch_push_tiles(f->planes[p], m->devBase + (uint64_t)slot * sh->slotStride + sh->planeOff[p], f->W, f->H, c->rank, c->world, f->localTiles, kPlaneBpp[p], s);
        if (e == cudaSuccess && (host_plane_mask >> p & 1u))
            e = launch_push_tiles(f->planes[p], m->hostBaseDev + (uint64_t)slot * sh->hostSlotStride + sh->hostPlaneOff[p], f->W, f->H, c->rank, c->world, f->localTiles, kPlaneBpp[p], s);
        if (f->localTiles > 0) f->launches += (plane_mask >> p & 1u) + (host_plane_mask >> p & 1u);
    }
    if (e == cudaSuccess) {
        uint32_t* devFlag = devFlagWanted ? (uint32_t*)(m->devBase + sh->arriveOff) + slot * kMaxWorld + c->rank : nullptr;
        uint32_t* hostFlag = hostFlagWanted ? m->shmDevArrive + slot * kMaxWorld + c->rank : nullptr;
        // arrival flags: stream memory operations where the driver has them (no kernel, no SM), else the one-thread kernel
        bool done = false;
        if (g_writeValue32 && !getenv("VRT_COMM_SIGNAL_KERNEL")) {
            int rc = 0;
            if (devFlag) rc = g_writeValue32(s, (unsigned long long)devFlag, (uint32_t)g, 0);
            if (rc == 0 && hostFlag) rc = g_writeValue32(s, (unsigned long long)hostFlag, (uint32_t)g, 0);
            done = rc == 0;
            if (!done) { g_writeValue32 = nullptr; cudaGetLastError(); }
        }
        if (!done) { comm_signal_kernel<<<1, 1, 0, s>>>(devFlag, hostFlag, (uint32_t)g); f->launches++; e = cudaGetLastError(); }
    }
    if (e == cudaSuccess && hop) {
        e = cudaEventRecord(f->copied, m->pushStream);
        f->copyPending = true;
    }
    if (e != cudaSuccess) { sh->error.store(1); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    return VRT_OK;
}

// This rank's tiles of the planes in plane_mask (device frame on rank 0) and host_plane_mask (shared host frame) of `slot`, then the
// arrival flags. Asynchronous (on `stream`, or behind the frame's kernels); the call itself may wait on the HOST until rank 0 has
// released what the slot held before (vrt_comm_release): with S slots the ranks run at most S-1 frames ahead of the consumer.
extern "C" int vrt_frame_exchange(vrt_frame* f, int slot, uint32_t plane_mask, uint32_t host_plane_mask, void* stream) {
    if (!f) return fail(VRT_E_INVALID, "null frame");
    uint64_t g = 0;
    const int rc = exchange_begin(f->ctx, f->W, f->H, slot, plane_mask, host_plane_mask, &g);
    if (rc) return rc;
    return exchange_finish(f, slot, g, plane_mask, host_plane_mask, plane_mask != 0, host_plane_mask != 0, stream);
}

// vrt_render + vrt_frame_exchange in one call, with the exchange FUSED into the pass where it can be: the colour and G-buffer planes
// named in plane_mask are stored by the pass's own kernels (primary / resolve) straight into rank 0's frame as they are produced —
// peer stores over NVLink beside the local ones — and the arrival flag follows the last kernel on the frame's stream as a stream
// memory operation. No push kernel, no extra pass over the tiles: the last kernel of every rank is the collective.
// Instrumentation planes and the host frame (host_plane_mask) still go through push kernels.
extern "C" int vrt_render_exchange(vrt_context* c, const vrt_scene* sc, const vrt_camera* cam, const vrt_render_settings* st, int slot,
                                   uint32_t plane_mask, uint32_t host_plane_mask, vrt_frame** frame) {
    if (!c || !cam || !st) return fail(VRT_E_INVALID, "null argument");
    uint64_t g = 0;
    int rc = exchange_begin(c, cam->screenSize[0], cam->screenSize[1], slot, plane_mask, host_plane_mask, &g);
    if (rc) return rc;
    vrt_comm* m = c->comm; CommShm* sh = m->shm;
    XPlanes x{};
    uint32_t fused = 0, fusedHost = 0;
    static const bool noFuse = getenv("VRT_COMM_NO_FUSE") != nullptr, fuseHost = getenv("VRT_COMM_FUSE_HOST") != nullptr;
    if (!noFuse && fuseHost && (host_plane_mask >> VRT_PLANE_COLOR_RGBA8 & 1u)) {
        x.colorHost = m->hostBaseDev + (uint64_t)slot * sh->hostSlotStride + sh->hostPlaneOff[VRT_PLANE_COLOR_RGBA8];
        fusedHost = 1u << VRT_PLANE_COLOR_RGBA8;
    }
    if (!noFuse) {
        auto at = [&](int p) { return m->devBase + (uint64_t)slot * sh->slotStride + sh->planeOff[p]; };
        const bool gb = (st->out_mask & VRT_OUT_GBUFFER) != 0;
        if (plane_mask >> VRT_PLANE_COLOR_RGBA8 & 1u) { x.color = at(VRT_PLANE_COLOR_RGBA8); fused |= 1u << VRT_PLANE_COLOR_RGBA8; }
        if (gb && (plane_mask >> VRT_PLANE_DEPTH_R32F & 1u)) { x.depth = (float*)at(VRT_PLANE_DEPTH_R32F); fused |= 1u << VRT_PLANE_DEPTH_R32F; }
        if (gb && (plane_mask >> VRT_PLANE_MOTION_RG32F & 1u)) { x.motion = (float*)at(VRT_PLANE_MOTION_RG32F); fused |= 1u << VRT_PLANE_MOTION_RG32F; }
        if (gb && (plane_mask >> VRT_PLANE_MASK_R8 & 1u)) { x.mask = at(VRT_PLANE_MASK_R8); fused |= 1u << VRT_PLANE_MASK_R8; }
        if (gb && (plane_mask >> VRT_PLANE_POSITION_RGBA32F & 1u)) { x.position = (float*)at(VRT_PLANE_POSITION_RGBA32F); fused |= 1u << VRT_PLANE_POSITION_RGBA32F; }
        if (gb && (plane_mask >> VRT_PLANE_NORMAL_RGBA8S & 1u)) { x.normal = (int8_t*)at(VRT_PLANE_NORMAL_RGBA8S); fused |= 1u << VRT_PLANE_NORMAL_RGBA8S; }
    }
    rc = vrt_render_impl(c, sc, cam, st, frame, (fused | fusedHost) ? &x : nullptr);
    if (rc) { sh->error.store(1); return rc; }
    return exchange_finish(*frame, slot, g, plane_mask & ~fused, host_plane_mask & ~fusedHost, plane_mask != 0, host_plane_mask != 0, nullptr);
}

// ---- strips: the denoiser of a sharded frame, distributed (SURVEY.md 8e "denoiser under sharding"; neither route (a) nor (b)) ----------
// Route (a) — all three input planes to rank 0, denoiser there — makes rank 0 the serial part (cfg3 at 8 GPUs: 3.4x). Route (b) — every
// rank re-traces a halo around its tiles — re-traces 56 % more pixels with 32x32 interleaved tiles. Here the pixels are traced once, by the
// rank that owns their tile, and the denoiser's inputs are stored (by the pass's own kernels, over NVLink peer memory) into the buffer of
// the rank that owns their ROWS: rank q denoises rows [q * stripRows, (q + 1) * stripRows) of the frame from a buffer that also holds
// `halo` rows either side, halo = the summed reach of all iterations' taps, so its rows come out byte-identical to the single-GPU
// frame (the rows of the halo itself come out wrong and are dropped; at the image's top and bottom the buffer ends where the image does,
// and the sampler's clamp is the reference's). Each rank then copies its rows into rank 0's denoised plane of the slot (and into the
// shared host frame) with plain row-range copies on the copy engines, and sets its arrival flag: vrt_comm_wait / _wait_host / _release
// on rank 0 work as for any exchange. Per frame every rank receives 24 B x (1 / world + halo share) of the pixels instead of rank 0
// receiving all of them.
static int denoise_reach_rows(const vrt_denoiser_settings* ds) {
    if (!ds || ds->iterations < 1 || ds->iterations > 10) return -1;
    int h = 0;
    for (int i = 0; i < ds->iterations; i++) {
        const float sw = fabsf((float)i * ds->stepWidth + 1.0f);           // denoiser_stage.cpp:143-154; tap offsets are -1, 0, 1
        if (!(sw < 4096.0f)) return -1;
        h += (int)ceilf(sw) + 1;                                           // + 1: a fractional tap position blends the texel beyond it
    }
    return h;
}
// the same as a utility of the ABI (pure host arithmetic: callable without a GPU): tests/test_sharding.py checks it against the oracle
extern "C" int vrt_denoise_reach_rows(const vrt_denoiser_settings* ds) { return denoise_reach_rows(ds); }
// `word` := g when `s` gets there: a stream memory operation where the driver has them (no SM), else the one-thread kernel
static cudaError_t stream_signal(cudaStream_t s, uint32_t* devFlag, uint32_t* hostFlag, uint32_t g, uint32_t* launches) {
    if (g_writeValue32 && !getenv("VRT_COMM_SIGNAL_KERNEL")) {
        int rc = 0;
        if (devFlag) rc = g_writeValue32(s, (unsigned long long)devFlag, g, 0);
        if (rc == 0 && hostFlag) rc = g_writeValue32(s, (unsigned long long)hostFlag, g, 0);
        if (rc == 0) return cudaSuccess;
        g_writeValue32 = nullptr; cudaGetLastError();
    }
    vrt::comm_signal_kernel<<<1, 1, 0, s>>>(devFlag, hostFlag, g);
    if (launches) (*launches)++;
    return cudaGetLastError();
}
static cudaError_t stream_wait_flags(vrt_comm* m, cudaStream_t s, const uint32_t* flags, int n, uint32_t g) {
    if (g_waitValue32) {
        bool ok = true;
        for (int r = 0; r < n && ok; r++) ok = g_waitValue32(s, (unsigned long long)(flags + r), g, 0x0 /* CU_STREAM_WAIT_VALUE_GEQ */) == 0;
        if (ok) return cudaSuccess;
        g_waitValue32 = nullptr; cudaGetLastError();
    }
    vrt::comm_wait_kernel<<<1, 32, 0, s>>>(flags, n, g, m->devError);
    return cudaGetLastError();
}

// Collective (same arguments on every rank), after vrt_comm_alloc_frames with VRT_PLANE_DENOISED_RGBA8 in plane_mask: every rank
// allocates the strip buffers of its rows for all slots and maps everybody else's. The denoiser settings fix the halo; they are the
// ones vrt_render_denoise_exchange runs. VRT_E_UNSUPPORTED when a strip is thinner than the denoiser's reach (use route (a) then).
extern "C" int vrt_comm_alloc_strips(vrt_context* c, const vrt_denoiser_settings* ds) {
    if (!c || !c->comm) return fail(VRT_E_INVALID, "no communicator");
    vrt_comm* m = c->comm; CommShm* sh = m->shm;
    const bool plain = ds == nullptr;                                      // no denoiser: the colour plane alone goes through the strips
    const uint32_t need = 1u << (plain ? VRT_PLANE_COLOR_RGBA8 : VRT_PLANE_DENOISED_RGBA8);
    if (!m->devBase || !((sh->planeMask | sh->hostPlaneMask) & need)) return fail(VRT_E_INVALID, "call vrt_comm_alloc_frames with the plane the strips deliver (denoised, or colour without a denoiser) first");
    if (!plain && (ds->iterations < 1 || ds->iterations > 10)) return fail(VRT_E_INVALID, "denoiser iterations must be 1..10 (MAX_DENOISER_PASSES)");
    if (!plain && ds->mode != VRT_DENOISE_AS_EXECUTED && ds->mode != VRT_DENOISE_AS_INTENDED) return fail(VRT_E_INVALID, "bad denoise mode");
    const int W = sh->w, H = sh->h, slots = sh->slots, world = m->world;
    const int SR = (H + world - 1) / world, halo = plain ? 0 : denoise_reach_rows(ds);
    if (halo < 0 || halo > SR) return fail(VRT_E_UNSUPPORTED, "a strip of " + std::to_string(SR) + " rows is thinner than the denoiser's reach: exchange the planes and denoise on rank 0 (vrt_denoise_on)");
    CU(cudaSetDevice(c->device));
    CU(cudaDeviceSynchronize());
    int rc = comm_barrier(m);                                              // nobody is still storing into the old strips
    if (rc) return rc;
    comm_unmap_strips(m);
    const int bufRows = SR + 2 * halo;
    auto up = [](uint64_t v) { return (v + 255u) & ~(uint64_t)255u; };
    const uint64_t colorBytes = up((uint64_t)bufRows * W * 4), normalBytes = plain ? 0 : colorBytes, positionBytes = plain ? 0 : up((uint64_t)bufRows * W * 16);
    m->stripSlotBytes = colorBytes + normalBytes + positionBytes;
    m->stripFlagsOff = m->stripSlotBytes * slots;
    const uint64_t total = m->stripFlagsOff + sizeof(uint32_t) * kMaxSlots * kMaxWorld;
    cudaError_t e = cudaMalloc(&m->stripBase[m->rank], total);
    if (e == cudaSuccess) e = cudaMemset(m->stripBase[m->rank], 0, total);
    if (e == cudaSuccess && world > 1) e = cudaIpcGetMemHandle(&sh->stripHandle[m->rank], m->stripBase[m->rank]);
    if (e == cudaSuccess && !plain) e = cudaMalloc(&m->dnOut, (size_t)bufRows * W * 4);
    if (e == cudaSuccess && !plain) e = cudaMalloc(&m->dnScratch, (size_t)bufRows * W * 4);
    if (e == cudaSuccess) e = cudaMalloc(&m->stripTables, sizeof(StripTable) * slots);
    if (e != cudaSuccess) { sh->error.store(1); return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e)); }
    rc = comm_barrier(m);                                                  // every handle is published
    if (rc) return rc;
    for (int r = 0; r < world; r++) {
        if (r == m->rank) continue;
        void* base = nullptr;
        e = cudaIpcOpenMemHandle(&base, sh->stripHandle[r], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { sh->error.store(1); return fail(VRT_E_CUDA, std::string("cudaIpcOpenMemHandle (strips): ") + cudaGetErrorString(e)); }
        m->stripBase[r] = (uint8_t*)base;
    }
    StripTable tables[kMaxSlots];
    memset(tables, 0, sizeof(tables));
    for (int s = 0; s < slots; s++) {
        StripTable& t = tables[s];
        for (int r = 0; r < world; r++) {
            uint8_t* b = m->stripBase[r] + (uint64_t)s * m->stripSlotBytes;
            t.color[r] = b;
            if (!plain) { t.normal[r] = (int8_t*)(b + colorBytes); t.position[r] = (float*)(b + colorBytes + normalBytes); }
        }
        t.world = world; t.stripRows = SR; t.halo = halo; t.W = W; t.H = H;
    }
    e = cudaMemcpy(m->stripTables, tables, sizeof(StripTable) * slots, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { sh->error.store(1); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    m->stripRows = SR; m->stripHalo = halo; m->stripBufRows = bufRows; m->stripPlain = plain;
    if (!plain) m->stripDs = *ds;
    return comm_barrier(m);
}

// vrt_render + the distributed denoiser in one call (every rank, same arguments): the pass's kernels store colour, normal and position
// of this rank's tiles into the strips of the ranks that own their rows; this rank then denoises its own rows as soon as every rank's
// contribution has arrived and copies them into the denoised plane of `slot` on rank 0 (to_host: and of the shared host frame, which
// must have been allocated with the denoised plane in host_plane_mask). st->denoise is ignored (the settings are those of
// vrt_comm_alloc_strips); the frame's own G-buffer planes are produced as usual. Asynchronous, like vrt_render_exchange.
extern "C" int vrt_render_denoise_exchange(vrt_context* c, const vrt_scene* sc, const vrt_camera* cam, const vrt_render_settings* st, int slot,
                                           int to_host, vrt_frame** frame) {
    if (!c || !cam || !st || !frame) return fail(VRT_E_INVALID, "null argument");
    if (to_host < 0 || to_host > 2) return fail(VRT_E_INVALID, "to_host is 0 (device frame), 1 (device + host frame) or 2 (host frame only)");
    if (!c->comm || !c->comm->stripRows) return fail(VRT_E_INVALID, "call vrt_comm_alloc_strips first");
    if (st->pipeline == VRT_PIPELINE_MEGAKERNEL) return fail(VRT_E_UNSUPPORTED, "strips are stored by the wavefront kernels");
    vrt_comm* m = c->comm; CommShm* sh = m->shm;
    const int W = cam->screenSize[0], H = cam->screenSize[1];
    const int outPlane = m->stripPlain ? VRT_PLANE_COLOR_RGBA8 : VRT_PLANE_DENOISED_RGBA8;
    const uint32_t dn = 1u << outPlane;
    const bool toDev = to_host != 2, toHost = to_host != 0;                // to_host: 0 rank 0's device frame, 1 both, 2 the shared host frame only
    uint64_t g64 = 0;
    int rc = exchange_begin(c, W, H, slot, toDev ? dn : 0u, toHost ? dn : 0u, &g64);
    if (rc) return rc;
    const uint32_t g = (uint32_t)g64;
    vrt_render_settings s2 = *st;
    s2.denoise.enable = 0;
    if (!m->stripPlain) s2.out_mask |= VRT_OUT_GBUFFER;
    XPlanes x{};
    x.strips = m->stripTables + slot;
    rc = vrt_render_impl(c, sc, cam, &s2, frame, &x);
    if (rc) { sh->error.store(1); return rc; }
    vrt_frame* f = *frame;
    cudaError_t e = cudaSuccess;
    // every owner learns that this rank's pixels of generation g are in its strip (behind the pass's last kernel on the frame's stream)
    for (int q = 0; q < m->world && e == cudaSuccess; q++)
        e = stream_signal(f->stream, (uint32_t*)(m->stripBase[q] + m->stripFlagsOff) + slot * kMaxWorld + m->rank, nullptr, g, &f->launches);
    // this rank's rows: wait for every rank, denoise, deliver
    if (e == cudaSuccess) e = stream_wait_flags(m, m->stripStream, (const uint32_t*)(m->stripBase[m->rank] + m->stripFlagsOff) + slot * kMaxWorld, m->world, g);
    const int SR = m->stripRows, y0 = m->rank * SR, y1 = y0 + SR < H ? y0 + SR : H;
    if (e == cudaSuccess && y0 < y1) {
        const int lo = strip_lo(m->rank, SR, m->stripHalo), hi = strip_hi(m->rank, SR, m->stripHalo, H);
        uint8_t* b = m->stripBase[m->rank] + (uint64_t)slot * m->stripSlotBytes;
        const uint64_t colorBytes = ((uint64_t)m->stripBufRows * W * 4 + 255u) & ~(uint64_t)255u;
        const size_t rowBytes = (size_t)W * 4;
        const uint8_t* src = b + (size_t)(y0 - lo) * rowBytes;             // without a denoiser: the rows as they arrived
        if (!m->stripPlain) {
            void* ping[2] = {m->dnOut, m->dnScratch};
            if (((m->stripDs.iterations - 1) % 2) == 1) { ping[0] = m->dnScratch; ping[1] = m->dnOut; }   // the last iteration lands in dnOut
            int last = 0; uint32_t n = 0;
            rc = run_denoise(c, m->stripStream, b, b + colorBytes, b + 2 * colorBytes, W, hi - lo, &m->stripDs, ping, &last, &n);
            if (rc) { sh->error.store(1); return rc; }
            f->launches += n;
            src = m->dnOut + (size_t)(y0 - lo) * rowBytes;
        }
        if (toDev)
            e = cudaMemcpyAsync(m->devBase + (uint64_t)slot * sh->slotStride + sh->planeOff[outPlane] + (size_t)y0 * rowBytes, src, (size_t)(y1 - y0) * rowBytes, cudaMemcpyDefault, m->stripStream);
        if (e == cudaSuccess && toHost)
            e = cudaMemcpyAsync(m->hostBase + (uint64_t)slot * sh->hostSlotStride + sh->hostPlaneOff[outPlane] + (size_t)y0 * rowBytes, src, (size_t)(y1 - y0) * rowBytes, cudaMemcpyDefault, m->stripStream);
    }
    if (e == cudaSuccess)
        e = stream_signal(m->stripStream, toDev ? (uint32_t*)(m->devBase + sh->arriveOff) + slot * kMaxWorld + m->rank : nullptr, toHost ? m->shmDevArrive + slot * kMaxWorld + m->rank : nullptr, g, &f->launches);
    if (e != cudaSuccess) { sh->error.store(1); return fail(VRT_E_CUDA, cudaGetErrorString(e)); }
    return VRT_OK;
}

// Rank 0: everything queued on `stream` (null: the communicator's stream) after this call runs once every rank's tiles of the latest
// exchange on `slot` are in the device frame.
extern "C" int vrt_comm_wait(vrt_context* c, int slot, void* stream) {
    if (!c || !c->comm || !c->comm->devBase) return fail(VRT_E_INVALID, "no communicator / frames");
    vrt_comm* m = c->comm; CommShm* sh = m->shm;
    if (m->rank != 0) return fail(VRT_E_INVALID, "the frame is assembled on rank 0");
    if (slot < 0 || slot >= sh->slots || m->gen[slot] == 0) return fail(VRT_E_INVALID, "nothing was exchanged on this slot");
    CU(cudaSetDevice(c->device));
    const cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    const uint32_t* flags = (const uint32_t*)(m->devBase + sh->arriveOff) + slot * kMaxWorld;
    if (g_waitValue32) {
        bool ok = true;
        for (int r = 0; r < m->world && ok; r++) ok = g_waitValue32(s, (unsigned long long)(flags + r), (uint32_t)m->gen[slot], 0x0 /* CU_STREAM_WAIT_VALUE_GEQ */) == 0;
        if (ok) return VRT_OK;
        g_waitValue32 = nullptr; cudaGetLastError();
    }
    comm_wait_kernel<<<1, 32, 0, s>>>(flags, m->world, (uint32_t)m->gen[slot], m->devError);
    CU(cudaGetLastError());
    return VRT_OK;
}

// Rank 0: the HOST waits until every rank's tiles of the latest exchange on `slot` are in the shared host frame.
extern "C" int vrt_comm_wait_host(vrt_context* c, int slot) {
    if (!c || !c->comm || !c->comm->hostBase) return fail(VRT_E_INVALID, "no communicator / host frames");
    vrt_comm* m = c->comm;
    if (m->rank != 0) return fail(VRT_E_INVALID, "the frame is assembled on rank 0");
    if (slot < 0 || slot >= m->shm->slots || m->gen[slot] == 0) return fail(VRT_E_INVALID, "nothing was exchanged on this slot");
    struct Arg { int slot; uint32_t g; } a{slot, (uint32_t)m->gen[slot]};
    return comm_spin(m, "the ranks' tiles in the host frame", [](vrt_comm* mm, const void* q) {
        const Arg* x = (const Arg*)q;
        for (int r = 0; r < mm->world; r++) if ((int32_t)(mm->shm->hostArrive[x->slot][r] - x->g) < 0) return false;
        std::atomic_thread_fence(std::memory_order_acquire);
        return true; }, &a);
}

// Rank 0: the consumer is done with what `slot` holds; the ranks may store the next frame into it. _on: when `stream` gets there.
extern "C" int vrt_comm_release(vrt_context* c, int slot) {
    if (!c || !c->comm || c->comm->rank != 0 || slot < 0 || slot >= kMaxSlots) return fail(VRT_E_INVALID, "bad argument");
    c->comm->shm->released[slot].store(c->comm->gen[slot], std::memory_order_release);
    return VRT_OK;
}
namespace { struct ReleaseNote { std::atomic<uint64_t>* word; uint64_t g; }; }
extern "C" int vrt_comm_release_on(vrt_context* c, int slot, void* stream) {
    if (!c || !c->comm || c->comm->rank != 0 || slot < 0 || slot >= kMaxSlots) return fail(VRT_E_INVALID, "bad argument");
    vrt_comm* m = c->comm;
    CU(cudaSetDevice(c->device));
    // a stream memory operation into the (mapped) rendezvous segment where the driver has them: no host callback thread in the loop
    if (g_writeValue32 && m->shmDevReleased && !getenv("VRT_COMM_RELEASE_HOSTFUNC")) {
        if (g_writeValue32(stream ? (cudaStream_t)stream : m->stream, (unsigned long long)(m->shmDevReleased + slot), (uint32_t)m->gen[slot], 0) == 0) return VRT_OK;
        cudaGetLastError();
    }
    ReleaseNote* n = new (std::nothrow) ReleaseNote{&m->shm->released[slot], m->gen[slot]};
    if (!n) return fail(VRT_E_OOM, "host allocation failed");
    CU(cudaLaunchHostFunc(stream ? (cudaStream_t)stream : m->stream, [](void* q) { ReleaseNote* x = (ReleaseNote*)q; x->word->store(x->g, std::memory_order_release); delete x; }, n));
    return VRT_OK;
}

// The assembled plane of a slot as this process sees it: device pointer (rank 0: its memory; others: the peer mapping) and the
// host pointer into the shared pinned frame (null if the plane has none).
extern "C" int vrt_comm_frame_plane(vrt_context* c, int slot, int plane, void** d_ptr, void** h_ptr, size_t* bytes) {
    if (!c || !c->comm || !c->comm->devBase) return fail(VRT_E_INVALID, "no communicator / frames");
    vrt_comm* m = c->comm; CommShm* sh = m->shm;
    if (slot < 0 || slot >= sh->slots || plane < 0 || plane >= VRT_PLANE_COUNT) return fail(VRT_E_INVALID, "bad slot / plane");
    if (d_ptr) *d_ptr = (sh->planeMask >> plane & 1u) ? m->devBase + (uint64_t)slot * sh->slotStride + sh->planeOff[plane] : nullptr;
    if (h_ptr) *h_ptr = (sh->hostPlaneMask >> plane & 1u) && m->hostBase ? m->hostBase + (uint64_t)slot * sh->hostSlotStride + sh->hostPlaneOff[plane] : nullptr;
    if (bytes) *bytes = (size_t)sh->w * sh->h * kPlaneBpp[plane];
    return VRT_OK;
}

// the context stream waits for everything queued on the communicator's stream so far (the end of a timed region on rank 0)
int vrt_comm_join_into(vrt_context* c) {
    vrt_comm* m = c->comm;
    if (!m) return VRT_OK;
    CU(cudaEventRecord(m->joined, m->stream));
    CU(cudaStreamWaitEvent(c->stream, m->joined, 0));
    return VRT_OK;
}
int vrt_comm_sync(vrt_context* c) {
    vrt_comm* m = c->comm;
    if (!m) return VRT_OK;
    CU(cudaStreamSynchronize(m->stream));
    uint32_t err = 0;
    CU(cudaMemcpy(&err, m->devError, sizeof(err), cudaMemcpyDeviceToHost));
    if (err) { m->shm->error.store(1); return fail(VRT_E_COMM, "a wait kernel gave up: a rank's tiles never arrived"); }
    return VRT_OK;
}
