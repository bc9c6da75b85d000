// vrt_launch.h — host-callable launchers of the CUDA kernels (internal; the public surface is include/vrt.h).
#pragma once
#include <cuda_runtime.h>
#include "vrt_types.h"
#include "vrt_flatten.h"

namespace vrt {

// voxel_volume pass (vrt_render.cu)
cudaError_t launch_render(const SceneDev& sc, const FrameParams& fp, const PlanesDev& pl, Counters* ctr, cudaStream_t stream);
// the same pass as ray waves over compact hit records (vrt_wave.cu); `sms` sizes the persistent grids
// `side` (may be null): a low-priority stream + two untimed events. The AO / shadow rays of the PRIMARY hits do not depend on the
// reflection waves, so they run beside them as many short-lived blocks that fill the latency-bound tails of those launches.
struct WaveSide { cudaStream_t stream; cudaEvent_t fork, join; };
cudaError_t launch_wavefront(const SceneDev& sc, const FrameParams& fp, const PlanesDev& pl, Counters* ctr, const WaveBuf& wb, int sms,
                             cudaStream_t stream, const WaveSide* side, uint32_t* launches,
                             cudaEvent_t* stageEvents /* 3: after primary, bounces, any-hit; or null */);
cudaError_t launch_detile(const void* src, void* dst, int W, int H, int world, size_t rankStridePixels, int bpp, cudaStream_t stream);
cudaError_t launch_push_tiles(const void* src, void* dst, int W, int H, int rank, int world, int localTiles, int bpp, cudaStream_t stream);

// brick acceleration build (vrt_accel.cu): bmask + bdist from the dense volume; tmpA/tmpB are scratch of Bx*By*Bz bytes
cudaError_t launch_build_accel(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, unsigned long long* bmask, uint8_t* bdist,
                               uint8_t* tmpA, uint8_t* tmpB, cudaStream_t stream, int* launches);

// per-voxel safe-step field (bricked fdist + linear aproned ldist) from the volume and bdist; a/t1/t2 are scratch of W*H*D bytes each
cudaError_t launch_build_fdist(const uint8_t* vol, const uint8_t* bdist, uint32_t W, uint32_t H, uint32_t D, uint8_t* fdist, uint8_t* ldist,
                               uint8_t* a, uint8_t* t1, uint8_t* t2, cudaStream_t stream, int* launches);

// directional safe-step fields (vrt_octant.cu): odist[0..7] of (W+2)(H+2)(D+2) bytes each, scratch of octant_scratch_bytes()
size_t octant_scratch_bytes(uint32_t W, uint32_t H, uint32_t D);
cudaError_t launch_build_odist(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, uint8_t* const odist[8], uint32_t* scratch,
                               cudaStream_t stream, int* launches);

// the same build one octant at a time, and the two-level (sparse) layout made from an octant's linear field (SceneDev::sdist)
cudaError_t launch_odist_init(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, uint32_t* scratch, cudaStream_t stream, int* launches);
cudaError_t launch_odist_octant(uint32_t W, uint32_t H, uint32_t D, int octant, uint8_t* out, uint32_t* scratch, cudaStream_t stream, int* launches);
void sparse_top_dims(uint32_t W, uint32_t H, uint32_t D, uint32_t tb[3]);
cudaError_t launch_sparse_classify(const uint8_t* lin, uint32_t W, uint32_t H, uint32_t D, uint32_t* top, uint32_t* fineCount, cudaStream_t stream);
cudaError_t launch_sparse_fill(const uint8_t* lin, uint32_t W, uint32_t H, uint32_t D, const uint32_t* top, uint8_t* fine, cudaStream_t stream);
cudaError_t launch_sparse_volume_classify(const uint8_t* vol, int generator, uint32_t seed, uint32_t W, uint32_t H, uint32_t D, uint32_t* top, uint32_t* fineCount, cudaStream_t stream);
cudaError_t launch_sparse_volume_fill(const uint8_t* vol, int generator, uint32_t seed, uint32_t W, uint32_t H, uint32_t D, const uint32_t* top, uint8_t* fine, cudaStream_t stream);
cudaError_t launch_sparse_volume_expand(const uint32_t* top, const uint8_t* fine, uint32_t W, uint32_t H, uint32_t D, uint32_t zBase, uint32_t slabD, uint8_t* out, cudaStream_t stream);
cudaError_t launch_odist_init_sparse(const uint32_t* top, const uint8_t* fine, uint32_t W, uint32_t H, uint32_t D, uint32_t* scratch, cudaStream_t stream, int* launches);
cudaError_t launch_sparse_expand(const uint32_t* top, const uint8_t* fine, uint32_t W, uint32_t H, uint32_t D, uint8_t* out, cudaStream_t stream);

// procedural volumes (vrt_generate.cu)
cudaError_t launch_generate(int generator, uint32_t seed, uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, cudaStream_t stream);
void generator_palette(int generator, uint32_t seed, vrt_material palette[256]);   // host

// denoiser (vrt_denoise.cu)
struct DenoiseTaps { int n; int ox[9], oy[9]; float k[9]; int dx[9], dy[9]; float fx[9], fy[9]; };   // taps with non-zero kernel weight, in loop order; per iteration (launch_denoise): fx/fy = offset*stepWidth, dx/dy the same as integers when all are integral
struct DenoiseIter { float phiColor, phiNormal, phiPos, stepWidth; };
cudaError_t launch_denoise(const uint8_t* color, const int8_t* normal, const float* position, uint8_t* out, int W, int H,
                           const DenoiseTaps& taps, const DenoiseIter& it, cudaStream_t stream);

// scene ingestion (vrt_flatten.cu): one instance of a .vox scene into the dense volume (cells: the model's voxels on the device)
cudaError_t launch_flatten_instance(const uint8_t* cells, uint8_t* vol, const FlattenInstance& f, cudaStream_t stream);

// output stage (vrt_output.cu): blit.frag; flags = VRT_BLIT_*
cudaError_t launch_blit(const void* src, int sw, int sh, void* dst, int tw, int th, uint32_t flags, cudaStream_t stream);

}  // namespace vrt
