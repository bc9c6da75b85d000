// vrt_flatten.cu — scene ingestion on the GPU: the instance flatten of VoxelScene::VoxelScene
// (source/voxels/resource/voxel_scene.cpp:83-105) as a kernel. The host uploads every model ONCE (a .vox places the same model many
// times) and launches one kernel per instance, in the reference's instance order on one stream, so "a later instance overwrites an
// earlier one" holds without atomics; within an instance a signed-permutation transform maps distinct voxels to distinct cells.
// HBM-streaming: nx*ny*nz model bytes read (coalesced along x), one scattered byte store per solid voxel.
#include <cuda_runtime.h>
#include "vrt_flatten.h"
#include "vrt_launch.h"

namespace vrt {

__global__ void __launch_bounds__(256) flatten_instance_kernel(const uint8_t* __restrict__ cells, uint8_t* __restrict__ vol, const FlattenInstance f) {
    // 4 model voxels (consecutive in x) per thread: one 32-bit load when the row length allows it
    const size_t n = (size_t)f.nx * f.ny * f.nz;
    const size_t i0 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= n) return;
    uint32_t word = 0;
    if (i0 + 4 <= n && ((uintptr_t)(cells + i0) & 3u) == 0) word = *reinterpret_cast<const uint32_t*>(cells + i0);
    else for (int k = 0; k < 4 && i0 + k < n; k++) word |= (uint32_t)cells[i0 + k] << (8 * k);
    if (!word) return;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint8_t v = (uint8_t)(word >> (8 * k));
        if (!v) continue;
        const size_t i = i0 + k;
        const int x = (int)(i % f.nx), y = (int)((i / f.nx) % f.ny), z = (int)(i / ((size_t)f.nx * f.ny));
        const long long at = flatten_cell(f, x, y, z);
        if (at >= 0) vol[at] = v;
    }
}

cudaError_t launch_flatten_instance(const uint8_t* cells, uint8_t* vol, const FlattenInstance& f, cudaStream_t stream) {
    const size_t n = (size_t)f.nx * f.ny * f.nz;
    if (!n) return cudaSuccess;
    const size_t threads = (n + 3) / 4;
    flatten_instance_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(cells, vol, f);
    return cudaGetLastError();
}

}  // namespace vrt
