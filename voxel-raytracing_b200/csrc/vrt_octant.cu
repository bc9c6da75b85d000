// vrt_octant.cu — builds the directional safe-step fields (SceneDev::odist, vrt_types.h) on the GPU.
//
// odist[o][v] = largest level n of VRT_OCT_LEVEL_LIST such that the cube of side n anchored at v and extending towards
// octant o (offsets 0..n-1 per axis, in the octant's sign direction) contains no solid and no out-of-bounds voxel; 0 for a
// solid voxel; apron = VRT_LIN_OOB. Computed exactly on bit-volumes (1 bit per voxel, 32 voxels along x per word):
//   E_1(v)      = voxel v is empty (out of bounds reads 0)
//   E_{m+d}(v)  = AND over (a,b,c) in {0,d}^3 of E_m(v + sign_o * (a,b,c))        for d <= m  (8 cubes of side m cover side m+d)
// one fused kernel per level (x: funnel shifts inside the row, y/z: row offsets), then one pass that counts the levels each
// voxel satisfies (bit-sliced adder) and writes the byte field. Load-time only: HBM-streaming, ~200 launches per scene.
#include <cuda_runtime.h>
#include "vrt_launch.h"
#include "vrt_generate.cuh"

namespace vrt {

__constant__ uint8_t c_octLevels[VRT_OCT_NLEVELS + 1];

// level 0: emptiness bits
__global__ void oct_init_kernel(const uint8_t* __restrict__ vol, uint32_t* __restrict__ bits, uint32_t W, uint32_t H, uint32_t WW) {
    const uint32_t wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (wx >= WW || y >= H) return;
    const uint8_t* row = vol + (size_t)y * W + (size_t)z * W * H;
    uint32_t m = 0;
    const uint32_t x0 = wx * 32u;
    if (x0 + 32u <= W && (W & 15u) == 0) {
        const uint4 a = *reinterpret_cast<const uint4*>(row + x0), b = *reinterpret_cast<const uint4*>(row + x0 + 16);
        const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t v = w[k];
            m |= (((v & 0xffu) ? 0u : 1u) | ((v & 0xff00u) ? 0u : 2u) | ((v & 0xff0000u) ? 0u : 4u) | ((v & 0xff000000u) ? 0u : 8u)) << (4 * k);
        }
    } else {
        for (uint32_t b = 0; b < 32u && x0 + b < W; b++) if (row[x0 + b] == 0) m |= 1u << b;
    }
    bits[(size_t)wx + (size_t)y * WW + (size_t)z * WW * H] = m;
}

// row word wx of E shifted so that bit x holds E(x + sx*d): sx = +1 -> towards higher x
__device__ __forceinline__ uint32_t oct_shifted(const uint32_t* __restrict__ row, int wx, int WW, int d, int sx) {
    const int q = d >> 5, r = d & 31;
    if (sx > 0) {
        const int i0 = wx + q, i1 = wx + q + 1;
        const uint32_t lo = i0 < WW ? row[i0] : 0u, hi = (r && i1 < WW) ? row[i1] : 0u;
        return __funnelshift_r(lo, hi, r);
    }
    const int i1 = wx - q, i0 = wx - q - 1;
    const uint32_t hi = i1 >= 0 ? row[i1] : 0u, lo = (r && i0 >= 0) ? row[i0] : 0u;
    return __funnelshift_l(lo, hi, r);
}

// next = level m+d from prev = level m
__global__ void oct_level_kernel(const uint32_t* __restrict__ prev, uint32_t* __restrict__ next, int WW, int H, int D, int d, int sx, int sy, int sz) {
    const int wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (wx >= WW || y >= H) return;
    uint32_t acc = 0xffffffffu;
#pragma unroll
    for (int c = 0; c < 2; c++) {
        const int zz = z + (c ? sz * d : 0);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int yy = y + (b ? sy * d : 0);
            if (zz < 0 || zz >= D || yy < 0 || yy >= H) { acc = 0u; continue; }
            const uint32_t* row = prev + (size_t)yy * WW + (size_t)zz * WW * H;
            acc &= row[wx] & oct_shifted(row, wx, WW, d, sx);
        }
    }
    next[(size_t)wx + (size_t)y * WW + (size_t)z * WW * H] = acc;
}

// count the levels each voxel satisfies (levels are nested: E_n subset of E_m for n > m) and write the aproned byte field
__global__ void oct_final_kernel(const uint32_t* __restrict__ levels, size_t levelWords, uint8_t* __restrict__ out, uint32_t W, uint32_t H, uint32_t WW) {
    const uint32_t wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (wx >= WW || y >= H) return;
    const size_t i = (size_t)wx + (size_t)y * WW + (size_t)z * WW * H;
    uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
    for (int l = 0; l < VRT_OCT_NLEVELS; l++) {
        uint32_t w = levels[(size_t)l * levelWords + i], t;
        t = c0 & w; c0 ^= w; w = t;
        t = c1 & w; c1 ^= w; w = t;
        t = c2 & w; c2 ^= w; w = t;
        t = c3 & w; c3 ^= w; w = t;
        c4 ^= w;
    }
    uint8_t* dst = out + (size_t)(wx * 32u + 1u) + (size_t)(y + 1u) * (W + 2u) + (size_t)(z + 1u) * (W + 2u) * (H + 2u);
    for (uint32_t b = 0; b < 32u && wx * 32u + b < W; b++) {
        const uint32_t k = ((c0 >> b) & 1u) | (((c1 >> b) & 1u) << 1) | (((c2 >> b) & 1u) << 2) | (((c3 >> b) & 1u) << 3) | (((c4 >> b) & 1u) << 4);
        dst[b] = c_octLevels[k];
    }
}

size_t octant_scratch_bytes(uint32_t W, uint32_t H, uint32_t D) {
    const size_t WW = (W + 31u) / 32u;
    return (size_t)VRT_OCT_NLEVELS * WW * H * D * sizeof(uint32_t);
}

static cudaError_t upload_level_table(cudaStream_t stream) {
    static const int lv[VRT_OCT_NLEVELS] = VRT_OCT_LEVEL_LIST;
    uint8_t table[VRT_OCT_NLEVELS + 1];
    table[0] = 0;
    for (int l = 0; l < VRT_OCT_NLEVELS; l++) table[l + 1] = (uint8_t)safe_encode((uint32_t)lv[l]);
    return cudaMemcpyToSymbolAsync(c_octLevels, table, sizeof(table), 0, cudaMemcpyHostToDevice, stream);
}

// level 0 of the scratch (emptiness bits), shared by the eight octants
cudaError_t launch_odist_init(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, uint32_t* scratch, cudaStream_t stream, int* launches) {
    cudaError_t e = upload_level_table(stream);
    if (e != cudaSuccess) return e;
    const uint32_t WW = (W + 31u) / 32u;
    const dim3 block(32, 8), grid((WW + 31u) / 32u, (H + 7u) / 8u, D);
    oct_init_kernel<<<grid, block, 0, stream>>>(vol, scratch, W, H, WW);
    if (launches) *launches += 1;
    return cudaGetLastError();
}
// the aproned linear field of ONE octant from the scratch prepared by launch_odist_init
cudaError_t launch_odist_octant(uint32_t W, uint32_t H, uint32_t D, int o, uint8_t* out, uint32_t* scratch, cudaStream_t stream, int* launches) {
    static const int lv[VRT_OCT_NLEVELS] = VRT_OCT_LEVEL_LIST;
    const uint32_t WW = (W + 31u) / 32u;
    const size_t levelWords = (size_t)WW * H * D;
    const dim3 block(32, 8), grid((WW + 31u) / 32u, (H + 7u) / 8u, D);
    const int sx = (o & 1) ? -1 : 1, sy = (o & 2) ? -1 : 1, sz = (o & 4) ? -1 : 1;
    cudaError_t e = cudaMemsetAsync(out, (int)VRT_LIN_OOB, (size_t)(W + 2u) * (H + 2u) * (D + 2u), stream);   // apron
    if (e != cudaSuccess) return e;
    for (int l = 1; l < VRT_OCT_NLEVELS; l++)
        oct_level_kernel<<<grid, block, 0, stream>>>(scratch + (size_t)(l - 1) * levelWords, scratch + (size_t)l * levelWords, (int)WW, (int)H, (int)D,
                                                     lv[l] - lv[l - 1], sx, sy, sz);
    oct_final_kernel<<<grid, block, 0, stream>>>(scratch, levelWords, out, W, H, WW);
    if (launches) *launches += VRT_OCT_NLEVELS;
    return cudaGetLastError();
}

// ---- two-level (sparse) layout of one octant's field: SceneDev::sdist (vrt_types.h) ----
// One block of 64 threads per 8x8x8 brick; thread t owns the 8 voxels of row (y, z) = (t & 7, t >> 3). `lin` is the aproned linear
// field; brick coordinates are shifted by one brick (voxel x of brick bx: bx*8 - 8 + i), voxels beyond the apron count as outside.
// where a brick's 512 byte codes come from
struct SrcField {     // an octant's aproned linear safe-step field; beyond the apron = outside
    const uint8_t* lin; int W, H, D;
    __device__ __forceinline__ uint32_t operator()(int x, int y, int z) const {
        if (x < -1 || y < -1 || z < -1 || x > W || y > H || z > D) return VRT_LIN_OOB;
        return lin[(size_t)(x + 1) + (size_t)(y + 1) * (size_t)(W + 2) + (size_t)(z + 1) * (size_t)(W + 2) * (size_t)(H + 2)];
    }
};
struct SrcDenseVol {  // the dense palette volume; outside = 0 (empty)
    const uint8_t* vol; int W, H, D;
    __device__ __forceinline__ uint32_t operator()(int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= W || y >= H || z >= D) return 0u;
        return vol[(size_t)x + (size_t)y * (size_t)W + (size_t)z * (size_t)W * (size_t)H];
    }
};
struct SrcGenerator { // a procedural volume, evaluated in place: the dense volume is never materialised
    int generator; uint32_t seed; int W, H, D;
    __device__ __forceinline__ uint32_t operator()(int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= W || y >= H || z >= D) return 0u;
        return gen_voxel(generator, seed, (uint32_t)x, (uint32_t)y, (uint32_t)z, (uint32_t)W, (uint32_t)H, (uint32_t)D);
    }
};
// pass 1: classify every brick; bricks that need per-voxel codes get a slot of the pool (atomic counter). FIELD: the codes are safe-step
// budgets, so a brick without solid / outside voxels whose minimum is at least VRT_SP_UNIFORM_MIN may stand for all of its voxels;
// a volume brick (palette indices) is uniform only if all 512 values are equal.
template <class SRC, bool FIELD>
__global__ void __launch_bounds__(64) sp_classify_kernel(const SRC src, uint32_t TBx, uint32_t TBy, uint32_t* __restrict__ top, uint32_t* __restrict__ fineCount) {
    const uint32_t bx = blockIdx.x, by = blockIdx.y, bz = blockIdx.z, t = threadIdx.x;
    const int y = (int)by * 8 - 8 + (int)(t & 7u), z = (int)bz * 8 - 8 + (int)(t >> 3);
    uint32_t mn = 0xffffffffu, first = 0, same = 1, special = 0;
    for (int i = 0; i < 8; i++) {
        const uint32_t c = src((int)bx * 8 - 8 + i, y, z);
        if (i == 0) first = c;
        same &= c == first ? 1u : 0u;
        special |= (c == 0u || c == VRT_LIN_OOB) ? 1u : 0u;
        mn = min(mn, safe_decode(c));
    }
    __shared__ uint32_t sFirst[64], sMin[2], sFlags[2];
    sFirst[t] = first;
    const uint32_t wmin = __reduce_min_sync(0xffffffffu, mn), wsame = __reduce_and_sync(0xffffffffu, same), wspecial = __reduce_or_sync(0xffffffffu, special);
    if ((t & 31u) == 0) { sMin[t >> 5] = wmin; sFlags[t >> 5] = wsame | (wspecial << 1); }
    __syncthreads();
    if (t == 0) {
        bool allSame = (sFlags[0] & 1u) && (sFlags[1] & 1u);
        for (int k = 1; k < 64 && allSame; k++) allSame = sFirst[k] == sFirst[0];
        const bool anySpecial = ((sFlags[0] | sFlags[1]) >> 1) != 0;
        const uint32_t m = min(sMin[0], sMin[1]);
        uint32_t e;
        if (allSame) e = sFirst[0];
        else if (FIELD && !anySpecial && m >= VRT_SP_UNIFORM_MIN) e = safe_encode(m);   // the minimum is a valid lower bound for every voxel of the brick
        else e = VRT_SP_FINE | atomicAdd(fineCount, 1u);
        top[bx + (by + (size_t)bz * TBy) * TBx] = e;
    }
}
// pass 2: per-voxel codes of the bricks that were given a slot
template <class SRC>
__global__ void __launch_bounds__(64) sp_fill_kernel(const SRC src, uint32_t TBx, uint32_t TBy, const uint32_t* __restrict__ top, uint8_t* __restrict__ fine) {
    const uint32_t bx = blockIdx.x, by = blockIdx.y, bz = blockIdx.z, t = threadIdx.x;
    const uint32_t e = top[bx + (by + (size_t)bz * TBy) * TBx];
    if (!(e & VRT_SP_FINE)) return;
    const int y = (int)by * 8 - 8 + (int)(t & 7u), z = (int)bz * 8 - 8 + (int)(t >> 3);
    unsigned long long row = 0ull;
    for (int i = 0; i < 8; i++) row |= (unsigned long long)src((int)bx * 8 - 8 + i, y, z) << (8 * i);
    reinterpret_cast<unsigned long long*>(fine + (size_t)(e & ~VRT_SP_FINE) * 512u)[t] = row;   // (x&7) | (y&7)<<3 | (z&7)<<6
}
__device__ __forceinline__ uint32_t sp_fetch(const uint32_t* __restrict__ top, const uint8_t* __restrict__ fine, uint32_t TBx, uint32_t TBy, int x, int y, int z) {
    const uint32_t X = (uint32_t)(x + 8), Y = (uint32_t)(y + 8), Z = (uint32_t)(z + 8);
    const uint32_t e = top[(X >> 3) + ((Y >> 3) + (size_t)(Z >> 3) * TBy) * TBx];
    return (e & VRT_SP_FINE) ? fine[(size_t)(e & ~VRT_SP_FINE) * 512u + ((X & 7u) | ((Y & 7u) << 3) | ((Z & 7u) << 6))] : e;
}
// emptiness bits (level 0 of the octant build) and a dense copy (tests) from a two-level volume
__global__ void oct_init_sparse_kernel(const uint32_t* __restrict__ top, const uint8_t* __restrict__ fine, uint32_t TBx, uint32_t TBy, uint32_t* __restrict__ bits,
                                       uint32_t W, uint32_t H, uint32_t WW) {
    const uint32_t wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (wx >= WW || y >= H) return;
    uint32_t m = 0;
    for (uint32_t b = 0; b < 32u && wx * 32u + b < W; b++) if (sp_fetch(top, fine, TBx, TBy, (int)(wx * 32u + b), (int)y, (int)z) == 0u) m |= 1u << b;
    bits[(size_t)wx + (size_t)y * WW + (size_t)z * WW * H] = m;
}
__global__ void sp_expand_volume_kernel(const uint32_t* __restrict__ top, const uint8_t* __restrict__ fine, uint32_t TBx, uint32_t TBy, uint8_t* __restrict__ out,
                                        uint32_t W, uint32_t H, uint32_t zBase) {
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
    if (x >= W) return;
    out[(size_t)x + (size_t)y * W + (size_t)z * W * H] = (uint8_t)sp_fetch(top, fine, TBx, TBy, (int)x, (int)y, (int)(z + zBase));
}
// expand a sparse field back to the aproned linear layout (tests)
__global__ void sp_expand_kernel(const uint32_t* __restrict__ top, const uint8_t* __restrict__ fine, uint32_t TBx, uint32_t TBy, uint8_t* __restrict__ out, int W, int H, int D) {
    const int x = (int)(blockIdx.x * blockDim.x + threadIdx.x) - 1, y = (int)blockIdx.y - 1, z = (int)blockIdx.z - 1;
    if (x > W) return;
    const uint32_t X = (uint32_t)(x + 8), Y = (uint32_t)(y + 8), Z = (uint32_t)(z + 8);
    const uint32_t e = top[(X >> 3) + ((Y >> 3) + (size_t)(Z >> 3) * TBy) * TBx];
    const uint32_t c = (e & VRT_SP_FINE) ? fine[(size_t)(e & ~VRT_SP_FINE) * 512u + ((X & 7u) | ((Y & 7u) << 3) | ((Z & 7u) << 6))] : e;
    out[(size_t)(x + 1) + (size_t)(y + 1) * (size_t)(W + 2) + (size_t)(z + 1) * (size_t)(W + 2) * (size_t)(H + 2)] = (uint8_t)c;
}

void sparse_top_dims(uint32_t W, uint32_t H, uint32_t D, uint32_t tb[3]) { tb[0] = ((W + 8u) >> 3) + 1u; tb[1] = ((H + 8u) >> 3) + 1u; tb[2] = ((D + 8u) >> 3) + 1u; }
cudaError_t launch_sparse_classify(const uint8_t* lin, uint32_t W, uint32_t H, uint32_t D, uint32_t* top, uint32_t* fineCount, cudaStream_t stream) {
    uint32_t tb[3]; sparse_top_dims(W, H, D, tb);
    sp_classify_kernel<SrcField, true><<<dim3(tb[0], tb[1], tb[2]), 64, 0, stream>>>(SrcField{lin, (int)W, (int)H, (int)D}, tb[0], tb[1], top, fineCount);
    return cudaGetLastError();
}
cudaError_t launch_sparse_fill(const uint8_t* lin, uint32_t W, uint32_t H, uint32_t D, const uint32_t* top, uint8_t* fine, cudaStream_t stream) {
    uint32_t tb[3]; sparse_top_dims(W, H, D, tb);
    sp_fill_kernel<SrcField><<<dim3(tb[0], tb[1], tb[2]), 64, 0, stream>>>(SrcField{lin, (int)W, (int)H, (int)D}, tb[0], tb[1], top, fine);
    return cudaGetLastError();
}
// the palette volume itself in the two-level layout: from a dense volume (vol != null) or straight from a generator (vol == null)
cudaError_t launch_sparse_volume_classify(const uint8_t* vol, int generator, uint32_t seed, uint32_t W, uint32_t H, uint32_t D, uint32_t* top, uint32_t* fineCount, cudaStream_t stream) {
    uint32_t tb[3]; sparse_top_dims(W, H, D, tb);
    const dim3 grid(tb[0], tb[1], tb[2]);
    if (vol) sp_classify_kernel<SrcDenseVol, false><<<grid, 64, 0, stream>>>(SrcDenseVol{vol, (int)W, (int)H, (int)D}, tb[0], tb[1], top, fineCount);
    else sp_classify_kernel<SrcGenerator, false><<<grid, 64, 0, stream>>>(SrcGenerator{generator, seed, (int)W, (int)H, (int)D}, tb[0], tb[1], top, fineCount);
    return cudaGetLastError();
}
cudaError_t launch_sparse_volume_fill(const uint8_t* vol, int generator, uint32_t seed, uint32_t W, uint32_t H, uint32_t D, const uint32_t* top, uint8_t* fine, cudaStream_t stream) {
    uint32_t tb[3]; sparse_top_dims(W, H, D, tb);
    const dim3 grid(tb[0], tb[1], tb[2]);
    if (vol) sp_fill_kernel<SrcDenseVol><<<grid, 64, 0, stream>>>(SrcDenseVol{vol, (int)W, (int)H, (int)D}, tb[0], tb[1], top, fine);
    else sp_fill_kernel<SrcGenerator><<<grid, 64, 0, stream>>>(SrcGenerator{generator, seed, (int)W, (int)H, (int)D}, tb[0], tb[1], top, fine);
    return cudaGetLastError();
}
// z-slab [zBase, zBase + slabD) of a two-level volume as dense bytes (vrt_scene_download_volume)
cudaError_t launch_sparse_volume_expand(const uint32_t* top, const uint8_t* fine, uint32_t W, uint32_t H, uint32_t D, uint32_t zBase, uint32_t slabD, uint8_t* out, cudaStream_t stream) {
    uint32_t tb[3]; sparse_top_dims(W, H, D, tb);
    sp_expand_volume_kernel<<<dim3((W + 127u) / 128u, H, slabD), 128, 0, stream>>>(top, fine, tb[0], tb[1], out, W, H, zBase);
    return cudaGetLastError();
}
// level 0 of the octant build from a two-level volume
cudaError_t launch_odist_init_sparse(const uint32_t* top, const uint8_t* fine, uint32_t W, uint32_t H, uint32_t D, uint32_t* scratch, cudaStream_t stream, int* launches) {
    cudaError_t e = upload_level_table(stream);
    if (e != cudaSuccess) return e;
    uint32_t tb[3]; sparse_top_dims(W, H, D, tb);
    const uint32_t WW = (W + 31u) / 32u;
    const dim3 block(32, 8), grid((WW + 31u) / 32u, (H + 7u) / 8u, D);
    oct_init_sparse_kernel<<<grid, block, 0, stream>>>(top, fine, tb[0], tb[1], scratch, W, H, WW);
    if (launches) *launches += 1;
    return cudaGetLastError();
}
cudaError_t launch_sparse_expand(const uint32_t* top, const uint8_t* fine, uint32_t W, uint32_t H, uint32_t D, uint8_t* out, cudaStream_t stream) {
    uint32_t tb[3]; sparse_top_dims(W, H, D, tb);
    sp_expand_kernel<<<dim3((W + 2u + 127u) / 128u, H + 2u, D + 2u), 128, 0, stream>>>(top, fine, tb[0], tb[1], out, (int)W, (int)H, (int)D);
    return cudaGetLastError();
}

// odist[0..7]: (W+2)(H+2)(D+2) bytes each; scratch: octant_scratch_bytes()
cudaError_t launch_build_odist(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, uint8_t* const odist[8], uint32_t* scratch,
                               cudaStream_t stream, int* launches) {
    cudaError_t e = launch_odist_init(vol, W, H, D, scratch, stream, launches);
    for (int o = 0; o < 8 && e == cudaSuccess; o++) e = launch_odist_octant(W, H, D, o, odist[o], scratch, stream, launches);
    return e;
}

}  // namespace vrt
