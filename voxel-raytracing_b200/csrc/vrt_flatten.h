// vrt_flatten.h — apply_vox_transform of the reference's flatten (source/voxels/resource/voxel_scene.cpp:9-31): a model voxel's cell
// centre, relative to the model's pivot, through the instance transform (row-vector convention, translation in row 3), floored to
// the scene cell. Shared by the host flatten (vrt_vox.cpp) and the device flatten (vrt_flatten.cu); sums in the reference's order.
#pragma once
#include <math.h>
#include "vrt_types.h"

namespace vrt {

VRT_HD void vox_transform(const float* t, const int* p, int x, int y, int z, int o[3]) {
    const float v0 = ((float)x + 0.5f) - (float)p[0], v1 = ((float)y + 0.5f) - (float)p[1], v2 = ((float)z + 0.5f) - (float)p[2];
    for (int k = 0; k < 3; k++) {
        const float s = (t[k] * v0 + t[4 + k] * v1) + (t[8 + k] * v2 + t[12 + k] * 1.0f);
        const float fl = floorf(s);
        o[k] = (fl != fl) ? 0 : (fl >= 2147483648.0f ? 2147483647 : (fl <= -2147483648.0f ? (-2147483647 - 1) : (int)fl));
    }
}

struct FlattenInstance { float xf[16]; int pivot[3]; int lo[3]; uint32_t nx, ny, nz; uint32_t W, H, D; };

// scene cell (shader-space layout x + y*W + z*W*H with the reference's Y/Z swap) of model voxel (x, y, z), or -1 when outside
VRT_HD long long flatten_cell(const FlattenInstance& f, int x, int y, int z) {
    int t[3]; vox_transform(f.xf, f.pivot, x, y, z, t);
    const long long at = (long long)(t[0] - f.lo[0]) + (long long)(t[2] - f.lo[2]) * f.W + (long long)(t[1] - f.lo[1]) * f.W * f.H;
    return (at >= 0 && at < (long long)f.W * f.H * f.D) ? at : -1;   // the reference writes unchecked (UB past the end); dropped here
}

}  // namespace vrt
