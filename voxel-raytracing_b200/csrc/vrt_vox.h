// vrt_vox.h — host-side scene ingestion: MagicaVoxel .vox bytes -> dense palette-indexed volume + 256 materials,
// with the semantics of VoxelScene::VoxelScene (source/voxels/resource/voxel_scene.cpp:33-117) over the .vox
// container as the reference's parser reads it (thirdparty/opengametools/include/ogt_vox.h, read_flags = 0).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/vrt.h"

namespace vrt {

struct DenseScene {
    uint32_t w = 0, h = 0, d = 0;          // shader-space bounds: width = dx, height = dz(vox), depth = dy(vox)
    std::vector<uint8_t> volume;            // x + y*w + z*w*h
    vrt_material palette[256];
};

// Returns VRT_OK or a negative vrt_status; `err` receives a message on failure.
int load_vox_bytes(const uint8_t* bytes, size_t n, DenseScene& out, std::string& err);
int load_vox_file(const char* path, DenseScene& out, std::string& err);

}  // namespace vrt
