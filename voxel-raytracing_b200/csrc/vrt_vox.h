// vrt_vox.h — host-side scene ingestion: MagicaVoxel .vox bytes -> dense palette-indexed volume + 256 materials,
// with the semantics of VoxelScene::VoxelScene (source/voxels/resource/voxel_scene.cpp:33-117) over the .vox
// container as the reference's parser reads it (thirdparty/opengametools/include/ogt_vox.h, read_flags = 0).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/vrt.h"
#include "vrt_flatten.h"

namespace vrt {

struct DenseScene {
    uint32_t w = 0, h = 0, d = 0;          // shader-space bounds: width = dx, height = dz(vox), depth = dy(vox)
    std::vector<uint8_t> volume;            // x + y*w + z*w*h
    vrt_material palette[256];
};

// A parsed .vox before the flatten: unique models, the instances in the order the reference flattens them (a later instance overwrites
// an earlier one, voxel_scene.cpp:83-105), the scene bounds and the materials. This is what the GPU flatten (vrt_flatten.cu) uploads:
// each model once, however many instances place it.
struct VoxLayout {
    struct Model { uint32_t nx = 0, ny = 0, nz = 0; std::vector<uint8_t> cells; };   // x + y*nx + z*nx*ny, .vox axes
    struct Instance { uint32_t model = 0; float xf[16]; int pivot[3]; bool permutation = true; };
    std::vector<Model> models;
    std::vector<Instance> instances;
    int lo[3] = {0, 0, 0};                   // min corner in .vox axes (voxel_scene.cpp:53-72)
    uint32_t w = 0, h = 0, d = 0;
    vrt_material palette[256];
};
int parse_vox_bytes(const uint8_t* bytes, size_t n, VoxLayout& out, std::string& err);
int parse_vox_file(const char* path, VoxLayout& out, std::string& err);
FlattenInstance flatten_instance_params(const VoxLayout& lay, size_t instance);   // kernel / loop parameters of one instance
void flatten_host(const VoxLayout& lay, uint8_t* volume);   // volume: w*h*d bytes, zeroed by the callee

// Returns VRT_OK or a negative vrt_status; `err` receives a message on failure.
int load_vox_bytes(const uint8_t* bytes, size_t n, DenseScene& out, std::string& err);
int load_vox_file(const char* path, DenseScene& out, std::string& err);

}  // namespace vrt
