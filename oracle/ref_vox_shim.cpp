// ORACLE — TEST INFRASTRUCTURE ONLY. Compiles the reference's vendored .vox parser UNMODIFIED from where it
// lies (/root/reference/thirdparty/opengametools/include/ogt_vox.h, via -I) into oracle/_ref/libref_vox.so.
// No reference source is copied into this repository. Used to (1) pin oracle/orc_scene.cpp and the product
// loader against the real parser, (2) write .vox fixtures with the reference's own writer.
// The flatten below is voxel_scene.cpp:9-117 written against the real ogt structs exactly as the reference does
// (voxel_scene.cpp itself needs glm + Vulkan and cannot be compiled here: SURVEY.md F3).
#define OGT_VOX_IMPLEMENTATION
#include "ogt_vox.h"
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>

static void pivot_of(const ogt_vox_model* m, int pv[3]) {  // voxel_scene.cpp:9-16
    pv[0] = (int)floorf(m->size_x / 2.0f); pv[1] = (int)floorf(m->size_y / 2.0f); pv[2] = (int)floorf(m->size_z / 2.0f);
}
static void apply_xf(const ogt_vox_transform& t, const int pv[3], int x, int y, int z, int out[3]) {  // :18-31
    const float c0[4] = {t.m00, t.m01, t.m02, t.m03}, c1[4] = {t.m10, t.m11, t.m12, t.m13};
    const float c2[4] = {t.m20, t.m21, t.m22, t.m23}, c3[4] = {t.m30, t.m31, t.m32, t.m33};
    const float v[4] = {((float)x + 0.5f) - (float)pv[0], ((float)y + 0.5f) - (float)pv[1], ((float)z + 0.5f) - (float)pv[2], 1.0f};
    for (int k = 0; k < 3; k++) out[k] = (int)floorf((c0[k] * v[0] + c1[k] * v[1]) + (c2[k] * v[2] + c3[k] * v[3]));
}

extern "C" int ref_load_vox(const uint8_t* bytes, size_t n, uint8_t** volume, uint32_t whd[3], float palette[256 * 8]) {
    const ogt_vox_scene* sc = ogt_vox_read_scene(bytes, (uint32_t)n);
    if (!sc) return -1;
    if (sc->num_instances < 1) { ogt_vox_destroy_scene(sc); return -3; }
    int mn[3] = {100000, 100000, 100000}, mx[3] = {-100000, -100000, -100000};
    for (uint32_t i = 0; i < sc->num_instances; i++) {
        const ogt_vox_instance* in = &sc->instances[i];
        const ogt_vox_model* m = sc->models[in->model_index];
        ogt_vox_transform xf = ogt_vox_sample_instance_transform(in, 0, sc);
        int pv[3], c1[3], c2[3]; pivot_of(m, pv);
        apply_xf(xf, pv, 0, 0, 0, c1);
        apply_xf(xf, pv, (int)m->size_x, (int)m->size_y, (int)m->size_z, c2);
        for (int k = 0; k < 3; k++) { mn[k] = std::min(mn[k], std::min(c1[k], c2[k])); mx[k] = std::max(mx[k], std::max(c1[k], c2[k])); }
    }
    uint32_t width = mx[0] - mn[0], height = mx[2] - mn[2], depth = mx[1] - mn[1];
    size_t size = (size_t)width * height * depth;
    uint8_t* scene = (uint8_t*)calloc(size ? size : 1, 1);
    for (uint32_t i = 0; i < sc->num_instances; i++) {
        const ogt_vox_instance* in = &sc->instances[i];
        const ogt_vox_model* m = sc->models[in->model_index];
        int pv[3]; pivot_of(m, pv);
        for (uint32_t x = 0; x < m->size_x; x++)
            for (uint32_t y = 0; y < m->size_y; y++)
                for (uint32_t z = 0; z < m->size_z; z++) {
                    uint8_t vox = m->voxel_data[x + (y * m->size_x) + (z * m->size_x * m->size_y)];
                    if (vox != 0) {
                        ogt_vox_transform xf = ogt_vox_sample_instance_transform(in, 0, sc);
                        int t[3]; apply_xf(xf, pv, (int)x, (int)y, (int)z, t);
                        long long pos = (long long)(t[0] - mn[0]) + (long long)(t[2] - mn[2]) * width + (long long)(t[1] - mn[1]) * width * height;
                        if (pos >= 0 && (size_t)pos < size) scene[pos] = vox;  // reference writes unchecked; see orc_scene.cpp
                    }
                }
    }
    for (int m = 0; m < 256; m++) {
        const ogt_vox_rgba c = sc->palette.color[m];
        float* o = palette + 8 * m;
        o[0] = powf(c.r / 255.0f, 2.2f); o[1] = powf(c.g / 255.0f, 2.2f); o[2] = powf(c.b / 255.0f, 2.2f); o[3] = powf(c.a / 255.0f, 2.2f);
        o[4] = sc->materials.matl[m].metal; o[5] = o[6] = o[7] = 0.0f;
    }
    ogt_vox_destroy_scene(sc);
    *volume = scene; whd[0] = width; whd[1] = height; whd[2] = depth;
    return 0;
}

// Write a .vox with the reference's own writer (ogt_vox_write_scene, ogt_vox.h:2169).
// models: n_models x {sx,sy,sz}; voxels concatenated (x fastest, then y, then z).
// instances: n_inst x {model_index, rot[9] as m00 m01 m02 m10 ... m22, tx,ty,tz}.  palette_rgba: 256*4 (index i = colour of voxel value i)
// metal[256]: < 0 means "no MATL chunk for this index".
extern "C" uint8_t* ref_write_vox(uint32_t n_models, const uint32_t* sizes, const uint8_t* voxels, uint32_t n_inst,
                                  const uint32_t* inst_model, const float* inst_xf, const uint8_t* palette_rgba,
                                  const float* metal, uint32_t* out_size) {
    std::vector<ogt_vox_model> models(n_models);
    std::vector<const ogt_vox_model*> mptr(n_models);
    size_t off = 0;
    for (uint32_t i = 0; i < n_models; i++) {
        models[i] = {sizes[3 * i], sizes[3 * i + 1], sizes[3 * i + 2], 0, voxels + off};
        off += (size_t)sizes[3 * i] * sizes[3 * i + 1] * sizes[3 * i + 2];
        mptr[i] = &models[i];
    }
    std::vector<ogt_vox_instance> inst(n_inst);
    for (uint32_t i = 0; i < n_inst; i++) {
        memset(&inst[i], 0, sizeof(ogt_vox_instance));
        const float* x = inst_xf + 12 * i;
        ogt_vox_transform t = {x[0], x[1], x[2], 0, x[3], x[4], x[5], 0, x[6], x[7], x[8], 0, x[9], x[10], x[11], 1};
        inst[i].transform = t; inst[i].model_index = inst_model[i]; inst[i].layer_index = 0; inst[i].group_index = 0;
    }
    ogt_vox_layer layer; memset(&layer, 0, sizeof(layer));
    ogt_vox_group group; memset(&group, 0, sizeof(group));
    group.transform = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    group.parent_group_index = k_invalid_group_index;
    ogt_vox_scene sc; memset(&sc, 0, sizeof(sc));
    sc.num_models = n_models; sc.models = mptr.data();
    sc.num_instances = n_inst; sc.instances = inst.data();
    sc.num_layers = 1; sc.layers = &layer;
    sc.num_groups = 1; sc.groups = &group;
    memcpy(&sc.palette, palette_rgba, 1024);
    for (int i = 0; i < 256; i++)
        if (metal && metal[i] >= 0.0f) {
            sc.materials.matl[i].content_flags = k_ogt_vox_matl_have_metal;
            sc.materials.matl[i].type = ogt_matl_type_metal;
            sc.materials.matl[i].metal = metal[i];
        }
    return ogt_vox_write_scene(&sc, out_size);
}
extern "C" void ref_free(void* p) { free(p); }
extern "C" void ref_default_palette(uint8_t out[1024]) { memcpy(out, k_default_vox_palette, 1024); }
