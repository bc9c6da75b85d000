// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h). CPU restatement of the scene ingestion that feeds the
// shader: /root/reference/source/voxels/resource/voxel_scene.cpp:9-117 (pivot, transform, bounds, Y/Z swap,
// dense fill, palette) on top of an independent reader of the MagicaVoxel .vox container that follows the
// behaviour of the vendored third-party parser /root/reference/thirdparty/opengametools/include/ogt_vox.h
// (ogt_vox_read_scene with read_flags = 0: :1179-1990). That parser itself is compiled unmodified into
// oracle/_ref (see oracle/ref_vox_shim.cpp) and tests pin this file against it.
#include "vrt_oracle.h"
#include "orc_math.h"
#include <cstdlib>
#include <string>
#include <vector>

namespace {

struct Xf { float m[4][4]; };  // ogt_vox_transform: m[r][c] = m<r><c>, row-vector convention (translation in row 3)
Xf xf_identity() { Xf t{}; for (int i = 0; i < 4; i++) t.m[i][i] = 1.0f; return t; }
Xf xf_mul(const Xf& a, const Xf& b) {  // ogt_vox.h:749-768
    Xf r{};
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            r.m[i][j] = (a.m[i][0] * b.m[0][j]) + (a.m[i][1] * b.m[1][j]) + (a.m[i][2] * b.m[2][j]) + (a.m[i][3] * b.m[3][j]);
    return r;
}

struct Model { uint32_t sx = 0, sy = 0, sz = 0; std::vector<uint8_t> vox; bool null_model = true; };
enum NodeType { N_INVALID = 0, N_GROUP, N_TRANSFORM, N_SHAPE };
struct Node { NodeType type = N_INVALID; Xf xf = xf_identity(); uint32_t child = 0; std::vector<uint32_t> children; uint32_t model = 0; };
struct Instance { uint32_t model; Xf xf; };
struct Group { Xf xf; int parent; };

struct Reader {
    const uint8_t* p; size_t n, off = 0;
    bool ok = true;
    size_t remaining() const { return n - off; }
    uint32_t u32() { if (remaining() < 4) { ok = false; off = n; return 0; } uint32_t v; memcpy(&v, p + off, 4); off += 4; return v; }
    std::string str() { uint32_t len = u32(); if (remaining() < len) { ok = false; off = n; return ""; } std::string s((const char*)p + off, len); off += len; return s; }
    void skip(size_t k) { off = (k > remaining()) ? n : off + k; }
};
typedef std::vector<std::pair<std::string, std::string>> Dict;
Dict read_dict(Reader& r) {  // ogt_vox.h:781-825
    Dict d; uint32_t np = r.u32();
    for (uint32_t i = 0; i < np && r.ok; i++) { std::string k = r.str(); std::string v = r.str(); d.emplace_back(k, v); }
    return d;
}
const std::string* dict_get(const Dict& d, const char* key) {  // case-insensitive, ogt_vox.h:828-833
    for (auto& kv : d) {
        const std::string& k = kv.first; size_t i = 0;
        for (; i < k.size() && key[i]; i++) if (tolower((unsigned char)k[i]) != tolower((unsigned char)key[i])) break;
        if (i == k.size() && key[i] == 0) return &kv.second;
    }
    return nullptr;
}

// ogt_vox.h:861-907: packed rotation byte -> rows; rows become columns of the stored transform
Xf make_transform(const std::string* rot, const std::string* trans) {
    Xf t = xf_identity();
    if (rot) {
        static const float kv[4][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, 0}};
        static const int row2[8] = {-1, -1, -1, 2, -1, 1, 0, -1};
        uint32_t bits = (uint32_t)atoi(rot->c_str());
        uint32_t i0 = bits & 3, i1 = (bits >> 2) & 3;
        int i2 = row2[((1u << i0) | (1u << i1)) & 7];
        if (i2 < 0) i2 = 3;
        float r0[3], r1[3], r2[3];
        for (int k = 0; k < 3; k++) {
            r0[k] = (bits & 16) ? -kv[i0][k] : kv[i0][k];
            r1[k] = (bits & 32) ? -kv[i1][k] : kv[i1][k];
            r2[k] = (bits & 64) ? -kv[i2][k] : kv[i2][k];
        }
        for (int k = 0; k < 3; k++) { t.m[k][0] = r0[k]; t.m[k][1] = r1[k]; t.m[k][2] = r2[k]; }
    }
    if (trans) {
        int x = 0, y = 0, z = 0;
        sscanf(trans->c_str(), "%i %i %i", &x, &y, &z);
        t.m[3][0] = (float)x; t.m[3][1] = (float)y; t.m[3][2] = (float)z;
    }
    return t;
}

// ogt_vox.h:968-1064 (generate_instances_for_node), read_flags = 0
void gen_instances(const std::vector<Node>& nodes, uint32_t idx, std::vector<const Node*>& stack, const std::vector<Model>& models,
                   std::vector<Instance>& inst, std::vector<Group>& groups, int group_index, int depth) {
    if (idx >= nodes.size() || depth > 256) return;
    const Node* node = &nodes[idx];
    switch (node->type) {
        case N_TRANSFORM:
            stack.push_back(node);
            gen_instances(nodes, node->child, stack, models, inst, groups, group_index, depth + 1);
            stack.pop_back();
            break;
        case N_GROUP: {
            const Node* last = stack.empty() ? nullptr : stack.back();
            int next = (int)groups.size();
            groups.push_back({last ? last->xf : xf_identity(), group_index});
            stack.push_back(node);
            for (uint32_t c : node->children) gen_instances(nodes, c, stack, models, inst, groups, next, depth + 1);
            stack.pop_back();
            break;
        }
        case N_SHAPE:
            if (node->model < models.size() && !models[node->model].null_model) {
                const Node* last = stack.empty() ? nullptr : stack.back();
                Xf t = last ? last->xf : xf_identity();
                // flatten through the group chain (ogt_vox.h:1737-1747)
                int g = group_index;
                while (g >= 0) { t = xf_mul(t, groups[g].xf); g = groups[g].parent; }
                inst.push_back({node->model, t});
            }
            break;
        default: break;
    }
}

// MagicaVoxel default palette (file order, i.e. before the rotate of ogt_vox.h:1834-1840): the 6x6x6 colour
// cube without black, then red / green / blue / grey ramps, then black. Generated, and pinned against the
// table at ogt_vox.h:474 by tests (through oracle/_ref).
void default_palette(uint8_t pal[256 * 4]) {
    static const uint8_t ramp[10] = {0xee, 0xdd, 0xbb, 0xaa, 0x88, 0x77, 0x55, 0x44, 0x22, 0x11};
    int k = 0;
    for (int i = 0; i < 215; i++, k++) {
        pal[4 * k + 0] = (uint8_t)(0xff - 0x33 * (i / 36)); pal[4 * k + 1] = (uint8_t)(0xff - 0x33 * ((i / 6) % 6));
        pal[4 * k + 2] = (uint8_t)(0xff - 0x33 * (i % 6)); pal[4 * k + 3] = 0xff;
    }
    for (int c = 0; c < 4; c++)
        for (int j = 0; j < 10; j++, k++) {
            for (int q = 0; q < 3; q++) pal[4 * k + q] = (c == 3 || c == q) ? ramp[j] : 0;
            pal[4 * k + 3] = 0xff;
        }
    pal[4 * k + 0] = pal[4 * k + 1] = pal[4 * k + 2] = 0; pal[4 * k + 3] = 0xff;
}

}  // namespace

static inline uint32_t fourcc(const char* s) { return (uint32_t)s[0] | ((uint32_t)s[1] << 8) | ((uint32_t)s[2] << 16) | ((uint32_t)s[3] << 24); }

extern "C" void orc_free(void* p) { free(p); }

extern "C" int orc_load_vox(const uint8_t* bytes, size_t n, uint8_t** volume, uint32_t whd[3], float palette[256 * 8]) {
    if (!bytes || !volume || !whd || !palette) return -1;
    static const char lfs[] = "version https://git-lfs";
    if (n >= sizeof(lfs) - 1 && memcmp(bytes, lfs, sizeof(lfs) - 1) == 0) return -2;
    Reader r{bytes, n};
    uint32_t hdr = r.u32(), ver = r.u32();
    if (!r.ok || hdr != fourcc("VOX ") || (ver != 150 && ver != 200)) return -1;  // ogt_vox.h:1228

    std::vector<Model> models;
    std::vector<Node> nodes;
    uint8_t pal[256 * 4]; default_palette(pal);
    float metal[256] = {0};
    uint8_t imap[256]; bool have_imap = false;
    uint32_t sx = 0, sy = 0, sz = 0;
    auto node_at = [&](uint32_t id) -> Node& { if (id >= nodes.size()) nodes.resize(id + 1); return nodes[id]; };

    while (r.remaining() >= 12) {  // ogt_vox.h:1232-1671
        uint32_t id = r.u32(), size = r.u32(); (void)r.u32();
        size_t body = r.off;
        if (id == fourcc("MAIN")) { continue; }
        else if (id == fourcc("SIZE")) { sx = r.u32(); sy = r.u32(); sz = r.u32(); }
        else if (id == fourcc("XYZI")) {
            uint32_t nv = r.u32();
            Model m;
            if (nv != 0) {
                m.null_model = false; m.sx = sx; m.sy = sy; m.sz = sz;
                m.vox.assign((size_t)sx * sy * sz, 0);
                uint32_t can = (uint32_t)std::min<size_t>(r.remaining() / 4, nv);
                const uint8_t* q = bytes + r.off;
                for (uint32_t i = 0; i < can; i++) {
                    uint8_t x = q[4 * i], y = q[4 * i + 1], z = q[4 * i + 2], c = q[4 * i + 3];
                    if (x < sx && y < sy && z < sz) m.vox[x + (size_t)y * sx + (size_t)z * sx * sy] = c;
                }
            }
            models.push_back(std::move(m));
            r.off = body; r.skip(4 + (size_t)nv * 4);
            continue;
        }
        else if (id == fourcc("RGBA")) { if (r.remaining() >= 1024) memcpy(pal, bytes + r.off, 1024); r.skip(1024); continue; }
        else if (id == fourcc("nTRN")) {
            uint32_t nid = r.u32(); (void)read_dict(r);
            uint32_t child = r.u32(); (void)r.u32(); (void)r.u32(); uint32_t nf = r.u32();
            Xf first = xf_identity();
            for (uint32_t f = 0; f < nf && r.ok; f++) { Dict d = read_dict(r); if (f == 0) first = make_transform(dict_get(d, "_r"), dict_get(d, "_t")); }
            Node& nd = node_at(nid); nd.type = N_TRANSFORM; nd.child = child; nd.xf = first;
            continue;  // cursor is already past the chunk (ogt does not re-seek either)
        }
        else if (id == fourcc("nGRP")) {
            uint32_t nid = r.u32(); (void)read_dict(r);
            uint32_t nc = r.u32();
            std::vector<uint32_t> ch; for (uint32_t i = 0; i < nc && r.ok; i++) ch.push_back(r.u32());
            Node& nd = node_at(nid); nd.type = N_GROUP; nd.children = ch;
            continue;
        }
        else if (id == fourcc("nSHP")) {
            uint32_t nid = r.u32(); (void)read_dict(r);
            uint32_t nm = r.u32(); uint32_t first = 0;
            for (uint32_t i = 0; i < nm && r.ok; i++) { uint32_t mi = r.u32(); (void)read_dict(r); if (i == 0) first = mi; }
            Node& nd = node_at(nid); nd.type = N_SHAPE; nd.model = first;
            continue;
        }
        else if (id == fourcc("IMAP")) { if (r.remaining() >= 256) { memcpy(imap, bytes + r.off, 256); have_imap = true; } r.skip(256); continue; }
        else if (id == fourcc("LAYR")) { (void)r.u32(); (void)read_dict(r); (void)r.u32(); continue; }
        else if (id == fourcc("MATL")) {
            uint32_t mid = r.u32() & 0xFF; Dict d = read_dict(r);
            if (const std::string* v = dict_get(d, "_metal")) metal[mid] = (float)atof(v->c_str());
            continue;
        }
        else if (id == fourcc("MATT")) {  // ogt_vox.h:1570-1622
            uint32_t mid = r.u32() & 0xFF; uint32_t type = r.u32(); float wgt; uint32_t wb = r.u32(); memcpy(&wgt, &wb, 4); (void)r.u32();
            if (type == 1) metal[mid] = wgt;
            r.skip(size >= 16 ? size - 16 : 0);
            continue;
        }
        else if (id == fourcc("rCAM")) { (void)r.u32(); (void)read_dict(r); continue; }
        else { r.skip(size); continue; }
    }

    // instances (ogt_vox.h:1673-1774)
    std::vector<Instance> inst;
    if (!nodes.empty()) {
        std::vector<Group> groups; std::vector<const Node*> stack;
        gen_instances(nodes, 0, stack, models, inst, groups, -1, 0);
    } else if (models.size() == 1) {
        inst.push_back({0, xf_identity()});
    }
    // IMAP (ogt_vox.h:1800-1831)
    float metal2[256]; memcpy(metal2, metal, sizeof(metal));
    if (have_imap) {
        uint8_t inv[256]; for (int i = 0; i < 256; i++) inv[imap[i]] = (uint8_t)i;
        uint8_t old[1024]; memcpy(old, pal, 1024);
        for (int i = 0; i < 256; i++) memcpy(pal + 4 * i, old + 4 * ((imap[i] + 255) & 0xFF), 4);
        for (int i = 0; i < 256; i++) metal2[i] = metal[imap[(i + 255) & 0xFF]];
        for (auto& m : models) for (auto& v : m.vox) v = (uint8_t)(1 + inv[v]);
    }
    // palette rotate (ogt_vox.h:1834-1840)
    uint8_t rot[1024];
    memcpy(rot + 4, pal, 1020); memcpy(rot, pal + 1020, 4); rot[3] = 0;

    if (inst.empty()) return -3;  // voxel_scene.cpp:49-50

    // ---- voxel_scene.cpp:9-31: pivot and transform ----
    auto pivot = [](const Model& m, int pv[3]) {
        pv[0] = (int)floorf(m.sx / 2.0f); pv[1] = (int)floorf(m.sy / 2.0f); pv[2] = (int)floorf(m.sz / 2.0f);
    };
    auto apply = [](const Xf& t, const int pv[3], int x, int y, int z, int out[3]) {
        // glm column-major mat4{col0..col3} * vec4, columns = ogt rows (voxel_scene.cpp:25-29); glm order (c0*x + c1*y) + (c2*z + c3*w)
        float v[4] = {((float)x + 0.5f) - (float)pv[0], ((float)y + 0.5f) - (float)pv[1], ((float)z + 0.5f) - (float)pv[2], 1.0f - 0.0f};
        for (int k = 0; k < 3; k++) {
            float a = (t.m[0][k] * v[0] + t.m[1][k] * v[1]) + (t.m[2][k] * v[2] + t.m[3][k] * v[3]);
            out[k] = orc::f2i(floorf(a));
        }
    };
    // ---- voxel_scene.cpp:52-74: bounds from two corners per instance, Y/Z swap ----
    int mn[3] = {100000, 100000, 100000}, mx[3] = {-100000, -100000, -100000};
    for (auto& in : inst) {
        const Model& m = models[in.model];
        int pv[3]; pivot(m, pv);
        int c1[3], c2[3];
        apply(in.xf, pv, 0, 0, 0, c1);
        apply(in.xf, pv, (int)m.sx, (int)m.sy, (int)m.sz, c2);
        for (int k = 0; k < 3; k++) {
            mn[k] = std::min(mn[k], std::min(c1[k], c2[k]));
            mx[k] = std::max(mx[k], std::max(c1[k], c2[k]));
        }
    }
    const uint32_t width = (uint32_t)(mx[0] - mn[0]), height = (uint32_t)(mx[2] - mn[2]), depth = (uint32_t)(mx[1] - mn[1]);
    const size_t size = (size_t)width * height * depth;
    if (size == 0 || size > ((size_t)1 << 34)) return -4;
    uint8_t* scene = (uint8_t*)calloc(size, 1);
    if (!scene) return -5;
    // ---- voxel_scene.cpp:81-105: dense fill; later instances overwrite earlier ones ----
    for (auto& in : inst) {
        const Model& m = models[in.model];
        int pv[3]; pivot(m, pv);
        for (uint32_t x = 0; x < m.sx; x++)
            for (uint32_t y = 0; y < m.sy; y++)
                for (uint32_t z = 0; z < m.sz; z++) {
                    uint8_t vox = m.vox[x + (size_t)y * m.sx + (size_t)z * m.sx * m.sy];
                    if (vox != 0) {
                        int t[3]; apply(in.xf, pv, (int)x, (int)y, (int)z, t);
                        long long tx = t[0] - mn[0], ty = t[1] - mn[1], tz = t[2] - mn[2];
                        long long pos = tx + tz * (long long)width + ty * (long long)width * height;
                        // The reference writes sceneData[pos] unchecked (UB when a mirrored instance lands on x == width
                        // at the far corner); pinned: linear index as computed, writes outside the buffer are dropped.
                        if (pos >= 0 && (size_t)pos < size) scene[pos] = vox;
                    }
                }
    }
    // ---- voxel_scene.cpp:107-117: palette ----
    for (int m = 0; m < 256; m++) {
        float* o = palette + 8 * m;
        for (int k = 0; k < 4; k++) o[k] = powf((float)rot[4 * m + k] / 255.0f, 2.2f);
        o[4] = metal2[m]; o[5] = o[6] = o[7] = 0.0f;
    }
    *volume = scene; whd[0] = width; whd[1] = height; whd[2] = depth;
    return 0;
}
