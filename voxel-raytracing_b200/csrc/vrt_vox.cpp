// vrt_vox.cpp — see vrt_vox.h. Three stages: (1) chunk scan of the RIFF-like .vox container, (2) scene-graph walk
// producing flattened instance transforms, (3) the reference's flatten into one dense volume with the Y/Z swap.
#include "vrt_vox.h"
#include "vrt_flatten.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <map>
#include <new>
#include <stdexcept>

namespace vrt {
namespace {

struct Mat4 { float e[16]; };   // e[4*r+c], row-vector convention like ogt_vox_transform (translation in row 3)
Mat4 ident() { Mat4 m{}; m.e[0] = m.e[5] = m.e[10] = m.e[15] = 1.0f; return m; }
Mat4 mul(const Mat4& a, const Mat4& b) {
    Mat4 r{};
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            r.e[4 * i + j] = (a.e[4 * i] * b.e[j]) + (a.e[4 * i + 1] * b.e[4 + j]) + (a.e[4 * i + 2] * b.e[8 + j]) + (a.e[4 * i + 3] * b.e[12 + j]);
    return r;
}

struct Cursor {
    const uint8_t* base; size_t size, at;
    bool bad = false;
    size_t left() const { return size - at; }
    uint32_t word() { if (left() < 4) { bad = true; at = size; return 0; } uint32_t v; memcpy(&v, base + at, 4); at += 4; return v; }
    std::string text() { uint32_t k = word(); if (left() < k) { bad = true; at = size; return std::string(); } std::string s(reinterpret_cast<const char*>(base + at), k); at += k; return s; }
    void advance(size_t k) { at = k > left() ? size : at + k; }
};

// .vox DICT: case-insensitive keys (the reference parser compares with strcasecmp)
struct Props {
    std::map<std::string, std::string> kv;
    static std::string lower(std::string s) { for (auto& c : s) c = (char)tolower((unsigned char)c); return s; }
    void read(Cursor& c) { kv.clear(); uint32_t n = c.word(); for (uint32_t i = 0; i < n && !c.bad; i++) { std::string k = lower(c.text()); std::string v = c.text(); if (!kv.count(k)) kv[k] = v; } }
    const std::string* find(const char* k) const { auto it = kv.find(k); return it == kv.end() ? nullptr : &it->second; }
};

struct VoxModel { uint32_t nx = 0, ny = 0, nz = 0; std::vector<uint8_t> cells; bool present = false; };
struct GraphNode { int kind = 0 /*1 xform 2 group 3 shape*/; Mat4 xf = ident(); uint32_t child = 0; std::vector<uint32_t> kids; uint32_t model = 0; };
struct Placed { uint32_t model; Mat4 xf; };

Mat4 frame_transform(const Props& p) {
    Mat4 t = ident();
    if (const std::string* r = p.find("_r")) {
        const uint32_t bits = (uint32_t)atoi(r->c_str());
        const uint32_t a = bits & 3u, b = (bits >> 2) & 3u;
        uint32_t c = 3;   // remaining axis by elimination; 3 = invalid (zero row), as the reference's lookup table yields
        const uint32_t used = (1u << a) | (1u << b);
        if (used == 3u) c = 2; else if (used == 5u) c = 1; else if (used == 6u) c = 0;
        const float sa = (bits & 16u) ? -1.0f : 1.0f, sb = (bits & 32u) ? -1.0f : 1.0f, sc = (bits & 64u) ? -1.0f : 1.0f;
        for (int k = 0; k < 3; k++) {   // file stores rows; the transform keeps them as columns
            t.e[4 * k + 0] = (a < 3 && (int)a == k) ? sa : (sa < 0 ? -0.0f : 0.0f);
            t.e[4 * k + 1] = (b < 3 && (int)b == k) ? sb : (sb < 0 ? -0.0f : 0.0f);
            t.e[4 * k + 2] = (c < 3 && (int)c == k) ? sc : (sc < 0 ? -0.0f : 0.0f);
        }
    }
    if (const std::string* tr = p.find("_t")) {
        int x = 0, y = 0, z = 0;
        sscanf(tr->c_str(), "%i %i %i", &x, &y, &z);
        t.e[12] = (float)x; t.e[13] = (float)y; t.e[14] = (float)z;
    }
    return t;
}

// A crafted graph (a group listing itself, or the same child many times) would otherwise fan out exponentially: the walk visits at
// most kMaxVisits nodes and then stops (a real scene has a few thousand nodes).
constexpr uint32_t kMaxNodeId = 1u << 20, kMaxVisits = 1u << 22;
constexpr size_t kMaxModelCells = (size_t)1 << 31;
void walk(const std::vector<GraphNode>& g, uint32_t id, std::vector<const GraphNode*>& path, const std::vector<VoxModel>& models,
          std::vector<std::pair<Mat4, int>>& groups, int group, std::vector<Placed>& out, int depth, uint32_t& visits) {
    if (id >= g.size() || depth > 256 || ++visits > kMaxVisits) return;
    const GraphNode& n = g[id];
    if (n.kind == 1) {
        path.push_back(&n); walk(g, n.child, path, models, groups, group, out, depth + 1, visits); path.pop_back();
    } else if (n.kind == 2) {
        const int mine = (int)groups.size();
        groups.emplace_back(path.empty() ? ident() : path.back()->xf, group);
        path.push_back(&n);
        for (uint32_t k : n.kids) walk(g, k, path, models, groups, mine, out, depth + 1, visits);
        path.pop_back();
    } else if (n.kind == 3) {
        if (n.model < models.size() && models[n.model].present) {
            Mat4 t = path.empty() ? ident() : path.back()->xf;
            for (int gi = group; gi >= 0; gi = groups[gi].second) t = mul(t, groups[gi].first);
            out.push_back({n.model, t});
        }
    }
}

void stock_palette(uint8_t* p) {   // MagicaVoxel default palette in file order (colour of voxel value i+1 at slot i)
    static const uint8_t ramp[10] = {0xee, 0xdd, 0xbb, 0xaa, 0x88, 0x77, 0x55, 0x44, 0x22, 0x11};
    int k = 0;
    for (int r = 0; r < 6; r++) for (int g = 0; g < 6; g++) for (int b = 0; b < 6; b++) {
        if (r == 5 && g == 5 && b == 5) continue;
        p[4 * k] = (uint8_t)(255 - 51 * r); p[4 * k + 1] = (uint8_t)(255 - 51 * g); p[4 * k + 2] = (uint8_t)(255 - 51 * b); p[4 * k + 3] = 255; k++;
    }
    for (int ch = 0; ch < 4; ch++) for (int j = 0; j < 10; j++, k++) {
        for (int q = 0; q < 3; q++) p[4 * k + q] = (ch == 3 || ch == q) ? ramp[j] : 0;
        p[4 * k + 3] = 255;
    }
    p[4 * k] = p[4 * k + 1] = p[4 * k + 2] = 0; p[4 * k + 3] = 255;
}

constexpr uint32_t tag(char a, char b, char c, char d) { return (uint32_t)(uint8_t)a | ((uint32_t)(uint8_t)b << 8) | ((uint32_t)(uint8_t)c << 16) | ((uint32_t)(uint8_t)d << 24); }

}  // namespace

static int parse_vox_bytes_unchecked(const uint8_t* bytes, size_t n, VoxLayout& out, std::string& err);
// Nothing may throw across the C ABI: sizes taken from the file can make the containers throw (bad_alloc / length_error).
int parse_vox_bytes(const uint8_t* bytes, size_t n, VoxLayout& out, std::string& err) {
    try { return parse_vox_bytes_unchecked(bytes, n, out, err); }
    catch (const std::bad_alloc&) { err = "out of host memory while parsing the voxel scene"; return VRT_E_OOM; }
    catch (const std::exception&) { err = "Could not parse voxel scene"; return VRT_E_PARSE; }
}
static int parse_vox_bytes_unchecked(const uint8_t* bytes, size_t n, VoxLayout& out, std::string& err) {
    static const char lfs[] = "version https://git-lfs";
    if (bytes && n >= sizeof(lfs) - 1 && memcmp(bytes, lfs, sizeof(lfs) - 1) == 0) {
        err = "file is a git-LFS pointer stub, not .vox data"; return VRT_E_LFS_POINTER;
    }
    Cursor c{bytes, n, 0};
    const uint32_t magic = c.word(), version = c.word();
    if (!bytes || c.bad || magic != tag('V', 'O', 'X', ' ') || (version != 150 && version != 200)) { err = "Could not parse voxel scene"; return VRT_E_PARSE; }

    std::vector<VoxModel> models;
    std::vector<GraphNode> graph;
    uint8_t filePal[1024]; stock_palette(filePal);
    float metal[256]; for (float& m : metal) m = 0.0f;
    uint8_t imap[256]; bool haveImap = false;
    uint32_t nx = 0, ny = 0, nz = 0;
    Props props;
    bool badNode = false;
    GraphNode scrap;
    auto node = [&](uint32_t id) -> GraphNode& { if (id > kMaxNodeId) { badNode = true; return scrap; } if (id >= graph.size()) graph.resize((size_t)id + 1); return graph[id]; };

    // (1) chunk scan. Like the reference parser, structured chunks are consumed field by field (their declared
    // size is not used to re-seek) and unknown chunks are skipped by their declared size.
    while (c.left() >= 12) {
        const uint32_t id = c.word(), bytesInChunk = c.word(); c.word();
        switch (id) {
            case tag('M', 'A', 'I', 'N'): break;
            case tag('S', 'I', 'Z', 'E'): nx = c.word(); ny = c.word(); nz = c.word(); break;
            case tag('X', 'Y', 'Z', 'I'): {
                const uint32_t count = c.word();
                VoxModel m;
                if (count) {
                    // the sizes come from the file: the product must neither wrap nor ask for an absurd allocation
                    const unsigned __int128 cells = (unsigned __int128)nx * ny * nz;
                    if (cells == 0 || cells > kMaxModelCells) { err = "Could not parse voxel scene (model size)"; return VRT_E_PARSE; }
                    m.present = true; m.nx = nx; m.ny = ny; m.nz = nz;
                    m.cells.assign((size_t)cells, 0);
                    const uint32_t readable = (uint32_t)std::min<size_t>(c.left() / 4, count);
                    const uint8_t* q = bytes + c.at;
                    for (uint32_t i = 0; i < readable; i++, q += 4)
                        if (q[0] < nx && q[1] < ny && q[2] < nz) m.cells[q[0] + (size_t)q[1] * nx + (size_t)q[2] * nx * ny] = q[3];
                }
                models.push_back(std::move(m));
                c.advance((size_t)count * 4);
                break;
            }
            case tag('R', 'G', 'B', 'A'): if (c.left() >= 1024) memcpy(filePal, bytes + c.at, 1024); c.advance(1024); break;
            case tag('n', 'T', 'R', 'N'): {
                const uint32_t nid = c.word(); props.read(c);
                const uint32_t child = c.word(); c.word(); c.word();
                const uint32_t frames = c.word();
                Mat4 first = ident();
                for (uint32_t f = 0; f < frames && !c.bad; f++) { props.read(c); if (f == 0) first = frame_transform(props); }
                GraphNode& g = node(nid); g.kind = 1; g.child = child; g.xf = first;
                break;
            }
            case tag('n', 'G', 'R', 'P'): {
                const uint32_t nid = c.word(); props.read(c);
                const uint32_t kids = c.word();
                if (kids > c.left() / 4) { err = "Could not parse voxel scene (group size)"; return VRT_E_PARSE; }
                std::vector<uint32_t> ids; for (uint32_t i = 0; i < kids && !c.bad; i++) ids.push_back(c.word());
                GraphNode& g = node(nid); g.kind = 2; g.kids.swap(ids);
                break;
            }
            case tag('n', 'S', 'H', 'P'): {
                const uint32_t nid = c.word(); props.read(c);
                const uint32_t nm = c.word(); uint32_t first = 0;
                for (uint32_t i = 0; i < nm && !c.bad; i++) { const uint32_t mi = c.word(); props.read(c); if (i == 0) first = mi; }
                GraphNode& g = node(nid); g.kind = 3; g.model = first;
                break;
            }
            case tag('I', 'M', 'A', 'P'): if (c.left() >= 256) { memcpy(imap, bytes + c.at, 256); haveImap = true; } c.advance(256); break;
            case tag('L', 'A', 'Y', 'R'): c.word(); props.read(c); c.word(); break;
            case tag('M', 'A', 'T', 'L'): {
                const uint32_t mid = c.word() & 0xFFu; props.read(c);
                if (const std::string* v = props.find("_metal")) metal[mid] = (float)atof(v->c_str());
                break;
            }
            case tag('M', 'A', 'T', 'T'): {
                const uint32_t mid = c.word() & 0xFFu, type = c.word(), wbits = c.word(); c.word();
                float w; memcpy(&w, &wbits, 4);
                if (type == 1) metal[mid] = w;
                c.advance(bytesInChunk >= 16 ? bytesInChunk - 16 : 0);
                break;
            }
            case tag('r', 'C', 'A', 'M'): c.word(); props.read(c); break;
            default: c.advance(bytesInChunk); break;
        }
    }

    if (badNode) { err = "Could not parse voxel scene (node id)"; return VRT_E_PARSE; }
    // (2) instances
    std::vector<Placed> placed;
    if (!graph.empty()) {
        std::vector<const GraphNode*> path; std::vector<std::pair<Mat4, int>> groups;
        uint32_t visits = 0;
        walk(graph, 0, path, models, groups, -1, placed, 0, visits);
        if (visits > kMaxVisits) { err = "Could not parse voxel scene (scene graph too large)"; return VRT_E_PARSE; }
    } else if (models.size() == 1) {
        placed.push_back({0, ident()});
    }
    // display-order remap (IMAP) of palette, materials and voxel values, then the palette rotate that makes a
    // voxel value index the palette directly (value 0 = empty, alpha 0)
    float metalOut[256]; memcpy(metalOut, metal, sizeof(metal));
    if (haveImap) {
        uint8_t inv[256]; for (int i = 0; i < 256; i++) inv[imap[i]] = (uint8_t)i;
        uint8_t old[1024]; memcpy(old, filePal, 1024);
        for (int i = 0; i < 256; i++) memcpy(filePal + 4 * i, old + 4 * ((imap[i] + 255) & 0xFF), 4);
        for (int i = 0; i < 256; i++) metalOut[i] = metal[imap[(i + 255) & 0xFF]];
        for (VoxModel& m : models) for (uint8_t& v : m.cells) v = (uint8_t)(1 + inv[v]);
    }
    uint8_t pal[1024];
    memcpy(pal + 4, filePal, 1020); memcpy(pal, filePal + 1020, 4); pal[3] = 0;

    if (placed.empty()) { err = "Voxel scene does not contain an instance."; return VRT_E_NO_INSTANCE; }

    // (3) bounds of the flatten (voxel_scene.cpp:53-76); the copy itself is flatten_host / the device flatten
    out.models.clear(); out.instances.clear();
    for (const VoxModel& m : models) { VoxLayout::Model o; o.nx = m.nx; o.ny = m.ny; o.nz = m.nz; o.cells = m.cells; out.models.push_back(std::move(o)); }
    int lo[3] = {100000, 100000, 100000}, hi[3] = {-100000, -100000, -100000};
    for (const Placed& pl : placed) {
        const VoxModel& m = models[pl.model];
        VoxLayout::Instance in; in.model = pl.model; memcpy(in.xf, pl.xf.e, sizeof(in.xf));
        in.pivot[0] = (int)floorf(m.nx / 2.0f); in.pivot[1] = (int)floorf(m.ny / 2.0f); in.pivot[2] = (int)floorf(m.nz / 2.0f);
        // a signed permutation in the rotation part maps distinct voxels to distinct cells; anything else (the zero row an invalid _r yields)
        // can collide inside one instance, where the reference's loop order decides
        bool perm = true;
        for (int r = 0; r < 3 && perm; r++) {
            int nzRow = 0, nzCol = 0;
            for (int k = 0; k < 3; k++) {
                const float a = in.xf[4 * r + k], b = in.xf[4 * k + r];
                if (a != 0.0f) { nzRow++; if (a != 1.0f && a != -1.0f) perm = false; }
                if (b != 0.0f) nzCol++;
            }
            if (nzRow != 1 || nzCol != 1) perm = false;
        }
        in.permutation = perm;
        int a[3], b[3];
        vox_transform(in.xf, in.pivot, 0, 0, 0, a); vox_transform(in.xf, in.pivot, (int)m.nx, (int)m.ny, (int)m.nz, b);
        for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], std::min(a[k], b[k])); hi[k] = std::max(hi[k], std::max(a[k], b[k])); }
        out.instances.push_back(in);
    }
    for (int k = 0; k < 3; k++) {
        out.lo[k] = lo[k];
        const long long ext = (long long)hi[k] - (long long)lo[k];
        if (ext <= 0 || ext > 32767) { err = "degenerate scene bounds"; return VRT_E_PARSE; }   // vrt_scene bounds are 1..32767 per axis
    }
    out.w = (uint32_t)(hi[0] - lo[0]); out.h = (uint32_t)(hi[2] - lo[2]); out.d = (uint32_t)(hi[1] - lo[1]);
    const size_t total = (size_t)out.w * out.h * out.d;
    if (total == 0 || total > ((size_t)1 << 34)) { err = "degenerate scene bounds"; return VRT_E_PARSE; }
    for (int i = 0; i < 256; i++) {   // voxel_scene.cpp:107-117
        vrt_material& m = out.palette[i];
        for (int k = 0; k < 4; k++) m.diffuse[k] = powf((float)pal[4 * i + k] / 255.0f, 2.2f);
        m.metallic = metalOut[i]; m.pad1 = m.pad2 = m.pad3 = 0.0f;
    }
    return VRT_OK;
}

static FlattenInstance flatten_params(const VoxLayout& lay, const VoxLayout::Instance& in) {
    const VoxLayout::Model& m = lay.models[in.model];
    FlattenInstance f;
    memcpy(f.xf, in.xf, sizeof(f.xf));
    for (int k = 0; k < 3; k++) { f.pivot[k] = in.pivot[k]; f.lo[k] = lay.lo[k]; }
    f.nx = m.nx; f.ny = m.ny; f.nz = m.nz; f.W = lay.w; f.H = lay.h; f.D = lay.d;
    return f;
}
FlattenInstance flatten_instance_params(const VoxLayout& lay, size_t i) { return flatten_params(lay, lay.instances[i]); }

// voxel_scene.cpp:83-105, loop order included (x outermost, z innermost; instances in order)
void flatten_host(const VoxLayout& lay, uint8_t* volume) {
    memset(volume, 0, (size_t)lay.w * lay.h * lay.d);
    for (const VoxLayout::Instance& in : lay.instances) {
        const VoxLayout::Model& m = lay.models[in.model];
        const FlattenInstance f = flatten_params(lay, in);
        for (uint32_t x = 0; x < m.nx; x++) for (uint32_t y = 0; y < m.ny; y++) for (uint32_t z = 0; z < m.nz; z++) {
            const uint8_t v = m.cells[x + (size_t)y * m.nx + (size_t)z * m.nx * m.ny];
            if (!v) continue;
            const long long at = flatten_cell(f, (int)x, (int)y, (int)z);
            if (at >= 0) volume[(size_t)at] = v;
        }
    }
}

int load_vox_bytes(const uint8_t* bytes, size_t n, DenseScene& out, std::string& err) {
    try {
        VoxLayout lay;
        const int rc = parse_vox_bytes(bytes, n, lay, err);
        if (rc) return rc;
        out.w = lay.w; out.h = lay.h; out.d = lay.d;
        out.volume.resize((size_t)lay.w * lay.h * lay.d);
        flatten_host(lay, out.volume.data());
        memcpy(out.palette, lay.palette, sizeof(out.palette));
        return VRT_OK;
    } catch (const std::bad_alloc&) { err = "out of host memory while flattening the voxel scene"; return VRT_E_OOM; }
    catch (const std::exception&) { err = "Could not parse voxel scene"; return VRT_E_PARSE; }
}

static int read_file(const char* path, std::vector<uint8_t>& buf, std::string& err) {
    FILE* f = path ? fopen(path, "rb") : nullptr;
    if (!f) { err = "Failed to read voxel scene"; return VRT_E_IO; }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    if (sz < 0) { fclose(f); err = "Failed to read voxel scene"; return VRT_E_IO; }
    try { buf.resize((size_t)sz); } catch (const std::exception&) { fclose(f); err = "out of host memory"; return VRT_E_OOM; }
    const size_t got = sz ? fread(buf.data(), 1, (size_t)sz, f) : 0;
    fclose(f);
    if (got != (size_t)sz) { err = "Failed to read voxel scene"; return VRT_E_IO; }
    return VRT_OK;
}
int load_vox_file(const char* path, DenseScene& out, std::string& err) {
    std::vector<uint8_t> buf;
    const int rc = read_file(path, buf, err);
    return rc ? rc : load_vox_bytes(buf.data(), buf.size(), out, err);
}
int parse_vox_file(const char* path, VoxLayout& out, std::string& err) {
    std::vector<uint8_t> buf;
    const int rc = read_file(path, buf, err);
    return rc ? rc : parse_vox_bytes(buf.data(), buf.size(), out, err);
}

}  // namespace vrt
