// TEST INFRASTRUCTURE. Compiles the product's per-pixel device code (csrc/vrt_pixel.cuh, vrt_core.cuh) for the
// HOST with g++ so that the CPU-only test tier can compare the kernel logic bit-for-bit with the oracle
// without a GPU. This is not a product path and is not reachable from libvrt.
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include "../../voxel-raytracing_b200/csrc/vrt_wave.cuh"
#include "../../voxel-raytracing_b200/csrc/vrt_output.cuh"

using namespace vrt;

// CPU restatement of the GPU brick builder (csrc/vrt_accel.cu) for the harness only.
static void build_bricks(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, std::vector<unsigned long long>& bmask,
                         std::vector<uint8_t>& bdist, uint32_t& Bx, uint32_t& By, uint32_t& Bz) {
    Bx = (W + 3) / 4; By = (H + 3) / 4; Bz = (D + 3) / 4;
    bmask.assign((size_t)Bx * By * Bz, 0ull);
    bdist.assign((size_t)Bx * By * Bz, 0);
    for (uint32_t z = 0; z < D; z++) for (uint32_t y = 0; y < H; y++) for (uint32_t x = 0; x < W; x++)
        if (vol[x + (size_t)y * W + (size_t)z * W * H]) {
            size_t b = (size_t)(z >> 2) * By * Bx + (size_t)(y >> 2) * Bx + (x >> 2);
            bmask[b] |= 1ull << ((x & 3) | ((y & 3) << 2) | ((z & 3) << 4));
        }
    std::vector<uint8_t> cur((size_t)Bx * By * Bz);
    for (uint32_t z = 0; z < Bz; z++) for (uint32_t y = 0; y < By; y++) for (uint32_t x = 0; x < Bx; x++) {
        size_t b = (size_t)z * By * Bx + (size_t)y * Bx + x;
        bool straddle = (x * 4 + 4 > W) || (y * 4 + 4 > H) || (z * 4 + 4 > D);
        cur[b] = (bmask[b] != 0 || straddle) ? 0 : VRT_DIST_CAP;
    }
    // Chebyshev distance by repeated 3x3x3 min (+1), outside the grid counts as distance 0
    for (int it = 0; it < VRT_DIST_CAP; it++) {
        std::vector<uint8_t> nxt = cur; bool changed = false;
        for (int z = 0; z < (int)Bz; z++) for (int y = 0; y < (int)By; y++) for (int x = 0; x < (int)Bx; x++) {
            size_t b = (size_t)z * By * Bx + (size_t)y * Bx + x;
            int mn = 255;
            for (int dz = -1; dz <= 1; dz++) for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) {
                int X = x + dx, Y = y + dy, Z = z + dz;
                int v = (X < 0 || Y < 0 || Z < 0 || X >= (int)Bx || Y >= (int)By || Z >= (int)Bz) ? 0 : cur[(size_t)Z * By * Bx + (size_t)Y * Bx + X];
                mn = std::min(mn, v);
            }
            int nv = std::min((int)cur[b], mn + 1);
            if (nv != cur[b]) { nxt[b] = (uint8_t)nv; changed = true; }
        }
        cur.swap(nxt);
        if (!changed) break;
    }
    bdist = cur;
}

// CPU restatement of the per-voxel safe-step field builder (csrc/vrt_accel.cu launch_build_fdist)
static void build_fdist(const uint8_t* vol, int W, int H, int D, const std::vector<uint8_t>& bdist, uint32_t Bx, uint32_t By, uint32_t Bz,
                        std::vector<uint8_t>& fdist, std::vector<uint8_t>& ldist) {
    const size_t n = (size_t)W * H * D;
    std::vector<uint8_t> a(n), nx(n);
    for (size_t i = 0; i < n; i++) a[i] = vol[i] ? 0 : VRT_FINE_CAP;
    auto at = [&](const std::vector<uint8_t>& v, int x, int y, int z) -> int {
        return (x < 0 || y < 0 || z < 0 || x >= W || y >= H || z >= D) ? 0 : v[(size_t)x + (size_t)y * W + (size_t)z * W * H]; };
    for (int r = 1; r < VRT_FINE_CAP; r++) {
        for (int z = 0; z < D; z++) for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
            int mn = 255;
            for (int dz = -1; dz <= 1; dz++) for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) mn = std::min(mn, at(a, x + dx, y + dy, z + dz));
            const size_t v = (size_t)x + (size_t)y * W + (size_t)z * W * H;
            nx[v] = (uint8_t)std::min((int)a[v], mn + 1);
        }
        a.swap(nx);
    }
    fdist.assign((size_t)Bx * By * Bz * 64, 0);
    ldist.assign((size_t)(W + 2) * (H + 2) * (D + 2), (uint8_t)VRT_LIN_OOB);
    for (int z = 0; z < D; z++) for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
        const uint32_t f = a[(size_t)x + (size_t)y * W + (size_t)z * W * H];
        const size_t b = (size_t)(z >> 2) * By * Bx + (size_t)(y >> 2) * Bx + (x >> 2);
        const uint32_t d = bdist[b];
        uint32_t v = f;
        if (f != 0 && d != 0) v = std::max(f, (d - 1u) * 4u + 1u);
        fdist[b * 64 + ((x & 3) | ((y & 3) << 2) | ((z & 3) << 4))] = (uint8_t)std::min(v, VRT_SAFE_MAX);
        ldist[(size_t)(x + 1) + (size_t)(y + 1) * (W + 2) + (size_t)(z + 1) * (W + 2) * (H + 2)] = (uint8_t)std::min(v, VRT_SAFE_MAX);
    }
}

// CPU restatement of the directional safe-step fields (csrc/vrt_octant.cu): exact largest empty cube anchored at v towards
// the octant (7-neighbour dynamic programme), rounded down to a level of VRT_OCT_LEVEL_LIST
static void build_odist(const uint8_t* vol, int W, int H, int D, std::vector<uint8_t> odist[8]) {
    static const int lv[VRT_OCT_NLEVELS] = VRT_OCT_LEVEL_LIST;
    const size_t n = (size_t)W * H * D;
    std::vector<int> e(n);
    auto at = [&](int x, int y, int z) -> int { return (x < 0 || y < 0 || z < 0 || x >= W || y >= H || z >= D) ? 0 : e[(size_t)x + (size_t)y * W + (size_t)z * W * H]; };
    for (int o = 0; o < 8; o++) {
        const int sx = (o & 1) ? -1 : 1, sy = (o & 2) ? -1 : 1, sz = (o & 4) ? -1 : 1;
        odist[o].assign((size_t)(W + 2) * (H + 2) * (D + 2), (uint8_t)VRT_LIN_OOB);
        for (int kz = 0; kz < D; kz++) for (int ky = 0; ky < H; ky++) for (int kx = 0; kx < W; kx++) {
            const int x = sx > 0 ? W - 1 - kx : kx, y = sy > 0 ? H - 1 - ky : ky, z = sz > 0 ? D - 1 - kz : kz;   // far end of the octant first
            int v = 0;
            if (vol[(size_t)x + (size_t)y * W + (size_t)z * W * H] == 0) {
                int m = at(x + sx, y, z);
                m = std::min(m, at(x, y + sy, z)); m = std::min(m, at(x, y, z + sz));
                m = std::min(m, at(x + sx, y + sy, z)); m = std::min(m, at(x + sx, y, z + sz)); m = std::min(m, at(x, y + sy, z + sz));
                m = std::min(m, at(x + sx, y + sy, z + sz));
                v = std::min(1 + m, 1 << 20);
            }
            e[(size_t)x + (size_t)y * W + (size_t)z * W * H] = v;
            int q = 0;
            for (int l = 0; l < VRT_OCT_NLEVELS; l++) if (lv[l] <= v) q = lv[l];
            odist[o][(size_t)(x + 1) + (size_t)(y + 1) * (W + 2) + (size_t)(z + 1) * (W + 2) * (H + 2)] = (uint8_t)safe_encode((uint32_t)q);
        }
    }
}

// CPU restatement of the two-level layout (csrc/vrt_octant.cu sp_classify_kernel / sp_fill_kernel) from an octant's linear field
static void build_sdist(const std::vector<uint8_t>& lin, int W, int H, int D, const uint32_t tb[3], std::vector<uint32_t>& top, std::vector<uint8_t>& fine) {
    auto code = [&](int x, int y, int z) -> uint32_t {
        if (x < -1 || y < -1 || z < -1 || x > W || y > H || z > D) return VRT_LIN_OOB;
        return lin[(size_t)(x + 1) + (size_t)(y + 1) * (W + 2) + (size_t)(z + 1) * (W + 2) * (H + 2)]; };
    top.assign((size_t)tb[0] * tb[1] * tb[2], 0u); fine.clear();
    for (uint32_t bz = 0; bz < tb[2]; bz++) for (uint32_t by = 0; by < tb[1]; by++) for (uint32_t bx = 0; bx < tb[0]; bx++) {
        uint8_t c[512]; uint32_t mn = 0xffffffffu; bool same = true, special = false;
        for (int k = 0; k < 512; k++) {
            c[k] = (uint8_t)code((int)bx * 8 - 8 + (k & 7), (int)by * 8 - 8 + ((k >> 3) & 7), (int)bz * 8 - 8 + (k >> 6));
            same = same && c[k] == c[0]; special = special || c[k] == 0 || c[k] == VRT_LIN_OOB;
            mn = std::min(mn, safe_decode(c[k]));
        }
        uint32_t e;
        if (same) e = c[0];
        else if (!special && mn >= VRT_SP_UNIFORM_MIN) e = safe_encode(mn);
        else { e = VRT_SP_FINE | (uint32_t)(fine.size() / 512); fine.insert(fine.end(), c, c + 512); }
        top[bx + (by + (size_t)bz * tb[1]) * tb[0]] = e;
    }
}

struct HScene {
    SceneDev dev;
    std::vector<unsigned long long> bmask;
    std::vector<uint8_t> bdist, fdist, ldist, odist[8], sfine[8];
    std::vector<uint32_t> stop[8];
};

extern "C" void* hh_scene_create(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, const float* palette,
                                 const uint8_t* noise, const float* sky, uint32_t skyW, uint32_t skyH) {
    HScene* s = new HScene();
    memset(&s->dev, 0, sizeof(SceneDev));
    s->dev.vol = vol; s->dev.W = W; s->dev.H = H; s->dev.D = D;
    build_bricks(vol, W, H, D, s->bmask, s->bdist, s->dev.Bx, s->dev.By, s->dev.Bz);
    build_fdist(vol, (int)W, (int)H, (int)D, s->bdist, s->dev.Bx, s->dev.By, s->dev.Bz, s->fdist, s->ldist);
    s->dev.bmask = s->bmask.data(); s->dev.bdist = s->bdist.data(); s->dev.fdist = s->fdist.data(); s->dev.ldist = s->ldist.data();
    build_odist(vol, (int)W, (int)H, (int)D, s->odist);
    for (int o = 0; o < 8; o++) s->dev.odist[o] = s->odist[o].data();
    uint32_t tb[3] = {((W + 8u) >> 3) + 1u, ((H + 8u) >> 3) + 1u, ((D + 8u) >> 3) + 1u};
    s->dev.TBx = tb[0]; s->dev.TBy = tb[1]; s->dev.TBz = tb[2];
    for (int o = 0; o < 8; o++) {
        build_sdist(s->odist[o], (int)W, (int)H, (int)D, tb, s->stop[o], s->sfine[o]);
        s->dev.sdist[o].top = s->stop[o].data(); s->dev.sdist[o].fine = s->sfine[o].data();
    }
    s->dev.palette = reinterpret_cast<const vrt_material*>(palette);
    s->dev.noise = noise; s->dev.sky = sky; s->dev.skyW = skyW; s->dev.skyH = skyH;
    return s;
}
extern "C" void hh_scene_destroy(void* p) { delete (HScene*)p; }
extern "C" void hh_scene_bdist(void* p, uint8_t* out) { HScene* s = (HScene*)p; memcpy(out, s->bdist.data(), s->bdist.size()); }
extern "C" void hh_scene_fdist(void* p, uint8_t* out) { HScene* s = (HScene*)p; memcpy(out, s->fdist.data(), s->fdist.size()); }
extern "C" void hh_scene_bmask(void* p, unsigned long long* out) { HScene* s = (HScene*)p; memcpy(out, s->bmask.data(), 8 * s->bmask.size()); }

extern "C" int hh_render(void* scene, const vrt_camera* cam, const vrt_render_settings* st, int traversal, const PlanesDev* pl,
                         int row_begin, int row_end, int row_step) {
    HScene* s = (HScene*)scene;
    FrameParams fp; memset(&fp, 0, sizeof(fp));
    fp.cam = *cam; fp.pass_mask = st->pass_mask; fp.out_mask = st->out_mask; fp.traversal = traversal;
    fp.aoSamples = st->vol.aoSamples; fp.ambientIntensity = st->vol.ambientIntensity;
    for (int k = 0; k < 3; k++) { fp.lightDir[k] = st->light.direction[k]; fp.lightColor[k] = st->light.color[k]; }
    fp.lightIntensity = st->light.intensity; fp.rank = 0; fp.world = 1;
    const int W = cam->screenSize[0], H = cam->screenSize[1];
    if (row_end > H) row_end = H;
#pragma omp parallel for schedule(dynamic, 4)
    for (int py = row_begin; py < row_end; py += row_step)
        for (int px = 0; px < W; px++) {
            PixelOut o; memset(&o, 0, sizeof(o));
            if (traversal == VRT_TRAVERSAL_DENSE_U8) render_pixel<VRT_TRAVERSAL_DENSE_U8>(s->dev, fp, px, py, o);
            else if (traversal == VRT_TRAVERSAL_FINE) render_pixel<VRT_TRAVERSAL_FINE>(s->dev, fp, px, py, o);
            else if (traversal == VRT_TRAVERSAL_LINEAR) render_pixel<VRT_TRAVERSAL_LINEAR>(s->dev, fp, px, py, o);
            else if (traversal == VRT_TRAVERSAL_OCTANT) render_pixel<VRT_TRAVERSAL_OCTANT>(s->dev, fp, px, py, o);
            else if (traversal == VRT_TRAVERSAL_SPARSE) render_pixel<VRT_TRAVERSAL_SPARSE>(s->dev, fp, px, py, o);
            else render_pixel<VRT_TRAVERSAL_BRICK>(s->dev, fp, px, py, o);
            size_t p = (size_t)py * W + px;
            auto sat = [](int v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); };
            if (pl->color) { uint8_t* q = pl->color + 4 * p; q[0] = to_unorm8(o.color.x); q[1] = to_unorm8(o.color.y); q[2] = to_unorm8(o.color.z); q[3] = 0; }
            if (pl->depth) pl->depth[p] = o.depth;
            if (pl->motion) { pl->motion[2 * p] = 0; pl->motion[2 * p + 1] = 0; }
            if (pl->mask) pl->mask[p] = to_unorm8(o.mask);
            if (pl->position) { float* q = pl->position + 4 * p; q[0] = o.pos.x; q[1] = o.pos.y; q[2] = o.pos.z; q[3] = 0; }
            if (pl->normal) { int8_t* q = pl->normal + 4 * p; q[0] = to_snorm8(o.normal.x); q[1] = to_snorm8(o.normal.y); q[2] = to_snorm8(o.normal.z); q[3] = 0; }
            if (pl->hit) { int16_t* q = pl->hit + 4 * p; q[0] = sat(o.hx); q[1] = sat(o.hy); q[2] = sat(o.hz); q[3] = (int16_t)o.steps; }
            if (pl->hit_palette) pl->hit_palette[p] = (uint8_t)o.palette;
            if (pl->color_f32) { float* q = pl->color_f32 + 3 * p; q[0] = o.color.x; q[1] = o.color.y; q[2] = o.color.z; }
            if (pl->rays) pl->rays[p] = o.rays;
            if (pl->fetches) pl->fetches[p] = o.fetches;
        }
    return 0;
}

// ---- wavefront pipeline (csrc/vrt_wave.cuh) run through a serial emulation of the kernels' queues ----
template <int T>
static void render_wave(const SceneDev& sc, const FrameParams& fp, const PlanesDev* pl, int row_begin, int row_end, int row_step) {
    const int W = fp.cam.screenSize[0];
    const uint32_t aoN = (fp.pass_mask & VRT_PASS_AO) ? fp.aoSamples : 0u;
    const bool shadow = (fp.pass_mask & VRT_PASS_SHADOW) != 0;
    std::vector<RecQ> rec; std::vector<BounceEnt> q[2]; std::vector<uint32_t> roots, sq;
    auto sat = [](int v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); };
    auto put_color = [&](size_t p, V3 c) {
        if (pl->color) { uint8_t* d = pl->color + 4 * p; d[0] = to_unorm8(c.x); d[1] = to_unorm8(c.y); d[2] = to_unorm8(c.z); d[3] = 0; }
        if (pl->color_f32) { float* d = pl->color_f32 + 3 * p; d[0] = c.x; d[1] = c.y; d[2] = c.z; } };
    // primary wave
    for (int py = row_begin; py < row_end; py += row_step)
        for (int px = 0; px < W; px++) {
            PixelOut o; memset(&o, 0, sizeof(o)); V3 rd;
            const int action = wave_primary<T>(sc, fp, px, py, o, rd);
            const size_t p = (size_t)py * W + px;
            if (action == WAVE_PIXEL_FINAL) put_color(p, o.color);
            if (pl->depth) pl->depth[p] = o.depth;
            if (pl->motion) { pl->motion[2 * p] = 0; pl->motion[2 * p + 1] = 0; }
            if (pl->mask) pl->mask[p] = to_unorm8(o.mask);
            if (pl->position) { float* d = pl->position + 4 * p; d[0] = o.pos.x; d[1] = o.pos.y; d[2] = o.pos.z; d[3] = 0; }
            if (pl->normal) { int8_t* d = pl->normal + 4 * p; d[0] = to_snorm8(o.normal.x); d[1] = to_snorm8(o.normal.y); d[2] = to_snorm8(o.normal.z); d[3] = 0; }
            if (pl->hit) { int16_t* d = pl->hit + 4 * p; d[0] = sat(o.hx); d[1] = sat(o.hy); d[2] = sat(o.hz); d[3] = (int16_t)o.steps; }
            if (pl->hit_palette) pl->hit_palette[p] = (uint8_t)o.palette;
            if (pl->rays) pl->rays[p] = 1;
            if (pl->fetches) pl->fetches[p] = o.fetches;
            if (action != WAVE_PIXEL_FINAL) {
                const uint32_t idx = (uint32_t)(rec.size() / 3);
                rec.push_back(mkq(o.pos, (uint32_t)px | ((uint32_t)py << 16))); rec.push_back(mkq(o.normal, o.palette)); rec.push_back(mkq(rd, WAVE_NONE));
                roots.push_back(idx);
                if (action == WAVE_PIXEL_SHADE_AND_BOUNCE) { BounceEnt e; e.rec = idx; e.root = idx; q[0].push_back(e); }
            }
        }
    auto pix = [&](uint32_t pxy) { return (size_t)(pxy >> 16) * W + (pxy & 0xffffu); };
    // bounce waves
    for (int w = 0; w < MAX_REFLECTIONS; w++) {
        std::vector<BounceEnt>& in = q[w & 1]; std::vector<BounceEnt>& out = q[(w + 1) & 1]; out.clear();
        for (const BounceEnt& ent : in) {
            const RecQ p0 = rec[3 * (size_t)ent.rec], p1 = rec[3 * (size_t)ent.rec + 1], p2 = rec[3 * (size_t)ent.rec + 2];
            V3 rd; const TraceRes r = wave_bounce<T>(sc, q3(p0), q3(p1), q3(p2), rd);
            if (pl->rays) pl->rays[pix(p0.w)] += 1;
            if (pl->fetches) pl->fetches[pix(p0.w)] += r.fetches;
            const uint32_t child = (uint32_t)(rec.size() / 3);
            rec.push_back(mkq(r.pos, p0.w)); rec.push_back(mkq(r.normal, r.material | ((uint32_t)w << 8))); rec.push_back(mkq(rd, ent.rec));
            const bool terminal = r.material == 0 || sc.palette[r.material].metallic <= 0.0f;
            if (terminal) {
                rec[3 * (size_t)ent.root + 2].w = child;
                if (r.material != 0) sq.push_back(child);
                for (uint32_t k = ent.rec; k != ent.root; k = rec[3 * (size_t)k + 2].w) sq.push_back(k);
            } else if (w + 1 < MAX_REFLECTIONS) { BounceEnt e; e.rec = child; e.root = ent.root; out.push_back(e); }
        }
    }
    // any-hit wave (roots first, then queued bounce records; ray-kind-major like the kernel)
    const uint32_t nrec = (uint32_t)(rec.size() / 3), R = aoN + (shadow ? 1u : 0u);
    std::vector<uint8_t> res((size_t)nrec * (aoN + 1u), 0);
    WaveBuf wb; memset(&wb, 0, sizeof(wb)); wb.rec = rec.data(); wb.res = res.data(); wb.recCap = nrec;
    std::vector<uint32_t> todo = roots; todo.insert(todo.end(), sq.begin(), sq.end());
    for (uint32_t i = 0; i < R; i++)
        for (uint32_t idx : todo) {
            const RecQ r0 = rec[3 * (size_t)idx], r1 = rec[3 * (size_t)idx + 1];
            uint32_t f = 0;
            res[(size_t)i * nrec + idx] = wave_anyhit<T>(sc, fp, q3(r0), q3(r1), (int)(r0.w & 0xffffu), (int)(r0.w >> 16), r1.w >> 8, i, aoN, f) ? 1 : 0;
            if (pl->rays) pl->rays[pix(r0.w)] += 1;
            if (pl->fetches) pl->fetches[pix(r0.w)] += f;
        }
    // resolve
    for (uint32_t idx : roots) put_color(pix(rec[3 * (size_t)idx].w), wave_resolve(sc, fp, wb, idx, aoN));
}

extern "C" int hh_render_wave(void* scene, const vrt_camera* cam, const vrt_render_settings* st, int traversal, const PlanesDev* pl,
                              int row_begin, int row_end, int row_step) {
    HScene* s = (HScene*)scene;
    FrameParams fp; memset(&fp, 0, sizeof(fp));
    fp.cam = *cam; fp.pass_mask = st->pass_mask; fp.out_mask = st->out_mask; fp.traversal = traversal;
    fp.aoSamples = st->vol.aoSamples; fp.ambientIntensity = st->vol.ambientIntensity;
    for (int k = 0; k < 3; k++) { fp.lightDir[k] = st->light.direction[k]; fp.lightColor[k] = st->light.color[k]; }
    fp.lightIntensity = st->light.intensity; fp.rank = 0; fp.world = 1;
    if (row_end > cam->screenSize[1]) row_end = cam->screenSize[1];
    if (traversal == VRT_TRAVERSAL_DENSE_U8) render_wave<VRT_TRAVERSAL_DENSE_U8>(s->dev, fp, pl, row_begin, row_end, row_step);
    else if (traversal == VRT_TRAVERSAL_FINE) render_wave<VRT_TRAVERSAL_FINE>(s->dev, fp, pl, row_begin, row_end, row_step);
    else if (traversal == VRT_TRAVERSAL_LINEAR) render_wave<VRT_TRAVERSAL_LINEAR>(s->dev, fp, pl, row_begin, row_end, row_step);
    else if (traversal == VRT_TRAVERSAL_OCTANT) render_wave<VRT_TRAVERSAL_OCTANT>(s->dev, fp, pl, row_begin, row_end, row_step);
    else if (traversal == VRT_TRAVERSAL_SPARSE) render_wave<VRT_TRAVERSAL_SPARSE>(s->dev, fp, pl, row_begin, row_end, row_step);
    else render_wave<VRT_TRAVERSAL_BRICK>(s->dev, fp, pl, row_begin, row_end, row_step);
    return 0;
}

// ---- tile-sharded rendering (N>1 ranks): this rank's interleaved tiles into a compact tile-major colour buffer, exactly
// the layout vrt_render produces under vrt_context_set_shard (plane_index), and the assembly of vrt_assemble ----
extern "C" int hh_render_shard(void* scene, const vrt_camera* cam, const vrt_render_settings* st, int traversal, int rank, int world,
                               uint8_t* color_tiles, size_t shard_pixels) {
    HScene* s = (HScene*)scene;
    FrameParams fp; memset(&fp, 0, sizeof(fp));
    fp.cam = *cam; fp.pass_mask = st->pass_mask; fp.out_mask = st->out_mask; fp.traversal = traversal;
    fp.aoSamples = st->vol.aoSamples; fp.ambientIntensity = st->vol.ambientIntensity;
    for (int k = 0; k < 3; k++) { fp.lightDir[k] = st->light.direction[k]; fp.lightColor[k] = st->light.color[k]; }
    fp.lightIntensity = st->light.intensity; fp.rank = rank; fp.world = world;
    const int W = cam->screenSize[0], H = cam->screenSize[1];
    fp.tilesX = (W + VRT_TILE - 1) / VRT_TILE; fp.tilesY = (H + VRT_TILE - 1) / VRT_TILE;
    const int tiles = fp.tilesX * fp.tilesY;
    fp.localTiles = tiles > rank ? (tiles - rank + world - 1) / world : 0;
    if ((size_t)((tiles + world - 1) / world) * VRT_TILE * VRT_TILE != shard_pixels) return -1;
#pragma omp parallel for schedule(dynamic, 1)
    for (int j = 0; j < fp.localTiles; j++) {
        const int t = j * world + rank, tx = t % fp.tilesX, ty = t / fp.tilesX;
        for (int ly = 0; ly < VRT_TILE; ly++)
            for (int lx = 0; lx < VRT_TILE; lx++) {
                const int px = tx * VRT_TILE + lx, py = ty * VRT_TILE + ly;
                if (px >= W || py >= H) continue;
                PixelOut o; memset(&o, 0, sizeof(o));
                if (traversal == VRT_TRAVERSAL_LINEAR) render_pixel<VRT_TRAVERSAL_LINEAR>(s->dev, fp, px, py, o);
                else if (traversal == VRT_TRAVERSAL_OCTANT) render_pixel<VRT_TRAVERSAL_OCTANT>(s->dev, fp, px, py, o);
                else if (traversal == VRT_TRAVERSAL_SPARSE) render_pixel<VRT_TRAVERSAL_SPARSE>(s->dev, fp, px, py, o);
            else if (traversal == VRT_TRAVERSAL_SPARSE) render_pixel<VRT_TRAVERSAL_SPARSE>(s->dev, fp, px, py, o);
                else render_pixel<VRT_TRAVERSAL_FINE>(s->dev, fp, px, py, o);
                uint8_t* q = color_tiles + 4 * plane_index(fp, px, py);
                q[0] = to_unorm8(o.color.x); q[1] = to_unorm8(o.color.y); q[2] = to_unorm8(o.color.z); q[3] = 0;
            }
    }
    return 0;
}
extern "C" void hh_assemble(const uint8_t* gathered, int world, int W, int H, size_t shard_pixels, uint8_t* out) {
    const int tilesX = (W + VRT_TILE - 1) / VRT_TILE;
    for (int py = 0; py < H; py++)
        for (int px = 0; px < W; px++)
            memcpy(out + 4 * ((size_t)py * W + px), gathered + 4 * gathered_index(px, py, tilesX, world, shard_pixels), 4);
}

// the sparse field of an octant expanded to the aproned linear layout (what vrt_scene_download_sdist returns)
extern "C" void hh_scene_sdist(void* p, int octant, uint8_t* out) {
    HScene* s = (HScene*)p;
    const int W = (int)s->dev.W, H = (int)s->dev.H, D = (int)s->dev.D;
    for (int z = -1; z <= D; z++) for (int y = -1; y <= H; y++) for (int x = -1; x <= W; x++)
        out[(size_t)(x + 1) + (size_t)(y + 1) * (W + 2) + (size_t)(z + 1) * (W + 2) * (H + 2)] = (uint8_t)sparse_fetch(s->dev, s->dev.sdist[octant], x, y, z);
}
extern "C" void hh_scene_odist(void* p, int octant, uint8_t* out) { HScene* s = (HScene*)p; memcpy(out, s->odist[octant].data(), s->odist[octant].size()); }

// the output stage's per-pixel code (csrc/vrt_output.cuh) in the loop structure of blit_kernel
extern "C" void hh_blit(const uint8_t* src, int sw, int sh, uint8_t* dst, int tw, int th, uint32_t flags) {
    BlitParams p; p.sw = sw; p.sh = sh; p.tw = tw; p.th = th; p.flags = flags;
    float thr[256]; srgb_thresholds(thr);
    float lutU[256]; for (int i = 0; i < 256; i++) lutU[i] = (float)i / 255.0f;
    const float scale = blit_scale(p);
    for (int py = 0; py < th; py++) {
        const BlitAxis ry = blit_axis(py, sh, th, scale);
        for (int px = 0; px < tw; px++) {
            const BlitAxis cx = blit_axis(px, sw, tw, scale);
            const BlitTap t = blit_weights(cx.a, ry.a);
            const uint8_t* c00 = src + 4 * ((size_t)ry.i0 * sw + cx.i0); const uint8_t* c10 = src + 4 * ((size_t)ry.i0 * sw + cx.i1);
            const uint8_t* c01 = src + 4 * ((size_t)ry.i1 * sw + cx.i0); const uint8_t* c11 = src + 4 * ((size_t)ry.i1 * sw + cx.i1);
            uint8_t o[4];
            for (int c = 0; c < 4; c++) {
                const float v = blit_mix(t, lutU[c00[c]], lutU[c10[c]], lutU[c01[c]], lutU[c11[c]]);
                o[c] = ((flags & VRT_BLIT_SRGB) && c < 3) ? srgb8_lookup(thr, v) : to_unorm8(v);
            }
            if (flags & VRT_BLIT_BGRA) std::swap(o[0], o[2]);
            memcpy(dst + 4 * ((size_t)py * tw + px), o, 4);
        }
    }
}

// the strip layout of the distributed denoiser (csrc/vrt_types.h, used by the kernels' strip_store): for image row py, the ranks whose
// buffer holds it and the row's index there; returns how many (out arrays of 4)
extern "C" int hh_strip_holders(int py, int world, int stripRows, int halo, int H, int* q_out, int* row_out) {
    int n = 0;
    strip_for_each_holder(py, world, stripRows, halo, H, [&](int q, int row) { if (n < 4) { q_out[n] = q; row_out[n] = row; } n++; });
    return n;
}
