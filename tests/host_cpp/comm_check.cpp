// comm_check — the multi-GPU exchange of libvrt driven from plain C++ (no torch, no NCCL, no Python): forks `world` processes,
// one per rank (rank r on GPU r % ndev, so it also runs on a single GPU), each renders its interleaved tile shard of a few frames
// and exchanges colour + normal + position into rank 0's frame (peer memory) and colour into the shared pinned host frame. Rank 0
// denoises the assembled planes and compares everything, byte for byte, with the same frames rendered by a single-rank context.
// Second part: the same frames through vrt_render_denoise_exchange (every rank denoises its own rows), compared the same way.
//   usage: comm_check <world> <W> <H> <ndev> [frames]      exit code 0 = identical
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <ctime>
#include <string>
#include <vector>
#include <sys/wait.h>
#include <unistd.h>
#include "vrt.h"

#define CK(call) do { int rc_ = (call); if (rc_ != VRT_OK) { fprintf(stderr, "[rank %d] %s -> %d: %s\n", g_rank, #call, rc_, vrt_last_error()); return 10; } } while (0)
static int g_rank = -1;

static vrt_camera make_camera(const uint32_t whd[3], int W, int H, int k) {
    vrt_camera c{};
    const float a = 0.6f + 0.37f * (float)k;
    const float pos[3] = {0.5f * whd[0] + 1.1f * whd[0] * cosf(a), 0.9f * whd[1] + 0.13f, 0.5f * whd[2] + 1.1f * whd[2] * sinf(a)};
    const float v[3] = {0.5f * whd[0] - pos[0], 0.4f * whd[1] - pos[1], 0.5f * whd[2] - pos[2]};
    const float yaw = atan2f(v[2], v[0]) * 57.29577951f, pitch = atan2f(v[1], sqrtf(v[0] * v[0] + v[2] * v[2])) * 57.29577951f;
    float d[3], r[3], u[3];
    vrt_camera_vectors(yaw, pitch, 1.920982f, d, r, u);
    for (int i = 0; i < 3; i++) { c.camPos[i] = pos[i]; c.camDir[i] = d[i]; c.camRight[i] = r[i]; c.camUp[i] = u[i]; c.volumeBounds[i] = whd[i]; }
    c.camPos[3] = 1.0f; c.frame = (uint32_t)k; c.screenSize[0] = W; c.screenSize[1] = H;
    return c;
}

static int make_scene(vrt_context* ctx, vrt_scene** out) {
    CK(vrt_scene_generate(ctx, 2, 1234, 96, 80, 112, out));
    std::vector<uint8_t> noise(512 * 512 * 4);
    for (size_t i = 0; i < noise.size(); i++) { uint32_t x = (uint32_t)i * 2654435761u; x ^= x >> 15; x *= 2246822519u; x ^= x >> 13; noise[i] = (uint8_t)(x >> 24); }
    std::vector<float> sky(64 * 32 * 4);
    for (int y = 0; y < 32; y++) for (int x = 0; x < 64; x++) { float* p = &sky[(y * 64 + x) * 4]; p[0] = x / 64.0f; p[1] = y / 32.0f; p[2] = 0.75f; p[3] = 1.0f; }
    CK(vrt_scene_set_noise(*out, noise.data(), 512, 512));
    CK(vrt_scene_set_skybox(*out, sky.data(), 64, 32));
    return 0;
}

static int run_rank(int rank, int world, int W, int H, int ndev, int frames, const std::string& key) {
    g_rank = rank;
    vrt_context* ctx = nullptr;
    CK(vrt_context_create(rank % ndev, &ctx));
    CK(vrt_comm_create(ctx, rank, world, key.c_str()));
    vrt_scene* scene = nullptr;
    if (int rc = make_scene(ctx, &scene)) return rc;
    uint32_t whd[3]; CK(vrt_scene_bounds(scene, whd));
    const uint32_t planes = (1u << VRT_PLANE_COLOR_RGBA8) | (1u << VRT_PLANE_NORMAL_RGBA8S) | (1u << VRT_PLANE_POSITION_RGBA32F);
    const int slots = 2;
    const uint32_t dnPlane = 1u << VRT_PLANE_DENOISED_RGBA8;     // filled by the distributed denoiser (strips) in the second part
    CK(vrt_comm_alloc_frames(ctx, W, H, slots, planes | dnPlane, (1u << VRT_PLANE_COLOR_RGBA8) | dnPlane));
    vrt_render_settings st; vrt_default_settings(&st);
    st.denoise.enable = 0;                                   // sharded ranks do not denoise: rank 0 does, on the assembled planes
    vrt_render_settings st1 = st; st1.denoise.enable = 1;

    // rank 0: the reference frames from a single-rank context of its own
    vrt_context* solo = nullptr; vrt_scene* soloScene = nullptr; vrt_frame* soloFrame = nullptr;
    void *dOut = nullptr, *dScratch = nullptr;
    const size_t px = (size_t)W * H;
    std::vector<uint8_t> want(px * 4), wantDn(px * 4), got(px * 4);
    std::vector<uint8_t> wantN(px * 4), wantP(px * 16), gotBig(px * 16);
    if (rank == 0) {
        CK(vrt_context_create(0, &solo));
        if (int rc = make_scene(solo, &soloScene)) return rc;
        CK(vrt_buffer_alloc(ctx, px * 4, &dOut));
        CK(vrt_buffer_alloc(ctx, px * 4, &dScratch));
    }
    vrt_frame* fr[2] = {nullptr, nullptr};
    int bad = 0;
    for (int k = 0; k < frames; k++) {
        const int slot = k % slots;
        const vrt_camera cam = make_camera(whd, W, H, k);
        if (k & 1) {                                          // fused: the pass's own kernels store into rank 0's frame
            CK(vrt_render_exchange(ctx, scene, &cam, &st, slot, planes, 1u << VRT_PLANE_COLOR_RGBA8, &fr[slot]));
        } else {                                              // separate: push kernels after the pass
            CK(vrt_render(ctx, scene, &cam, &st, &fr[slot]));
            CK(vrt_frame_exchange(fr[slot], slot, planes, 1u << VRT_PLANE_COLOR_RGBA8, nullptr));
        }
        if (rank != 0) continue;
        void* cs = vrt_comm_stream(ctx);
        void *dc, *dn, *dp, *hc; size_t nb;
        CK(vrt_comm_frame_plane(ctx, slot, VRT_PLANE_COLOR_RGBA8, &dc, &hc, &nb));
        CK(vrt_comm_frame_plane(ctx, slot, VRT_PLANE_NORMAL_RGBA8S, &dn, nullptr, nullptr));
        CK(vrt_comm_frame_plane(ctx, slot, VRT_PLANE_POSITION_RGBA32F, &dp, nullptr, nullptr));
        CK(vrt_comm_wait(ctx, slot, nullptr));
        CK(vrt_denoise_on(ctx, dc, dn, dp, W, H, &st1.denoise, dOut, dScratch, cs));
        CK(vrt_context_join(ctx)); CK(vrt_context_sync(ctx));
        CK(vrt_render(solo, soloScene, &cam, &st1, &soloFrame));
        CK(vrt_frame_download(soloFrame, VRT_PLANE_COLOR_RGBA8, want.data(), want.size()));
        CK(vrt_frame_download(soloFrame, VRT_PLANE_DENOISED_RGBA8, wantDn.data(), wantDn.size()));
        CK(vrt_frame_download(soloFrame, VRT_PLANE_NORMAL_RGBA8S, wantN.data(), wantN.size()));
        CK(vrt_frame_download(soloFrame, VRT_PLANE_POSITION_RGBA32F, wantP.data(), wantP.size()));
        CK(vrt_buffer_download(ctx, dc, got.data(), got.size()));
        if (memcmp(got.data(), want.data(), got.size())) { fprintf(stderr, "frame %d: assembled colour differs\n", k); bad++; }
        CK(vrt_buffer_download(ctx, dn, got.data(), got.size()));
        if (memcmp(got.data(), wantN.data(), got.size())) { fprintf(stderr, "frame %d: assembled normal differs\n", k); bad++; }
        CK(vrt_buffer_download(ctx, dp, gotBig.data(), gotBig.size()));
        if (memcmp(gotBig.data(), wantP.data(), gotBig.size())) { fprintf(stderr, "frame %d: assembled position differs\n", k); bad++; }
        CK(vrt_buffer_download(ctx, dOut, got.data(), got.size()));
        if (memcmp(got.data(), wantDn.data(), got.size())) { fprintf(stderr, "frame %d: denoised frame differs\n", k); bad++; }
        CK(vrt_comm_wait_host(ctx, slot));
        if (memcmp(hc, want.data(), want.size())) { fprintf(stderr, "frame %d: shared host frame differs\n", k); bad++; }
        CK(vrt_comm_release(ctx, slot));
    }
    CK(vrt_context_sync(ctx));
    CK(vrt_comm_barrier(ctx));
    // second part: the denoiser distributed over the ranks (strips). Every rank denoises the rows it owns; rank 0 finds the finished
    // denoised plane in the slot (device frame and shared host frame) and compares it with the single-rank denoised frame.
    int stripsState = 0;                                     // 1 ran, 2 unsupported for this size (strips thinner than the denoiser's reach)
    {
        const int rc = vrt_comm_alloc_strips(ctx, &st1.denoise);
        if (rc == VRT_E_UNSUPPORTED) stripsState = 2;
        else if (rc != VRT_OK) { fprintf(stderr, "[rank %d] vrt_comm_alloc_strips -> %d: %s\n", rank, rc, vrt_last_error()); return 10; }
        else stripsState = 1;
    }
    for (int k = 0; stripsState == 1 && k < frames; k++) {
        const int slot = k % slots;
        const vrt_camera cam = make_camera(whd, W, H, k + 3);
        CK(vrt_render_denoise_exchange(ctx, scene, &cam, &st, slot, 1, &fr[slot]));
        if (rank != 0) continue;
        void *dd, *hd; size_t nb;
        CK(vrt_comm_frame_plane(ctx, slot, VRT_PLANE_DENOISED_RGBA8, &dd, &hd, &nb));
        CK(vrt_comm_wait(ctx, slot, nullptr));
        CK(vrt_context_join(ctx)); CK(vrt_context_sync(ctx));
        CK(vrt_render(solo, soloScene, &cam, &st1, &soloFrame));
        CK(vrt_frame_download(soloFrame, VRT_PLANE_DENOISED_RGBA8, wantDn.data(), wantDn.size()));
        CK(vrt_buffer_download(ctx, dd, got.data(), got.size()));
        if (memcmp(got.data(), wantDn.data(), got.size())) {
            size_t first = 0; while (first < got.size() && got[first] == wantDn[first]) first++;
            fprintf(stderr, "strips, frame %d: denoised frame differs (first at row %zu)\n", k, first / 4 / (size_t)W); bad++;
        }
        CK(vrt_comm_wait_host(ctx, slot));
        if (memcmp(hd, wantDn.data(), wantDn.size())) { fprintf(stderr, "strips, frame %d: shared host denoised frame differs\n", k); bad++; }
        CK(vrt_comm_release(ctx, slot));
    }
    CK(vrt_context_sync(ctx));
    CK(vrt_comm_barrier(ctx));
    // third part: strips without a denoiser — the colour plane crosses into row strips and every rank copies its rows into rank 0's
    // frame / the shared host frame (the read-back path): alternately both targets and the host frame alone
    CK(vrt_comm_alloc_strips(ctx, nullptr));
    for (int k = 0; k < frames; k++) {
        const int slot = k % slots, mode = (k & 1) ? 2 : 1;
        const vrt_camera cam = make_camera(whd, W, H, k + 7);
        CK(vrt_render_denoise_exchange(ctx, scene, &cam, &st, slot, mode, &fr[slot]));
        if (rank != 0) continue;
        void *dc, *hc; size_t nb;
        CK(vrt_comm_frame_plane(ctx, slot, VRT_PLANE_COLOR_RGBA8, &dc, &hc, &nb));
        CK(vrt_render(solo, soloScene, &cam, &st, &soloFrame));
        CK(vrt_frame_download(soloFrame, VRT_PLANE_COLOR_RGBA8, want.data(), want.size()));
        if (mode == 1) {
            CK(vrt_comm_wait(ctx, slot, nullptr));
            CK(vrt_context_join(ctx)); CK(vrt_context_sync(ctx));
            CK(vrt_buffer_download(ctx, dc, got.data(), got.size()));
            if (memcmp(got.data(), want.data(), got.size())) { fprintf(stderr, "plain strips, frame %d: colour frame differs\n", k); bad++; }
        }
        CK(vrt_comm_wait_host(ctx, slot));
        if (memcmp(hc, want.data(), want.size())) { fprintf(stderr, "plain strips, frame %d: shared host colour frame differs\n", k); bad++; }
        CK(vrt_comm_release(ctx, slot));
    }
    CK(vrt_context_sync(ctx));
    CK(vrt_comm_barrier(ctx));
    for (vrt_frame* f : fr) if (f) vrt_frame_release(f);
    if (rank == 0) {
        printf("comm_check: world %d, %dx%d, %d frames: %s\n", world, W, H, frames, bad ? "MISMATCH" : "assembled colour / normal / position, denoised and host frames identical to the single-rank frames");
        printf("comm_check: strips without a denoiser (colour rows to rank 0 / the host frame): %s\n", bad ? "MISMATCH" : "colour frame identical");
        printf("comm_check: strips (denoiser distributed over the ranks): %s\n", stripsState == 2 ? "unsupported at this size (strips thinner than the denoiser's reach)" : bad ? "MISMATCH" : "denoised frame identical");
        vrt_frame_release(soloFrame); vrt_scene_destroy(soloScene); vrt_context_destroy(solo);
        vrt_buffer_free(ctx, dOut); vrt_buffer_free(ctx, dScratch);
    }
    vrt_scene_destroy(scene);
    vrt_context_destroy(ctx);                                // destroys the communicator (collective)
    return bad ? 20 : 0;
}

int main(int argc, char** argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s world W H ndev [frames]\n", argv[0]); return 1; }
    const int world = atoi(argv[1]), W = atoi(argv[2]), H = atoi(argv[3]), ndev = atoi(argv[4]) > 0 ? atoi(argv[4]) : 1, frames = argc > 5 ? atoi(argv[5]) : 5;
    const std::string key = "cc" + std::to_string((long)getpid()) + "_" + std::to_string((long)time(nullptr));
    std::vector<pid_t> kids;
    for (int r = 0; r < world; r++) {                        // fork BEFORE any CUDA call: a CUDA context does not survive fork
        const pid_t p = fork();
        if (p == 0) { const int rc = run_rank(r, world, W, H, ndev, frames, key); fflush(nullptr); _exit(rc); }
        kids.push_back(p);
    }
    int worst = 0;
    for (pid_t p : kids) { int st = 0; waitpid(p, &st, 0); const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 99; if (code > worst) worst = code; }
    return worst;
}
