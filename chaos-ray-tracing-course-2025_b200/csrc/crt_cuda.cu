// C-ABI implementation (include/crt_cuda.h, crt_cuda_* group): device context, scene upload, kernel launches.
// There is no CPU fallback anywhere in this file: without a CUDA device every entry point fails loudly.

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <string>
#include <vector>

#include "crt_kernels.cuh"
#include "crt_wavefront.cuh"
#include "crt_persistent.cuh"
#include "crt_ppm.cuh"
#include "crt_push.cuh"
#include "crt_gi.h"

using namespace crtdev;

struct crt_cuda_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    std::string error;

    char *d_arena = nullptr; // every scene array, one allocation
    size_t d_arena_cap = 0;
    char *h_arena = nullptr; // pinned staging of the same layout
    size_t h_arena_cap = 0;
    DScene scene{};
    bool have_scene = false;
    bool has_secondary = false; // some material can spawn reflection / refraction rays
    int bucket_size = 0;        // Scene::bucket_size: only decides whether the frame gets any bucket at all (enqueue_frame)
    bool light_leaves = false;  // walk tree with a few triangles per leaf: k_wf_trace_lean (crt_trace_lean.cuh)
    size_t off_lights = 0, off_materials = 0, off_textures = 0; // arena offsets of the small shading arrays (crt_cuda_update_shading)
    const char *warm_base = nullptr; // what a k-d walk reads: triangles .. walk tree .. leaf references, one range of the arena
    size_t warm_bytes = 0;
    int warm_mode = -1;              // $CRT_WARM_L2: 0 never, 1 always (prefetches), 2 always (loads), -1 by scene size
    // $CRT_L2_PERSIST_MB > 0: that much of L2 set aside for persisting lines and the walk's working set (or only the
    // tree and its references: $CRT_L2_WINDOW=tree) marked persisting on the launch stream (experiment, step 26)
    size_t l2_window_offset = 0, l2_window_bytes = 0;
    cudaStream_t l2_policy_stream = nullptr;
    bool l2_policy_on = false;
    bool warm_loads = false;
    int packet_mode = -1;        // primary level as warp packets (crt_trace_packet.cuh): -1 by pixels per triangle, 0 / 1 = $CRT_PACKET
    bool packet_primary = false; // this frame's decision
    crt_render_params last_params{};
    bool last_wavefront = false;
    bool rerendered = false; // crt_cuda_synchronize had to render the frame again (a bounce level outgrew the ray pool)

    float *d_rgb = nullptr;
    uint8_t *d_rgb8 = nullptr;
    size_t fb_pixels = 0;
    Counters *d_counters = nullptr;
    Counters *h_counters = nullptr; // pinned
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    // crt_cuda_render starts the float frame's device-to-host copy right after the level-0 shade (ev_primary), on
    // copy_stream, under the deeper levels' kernels; rows holding non-terminal pixels are copied again after resolve
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_primary = nullptr;
    bool arm_early_copy = false, early_recorded = false;
    int last_width = 0, last_rows = 0, last_launches = 0;
    bool last_stats = false;

    PersistCtl *d_pctl = nullptr; // pixel cursor of the persistent pipeline (crt_persistent.cuh)
    WavefrontState wf; // ray queues of the wavefront pipeline (crt_wavefront.cuh)

    uint8_t *frame = nullptr; // full quantised frame the row bands are pushed into (own allocation on rank 0, peer mapping elsewhere)
    int frame_w = 0, frame_h = 0;
    int frame_kind = 0;       // 0 none, 1 own cudaMalloc, 2 cudaIpcOpenMemHandle mapping, 3 caller's pointer

    char *d_text = nullptr; // PPM text
    size_t d_text_cap = 0;
    unsigned long long *d_scan = nullptr;
    size_t d_scan_cap = 0;
};

namespace {

thread_local std::string g_create_error;

#define CRT_CK(ctx, call)                                                                                     \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess) {                                                                              \
            (ctx)->error = std::string(#call) + ": " + cudaGetErrorString(e_);                                \
            return CRT_ERR_CUDA;                                                                              \
        }                                                                                                     \
    } while (0)

int fail(crt_cuda_ctx *ctx, int code, const char *what) {
    ctx->error = what;
    return code;
}

void release_frame(crt_cuda_ctx *ctx) {
    if (ctx->frame_kind == 1)
        cudaFree(ctx->frame);
    else if (ctx->frame_kind == 2)
        cudaIpcCloseMemHandle(ctx->frame);
    ctx->frame = nullptr, ctx->frame_kind = 0, ctx->frame_w = 0, ctx->frame_h = 0;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int owned_rows(const crt_render_params &p, int height) {
    if (p.world <= 1)
        return height;
    const int band = p.band_rows > 0 ? p.band_rows : 8;
    const int bands = (height + band - 1) / band;
    int rows = 0;
    for (int b = p.rank; b < bands; b += p.world)
        rows += (b == bands - 1) ? height - b * band : band;
    return rows;
}

DParams device_params(const crt_render_params &p, int height) {
    DParams d{};
    d.max_ray_depth = p.max_ray_depth > uint32_t(kMaxDepthCap) ? uint32_t(kMaxDepthCap) : p.max_ray_depth;
    d.gi_rays = p.diffuse_reflection_ray_count;
    d.shadow_bias = p.shadow_bias;
    d.reflection_bias = p.reflection_bias;
    d.diffuse_reflection_bias = p.diffuse_reflection_bias;
    d.refraction_bias = p.refraction_bias;
    d.shadows = p.shadows;
    d.rank = p.rank;
    d.world = p.world;
    d.band_rows = p.band_rows > 0 ? p.band_rows : 8;
    d.rows = owned_rows(p, height);
    if (const char *v = std::getenv("CRT_COHERENT"); v && *v == '1')
        d.coherent = 1;
    return d;
}

int ensure_framebuffer(crt_cuda_ctx *ctx, size_t pixels) {
    if (pixels <= ctx->fb_pixels)
        return CRT_OK;
    if (ctx->d_rgb)
        cudaFree(ctx->d_rgb);
    if (ctx->d_rgb8)
        cudaFree(ctx->d_rgb8);
    ctx->d_rgb = nullptr, ctx->d_rgb8 = nullptr, ctx->fb_pixels = 0;
    CRT_CK(ctx, cudaMalloc(&ctx->d_rgb, pixels * 3 * sizeof(float)));
    CRT_CK(ctx, cudaMalloc(&ctx->d_rgb8, pixels * 3));
    ctx->fb_pixels = pixels;
    return CRT_OK;
}

void fill_stats(const crt_cuda_ctx *ctx, crt_render_stats *stats, float ms) {
    if (!stats)
        return;
    std::memset(stats, 0, sizeof *stats);
    stats->gpu_ms = ms;
    stats->kernel_launches = ctx->last_launches;
    stats->rows_rendered = ctx->last_rows;
    if (ctx->last_stats) {
        const Counters &c = *ctx->h_counters;
        stats->rays = c.rays;
        stats->shadow_rays = c.shadow_rays;
        stats->node_tests = c.node_tests;
        stats->tri_tests = c.tri_tests;
        stats->entered_rays = c.entered;
        for (int i = 0; i < 16; ++i)
            stats->rays_by_depth[i] = c.rays_by_depth[i];
    }
}

template <int TRAV>
void launch_mega(crt_cuda_ctx *ctx, const DParams &dp, bool stats, dim3 grid) {
    if (stats)
        k_render_mega<TRAV, true><<<grid, kBlock, 0, ctx->stream>>>(ctx->scene, dp, ctx->d_rgb, ctx->d_rgb8, ctx->d_counters);
    else
        k_render_mega<TRAV, false><<<grid, kBlock, 0, ctx->stream>>>(ctx->scene, dp, ctx->d_rgb, ctx->d_rgb8, ctx->d_counters);
}

template <bool SEEDED, bool SHADOWS>
void launch_persistent(crt_cuda_ctx *ctx, const DParams &dp, bool stats, bool gi, int blocks, int patches_x, int n_slots) {
#define CRT_LAUNCH_P(ST, G) \
    k_render_persistent<ST, G, SEEDED, SHADOWS><<<blocks, kTraceBlock, 0, ctx->stream>>>(ctx->scene, dp, ctx->d_pctl, patches_x, n_slots, ctx->d_rgb, ctx->d_rgb8, ctx->d_counters, ctx->wf.b)
    if (gi) // built in crt_gi.cu (vector divisions through binary64)
        crt_launch_persistent_gi(stats, SEEDED, SHADOWS, blocks, ctx->stream, &ctx->scene, sizeof ctx->scene, &dp, sizeof dp, ctx->d_pctl, patches_x, n_slots,
                                 ctx->d_rgb, ctx->d_rgb8, ctx->d_counters, &ctx->wf.b, sizeof ctx->wf.b);
    else if (stats)
        CRT_LAUNCH_P(true, false);
    else
        CRT_LAUNCH_P(false, false);
#undef CRT_LAUNCH_P
}

// $CRT_L2_PERSIST_MB: the access policy window of the stream the frame is launched on (set once per stream and scene)
int apply_l2_policy(crt_cuda_ctx *ctx) {
    const char *mb = std::getenv("CRT_L2_PERSIST_MB");
    const double want_mb = mb && *mb ? std::atof(mb) : 0.0;
    if (want_mb <= 0.0 || ctx->l2_window_bytes == 0)
        return CRT_OK;
    if (ctx->l2_policy_on && ctx->l2_policy_stream == ctx->stream)
        return CRT_OK;
    int max_persist = 0, max_window = 0;
    CRT_CK(ctx, cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, ctx->device));
    CRT_CK(ctx, cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device));
    const size_t set_aside = std::min(size_t(max_persist), size_t(want_mb * 1048576.0));
    CRT_CK(ctx, cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, set_aside));
    cudaStreamAttrValue attr{};
    attr.accessPolicyWindow.base_ptr = ctx->d_arena + ctx->l2_window_offset;
    attr.accessPolicyWindow.num_bytes = std::min(ctx->l2_window_bytes, size_t(max_window));
    double ratio = double(set_aside) / double(attr.accessPolicyWindow.num_bytes);
    if (const char *r = std::getenv("CRT_L2_HIT"); r && *r)
        ratio = std::atof(r);
    attr.accessPolicyWindow.hitRatio = float(std::min(1.0, ratio));
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
    if (const char *m = std::getenv("CRT_L2_MISS"); m && *m == 's')
        attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    CRT_CK(ctx, cudaStreamSetAttribute(ctx->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
    if (std::getenv("CRT_L2_VERBOSE"))
        std::fprintf(stderr, "crt: L2 policy: set aside %zu MB (max %d MB), window %zu MB (max %d MB), hit ratio %.2f\n", set_aside >> 20, max_persist >> 20,
                     size_t(attr.accessPolicyWindow.num_bytes) >> 20, max_window >> 20, attr.accessPolicyWindow.hitRatio);
    ctx->l2_policy_on = true, ctx->l2_policy_stream = ctx->stream;
    return CRT_OK;
}

// enqueue one frame on ctx->stream, bracketed by the timing events
int enqueue_frame(crt_cuda_ctx *ctx, const crt_render_params *params) {
    if (!ctx->have_scene)
        return fail(ctx, CRT_ERR_INVALID, "no scene uploaded");
    if (!params)
        return fail(ctx, CRT_ERR_INVALID, "params is NULL");
    if (params->traversal != CRT_TRAVERSAL_LITERAL && params->traversal != CRT_TRAVERSAL_KD)
        return fail(ctx, CRT_ERR_INVALID, "render supports CRT_TRAVERSAL_LITERAL or CRT_TRAVERSAL_KD (brute force is a primary-hit checker only)");
    if (params->traversal == CRT_TRAVERSAL_LITERAL && ctx->scene.n_nodes < 1)
        return fail(ctx, CRT_ERR_INVALID, "CRT_TRAVERSAL_LITERAL needs the reference's acceleration tree: flatten the scene with CRT_FLATTEN_REFERENCE_TREE");
    // the recursion state (frames of the per-pixel kernels, levels of the wavefront) is sized for depths 0 .. 15; the
    // reference has no limit, so a deeper request is refused rather than rendered differently
    if (params->max_ray_depth > uint32_t(kMaxDepthCap))
        return fail(ctx, CRT_ERR_INVALID, "max_ray_depth above 15 is not supported by this build (CRT_MAX_RAY_DEPTH_LIMIT)");
    if (params->world > 1 && (params->rank < 0 || params->rank >= params->world))
        return fail(ctx, CRT_ERR_INVALID, "rank outside [0, world)");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    const int W = ctx->scene.cam.width, H = ctx->scene.cam.height;
    const DParams dp = device_params(*params, H);
    if (int rc = ensure_framebuffer(ctx, size_t(W) * size_t(dp.rows > 0 ? dp.rows : 1)))
        return rc;
    const bool stats = params->collect_stats != 0;
    ctx->last_stats = stats;
    ctx->packet_primary = ctx->packet_mode >= 0 ? ctx->packet_mode == 1
                                                : ctx->scene.n_triangles > 0 && double(W) * double(H) / double(ctx->scene.n_triangles) >= 64.0;
    ctx->last_width = W;
    ctx->last_rows = dp.rows;
    ctx->last_launches = 0;
    int extra_launches = 0;
    ctx->last_wavefront = false;
    ctx->last_params = *params;
    if (int rc = apply_l2_policy(ctx))
        return rc;
    if (stats)
        CRT_CK(ctx, cudaMemsetAsync(ctx->d_counters, 0, sizeof(Counters), ctx->stream));
    CRT_CK(ctx, cudaEventRecord(ctx->ev_begin, ctx->stream));
    // the walk's working set on its way into L2 while the first kernels start (k_warm_l2, crt_push.cuh): pays when the
    // set is large enough to be fetched miss by miss otherwise, and small enough to stay
    if (params->traversal == CRT_TRAVERSAL_KD && ctx->warm_bytes > 0 &&
        ctx->warm_mode == 1) { // off unless $CRT_WARM_L2 asks: measured (r2e, r2m), it does not pay
        if (ctx->warm_loads)
            k_warm_l2_loads<<<148 * 8, 256, 0, ctx->stream>>>(reinterpret_cast<const uint4 *>(ctx->warm_base), ctx->warm_bytes / 16, nullptr);
        else
            k_warm_l2<<<148 * 4, 256, 0, ctx->stream>>>(ctx->warm_base, ctx->warm_bytes);
        ++extra_launches;
    }
    // render_image cuts the frame into int(float(extent) / bucket_size + 0.5) buckets per axis, the last one absorbing
    // the remainder (crt_renderer.cpp:160-170): an image narrower or lower than half a bucket gets NO bucket and
    // stays the zero-initialised Image -- reproduced; otherwise the bucket size has no effect on the pixels
    const auto bucket_count = [&](int extent) {
        return ctx->bucket_size > 0 ? static_cast<int>(double(float(extent) / float(ctx->bucket_size)) + 0.5) : 0;
    };
    if (dp.rows > 0 && W > 0 && (bucket_count(W) <= 0 || bucket_count(H) <= 0)) {
        CRT_CK(ctx, cudaMemsetAsync(ctx->d_rgb, 0, size_t(W) * size_t(dp.rows) * 3 * sizeof(float), ctx->stream));
        CRT_CK(ctx, cudaMemsetAsync(ctx->d_rgb8, 0, size_t(W) * size_t(dp.rows) * 3, ctx->stream));
    } else if (dp.rows > 0 && W > 0) {
        int pipeline = params->pipeline;
        const bool needs_mega = (ctx->scene.gi_on && dp.gi_rays > 0) || params->shadows;
        if (pipeline == CRT_PIPELINE_AUTO)
            // GI keeps a pixel's recursion on one lane (RNG order), and so do the light loops of the flagged shadow mode
            // (one shadow query per light): the persistent kernel; those under the verification traversal: the
            // per-pixel megakernel; everything else: the wavefront
            pipeline = needs_mega && params->traversal != CRT_TRAVERSAL_KD ? CRT_PIPELINE_MEGAKERNEL
                       : needs_mega                                                              ? CRT_PIPELINE_PERSISTENT
                                                                                                 : CRT_PIPELINE_WAVEFRONT;
            // deep recursions and band-sharded frames end in many small bounce levels, each lasting as long as its
            // longest walk: there everything below the primary level goes to ONE persistent kernel (measured on B200:
            // C4 depth 8 3.51 -> 3.36 ms, its 2-GPU band-sharded frame 2.47 -> 2.01 ms; depth-3 scenes prefer the levels)
            if (pipeline == CRT_PIPELINE_WAVEFRONT && params->traversal == CRT_TRAVERSAL_KD && ctx->light_leaves && ctx->has_secondary &&
                (dp.max_ray_depth >= 5 || params->world > 1))
                pipeline = CRT_PIPELINE_HYBRID;
        if (pipeline == CRT_PIPELINE_PERSISTENT) {
            if (params->traversal != CRT_TRAVERSAL_KD)
                return fail(ctx, CRT_ERR_INVALID, "the persistent pipeline walks the k-d walk tree only (CRT_TRAVERSAL_KD)");
            if (!ctx->d_pctl)
                CRT_CK(ctx, cudaMalloc(&ctx->d_pctl, sizeof(PersistCtl)));
            CRT_CK(ctx, cudaMemsetAsync(ctx->d_pctl, 0, sizeof(PersistCtl), ctx->stream));
            const int patches_x = (W + 7) / 8, patches_y = (dp.rows + 3) / 4;
            const int n_slots = patches_x * patches_y * 32;
            cudaDeviceProp prop;
            if (ctx->sm_count == 0) {
                CRT_CK(ctx, cudaGetDeviceProperties(&prop, ctx->device));
                ctx->sm_count = prop.multiProcessorCount;
            }
            const bool gi_frame = ctx->scene.gi_on && dp.gi_rays > 0;
            const int blocks = std::min(ctx->sm_count * (gi_frame ? CRT_P_MINBLOCKS_GI : CRT_P_MINBLOCKS), (n_slots + kTraceBlock - 1) / kTraceBlock);
            if (params->shadows)
                launch_persistent<false, true>(ctx, dp, stats, gi_frame, blocks, patches_x, n_slots);
            else
                launch_persistent<false, false>(ctx, dp, stats, gi_frame, blocks, patches_x, n_slots);
            ctx->last_launches = 1;
        } else if (pipeline == CRT_PIPELINE_WAVEFRONT && needs_mega)
            return fail(ctx, CRT_ERR_INVALID, "the wavefront pipeline does not handle GI or shadow rays; use CRT_PIPELINE_MEGAKERNEL");
        if (pipeline == CRT_PIPELINE_PERSISTENT) {
            // launched above
        } else if (pipeline == CRT_PIPELINE_HYBRID) {
            // primary level as wavefront kernels (dense, coherent), every deeper level in ONE persistent kernel
            if (needs_mega || params->traversal != CRT_TRAVERSAL_KD)
                return fail(ctx, CRT_ERR_INVALID, "the hybrid pipeline handles neither GI, shadow rays nor the literal traversal");
            int launches = 0;
            const bool deeper = ctx->has_secondary && dp.max_ray_depth > 0;
            // levels 0 .. seed-1 as wavefront kernels; $CRT_HYBRID_LEVEL moves the hand-over
            int seed = 1; // (round 1: C4 3.09 ms with 1, 3.01 ms with 2; with the lean visit of round 2, r3c: 2.25 with 1, 2.31 with 2, 2.39 with 3)
            if (const char *v = std::getenv("CRT_HYBRID_LEVEL"); v && *v)
                seed = std::max(1, std::atoi(v));
            seed = deeper ? std::min(seed, int(dp.max_ray_depth)) : 1;
            cudaError_t e = wavefront_render(ctx->wf, ctx->scene, dp, params->traversal, stats, ctx->has_secondary, ctx->d_rgb, ctx->d_rgb8,
                                             ctx->d_counters, ctx->stream, &launches, ctx->light_leaves, seed,
                                             ctx->arm_early_copy ? ctx->ev_primary : nullptr, ctx->packet_primary);
            ctx->early_recorded = ctx->arm_early_copy;
            ctx->last_wavefront = true;
            if (e != cudaSuccess)
                return fail(ctx, CRT_ERR_CUDA, cudaGetErrorString(e));
            if (deeper) {
                if (!ctx->d_pctl)
                    CRT_CK(ctx, cudaMalloc(&ctx->d_pctl, sizeof(PersistCtl)));
                CRT_CK(ctx, cudaMemsetAsync(ctx->d_pctl, 0, sizeof(PersistCtl), ctx->stream));
                launch_persistent<true, false>(ctx, dp, stats, false, ctx->wf.sm_count * CRT_P_MINBLOCKS, seed, 0);
                ++launches;
                for (int level = seed - 1; level >= 0; --level) {
                    k_wf_resolve<<<ctx->wf.sm_count * 4, 256, 0, ctx->stream>>>(ctx->scene, dp, ctx->wf.b, level, ctx->d_rgb, ctx->d_rgb8);
                    ++launches;
                }
            }
            ctx->last_launches = launches;
        } else if (pipeline == CRT_PIPELINE_WAVEFRONT) {
            int launches = 0;
            cudaError_t e = wavefront_render(ctx->wf, ctx->scene, dp, params->traversal, stats, ctx->has_secondary, ctx->d_rgb, ctx->d_rgb8,
                                             ctx->d_counters, ctx->stream, &launches, ctx->light_leaves, 0,
                                             ctx->arm_early_copy ? ctx->ev_primary : nullptr, ctx->packet_primary);
            ctx->early_recorded = ctx->arm_early_copy;
            ctx->last_wavefront = true;
            if (e != cudaSuccess)
                return fail(ctx, CRT_ERR_CUDA, cudaGetErrorString(e));
            ctx->last_launches = launches;
        } else {
            const dim3 grid((W + kTileW - 1) / kTileW, (dp.rows + kTileH - 1) / kTileH);
            if (params->traversal == CRT_TRAVERSAL_LITERAL)
                launch_mega<CRT_TRAVERSAL_LITERAL>(ctx, dp, stats, grid);
            else
                launch_mega<CRT_TRAVERSAL_KD>(ctx, dp, stats, grid);
            ctx->last_launches = 1;
        }
        CRT_CK(ctx, cudaGetLastError());
    }
    ctx->last_launches += extra_launches;
    CRT_CK(ctx, cudaEventRecord(ctx->ev_end, ctx->stream));
    if (stats)
        CRT_CK(ctx, cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, sizeof(Counters), cudaMemcpyDeviceToHost, ctx->stream));
    return CRT_OK;
}

} // namespace

extern "C" {

void crt_render_params_default(crt_render_params *p) {
    if (!p)
        return;
    std::memset(p, 0, sizeof *p);
    p->max_ray_depth = 3;                // DEFAULT_MAX_RAY_DEPTH, src/core/crt_renderer.h:10
    p->diffuse_reflection_ray_count = 4; // :11
    p->shadow_bias = p->reflection_bias = p->diffuse_reflection_bias = p->refraction_bias = 1e-2f; // :13-16
    p->traversal = CRT_TRAVERSAL_KD;
    p->pipeline = CRT_PIPELINE_AUTO;
    p->world = 1;
    p->band_rows = 8;
}

int crt_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
        return 0;
    return n;
}

int crt_cuda_create(int device, crt_cuda_ctx **out) {
    if (!out)
        return CRT_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        std::fprintf(stderr, "crt_cuda_create: no CUDA device (%s); this renderer has no CPU fallback\n",
                     e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return CRT_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n)
        return CRT_ERR_INVALID;
    crt_cuda_ctx *ctx = new crt_cuda_ctx;
    ctx->device = device;
    bool ok = cudaSetDevice(device) == cudaSuccess && cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreate(&ctx->ev_begin) == cudaSuccess && cudaEventCreate(&ctx->ev_end) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_primary, cudaEventDisableTiming) == cudaSuccess &&
              cudaMalloc(&ctx->d_counters, sizeof(Counters)) == cudaSuccess &&
              cudaHostAlloc(&ctx->h_counters, sizeof(Counters), cudaHostAllocDefault) == cudaSuccess;
    if (!ok) {
        std::fprintf(stderr, "crt_cuda_create: %s\n", cudaGetErrorString(cudaGetLastError()));
        crt_cuda_destroy(ctx);
        return CRT_ERR_CUDA;
    }
    std::memset(ctx->h_counters, 0, sizeof(Counters));
    ctx->stream = ctx->own_stream;
    *out = ctx;
    return CRT_OK;
}

void crt_cuda_destroy(crt_cuda_ctx *ctx) {
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    if (ctx->own_stream)
        cudaStreamSynchronize(ctx->own_stream);
    wavefront_free(ctx->wf);
    cudaFree(ctx->d_arena);
    cudaFreeHost(ctx->h_arena);
    cudaFree(ctx->d_rgb);
    cudaFree(ctx->d_rgb8);
    cudaFree(ctx->d_counters);
    cudaFreeHost(ctx->h_counters);
    release_frame(ctx);
    cudaFree(ctx->d_pctl);
    cudaFree(ctx->d_text);
    cudaFree(ctx->d_scan);
    if (ctx->ev_begin)
        cudaEventDestroy(ctx->ev_begin);
    if (ctx->ev_end)
        cudaEventDestroy(ctx->ev_end);
    if (ctx->ev_primary)
        cudaEventDestroy(ctx->ev_primary);
    if (ctx->copy_stream)
        cudaStreamDestroy(ctx->copy_stream);
    if (ctx->own_stream)
        cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

const char *crt_cuda_last_error(const crt_cuda_ctx *ctx) { return ctx ? ctx->error.c_str() : "null context"; }

int crt_cuda_set_stream(crt_cuda_ctx *ctx, void *stream) {
    if (!ctx)
        return CRT_ERR_INVALID;
    ctx->stream = stream ? static_cast<cudaStream_t>(stream) : ctx->own_stream;
    return CRT_OK;
}

// The walk tree in the device's 8-byte records (crt_device.cuh: kd_node_at).  Nodes are re-numbered so that the children
// of a node sit side by side (a, then b; an absent child takes no slot): the root stays node 0, a node's children are
// numbered when the node is reached, and child a's subtree is laid out before child b's, which keeps the pre-order
// locality of the builder's numbering along the a-chains.  Returns NULL, or what is wrong with the tree.
static const char *pack_walk_nodes(const crt_kd_node *src, int n, int2 *dst) {
    if (n > kNode8MaxNodes)
        return "walk tree has more nodes than the 8-byte device node can address (2^28)";
    std::vector<std::pair<int, int>> todo; // {builder's index, device index}
    todo.reserve(128);
    todo.push_back({ 0, 0 });
    int next = 1;
    while (!todo.empty()) {
        const auto [from, to] = todo.back();
        todo.pop_back();
        const crt_kd_node &k = src[from];
        if (k.axis == 3) {
            if (k.b < 0 || k.b >= kNode8MaxCount)
                return "walk tree leaf with a broken triangle count";
            dst[to] = make_int2(k.a, int(3u | (uint32_t(k.b) << 2)));
            continue;
        }
        if (k.axis < 0 || k.axis > 2 || k.a >= n || k.b >= n)
            return "walk tree node with a broken axis or child index";
        const int has_a = k.a >= 0, has_b = k.b >= 0;
        const int child = next;
        next += has_a + has_b;
        if (next > n)
            return "walk tree is not a tree (nodes reachable twice do not unfold into n records)";
        int split_bits;
        std::memcpy(&split_bits, &k.split, 4);
        dst[to] = make_int2(split_bits, int(uint32_t(k.axis) | uint32_t(has_a) << 2 | uint32_t(has_b) << 3 | uint32_t(child) << 4));
        if (has_b)
            todo.push_back({ k.b, child + has_a });
        if (has_a)
            todo.push_back({ k.a, child });
    }
    for (int i = next; i < n; ++i) // nodes nothing points to (none in the trees of core/): harmless, keep them inert
        dst[i] = make_int2(0, 3);
    return nullptr;
}

int crt_cuda_walk_node_bytes(void) { return int(sizeof(KdNodeRec)); }

int crt_cuda_pack_walk_tree(const crt_kd_node *nodes, int32_t n_nodes, uint64_t *out) {
    if (!nodes || !out || n_nodes < 1)
        return CRT_ERR_INVALID;
    static_assert(sizeof(int2) == sizeof(uint64_t), "8-byte node record");
    return pack_walk_nodes(nodes, n_nodes, reinterpret_cast<int2 *>(out)) ? CRT_ERR_INVALID : CRT_OK;
}

int crt_cuda_upload_scene(crt_cuda_ctx *ctx, const crt_flat_scene *s) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!s)
        return fail(ctx, CRT_ERR_INVALID, "scene is NULL");
    // n_nodes == 0: flattened without the reference's tree (crt_scene_flatten_ex); CRT_TRAVERSAL_LITERAL is then refused
    if (s->n_nodes < 0 || (s->n_nodes > 0 && !s->box_nodes))
        return fail(ctx, CRT_ERR_INVALID, "scene has a broken acceleration tree");
    if (s->n_fast_nodes < 1 || !s->fast_nodes)
        return fail(ctx, CRT_ERR_INVALID, "scene has no walk tree (fast_nodes)");
    if (s->max_depth + 2 > kStack || s->fast_max_depth + 2 > kStack)
        return fail(ctx, CRT_ERR_INVALID, "acceleration tree deeper than the traversal stack");
    if (s->camera.width < 0 || s->camera.height < 0)
        return fail(ctx, CRT_ERR_INVALID, "negative image size");
    CRT_CK(ctx, cudaSetDevice(ctx->device));

    struct Part {
        const void *src;
        size_t bytes;
        size_t offset;
    };
    Part parts[12] = {
        { s->tri_verts, size_t(s->n_triangles) * 3 * sizeof(crt_float4), 0 },
        { s->tri_meta, size_t(s->n_triangles) * 4 * sizeof(int32_t), 0 },
        { s->vert_normal, size_t(s->n_vertices) * sizeof(crt_float4), 0 },
        { s->vert_uv, size_t(s->n_vertices) * 2 * sizeof(float), 0 },
        { s->box_nodes, size_t(s->n_nodes) * sizeof(crt_box_node), 0 },
        { s->fast_nodes, size_t(s->n_fast_nodes) * sizeof(KdNodeRec), 0 }, // + the empty leaf; CRT_NODE8: packed below, not copied
        { s->refs, size_t(s->n_refs) * sizeof(uint32_t), 0 },
        { s->lights, size_t(s->n_lights) * sizeof(crt_float4), 0 },
        { s->materials, size_t(s->n_materials) * sizeof(crt_material), 0 },
        { s->textures, size_t(s->n_textures) * sizeof(crt_texture), 0 },
        { s->texels, size_t(s->n_texels) * sizeof(crt_float4), 0 },
        { s->fast_refs, size_t(s->n_fast_refs) * sizeof(uint32_t), 0 },
    };
    // arena order: what a k-d walk reads (triangles, walk tree, its leaf references) first and in one piece -- the range
    // k_warm_l2 brings into L2 at the start of a frame -- then everything the shading reads, then the reference's tree
    const int order[12] = { 0, 5, 11, 1, 2, 3, 7, 8, 9, 10, 4, 6 };
    size_t total = 0, warm_end = 0;
    for (int k = 0; k < 12; ++k) {
        Part &p = parts[order[k]];
        if (p.bytes && !p.src)
            return fail(ctx, CRT_ERR_INVALID, "scene array pointer is NULL");
        p.offset = total;
        total += align_up((p.bytes ? p.bytes : 16) + (&p == &parts[5] ? sizeof(KdNodeRec) : 0), 256);
        if (k == 2)
            warm_end = total;
    }
    if (total > ctx->d_arena_cap) {
        cudaFree(ctx->d_arena);
        ctx->d_arena = nullptr, ctx->d_arena_cap = 0;
        CRT_CK(ctx, cudaMalloc(&ctx->d_arena, total));
        ctx->d_arena_cap = total;
    }
    if (total > ctx->h_arena_cap) {
        cudaFreeHost(ctx->h_arena);
        ctx->h_arena = nullptr, ctx->h_arena_cap = 0;
        CRT_CK(ctx, cudaHostAlloc(&ctx->h_arena, total, cudaHostAllocDefault));
        ctx->h_arena_cap = total;
    }
    // make sure an earlier frame is not still reading the arena we are about to overwrite
    CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    for (const Part &p : parts)
        if (p.bytes && !(CRT_NODE8 && &p == &parts[5]))
            std::memcpy(ctx->h_arena + p.offset, p.src, p.bytes);
#if CRT_NODE8
    if (const char *why = pack_walk_nodes(s->fast_nodes, s->n_fast_nodes, reinterpret_cast<int2 *>(ctx->h_arena + parts[5].offset)))
        return fail(ctx, CRT_ERR_INVALID, why);
    reinterpret_cast<int2 *>(ctx->h_arena + parts[5].offset)[s->n_fast_nodes] = make_int2(0, 3);
#else
    // one node more than the builder made: a leaf without triangles.  A lane whose pending subtree a helper lane took
    // over finds this node in its stack instead (crt_trace_lean.cuh: lean_steal), so a visit never tests for theft
    reinterpret_cast<int4 *>(ctx->h_arena + parts[5].offset)[s->n_fast_nodes] = make_int4(0, 0, 0, 3);
    if (crtwalk::kLeafNeg) { // the device's leaves: {0, ~first, ~count, 3} (crt_walk_lean.h); the caller's crt_flat_scene stays plain
        int4 *recs = reinterpret_cast<int4 *>(ctx->h_arena + parts[5].offset);
        for (int i = 0; i <= s->n_fast_nodes; ++i)
            if (recs[i].w == 3)
                recs[i].y = ~recs[i].y, recs[i].z = ~recs[i].z;
    }
#endif
    CRT_CK(ctx, cudaMemcpyAsync(ctx->d_arena, ctx->h_arena, total, cudaMemcpyHostToDevice, ctx->stream));

    DScene &d = ctx->scene;
    d = DScene{};
    d.cam = s->camera;
    std::memcpy(d.bg, s->background, sizeof d.bg);
    d.gi_on = s->gi_on, d.reflections_on = s->reflections_on, d.refractions_on = s->refractions_on;
    std::memcpy(d.root_min, s->root_min, sizeof d.root_min);
    std::memcpy(d.root_max, s->root_max, sizeof d.root_max);
    d.n_triangles = s->n_triangles, d.n_nodes = s->n_nodes, d.n_lights = s->n_lights, d.n_materials = s->n_materials, d.n_textures = s->n_textures;
    d.kd_empty_leaf = s->n_fast_nodes;
    d.tri_verts = reinterpret_cast<const float4 *>(ctx->d_arena + parts[0].offset);
    d.tri_meta = reinterpret_cast<const int4 *>(ctx->d_arena + parts[1].offset);
    d.vert_normal = reinterpret_cast<const float4 *>(ctx->d_arena + parts[2].offset);
    d.vert_uv = reinterpret_cast<const float2 *>(ctx->d_arena + parts[3].offset);
    d.box_nodes = reinterpret_cast<const float4 *>(ctx->d_arena + parts[4].offset);
    d.kd_nodes = reinterpret_cast<const KdNodeRec *>(ctx->d_arena + parts[5].offset);
    d.refs = reinterpret_cast<const uint32_t *>(ctx->d_arena + parts[6].offset);
    d.lights = reinterpret_cast<const float4 *>(ctx->d_arena + parts[7].offset);
    d.materials = reinterpret_cast<const int4 *>(ctx->d_arena + parts[8].offset);
    d.textures = reinterpret_cast<const float4 *>(ctx->d_arena + parts[9].offset);
    d.texels = reinterpret_cast<const float4 *>(ctx->d_arena + parts[10].offset);
    d.kd_refs = reinterpret_cast<const uint32_t *>(ctx->d_arena + parts[11].offset);
    ctx->bucket_size = s->bucket_size;
    ctx->has_secondary = false;
    for (int i = 0; i < s->n_materials; ++i)
        if ((s->materials[i].type == 1 && s->reflections_on) || (s->materials[i].type == 2 && s->refractions_on))
            ctx->has_secondary = true;
    // which trace kernel suits the walk tree: mean triangles per leaf (CRT_TRACE_KERNEL=heavy|light overrides)
    size_t leaves = 0;
    for (int i = 0; i < s->n_fast_nodes; ++i)
        leaves += s->fast_nodes[i].axis == 3;
    ctx->light_leaves = leaves > 0 && double(s->n_fast_refs) / double(leaves) <= 6.0;
    if (const char *k = std::getenv("CRT_TRACE_KERNEL"); k && *k)
        ctx->light_leaves = std::strcmp(k, "light") == 0;
    // primary level as warp packets (crt_trace_packet.cuh): pays when a patch of 32 pixels sees few triangles -- the
    // rays of a warp then share their nodes and leaves.  Measured on B200 (tools/packet_sweep.py, profiles/
    // r1z_packet_sweep.txt: height fields covering a quarter of the frame): 5-25 % ahead of the per-lane walk from ~60
    // frame pixels per triangle upwards (C1 0.19 -> 0.17 ms, C3 0.82 -> 0.63 ms, C5 0.69 -> 0.63 ms), 1.3-3x BEHIND it
    // below 20 (C4: 8 pixels per triangle).  Threshold 64; $CRT_PACKET=0|1 overrides.
    ctx->off_lights = parts[7].offset, ctx->off_materials = parts[8].offset, ctx->off_textures = parts[9].offset;
    ctx->warm_base = ctx->d_arena, ctx->warm_bytes = warm_end;
    ctx->l2_window_offset = 0, ctx->l2_window_bytes = warm_end, ctx->l2_policy_on = false; // (a new scene: the window is set again)
    if (const char *k = std::getenv("CRT_L2_WINDOW"); k && std::strcmp(k, "tree") == 0)
        ctx->l2_window_offset = parts[5].offset, ctx->l2_window_bytes = warm_end - parts[5].offset;
    ctx->warm_mode = -1;
    if (const char *k = std::getenv("CRT_WARM_L2"); k && *k)
        ctx->warm_mode = (*k == '1' || *k == '2') ? 1 : 0, ctx->warm_loads = *k == '2';
    ctx->packet_mode = -1; // decided per frame from the image size (enqueue_frame)
    if (const char *k = std::getenv("CRT_PACKET"); k && *k)
        ctx->packet_mode = *k == '1' ? 1 : 0;
    ctx->wf.grow = 1.0; // the ray pool is re-sized per scene (it only grows while one scene is rendered)
    ctx->have_scene = true;
    return CRT_OK;
}

int crt_cuda_update_shading(crt_cuda_ctx *ctx, const crt_flat_scene *s) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!s || !ctx->have_scene)
        return fail(ctx, CRT_ERR_INVALID, "update_shading needs an uploaded scene and a flattened one");
    DScene &d = ctx->scene;
    if (s->n_lights != d.n_lights || s->n_materials != d.n_materials || s->n_textures != d.n_textures)
        return fail(ctx, CRT_ERR_INVALID, "update_shading: the number of lights, materials or textures changed -- upload the scene again");
    if ((s->n_lights && !s->lights) || (s->n_materials && !s->materials) || (s->n_textures && !s->textures))
        return fail(ctx, CRT_ERR_INVALID, "scene array pointer is NULL");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    struct Small {
        const void *src;
        size_t bytes, offset;
    } small[3] = { { s->lights, size_t(s->n_lights) * sizeof(crt_float4), ctx->off_lights },
                   { s->materials, size_t(s->n_materials) * sizeof(crt_material), ctx->off_materials },
                   { s->textures, size_t(s->n_textures) * sizeof(crt_texture), ctx->off_textures } };
    bool changed = false;
    for (const Small &p : small)
        changed = changed || (p.bytes && std::memcmp(ctx->h_arena + p.offset, p.src, p.bytes) != 0);
    if (changed) {
        CRT_CK(ctx, cudaStreamSynchronize(ctx->stream)); // the staging bytes of an earlier copy are about to be overwritten
        for (const Small &p : small)
            if (p.bytes) {
                std::memcpy(ctx->h_arena + p.offset, p.src, p.bytes);
                CRT_CK(ctx, cudaMemcpyAsync(ctx->d_arena + p.offset, ctx->h_arena + p.offset, p.bytes, cudaMemcpyHostToDevice, ctx->stream));
            }
    }
    // the scalar block travels with every launch (DScene is a kernel parameter)
    std::memcpy(d.bg, s->background, sizeof d.bg);
    d.gi_on = s->gi_on, d.reflections_on = s->reflections_on, d.refractions_on = s->refractions_on;
    ctx->bucket_size = s->bucket_size;
    ctx->has_secondary = false;
    for (int i = 0; i < s->n_materials; ++i)
        if ((s->materials[i].type == 1 && s->reflections_on) || (s->materials[i].type == 2 && s->refractions_on))
            ctx->has_secondary = true;
    return CRT_OK;
}

int crt_cuda_set_camera(crt_cuda_ctx *ctx, const crt_camera *camera) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!camera || !ctx->have_scene)
        return fail(ctx, CRT_ERR_INVALID, "set_camera needs an uploaded scene and a camera");
    // the camera block travels as a kernel parameter (__grid_constant__), i.e. with the launch itself
    ctx->scene.cam = *camera;
    return CRT_OK;
}

int crt_cuda_render_device(crt_cuda_ctx *ctx, const crt_render_params *params, crt_render_stats *stats) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (int rc = enqueue_frame(ctx, params))
        return rc;
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->kernel_launches = ctx->last_launches;
        stats->rows_rendered = ctx->last_rows;
    }
    return CRT_OK;
}

int crt_cuda_synchronize(crt_cuda_ctx *ctx, crt_render_stats *stats) {
    if (!ctx)
        return CRT_ERR_INVALID;
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    // A bounce level outgrew the ray pool: size the pool from what the discarded frame counted and render the frame
    // again (rare; the pool then stays large until the next scene upload).  The counts of an overflowed frame are a lower
    // bound -- rays that found no room spawned no children -- so this may take a few rounds: at least double every time,
    // and at least 1.25 x the counted demand.  A frame that still does not fit is an ERROR, never a silently black or
    // stale frame (a pool of P * (2^(d+1) - 2) rays always fits; cudaMalloc fails first for deep glass at large sizes).
    ctx->rerendered = false;
    constexpr int kPoolAttempts = 12;
    for (int attempt = 0; ctx->last_wavefront && ctx->wf.h_flags && *ctx->wf.h_flags; ++attempt) {
        *ctx->wf.h_flags = 0;
        if (attempt == kPoolAttempts)
            return fail(ctx, CRT_ERR_CUDA, "the wavefront ray pool still overflows after 12 re-sized attempts; the frame was NOT rendered");
        ctx->rerendered = true;
        double demand = 0.0;
        for (int l = 1; l < kMaxLevels; ++l)
            demand += double(ctx->wf.h_count[l]);
        const double pixels = double(std::max(1, ctx->wf.b.P));
        ctx->wf.grow = std::max(ctx->wf.grow * 2.0, 1.25 * demand / pixels);
        const crt_render_params again = ctx->last_params;
        if (int rc = enqueue_frame(ctx, &again))
            return rc;
        CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    }
    float ms = 0.0f;
    if (stats && ctx->last_width)
        cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end);
    fill_stats(ctx, stats, ms);
    return CRT_OK;
}

int crt_cuda_render(crt_cuda_ctx *ctx, const crt_render_params *params, float *rgb_out, uint8_t *rgb8_out, crt_render_stats *stats) {
    if (!ctx)
        return CRT_ERR_INVALID;
    // $CRT_EARLY_COPY=1 (wavefront and hybrid pipelines): every pixel whose primary ray ends on a terminal surface is
    // final after the level-0 shade, so the float frame can start travelling there, on the copy engine, under the
    // deeper levels' kernels; the rows that hold reflective / refractive pixels are then copied a second time after
    // resolve.  Off by default: on B200 the overlapped copy (1.8 ms for a 4K float frame) stretches the walk kernels
    // of C4 from 3.05 to 4.2 ms, so the frame reaches the host LATER than with the copy queued behind the kernels
    // (5.03 vs 4.85 ms, tools/e2e_probe.py).
    // ($CRT_EARLY_COPY=2, measurement aid: the row tracking of mode 1 -- its atomics -- WITHOUT the overlapped copy: what of
    // the slowdown is the atomics, what is a 100 MB copy streaming through L2 beside kernels that live on their L2 hits)
    const char *early_env = std::getenv("CRT_EARLY_COPY");
    const bool early_enabled = early_env && (*early_env == '1' || *early_env == '2');
    const bool overlap_copy = early_env && *early_env == '1';
    ctx->arm_early_copy = rgb_out != nullptr && early_enabled;
    ctx->early_recorded = false;
    if (ctx->wf.h_flags)
        ctx->wf.h_flags[1] = ctx->wf.h_flags[2] = 0;
    const int enq = enqueue_frame(ctx, params);
    ctx->arm_early_copy = false;
    if (enq)
        return enq;
    const size_t pixels = size_t(ctx->last_width) * size_t(ctx->last_rows);
    const size_t row_floats = size_t(ctx->last_width) * 3;
    const bool early = ctx->early_recorded && pixels && overlap_copy;
    if (rgb_out && pixels) {
        if (early) {
            CRT_CK(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_primary, 0));
            CRT_CK(ctx, cudaMemcpyAsync(rgb_out, ctx->d_rgb, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->copy_stream));
        } else {
            CRT_CK(ctx, cudaMemcpyAsync(rgb_out, ctx->d_rgb, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        }
    }
    if (rgb8_out && pixels)
        CRT_CK(ctx, cudaMemcpyAsync(rgb8_out, ctx->d_rgb8, pixels * 3, cudaMemcpyDeviceToHost, ctx->stream));
    const int rc = crt_cuda_synchronize(ctx, stats);
    if (early)
        cudaStreamSynchronize(ctx->copy_stream);
    if (rc)
        return rc;
    if (ctx->rerendered) { // the copies above carried the discarded frame
        if (rgb_out && pixels)
            CRT_CK(ctx, cudaMemcpyAsync(rgb_out, ctx->d_rgb, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        if (rgb8_out && pixels)
            CRT_CK(ctx, cudaMemcpyAsync(rgb8_out, ctx->d_rgb8, pixels * 3, cudaMemcpyDeviceToHost, ctx->stream));
        CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    } else if (early && ctx->wf.h_flags && ctx->wf.h_flags[1] > 0) {
        const int hi = std::min(ctx->wf.h_flags[1], ctx->last_rows), lo = std::max(0, 0x7fffffff - ctx->wf.h_flags[2]);
        if (hi > lo) {
            CRT_CK(ctx, cudaMemcpyAsync(rgb_out + size_t(lo) * row_floats, ctx->d_rgb + size_t(lo) * row_floats, size_t(hi - lo) * row_floats * sizeof(float),
                                        cudaMemcpyDeviceToHost, ctx->stream));
            CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
        }
    }
    return CRT_OK;
}

// this rank's row bands -> their place in a full frame in host memory: one strided copy for the whole bands, one more
// for a shorter last band of the image when this rank owns it
static int copy_bands_to_frame(crt_cuda_ctx *ctx, const void *dev_rows, void *frame, size_t pixel_bytes) {
    const crt_render_params &p = ctx->last_params;
    const int H = ctx->scene.cam.height, world = p.world > 1 ? p.world : 1, rank = p.world > 1 ? p.rank : 0;
    const int band = p.band_rows > 0 ? p.band_rows : 8;
    const size_t row_bytes = size_t(ctx->last_width) * pixel_bytes;
    if (world == 1) {
        CRT_CK(ctx, cudaMemcpyAsync(frame, dev_rows, size_t(ctx->last_rows) * row_bytes, cudaMemcpyDeviceToHost, ctx->stream));
        return CRT_OK;
    }
    const int bands = (H + band - 1) / band;
    const int owned = rank < bands ? (bands - 1 - rank) / world + 1 : 0;
    const int last_rows = H - (bands - 1) * band; // rows of the image's last band
    const bool short_last = owned > 0 && last_rows < band && (bands - 1) % world == rank;
    const int whole = owned - (short_last ? 1 : 0);
    char *dst = static_cast<char *>(frame);
    const char *src = static_cast<const char *>(dev_rows);
    if (whole > 0)
        CRT_CK(ctx, cudaMemcpy2DAsync(dst + size_t(rank) * band * row_bytes, size_t(world) * band * row_bytes, src, size_t(band) * row_bytes,
                                      size_t(band) * row_bytes, size_t(whole), cudaMemcpyDeviceToHost, ctx->stream));
    if (short_last)
        CRT_CK(ctx, cudaMemcpyAsync(dst + size_t(bands - 1) * band * row_bytes, src + size_t(whole) * band * row_bytes, size_t(last_rows) * row_bytes,
                                    cudaMemcpyDeviceToHost, ctx->stream));
    return CRT_OK;
}

int crt_cuda_render_frame(crt_cuda_ctx *ctx, const crt_render_params *params, float *rgb_frame, uint8_t *rgb8_frame, crt_render_stats *stats) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (int rc = enqueue_frame(ctx, params))
        return rc;
    for (int pass = 0; pass < 2; ++pass) {
        if (ctx->last_rows > 0 && ctx->last_width > 0) {
            if (rgb_frame)
                if (int rc = copy_bands_to_frame(ctx, ctx->d_rgb, rgb_frame, 3 * sizeof(float)))
                    return rc;
            if (rgb8_frame)
                if (int rc = copy_bands_to_frame(ctx, ctx->d_rgb8, rgb8_frame, 3))
                    return rc;
        }
        if (pass == 0) {
            if (int rc = crt_cuda_synchronize(ctx, stats))
                return rc;
            if (!ctx->rerendered) // (a re-rendered frame -- ray pool grown -- is copied once more)
                break;
        } else {
            CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
        }
    }
    return CRT_OK;
}

int crt_cuda_device_framebuffer(crt_cuda_ctx *ctx, void **rgb_dev, void **rgb8_dev, int *width, int *rows) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (rgb_dev)
        *rgb_dev = ctx->d_rgb;
    if (rgb8_dev)
        *rgb8_dev = ctx->d_rgb8;
    if (width)
        *width = ctx->last_width;
    if (rows)
        *rows = ctx->last_rows;
    return CRT_OK;
}

int crt_cuda_primary_rays(crt_cuda_ctx *ctx, float *rays_out) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!ctx->have_scene || !rays_out)
        return fail(ctx, CRT_ERR_INVALID, "primary_rays needs an uploaded scene and an output buffer");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    const int W = ctx->scene.cam.width, H = ctx->scene.cam.height;
    const size_t n = size_t(W) * H;
    if (!n)
        return CRT_OK;
    float *d = nullptr;
    CRT_CK(ctx, cudaMalloc(&d, n * 6 * sizeof(float)));
    k_primary_rays<<<dim3((W + kTileW - 1) / kTileW, (H + kTileH - 1) / kTileH), kBlock, 0, ctx->stream>>>(ctx->scene, d);
    cudaError_t e = cudaMemcpyAsync(rays_out, d, n * 6 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    CRT_CK(ctx, e);
    return CRT_OK;
}

int crt_cuda_primary_hits(crt_cuda_ctx *ctx, int traversal, int32_t *tri_id, float *t, int32_t *node_tests, int32_t *tri_tests, crt_render_stats *stats) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!ctx->have_scene)
        return fail(ctx, CRT_ERR_INVALID, "no scene uploaded");
    if (traversal < CRT_TRAVERSAL_LITERAL || traversal > CRT_TRAVERSAL_BRUTE)
        return fail(ctx, CRT_ERR_INVALID, "unknown traversal");
    if (traversal == CRT_TRAVERSAL_LITERAL && ctx->scene.n_nodes < 1)
        return fail(ctx, CRT_ERR_INVALID, "CRT_TRAVERSAL_LITERAL needs the reference's acceleration tree: flatten the scene with CRT_FLATTEN_REFERENCE_TREE");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    const int W = ctx->scene.cam.width, H = ctx->scene.cam.height;
    const size_t n = size_t(W) * H;
    if (!n)
        return CRT_OK;
    int32_t *d_tri = nullptr, *d_nn = nullptr, *d_nt = nullptr;
    float *d_t = nullptr;
    cudaError_t e = cudaMalloc(&d_tri, n * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_t, n * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_nn, n * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_nt, n * 4);
    if (e == cudaSuccess) e = cudaMemsetAsync(ctx->d_counters, 0, sizeof(Counters), ctx->stream);
    if (e == cudaSuccess) {
        const dim3 grid((W + kTileW - 1) / kTileW, (H + kTileH - 1) / kTileH);
        cudaEventRecord(ctx->ev_begin, ctx->stream);
        if (traversal == CRT_TRAVERSAL_LITERAL)
            k_primary_hits<CRT_TRAVERSAL_LITERAL><<<grid, kBlock, 0, ctx->stream>>>(ctx->scene, d_tri, d_t, d_nn, d_nt, ctx->d_counters);
        else if (traversal == CRT_TRAVERSAL_KD)
            k_primary_hits<CRT_TRAVERSAL_KD><<<grid, kBlock, 0, ctx->stream>>>(ctx->scene, d_tri, d_t, d_nn, d_nt, ctx->d_counters);
        else
            k_primary_hits<CRT_TRAVERSAL_BRUTE><<<grid, kBlock, 0, ctx->stream>>>(ctx->scene, d_tri, d_t, d_nn, d_nt, ctx->d_counters);
        cudaEventRecord(ctx->ev_end, ctx->stream);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess && tri_id) e = cudaMemcpyAsync(tri_id, d_tri, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && t) e = cudaMemcpyAsync(t, d_t, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && node_tests) e = cudaMemcpyAsync(node_tests, d_nn, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && tri_tests) e = cudaMemcpyAsync(tri_tests, d_nt, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, sizeof(Counters), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_tri), cudaFree(d_t), cudaFree(d_nn), cudaFree(d_nt);
    CRT_CK(ctx, e);
    ctx->last_stats = true;
    ctx->last_launches = 1;
    ctx->last_rows = H;
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end);
    fill_stats(ctx, stats, ms);
    return CRT_OK;
}

int crt_cuda_frame_create(crt_cuda_ctx *ctx, int width, int height, unsigned char handle_out[CRT_IPC_HANDLE_BYTES]) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (width <= 0 || height <= 0)
        return fail(ctx, CRT_ERR_INVALID, "frame_create: non-positive size");
    static_assert(sizeof(cudaIpcMemHandle_t) == CRT_IPC_HANDLE_BYTES, "handle size");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    release_frame(ctx);
    CRT_CK(ctx, cudaMalloc(&ctx->frame, size_t(width) * size_t(height) * 3));
    ctx->frame_kind = 1, ctx->frame_w = width, ctx->frame_h = height;
    if (handle_out) {
        cudaIpcMemHandle_t h;
        CRT_CK(ctx, cudaIpcGetMemHandle(&h, ctx->frame));
        std::memcpy(handle_out, &h, sizeof h);
    }
    return CRT_OK;
}

int crt_cuda_frame_open(crt_cuda_ctx *ctx, const unsigned char handle[CRT_IPC_HANDLE_BYTES], int width, int height) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!handle || width <= 0 || height <= 0)
        return fail(ctx, CRT_ERR_INVALID, "frame_open: handle and a positive size are required");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    release_frame(ctx);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    void *mapped = nullptr;
    CRT_CK(ctx, cudaIpcOpenMemHandle(&mapped, h, cudaIpcMemLazyEnablePeerAccess));
    ctx->frame = static_cast<uint8_t *>(mapped);
    ctx->frame_kind = 2, ctx->frame_w = width, ctx->frame_h = height;
    return CRT_OK;
}

int crt_cuda_frame_attach(crt_cuda_ctx *ctx, void *frame_dev, int width, int height) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!frame_dev || width <= 0 || height <= 0)
        return fail(ctx, CRT_ERR_INVALID, "frame_attach: pointer and a positive size are required");
    release_frame(ctx);
    ctx->frame = static_cast<uint8_t *>(frame_dev);
    ctx->frame_kind = 3, ctx->frame_w = width, ctx->frame_h = height;
    return CRT_OK;
}

int crt_cuda_frame_pointer(crt_cuda_ctx *ctx, void **frame_dev, int *width, int *height) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (frame_dev)
        *frame_dev = ctx->frame;
    if (width)
        *width = ctx->frame_w;
    if (height)
        *height = ctx->frame_h;
    return CRT_OK;
}

int crt_cuda_push_bands(crt_cuda_ctx *ctx) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!ctx->frame)
        return fail(ctx, CRT_ERR_INVALID, "push_bands: no frame (crt_cuda_frame_create / _open / _attach first)");
    if (!ctx->d_rgb8 || ctx->last_width != ctx->frame_w || ctx->scene.cam.height != ctx->frame_h)
        return fail(ctx, CRT_ERR_INVALID, "push_bands: the last rendered frame does not match the frame's size");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    const DParams dp = device_params(ctx->last_params, ctx->frame_h);
    if (dp.rows <= 0)
        return CRT_OK;
    const size_t row_bytes = size_t(ctx->frame_w) * 3;
    const bool vec = row_bytes % 16 == 0 && reinterpret_cast<uintptr_t>(ctx->frame) % 16 == 0 && reinterpret_cast<uintptr_t>(ctx->d_rgb8) % 16 == 0;
    const size_t units = vec ? row_bytes / 16 : row_bytes;
    const int blocks = int(std::min<size_t>((size_t(dp.rows) * units + 255) / 256, size_t(148 * 8)));
    if (vec)
        k_push_bands<uint4><<<blocks, 256, 0, ctx->stream>>>(reinterpret_cast<const uint4 *>(ctx->d_rgb8), reinterpret_cast<uint4 *>(ctx->frame), dp.rows, int(units), dp);
    else
        k_push_bands<uint8_t><<<blocks, 256, 0, ctx->stream>>>(ctx->d_rgb8, ctx->frame, dp.rows, int(units), dp);
    CRT_CK(ctx, cudaGetLastError());
    return CRT_OK;
}

// ---- one host thread driving several GPUs (crt_cuda_group_*): a context per device, the scene
// replicated, the frame cut into interleaved row bands, every device pushing its bands into the frame on devices[0]
struct crt_cuda_group {
    std::vector<crt_cuda_ctx *> ranks;
    std::string error;
};

int crt_cuda_group_create(const int *devices, int n, crt_cuda_group **out) {
    if (!out)
        return CRT_ERR_INVALID;
    *out = nullptr;
    if (!devices || n < 1)
        return CRT_ERR_INVALID;
    crt_cuda_group *g = new crt_cuda_group;
    for (int i = 0; i < n; ++i) {
        crt_cuda_ctx *ctx = nullptr;
        const int rc = crt_cuda_create(devices[i], &ctx);
        if (rc != CRT_OK) {
            crt_cuda_group_destroy(g);
            return rc;
        }
        g->ranks.push_back(ctx);
    }
    // peer stores into devices[0]'s frame need peer access from every other device (same-device contexts need nothing)
    for (int i = 1; i < n; ++i) {
        if (devices[i] == devices[0])
            continue;
        int can = 0;
        cudaDeviceCanAccessPeer(&can, devices[i], devices[0]);
        if (!can) {
            crt_cuda_group_destroy(g);
            return CRT_ERR_CUDA;
        }
        cudaSetDevice(devices[i]);
        const cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
            crt_cuda_group_destroy(g);
            return CRT_ERR_CUDA;
        }
        cudaGetLastError();
    }
    *out = g;
    return CRT_OK;
}

void crt_cuda_group_destroy(crt_cuda_group *g) {
    if (!g)
        return;
    for (size_t i = g->ranks.size(); i-- > 0;) // rank 0 owns the frame the others point into: release it last
        crt_cuda_destroy(g->ranks[i]);
    delete g;
}

const char *crt_cuda_group_last_error(const crt_cuda_group *g) { return g ? g->error.c_str() : "null group"; }

int crt_cuda_group_size(const crt_cuda_group *g) { return g ? int(g->ranks.size()) : 0; }

int crt_cuda_group_upload_scene(crt_cuda_group *g, const crt_flat_scene *scene) {
    if (!g)
        return CRT_ERR_INVALID;
    for (crt_cuda_ctx *ctx : g->ranks)
        if (int rc = crt_cuda_upload_scene(ctx, scene)) {
            g->error = ctx->error;
            return rc;
        }
    return CRT_OK;
}

int crt_cuda_group_update_shading(crt_cuda_group *g, const crt_flat_scene *scene) {
    if (!g)
        return CRT_ERR_INVALID;
    for (crt_cuda_ctx *ctx : g->ranks)
        if (int rc = crt_cuda_update_shading(ctx, scene)) {
            g->error = ctx->error;
            return rc;
        }
    return CRT_OK;
}

int crt_cuda_group_set_camera(crt_cuda_group *g, const crt_camera *camera) {
    if (!g)
        return CRT_ERR_INVALID;
    for (crt_cuda_ctx *ctx : g->ranks)
        if (int rc = crt_cuda_set_camera(ctx, camera)) {
            g->error = ctx->error;
            return rc;
        }
    return CRT_OK;
}

int crt_cuda_group_render(crt_cuda_group *g, const crt_render_params *params, float *rgb_out, uint8_t *rgb8_out, crt_render_stats *stats) {
    if (!g || !params)
        return CRT_ERR_INVALID;
    const int n = int(g->ranks.size());
    crt_cuda_ctx *root = g->ranks[0];
    if (!root->have_scene) {
        g->error = "no scene uploaded";
        return CRT_ERR_INVALID;
    }
    const int W = root->scene.cam.width, H = root->scene.cam.height;
#define CRT_GROUP_CK(ctx, call)            \
    if (int rc_ = (call)) {                \
        g->error = (ctx)->error;           \
        return rc_;                        \
    }
    if (W > 0 && H > 0 && (root->frame_kind != 1 || root->frame_w != W || root->frame_h != H)) {
        CRT_GROUP_CK(root, crt_cuda_frame_create(root, W, H, nullptr));
        for (int i = 1; i < n; ++i)
            CRT_GROUP_CK(g->ranks[i], crt_cuda_frame_attach(g->ranks[i], root->frame, W, H));
    }
    const int band = params->band_rows > 0 ? params->band_rows : 8;
    // a device's outputs: its bands into the frame on devices[0] (peer stores) and / or its float rows straight to their
    // place in the caller's frame, band by band
    auto issue_outputs = [&](crt_cuda_ctx *ctx, int i) -> int {
        if (W <= 0 || H <= 0 || ctx->last_rows <= 0)
            return CRT_OK;
        if (rgb8_out)
            if (int rc = crt_cuda_push_bands(ctx))
                return rc;
        if (rgb_out) {
            const size_t row_bytes = size_t(W) * 3 * sizeof(float);
            int local = 0;
            for (int b = i; b * band < H; b += n) {
                const int rows = std::min(band, H - b * band);
                const cudaError_t e = cudaMemcpyAsync(reinterpret_cast<char *>(rgb_out) + size_t(b) * band * row_bytes,
                                                      reinterpret_cast<const char *>(ctx->d_rgb) + size_t(local) * row_bytes, size_t(rows) * row_bytes,
                                                      cudaMemcpyDeviceToHost, ctx->stream);
                if (e != cudaSuccess) {
                    ctx->error = cudaGetErrorString(e);
                    return CRT_ERR_CUDA;
                }
                local += rows;
            }
        }
        return CRT_OK;
    };
    // enqueue the kernels on every device first, so that all of them are working before any copy is issued: a
    // device-to-host copy into PAGEABLE memory (crt::render_image passes a std::vector) blocks the calling thread until
    // that device is done, which would serialise the devices if it sat in the loop that launches them
    for (int i = 0; i < n; ++i) {
        crt_cuda_ctx *ctx = g->ranks[i];
        crt_render_params p = *params;
        p.rank = i, p.world = n, p.band_rows = band;
        CRT_GROUP_CK(ctx, enqueue_frame(ctx, &p));
    }
    for (int i = 0; i < n; ++i)
        CRT_GROUP_CK(g->ranks[i], issue_outputs(g->ranks[i], i));
    crt_render_stats total{};
    for (int i = 0; i < n; ++i) {
        crt_render_stats st{};
        CRT_GROUP_CK(g->ranks[i], crt_cuda_synchronize(g->ranks[i], &st));
        if (g->ranks[i]->rerendered) { // the frame was rendered again with a larger ray pool: its outputs once more
            CRT_GROUP_CK(g->ranks[i], issue_outputs(g->ranks[i], i));
            if (cudaStreamSynchronize(g->ranks[i]->stream) != cudaSuccess) {
                g->error = "cudaStreamSynchronize failed";
                return CRT_ERR_CUDA;
            }
        }
        total.gpu_ms = std::max(total.gpu_ms, st.gpu_ms);
        total.kernel_launches += st.kernel_launches;
        total.rows_rendered += st.rows_rendered;
        total.rays += st.rays, total.shadow_rays += st.shadow_rays, total.node_tests += st.node_tests, total.tri_tests += st.tri_tests;
        for (int k = 0; k < 16; ++k)
            total.rays_by_depth[k] += st.rays_by_depth[k];
    }
    if (rgb8_out && W > 0 && H > 0) { // every device's push has completed: one copy of the assembled frame
        cudaSetDevice(root->device);
        const cudaError_t e = cudaMemcpy(rgb8_out, root->frame, size_t(W) * size_t(H) * 3, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) {
            g->error = cudaGetErrorString(e);
            return CRT_ERR_CUDA;
        }
    }
#undef CRT_GROUP_CK
    if (stats)
        *stats = total;
    return CRT_OK;
}

int crt_cuda_group_encode_ppm(crt_cuda_group *g, char *text_out, size_t capacity, size_t *length) {
    if (!g)
        return CRT_ERR_INVALID;
    crt_cuda_ctx *root = g->ranks[0];
    const int rc = crt_cuda_encode_ppm_device(root, root->frame, root->frame_w, root->frame_h, text_out, capacity, length);
    if (rc)
        g->error = root->error;
    return rc;
}

int crt_cuda_group_frame(crt_cuda_group *g, void **frame_dev, int *width, int *height) {
    if (!g)
        return CRT_ERR_INVALID;
    return crt_cuda_frame_pointer(g->ranks[0], frame_dev, width, height);
}

int crt_cuda_encode_ppm_device(crt_cuda_ctx *ctx, const void *rgb8_dev, int width, int height, char *text_out, size_t capacity, size_t *length) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!rgb8_dev || width <= 0 || height <= 0)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm needs a device frame with positive dimensions");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    PpmHeader header{};
    header.length = std::snprintf(header.text, sizeof header.text, "P3\n%d %d\n255\n", width, height);
    const size_t pixels = size_t(width) * size_t(height);
    const int n_blocks = int((pixels + kPpmBlockPixels - 1) / kPpmBlockPixels);
    if (size_t(n_blocks) + 1 > ctx->d_scan_cap) {
        cudaFree(ctx->d_scan);
        ctx->d_scan = nullptr, ctx->d_scan_cap = 0;
        CRT_CK(ctx, cudaMalloc(&ctx->d_scan, (size_t(n_blocks) + 1) * sizeof(unsigned long long)));
        ctx->d_scan_cap = size_t(n_blocks) + 1;
    }
    const uint8_t *rgb8 = static_cast<const uint8_t *>(rgb8_dev);
    k_ppm_lengths<<<n_blocks, kPpmThreads, 0, ctx->stream>>>(rgb8, pixels, width, ctx->d_scan);
    k_ppm_scan<<<1, 1024, 0, ctx->stream>>>(ctx->d_scan, n_blocks);
    unsigned long long body = 0;
    CRT_CK(ctx, cudaMemcpyAsync(&body, ctx->d_scan + n_blocks, sizeof body, cudaMemcpyDeviceToHost, ctx->stream));
    CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    const size_t total = size_t(header.length) + size_t(body);
    if (length)
        *length = total;
    if (!text_out)
        return CRT_OK;
    if (capacity < total)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm: text buffer too small (query the length with text_out = NULL)");
    if (total + 16 > ctx->d_text_cap) {
        cudaFree(ctx->d_text);
        ctx->d_text = nullptr, ctx->d_text_cap = 0;
        CRT_CK(ctx, cudaMalloc(&ctx->d_text, total + 16));
        ctx->d_text_cap = total + 16;
    }
    k_ppm_format<<<n_blocks, kPpmThreads, 0, ctx->stream>>>(rgb8, pixels, width, ctx->d_scan, header, ctx->d_text);
    CRT_CK(ctx, cudaGetLastError());
    CRT_CK(ctx, cudaMemcpyAsync(text_out, ctx->d_text, total, cudaMemcpyDeviceToHost, ctx->stream));
    CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

int crt_cuda_encode_ppm(crt_cuda_ctx *ctx, char *text_out, size_t capacity, size_t *length) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!ctx->d_rgb8 || ctx->last_width <= 0 || ctx->last_rows <= 0)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm: no frame has been rendered on this context");
    if (ctx->last_params.world > 1)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm: this context holds only its own row bands; encode the gathered frame with crt_cuda_encode_ppm_device");
    return crt_cuda_encode_ppm_device(ctx, ctx->d_rgb8, ctx->last_width, ctx->last_rows, text_out, capacity, length);
}

#ifdef CRT_DEBUG_TAIL
int crt_debug_tail(unsigned long long *out /* 16*40 */, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, crtdev::g_tail, sizeof(unsigned long long) * 16 * 40);
    if (reset) {
        static unsigned long long zero[16 * 40];
        cudaMemcpyToSymbol(crtdev::g_tail, zero, sizeof zero);
    }
    return 0;
}
#endif

} // extern "C"

// ---- crt_cuda_encode_ppm_noalloc: the P3 encoder for callers that must not allocate ---------------------------------
// (no counterpart in the reference -- include/crt_cuda.h)
//
// The encoder's scratch -- the text bytes of every block of kPpmBlockPixels pixels, then their prefix sums -- lives in
// the file-scope device array below instead of ctx->d_scan, and the text goes straight into the CALLER's device buffer
// instead of ctx->d_text: the call issues three kernels and one 8-byte copy and allocates nothing.  Every GPU that
// loads the library holds the array (64 KiB + 64 B) until the process exits.  All calls on a device share it and it
// keeps the last call's contents: nothing in it has to start at zero (k_ppm_lengths writes entries [0, n_blocks),
// k_ppm_scan entry [n_blocks], before anything reads them -- so there is nothing to clear), but two calls that overlap
// on the GPU (different streams, different threads) race on it.
namespace {

constexpr int kPpmNoAllocMaxBlocks = 8192;                                            // of kPpmBlockPixels (1024) pixels each
constexpr size_t kPpmNoAllocMaxPixels = size_t(kPpmNoAllocMaxBlocks) * kPpmBlockPixels; // 8 388 608: a 3840x2160 frame fits
constexpr int kNoAllocMaxDevices = 64;
static_assert(kPpmNoAllocMaxPixels == CRT_PPM_NOALLOC_MAX_PIXELS, "the limit include/crt_cuda.h states");

static __device__ __align__(256) unsigned long long g_ppm_noalloc_scratch[kPpmNoAllocMaxBlocks + 8];

// In host code `g_ppm_noalloc_scratch` is the host-side stand-in of the symbol, not an address a kernel can use: the
// device address is asked for once per device, with that device current, and kept here.
unsigned long long *g_ppm_noalloc_ptr[kNoAllocMaxDevices] = {};

} // namespace

extern "C" int crt_cuda_encode_ppm_noalloc(crt_cuda_ctx *ctx, const void *rgb8_dev, int width, int height, void *text_dev, size_t capacity,
                                           size_t *length) {
    if (!ctx)
        return CRT_ERR_INVALID;
    if (!rgb8_dev || width <= 0 || height <= 0)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm_noalloc needs a device frame with positive dimensions");
    const size_t pixels = size_t(width) * size_t(height);
    if (pixels > kPpmNoAllocMaxPixels)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm_noalloc: frame larger than the static scratch allows (8388608 pixels); use crt_cuda_encode_ppm_device");
    if (ctx->device < 0 || ctx->device >= kNoAllocMaxDevices)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm_noalloc: device index beyond the 64 this entry point keeps an address for");
    CRT_CK(ctx, cudaSetDevice(ctx->device));
    unsigned long long *scratch = g_ppm_noalloc_ptr[ctx->device];
    if (!scratch) { // set-up, once per device (the same address whichever thread asks first)
        CRT_CK(ctx, cudaGetSymbolAddress(reinterpret_cast<void **>(&scratch), g_ppm_noalloc_scratch));
        g_ppm_noalloc_ptr[ctx->device] = scratch;
    }
    PpmHeader header{};
    header.length = std::snprintf(header.text, sizeof header.text, "P3\n%d %d\n255\n", width, height);
    const int n_blocks = int((pixels + kPpmBlockPixels - 1) / kPpmBlockPixels);
    const uint8_t *rgb8 = static_cast<const uint8_t *>(rgb8_dev);
    k_ppm_lengths<<<n_blocks, kPpmThreads, 0, ctx->stream>>>(rgb8, pixels, width, scratch);
    k_ppm_scan<<<1, 1024, 0, ctx->stream>>>(scratch, n_blocks);
    CRT_CK(ctx, cudaGetLastError());
    unsigned long long body = 0;
    CRT_CK(ctx, cudaMemcpyAsync(&body, scratch + n_blocks, sizeof body, cudaMemcpyDeviceToHost, ctx->stream));
    CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    const size_t total = size_t(header.length) + size_t(body);
    if (length)
        *length = total;
    if (!text_dev)
        return CRT_OK; // length query
    if (capacity < total)
        return fail(ctx, CRT_ERR_INVALID, "encode_ppm_noalloc: text buffer too small (query the length with text_dev = NULL)");
    // k_ppm_format stores exactly text_dev[0, total): single bytes up to the first 16-byte boundary and after the last
    k_ppm_format<<<n_blocks, kPpmThreads, 0, ctx->stream>>>(rgb8, pixels, width, scratch, header, static_cast<char *>(text_dev));
    CRT_CK(ctx, cudaGetLastError());
    CRT_CK(ctx, cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}
