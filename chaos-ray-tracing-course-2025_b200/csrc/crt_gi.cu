// The GI instantiations of k_render_persistent (one PCG32 stream per pixel, depth-first, crt_renderer.cpp:60-79), built
// as their own translation unit so that THEIR vector divisions go through binary64 (crt_device.cuh: CRT_DIV_VIA_DOUBLE,
// the same bits as the binary32 division, without its fix-up calls on zero numerators: 15-01/scene2 211 -> 186 ms)
// while every other kernel keeps the binary32 division, which is faster where numerators are never zero.  The whole
// device code is included under another namespace name, so nothing here collides with crt_cuda.cu's copy; the record
// types are the same declarations and therefore the same layout.
#define CRT_DIV_VIA_DOUBLE 1
#define crtdev crtdev_gi
#include "crt_kernels.cuh"
#include "crt_wavefront.cuh"
#include "crt_persistent.cuh"
#undef crtdev

#include "crt_gi.h"

namespace {

template <bool SEEDED, bool SHADOWS>
void launch(bool stats, int blocks, cudaStream_t stream, const crtdev_gi::DScene &scene, const crtdev_gi::DParams &dp, crtdev_gi::PersistCtl *pctl,
            int patches_x, int n_slots, float *rgb, uint8_t *rgb8, crtdev_gi::Counters *counters, const crtdev_gi::WfBuffers &wf) {
    if (stats)
        crtdev_gi::k_render_persistent<true, true, SEEDED, SHADOWS><<<blocks, crtdev_gi::kTraceBlock, 0, stream>>>(scene, dp, pctl, patches_x, n_slots, rgb, rgb8, counters, wf);
    else
        crtdev_gi::k_render_persistent<false, true, SEEDED, SHADOWS><<<blocks, crtdev_gi::kTraceBlock, 0, stream>>>(scene, dp, pctl, patches_x, n_slots, rgb, rgb8, counters, wf);
}

} // namespace

void crt_launch_persistent_gi(bool stats, bool seeded, bool shadows, int blocks, void *stream, const void *scene, size_t scene_bytes, const void *dparams,
                              size_t dparams_bytes, void *pctl, int patches_x, int n_slots, float *rgb, uint8_t *rgb8, void *counters, const void *wf,
                              size_t wf_bytes) {
    // the caller's records are crtdev::DScene / DParams / WfBuffers: same headers, same layout -- checked, not assumed
    if (scene_bytes != sizeof(crtdev_gi::DScene) || dparams_bytes != sizeof(crtdev_gi::DParams) || wf_bytes != sizeof(crtdev_gi::WfBuffers))
        abort();
    const auto &s = *static_cast<const crtdev_gi::DScene *>(scene);
    const auto &p = *static_cast<const crtdev_gi::DParams *>(dparams);
    const auto &w = *static_cast<const crtdev_gi::WfBuffers *>(wf);
    auto *ctl = static_cast<crtdev_gi::PersistCtl *>(pctl);
    auto *cn = static_cast<crtdev_gi::Counters *>(counters);
    auto st = static_cast<cudaStream_t>(stream);
    if (seeded && shadows)
        launch<true, true>(stats, blocks, st, s, p, ctl, patches_x, n_slots, rgb, rgb8, cn, w);
    else if (seeded)
        launch<true, false>(stats, blocks, st, s, p, ctl, patches_x, n_slots, rgb, rgb8, cn, w);
    else if (shadows)
        launch<false, true>(stats, blocks, st, s, p, ctl, patches_x, n_slots, rgb, rgb8, cn, w);
    else
        launch<false, false>(stats, blocks, st, s, p, ctl, patches_x, n_slots, rgb, rgb8, cn, w);
}
