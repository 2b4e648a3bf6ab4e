// crt_gi.cu <-> crt_cuda.cu: launch of the GI instantiations of k_render_persistent (see crt_gi.cu)
#pragma once
#include <cstddef>
#include <cstdint>

void crt_launch_persistent_gi(bool stats, bool seeded, bool shadows, int blocks, void *stream, const void *scene, size_t scene_bytes, const void *dparams,
                              size_t dparams_bytes, void *pctl, int patches_x, int n_slots, float *rgb, uint8_t *rgb8, void *counters, const void *wf,
                              size_t wf_bytes);
