// Wavefront pipeline (ray queues + compaction between bounce stages).  Placeholder interface: filled in next.
#pragma once
#include "crt_kernels.cuh"

namespace crtdev {

constexpr bool kWavefrontReady = false;

struct WavefrontState {};

inline void wavefront_free(WavefrontState &) {}

inline cudaError_t wavefront_render(WavefrontState &, const DScene &, const DParams &, int, bool, float *, uint8_t *, Counters *, cudaStream_t, int *) {
    return cudaErrorNotSupported;
}

} // namespace crtdev
