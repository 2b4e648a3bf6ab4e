// Multi-GPU stills, the exchange step: every rank PUSHES its row bands straight into the frame that
// lives in rank 0's memory -- plain 16-byte stores through a peer mapping (NVLink / NVSwitch), with the band
// un-interleave folded into the address computation.  There is no staging buffer, no gather collective and no
// separate scatter kernel on rank 0; the only collective left is the completion barrier the caller issues.
#pragma once

#include <cstdint>

#include "crt_device.cuh"

namespace crtdev {

// src: this rank's compact rows (rows_local x row_bytes); dst: the full frame (height x row_bytes) on rank 0
template <typename Unit>
__global__ void __launch_bounds__(256) k_push_bands(const Unit *__restrict__ src, Unit *__restrict__ dst, int rows_local, int units_per_row, DParams p) {
    const size_t total = size_t(rows_local) * size_t(units_per_row);
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
        const int ly = int(i / size_t(units_per_row));
        const int u = int(i - size_t(ly) * size_t(units_per_row));
        dst[size_t(global_row(p, ly)) * size_t(units_per_row) + u] = src[i];
    }
}

// Brings a byte range into L2 at streaming speed (one prefetch per 128-byte line, no register traffic).  A frame that
// starts with a cold L2 otherwise fetches the walk tree and the triangles one dependent cache miss at a time: DRAM
// latency per node visit instead of L2 latency, which is what a frame cut over 8 GPUs is made of (r2d: 1/8 of the C4
// frame took 0.91 ms, of which 0.72 ms one kernel whose critical path is a chain of dependent walks).
__global__ void __launch_bounds__(256) k_warm_l2(const char *__restrict__ base, size_t bytes) {
    const size_t lines = (bytes + 127) / 128;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < lines; i += size_t(gridDim.x) * blockDim.x)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(base + i * 128));
}

// The same with real loads (ld.global.cg: cached in L2, not in L1): a prefetch may be dropped when the memory system is
// busy, a load cannot.  `sink` is never written (the xor of 16-byte words of scene data is not the magic value).
__global__ void __launch_bounds__(256) k_warm_l2_loads(const uint4 *__restrict__ base, size_t words, unsigned *__restrict__ sink) {
    unsigned acc = 0;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < words; i += size_t(gridDim.x) * blockDim.x) {
        const uint4 v = __ldcg(base + i);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    if (acc == 0x9e3779b9u && sink)
        *sink = acc;
}

} // namespace crtdev
