// P3 text encoding of a quantised frame on the device: byte for byte the stream the reference's
// write_ppm produces (src/core/crt_image_ppm.cpp:9-23): "P3\nW H\n255\n", then "r g b\t" per pixel and "\n" per row.
// The channel bytes are already clamp(int(c * 255), 0, 255) (quantize(), crt_kernels.cuh), so all that is left is
// variable-length decimal formatting: lengths -> prefix sum -> formatted text staged through shared memory and
// written out with aligned 16-byte stores.
#pragma once

#include <cstdint>

namespace crtdev {

constexpr int kPpmThreads = 256;
constexpr int kPpmPerThread = 4;
constexpr int kPpmBlockPixels = kPpmThreads * kPpmPerThread;
constexpr int kPpmMaxPixelBytes = 13; // "255 255 255\t" + "\n"

struct PpmHeader {
    char text[48];
    int length;
};

__device__ __forceinline__ int ppm_digits(unsigned v) { return 1 + (v >= 10u) + (v >= 100u); }

__device__ __forceinline__ int ppm_pixel_length(const uint8_t *__restrict__ rgb8, size_t pixel, int width) {
    const uint8_t *p = rgb8 + pixel * 3;
    return ppm_digits(p[0]) + ppm_digits(p[1]) + ppm_digits(p[2]) + 3 + (int(pixel % size_t(width)) == width - 1 ? 1 : 0);
}

__device__ __forceinline__ char *ppm_put(char *out, unsigned v) {
    if (v >= 100u)
        *out++ = char('0' + v / 100u);
    if (v >= 10u)
        *out++ = char('0' + (v / 10u) % 10u);
    *out++ = char('0' + v % 10u);
    return out;
}

// block-wide exclusive prefix sum of one int per thread (kPpmThreads threads); returns the block total in `total`
__device__ __forceinline__ int ppm_block_scan(int value, int &total) {
    __shared__ int warp_sums[kPpmThreads / 32];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    int incl = value;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (int(lane) >= o)
            incl += v;
    }
    if (lane == 31)
        warp_sums[warp] = incl;
    __syncthreads();
    int before = 0, all = 0;
#pragma unroll
    for (int i = 0; i < kPpmThreads / 32; ++i) {
        const int s = warp_sums[i];
        before += i < int(warp) ? s : 0;
        all += s;
    }
    __syncthreads();
    total = all;
    return before + incl - value;
}

// pass 1: text bytes of every block of kPpmBlockPixels pixels
__global__ void __launch_bounds__(kPpmThreads) k_ppm_lengths(const uint8_t *__restrict__ rgb8, size_t pixels, int width, unsigned long long *__restrict__ block_bytes) {
    const size_t first = size_t(blockIdx.x) * kPpmBlockPixels + size_t(threadIdx.x) * kPpmPerThread;
    int mine = 0;
#pragma unroll
    for (int k = 0; k < kPpmPerThread; ++k)
        if (first + k < pixels)
            mine += ppm_pixel_length(rgb8, first + k, width);
    int total;
    ppm_block_scan(mine, total);
    if (threadIdx.x == 0)
        block_bytes[blockIdx.x] = (unsigned long long)total;
}

// pass 2: exclusive prefix sum over the blocks (one CTA; a 4K frame has 8100 blocks); block_bytes[n_blocks] = total
__global__ void __launch_bounds__(1024) k_ppm_scan(unsigned long long *__restrict__ block_bytes, int n_blocks) {
    __shared__ unsigned long long warp_sums[32];
    __shared__ unsigned long long carry;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0)
        carry = 0ull;
    __syncthreads();
    for (int base = 0; base < n_blocks; base += 1024) {
        const int i = base + int(threadIdx.x);
        const unsigned long long value = i < n_blocks ? block_bytes[i] : 0ull;
        unsigned long long incl = value;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long v = __shfl_up_sync(0xffffffffu, incl, o);
            if (int(lane) >= o)
                incl += v;
        }
        if (lane == 31)
            warp_sums[warp] = incl;
        __syncthreads();
        unsigned long long before = carry, all = 0ull;
        for (int w = 0; w < 32; ++w) {
            before += w < int(warp) ? warp_sums[w] : 0ull;
            all += warp_sums[w];
        }
        if (i < n_blocks)
            block_bytes[i] = before + incl - value;
        __syncthreads();
        if (threadIdx.x == 0)
            carry += all;
        __syncthreads();
    }
    if (threadIdx.x == 0)
        block_bytes[n_blocks] = carry;
}

// pass 3: format into shared memory at the destination's 16-byte phase, then aligned vector stores
__global__ void __launch_bounds__(kPpmThreads) k_ppm_format(const uint8_t *__restrict__ rgb8, size_t pixels, int width,
                                                            const unsigned long long *__restrict__ block_offsets, const PpmHeader header,
                                                            char *__restrict__ text) {
    __shared__ __align__(16) char stage[kPpmBlockPixels * kPpmMaxPixelBytes + 32];
    const size_t first = size_t(blockIdx.x) * kPpmBlockPixels + size_t(threadIdx.x) * kPpmPerThread;
    int mine = 0;
#pragma unroll
    for (int k = 0; k < kPpmPerThread; ++k)
        if (first + k < pixels)
            mine += ppm_pixel_length(rgb8, first + k, width);
    int total;
    const int before = ppm_block_scan(mine, total);
    char *dst = text + header.length + block_offsets[blockIdx.x];
    const int phase = int(reinterpret_cast<uintptr_t>(dst) & 15u);
    char *out = stage + phase + before;
#pragma unroll
    for (int k = 0; k < kPpmPerThread; ++k) {
        const size_t pixel = first + k;
        if (pixel < pixels) {
            const uint8_t *p = rgb8 + pixel * 3;
            out = ppm_put(out, p[0]);
            *out++ = ' ';
            out = ppm_put(out, p[1]);
            *out++ = ' ';
            out = ppm_put(out, p[2]);
            *out++ = '\t';
            if (int(pixel % size_t(width)) == width - 1)
                *out++ = '\n';
        }
    }
    if (blockIdx.x == 0 && int(threadIdx.x) < header.length)
        text[threadIdx.x] = header.text[threadIdx.x];
    __syncthreads();
    // stage[phase .. phase + total) -> dst[0 .. total): bytes up to the first 16-byte boundary, vectors, tail bytes
    const int head = min(total, (16 - phase) & 15);
    const int vectors = (total - head) / 16;
    const int tail_at = head + vectors * 16;
    if (int(threadIdx.x) < head)
        dst[threadIdx.x] = stage[phase + threadIdx.x];
    const uint4 *src16 = reinterpret_cast<const uint4 *>(stage + phase + head);
    uint4 *dst16 = reinterpret_cast<uint4 *>(dst + head);
    for (int i = threadIdx.x; i < vectors; i += kPpmThreads)
        dst16[i] = src16[i];
    if (int(threadIdx.x) < total - tail_at)
        dst[tail_at + threadIdx.x] = stage[phase + tail_at + threadIdx.x];
}

} // namespace crtdev
