// L2 read microbenchmarks: the roofline the L2-resident kernels (edge sweep, sphere check, grid kNN)
// are held to.  Their maps / node index (<= 32 MiB) live in the 126 MB L2, so HBM bandwidth is not
// what bounds them; these kernels measure, on the box the bench runs on, what the L2 delivers for
// the three access shapes those kernels use:
//   mode 0  stream : coalesced LDG.128, every warp reads consecutive 512-byte rows of the buffer
//   mode 1  sector : every thread reads 8 bytes at a random 32-byte-sector-aligned address
//                    (the DDA's brick-word probes); 32 bytes of L2 traffic per load
//   mode 2  row    : every warp reads a random 512-byte-aligned row with one LDG.128 per lane
//                    (the grid kNN's cell rows)
// All loads bypass L1 (ld.global.cg): the figure is L2 -> SM bandwidth, not L1 hits.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bgr {

__device__ __forceinline__ uint32_t mb_hash(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

constexpr int kMbUnroll = 8;

// bytes = buffer size (power of two), iters = loads per thread / kMbUnroll
__global__ void __launch_bounds__(256) l2_read_kernel(const uint4* __restrict__ buf, uint64_t bytes, int iters, int mode,
                                                      unsigned long long* __restrict__ sink) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    const uint32_t lane = threadIdx.x & 31, warp = tid >> 5, nwarps = nthreads >> 5;
    const uint64_t n16 = bytes / 16, nrows = bytes / 512, nsect = bytes / 32;
    unsigned long long acc = 0;
    if (mode == 0) {
        for (int it = 0; it < iters; ++it) {
            uint4 v[kMbUnroll];
#pragma unroll
            for (int u = 0; u < kMbUnroll; ++u) {
                const uint64_t row = ((uint64_t)(it * kMbUnroll + u) * nwarps + warp) & (nrows - 1);
                v[u] = __ldcg(buf + row * 32 + lane);
            }
#pragma unroll
            for (int u = 0; u < kMbUnroll; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
        }
    } else if (mode == 1) {
        const unsigned long long* b8 = reinterpret_cast<const unsigned long long*>(buf);
        for (int it = 0; it < iters; ++it) {
            unsigned long long v[kMbUnroll];
#pragma unroll
            for (int u = 0; u < kMbUnroll; ++u) {
                const uint64_t s = mb_hash(tid * 2654435761u + (uint32_t)(it * kMbUnroll + u) * 40503u) & (nsect - 1);
                v[u] = __ldcg(b8 + s * 4);
            }
#pragma unroll
            for (int u = 0; u < kMbUnroll; ++u) acc += v[u];
        }
    } else {
        for (int it = 0; it < iters; ++it) {
            uint4 v[kMbUnroll];
#pragma unroll
            for (int u = 0; u < kMbUnroll; ++u) {
                const uint64_t row = mb_hash(warp * 2654435761u + (uint32_t)(it * kMbUnroll + u) * 40503u) & (nrows - 1);
                v[u] = __ldcg(buf + row * 32 + lane);
            }
#pragma unroll
            for (int u = 0; u < kMbUnroll; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
        }
    }
    (void)n16;
    if (acc == 0x123456789abcdefull) *sink = acc;   // keeps the loads alive
}

// L2 bytes one launch moves (sector granularity for mode 1)
inline double l2_read_bytes(int mode, int grid, int iters) {
    const double loads = (double)grid * 256.0 * (double)iters * kMbUnroll;
    return loads * (mode == 1 ? 32.0 : 16.0);
}

}  // namespace bgr
