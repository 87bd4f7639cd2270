// Standalone form of the L2 read microbenchmark (the library exports the same kernels as
// bgr_microbench_l2, which bench.py calls live):
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/micro/l2_read_bench scripts/micro/l2_read_bench.cu
//   ./scripts/micro/l2_read_bench            -> one line per (buffer size, access shape)
#include <cstdio>
#include "../../bottleneck_guided_rrt_b200/csrc/microbench.cuh"

int main() {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const char* names[3] = {"stream LDG.128 rows", "random 32 B sectors (8 B loads)", "random 512 B rows (LDG.128)"};
    unsigned long long* sink;
    cudaMalloc(&sink, 8);
    for (long long mib : {16LL, 32LL, 64LL, 256LL}) {
        uint4* buf;
        const long long bytes = mib << 20;
        if (cudaMalloc(&buf, bytes) != cudaSuccess) return 1;
        cudaMemset(buf, 0x5a, bytes);
        for (int mode = 0; mode < 3; ++mode) {
            const int grid = sms * 8, iters = 64;
            cudaEvent_t a, b;
            cudaEventCreate(&a); cudaEventCreate(&b);
            float best = 1e30f;
            for (int r = 0; r < 12; ++r) {
                cudaEventRecord(a);
                bgr::l2_read_kernel<<<grid, 256>>>(buf, (uint64_t)bytes, iters, mode, sink);
                cudaEventRecord(b);
                if (cudaEventSynchronize(b) != cudaSuccess) return 2;
                float ms;
                cudaEventElapsedTime(&ms, a, b);
                if (r >= 2 && ms < best) best = ms;
            }
            printf("%4lld MiB  %-34s %9.1f GB/s  (%.3f ms)\n", mib, names[mode], bgr::l2_read_bytes(mode, grid, iters) / (best * 1e-3) / 1e9, best);
            cudaEventDestroy(a); cudaEventDestroy(b);
        }
        cudaFree(buf);
    }
    return 0;
}
