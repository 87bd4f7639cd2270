// bgr_microbench_l2 — on-box L2 read bandwidth (measurement aid of bench.py: the denominator of
// roofline.l2 for the L2-resident kernels; no counterpart in the reference).
#include "common.cuh"
#include "microbench.cuh"

using namespace bgr;

extern "C" int bgr_microbench_l2(int64_t bytes, int mode, int reps, double* gb_per_s) {
    BGR_ARG(gb_per_s != nullptr && mode >= 0 && mode <= 2 && reps >= 1, "bad microbench arguments");
    BGR_ARG(bytes >= (1 << 16) && (bytes & (bytes - 1)) == 0, "buffer size must be a power of two >= 64 KiB");
    uint4* buf = nullptr;
    unsigned long long* sink = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = BGR_OK;
    auto fail = [&](cudaError_t e) {
        set_error("microbench: %s", cudaGetErrorString(e));
        rc = BGR_ERR_CUDA;
    };
    cudaError_t e;
    if ((e = cudaMalloc(&buf, (size_t)bytes)) != cudaSuccess) fail(e);
    if (rc == BGR_OK && (e = cudaMalloc(&sink, 8)) != cudaSuccess) fail(e);
    if (rc == BGR_OK && (e = cudaMemset(buf, 0x5a, (size_t)bytes)) != cudaSuccess) fail(e);
    if (rc == BGR_OK && (e = cudaEventCreate(&e0)) != cudaSuccess) fail(e);
    if (rc == BGR_OK && (e = cudaEventCreate(&e1)) != cudaSuccess) fail(e);
    if (rc == BGR_OK) {
        const int grid = num_sms() * 8, iters = 64;
        float best = 1e30f;
        for (int r = 0; r < reps + 2 && rc == BGR_OK; ++r) {          // 2 warm-up launches load the buffer into L2
            cudaEventRecord(e0, 0);
            l2_read_kernel<<<grid, 256>>>(buf, (uint64_t)bytes, iters, mode, sink);
            note_launch();
            cudaEventRecord(e1, 0);
            if ((e = cudaEventSynchronize(e1)) != cudaSuccess) { fail(e); break; }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            if (r >= 2 && ms < best) best = ms;
        }
        if (rc == BGR_OK) *gb_per_s = l2_read_bytes(mode, grid, iters) / ((double)best * 1e-3) / 1e9;
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(buf);
    cudaFree(sink);
    return rc;
}
