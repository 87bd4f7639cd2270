// Microbenchmark: issue-slot cost of packed FP32 (FADD2/FMUL2/FFMA2, sm_100a) vs scalar
// FADD/FMUL on B200.  Prints elementwise FP32 ops per cycle per SM for each variant.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32x2_bench fp32x2_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint64_t pack(float a, float b){ uint64_t r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void unpack(uint64_t v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b){ uint64_t r; asm volatile("add.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b){ uint64_t r; asm volatile("mul.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c){ uint64_t r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(r):"l"(a),"l"(b),"l"(c)); return r;}

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed, long long* cyc) {
    const int N = 8;
    float s[2 * N];
    uint64_t p[N];
#pragma unroll
    for (int i = 0; i < 2 * N; ++i) s[i] = seed + (float)(threadIdx.x + i);
#pragma unroll
    for (int i = 0; i < N; ++i) p[i] = pack(s[2 * i], s[2 * i + 1]);
    const uint64_t c = pack(seed, seed * 0.5f);
    unsigned acc = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (MODE == 0) {          // scalar FADD, 16 independent chains
#pragma unroll
                for (int i = 0; i < 2 * N; ++i) s[i] = __fadd_rn(s[i], seed);
            } else if (MODE == 1) {   // FADD2, 8 chains (16 element ops)
#pragma unroll
                for (int i = 0; i < N; ++i) p[i] = add2(p[i], c);
            } else if (MODE == 2) {   // FMUL2
#pragma unroll
                for (int i = 0; i < N; ++i) p[i] = mul2(p[i], c);
            } else if (MODE == 3) {   // FFMA2
#pragma unroll
                for (int i = 0; i < N; ++i) p[i] = fma2(p[i], c, c);
            } else if (MODE == 4) {   // scalar FFMA
#pragma unroll
                for (int i = 0; i < 2 * N; ++i) s[i] = __fmaf_rn(s[i], seed, seed);
            } else if (MODE == 5) {   // FADD2 + one ALU op (LOP3/IADD) per packed op
#pragma unroll
                for (int i = 0; i < N; ++i) { p[i] = add2(p[i], c); acc = (acc ^ (unsigned)it) + (unsigned)i; }
            } else if (MODE == 6) {   // scalar FADD + one ALU op per 2 scalar ops (same element count as mode 5)
#pragma unroll
                for (int i = 0; i < N; ++i) { s[2*i] = __fadd_rn(s[2*i], seed); s[2*i+1] = __fadd_rn(s[2*i+1], seed); acc = (acc ^ (unsigned)it) + (unsigned)i; }
            } else if (MODE == 7) {   // FADD2 + 1 FSETP-like compare per element (as in the kNN scan)
#pragma unroll
                for (int i = 0; i < N; ++i) { p[i] = add2(p[i], c); float a, b; unpack(p[i], a, b); acc += (a <= seed) | (b <= seed); }
            }
        }
    }
    long long t1 = clock64();
    float r = 0;
    if (MODE == 0 || MODE == 4 || MODE == 6) { for (int i = 0; i < 2 * N; ++i) r += s[i]; }
    else { for (int i = 0; i < N; ++i) { float a, b; unpack(p[i], a, b); r += a + b; } }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + (float)acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>
void run(const char* name, float* out, long long* dcyc, int sms) {
    const int iters = 4096, blocks = sms * 8;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    k<MODE><<<blocks, 256>>>(out, iters, 1.0001f, dcyc);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    k<MODE><<<blocks, 256>>>(out, iters, 1.0001f, dcyc);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, a, b);
    long long cyc; cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
    const double elem_ops = (double)blocks * 256 * iters * 4 * 16;
    // per-SM per-cycle using block 0's cycle count: 8 blocks/SM resident at once (256 thr, few regs)
    printf("%-34s %8.3f ms  %7.2f Gelem-op/s  block0 cycles %lld  -> %.1f elem-ops/cycle/SM\n", name, ms,
           elem_ops / ms / 1e6, cyc, (double)8 * 256 * iters * 4 * 16 / (double)cyc);
}

int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out; long long* dcyc;
    cudaMalloc(&out, (size_t)sms * 8 * 256 * 4); cudaMalloc(&dcyc, 8);
    run<0>("scalar FADD", out, dcyc, sms);
    run<4>("scalar FFMA", out, dcyc, sms);
    run<1>("FADD2", out, dcyc, sms);
    run<2>("FMUL2", out, dcyc, sms);
    run<3>("FFMA2", out, dcyc, sms);
    run<6>("scalar FADD + 1 ALU per 2", out, dcyc, sms);
    run<5>("FADD2 + 1 ALU per packed op", out, dcyc, sms);
    run<7>("FADD2 + 2 FSETP per packed op", out, dcyc, sms);
    return 0;
}
