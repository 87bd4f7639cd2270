// Internal (non-ABI) accessors between the translation units of libbgrrt.
#pragma once
#include "common.cuh"

namespace bgr {
// node store of a bgr_nns handle (nns.cu)
int nns_reserve(bgr_nns* h, int64_t n);                       // capacity only, NaN padded, size unchanged
void nns_soa(bgr_nns* h, float** x, float** y, float** z, float4** f4);   // current arrays (valid until the next reserve)
void nns_set_size(bgr_nns* h, int64_t n);                     // after a device-side append
int nns_reset(bgr_nns* h);                                    // empty store again (buffers kept, NaN padding restored, index dropped)
cudaStream_t nns_stream(bgr_nns* h);
// Nearest over a query array whose valid length (count_base + *d_count <= nq_cap) is known only on the device
int nns_nearest_counted(bgr_nns* h, const float* d_q, int64_t nq_cap, int64_t nq_hint, const long long* d_count,
                        int64_t count_base, int32_t* d_idx, float* d_dist);
// plain scans laid out for a store of at most n_cap nodes whose current size the kernels read from *n_dev: the same
// launch parameters serve a growing tree (CUDA-graph replays of a planner step).  kNN: flags 0, no radius limit.
int nns_nearest_scan_cap(bgr_nns* h, const float* d_q, int64_t nq_cap, const long long* d_count, int64_t count_base,
                         int32_t* d_idx, float* d_dist, int64_t n_cap, const long long* n_dev);
int nns_knn_brute_cap(bgr_nns* h, const float* d_q, int64_t nq, int k, int32_t* d_idx, float* d_d2, int64_t n_cap,
                      const long long* n_dev);
void nns_scan_scratch(bgr_nns* h, const void** a, const void** b);   // scratch pointers the scans use (graph signature)
int nns_method(const bgr_nns* h);
}  // namespace bgr
