// Warp-level top-k (k <= 32) kept as a lane-distributed ascending list, plus the
// mbarrier / bulk-copy (TMA, SASS UBLKCP) wrappers used to stage node tiles.
//
// Ordering contract (BASELINE.json north_star): neighbours ascend by (d2, index); a key is
// the pair (float bits of d2, index) compared lexicographically as unsigned integers —
// d2 is a sum of squares, so it is never negative and its bit pattern is monotone.
#pragma once
#include "common.cuh"

namespace bgr {

__device__ __forceinline__ bool key_lt(uint32_t ad, uint32_t ai, uint32_t bd, uint32_t bi) {
    return ad < bd || (ad == bd && ai < bi);
}

constexpr uint32_t kKeyMax = 0xffffffffu;
constexpr uint32_t kInfBits = 0x7f800000u;

// Lane `l` holds the l-th smallest key.  Entries >= k are don't-care.
struct WarpList {
    uint32_t d, i;

    __device__ __forceinline__ void init(uint32_t sentinel_d) {
        d = sentinel_d;
        i = 0u;
    }
    // insert a warp-uniform candidate that is known to be < entry k-1
    __device__ __forceinline__ void insert(uint32_t cd, uint32_t ci, int lane) {
        const bool gt = key_lt(cd, ci, d, i);  // my entry is greater -> shifts up by one
        const unsigned m = __ballot_sync(kFull, gt);
        const uint32_t pd = __shfl_up_sync(kFull, d, 1);
        const uint32_t pi = __shfl_up_sync(kFull, i, 1);
        if (gt) {
            const bool prev_gt = lane > 0 && ((m >> (lane - 1)) & 1u);
            d = prev_gt ? pd : cd;
            i = prev_gt ? pi : ci;
        }
    }
    // merge a lane-distributed ASCENDING batch (bd,bi), keep the 32 smallest
    __device__ __forceinline__ void merge_sorted(uint32_t bd, uint32_t bi, int lane) {
        const uint32_t rd = __shfl_sync(kFull, bd, 31 - lane);
        const uint32_t ri = __shfl_sync(kFull, bi, 31 - lane);
        if (key_lt(rd, ri, d, i)) {
            d = rd;
            i = ri;
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            const uint32_t od = __shfl_xor_sync(kFull, d, s);
            const uint32_t oi = __shfl_xor_sync(kFull, i, s);
            const bool lower = (lane & s) == 0;
            const bool o_lt = key_lt(od, oi, d, i);
            const bool m_lt = key_lt(d, i, od, oi);
            if (lower ? o_lt : m_lt) {
                d = od;
                i = oi;
            }
        }
    }
};

// bitonic sort of one key per lane, ascending across lanes
__device__ __forceinline__ void warp_sort32(uint32_t& d, uint32_t& i, int lane) {
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int s = size >> 1; s > 0; s >>= 1) {
            const uint32_t od = __shfl_xor_sync(kFull, d, s);
            const uint32_t oi = __shfl_xor_sync(kFull, i, s);
            const bool up = (lane & size) == 0;  // size==32: always ascending
            const bool lower = (lane & s) == 0;
            const bool take_min = (lower == up);
            const bool o_lt = key_lt(od, oi, d, i);
            const bool m_lt = key_lt(d, i, od, oi);
            if (take_min ? o_lt : m_lt) {
                d = od;
                i = oi;
            }
        }
    }
}

// ---- 64-bit packed keys: (d2 bits << 32) | index, compared as one unsigned integer.
// Keys of real candidates are distinct (indices are unique); the all-ones sentinel may repeat,
// which no compare-exchange below is sensitive to.
constexpr unsigned long long kKey64Max = ~0ull;
__device__ __forceinline__ unsigned long long pack_key(uint32_t d, uint32_t i) {
    return ((unsigned long long)d << 32) | (unsigned long long)i;
}
__device__ __forceinline__ unsigned long long shfl_xor_u64(unsigned long long v, int m) {
    const uint32_t lo = __shfl_xor_sync(kFull, (uint32_t)v, m), hi = __shfl_xor_sync(kFull, (uint32_t)(v >> 32), m);
    return ((unsigned long long)hi << 32) | lo;
}
// bitonic sort, one key per lane, ascending across lanes (15 compare-exchange stages)
__device__ __forceinline__ void warp_sort32_u64(unsigned long long& key, int lane) {
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int s = size >> 1; s > 0; s >>= 1) {
            const unsigned long long o = shfl_xor_u64(key, s);
            const bool take_min = ((lane & size) == 0) == ((lane & s) == 0);
            if ((o < key) == take_min) key = o;
        }
    }
}
// `key` ascending across lanes, `other` ascending across lanes: keep the 32 smallest of the
// union in `key`, ascending (reverse + bitonic merge, 6 stages)
__device__ __forceinline__ void warp_merge32_u64(unsigned long long& key, unsigned long long other, int lane) {
    const uint32_t lo = __shfl_sync(kFull, (uint32_t)other, 31 - lane), hi = __shfl_sync(kFull, (uint32_t)(other >> 32), 31 - lane);
    const unsigned long long r = ((unsigned long long)hi << 32) | lo;
    if (r < key) key = r;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        const unsigned long long o = shfl_xor_u64(key, s);
        const bool take_min = (lane & s) == 0;
        if ((o < key) == take_min) key = o;
    }
}

// Offer one candidate per lane (pass == this lane's candidate may belong to the top-k) to
// the list; tau_* is the warp-uniform key of entry k-1 and is updated.  Few candidates:
// serial ballot/insert; many: bitonic sort + merge.  `eligible` must already contain the
// radius / self-match rules; the exact (d2,index) test against tau is done here.
__device__ __forceinline__ void warp_offer(WarpList& L, bool pass, uint32_t cd, uint32_t ci,
                                           uint32_t& tau_d, uint32_t& tau_i, int k, int lane) {
    pass = pass && key_lt(cd, ci, tau_d, tau_i);
    unsigned pm = __ballot_sync(kFull, pass);
    if (pm == 0u) return;
    if (__popc(pm) >= 8) {
        uint32_t bd = pass ? cd : kKeyMax, bi = pass ? ci : kKeyMax;
        warp_sort32(bd, bi, lane);
        L.merge_sorted(bd, bi, lane);
    } else {
        while (pm) {
            const int src = __ffs(pm) - 1;
            pm &= pm - 1;
            const uint32_t sd = __shfl_sync(kFull, cd, src);
            const uint32_t si = __shfl_sync(kFull, ci, src);
            if (key_lt(sd, si, tau_d, tau_i)) {
                L.insert(sd, si, lane);
                tau_d = __shfl_sync(kFull, L.d, k - 1);
                tau_i = __shfl_sync(kFull, L.i, k - 1);
            }
        }
        return;
    }
    tau_d = __shfl_sync(kFull, L.d, k - 1);
    tau_i = __shfl_sync(kFull, L.i, k - 1);
}

// ---------------------------------------------------------------- mbarrier + bulk copy
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0u;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// 1-D bulk copy global -> shared (TMA engine), completion reported on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

}  // namespace bgr
