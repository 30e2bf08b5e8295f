// common.cuh -- device helpers shared by the fusion-stage kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cf {

constexpr uint32_t FULL = 0xffffffffu;

// ---- order-preserving float <-> uint32 key ----------------------------------------------------------
// key(a) < key(b)  <=>  a < b for all non-NaN floats; -0 is canonicalised to +0 first so that the two
// zeros tie exactly like the float compare the reference (torch.topk) performs.
__device__ __forceinline__ uint32_t f2key(float v) {
    v = __fadd_rn(v, 0.0f);
    uint32_t b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
    uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(b);
}
constexpr uint32_t KEY_ZERO = 0x80000000u;  // f2key(0.0f)

// A candidate is one 64-bit composite: (key << 32) | ~gidx, gidx = class*H*W + y*W + x.  Descending
// composite order == (score desc, class asc, flat index asc): the total order of the oracle
// (oracle/cfd3d_oracle.c cand_before).  Composites of distinct pixels are distinct; 0 = "no element".
__device__ __forceinline__ unsigned long long make_comp(uint32_t key, uint32_t gidx) {
    return ((unsigned long long)key << 32) | (unsigned long long)(~gidx);
}
__device__ __forceinline__ uint32_t comp_key(unsigned long long c) { return (uint32_t)(c >> 32); }
__device__ __forceinline__ uint32_t comp_gidx(unsigned long long c) { return ~(uint32_t)c; }

__device__ __forceinline__ float fmax3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm volatile("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ---- shared-memory addresses, mbarrier and 1-D TMA bulk copy (cp.async.bulk -> SASS UBLKCP) ---------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0u;
}
// try_wait suspends the thread for a hardware time slice per attempt; a copy that never lands (a protocol bug) ends
// in a trap -- an error the host sees -- instead of a hung GPU
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins)
        if (spins > (1u << 22)) __trap();
}
// contiguous global -> shared bulk copy; dst, src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---- Python slice semantics: slice(lo, hi).indices(n) with step 1 -----------------------------------
__device__ __forceinline__ void pyslice(long long lo, long long hi, long long n, int& a, int& b) {
    if (lo < 0) {
        lo += n;
        if (lo < 0) lo = 0;
    } else if (lo > n)
        lo = n;
    if (hi < 0) {
        hi += n;
        if (hi < 0) hi = 0;
    } else if (hi > n)
        hi = n;
    a = (int)lo;
    b = (int)hi;  // empty when a >= b
}

}  // namespace cf
