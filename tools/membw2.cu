// membw2.cu -- staging experiments: TMA bulk ring vs cp.async ring vs direct LDG for a cold 28.7 MB read.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <vector>
#include <algorithm>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mb_expect(uint64_t* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mb_wait(uint64_t* b, uint32_t ph) {
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(s32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void bulk(void* d, const void* s, uint32_t n, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(d)), "l"(s), "r"(n), "r"(s32(b)) : "memory");
}
// persistent TMA ring: each CTA owns a contiguous byte range, tiles of TB bytes, S stages
template <int S>
__global__ void __launch_bounds__(256) tma_ring(const char* p, size_t bytes, int TB, float* out) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ uint64_t full[S];
    const int tid = threadIdx.x;
    size_t per = (bytes / gridDim.x + 127) & ~(size_t)127;
    size_t beg = per * blockIdx.x, end = min(bytes, beg + per);
    if (tid == 0) { for (int s = 0; s < S; ++s) mb_init(&full[s], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    size_t nxt = beg;
    if (tid == 0) for (int s = 0; s < S && nxt < end; ++s) { uint32_t n = (uint32_t)min((size_t)TB, end - nxt); mb_expect(&full[s], n); bulk(sm + (size_t)s * TB, p + nxt, n, &full[s]); nxt += n; }
    float m = -1e30f; int q = 0;
    for (size_t cur = beg; cur < end; cur += TB, ++q) {
        int s = q % S; uint32_t n = (uint32_t)min((size_t)TB, end - cur);
        mb_wait(&full[s], (q / S) & 1);
        const float4* t = (const float4*)(sm + (size_t)s * TB);
        for (uint32_t i = tid; i < n / 16; i += 256) { float4 v = t[i]; m = fmaxf(m, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w))); }
        __syncthreads();
        if (tid == 0 && nxt < end) { uint32_t n2 = (uint32_t)min((size_t)TB, end - nxt); mb_expect(&full[s], n2); bulk(sm + (size_t)s * TB, p + nxt, n2, &full[s]); nxt += n2; }
    }
    if (m == 12345.f) out[0] = m;
}
// same ring with per-thread 16-byte cp.async
template <int S>
__global__ void __launch_bounds__(256) cpasync_ring(const char* p, size_t bytes, int TB, float* out) {
    extern __shared__ __align__(128) unsigned char sm[];
    const int tid = threadIdx.x;
    size_t per = (bytes / gridDim.x + 127) & ~(size_t)127;
    size_t beg = per * blockIdx.x, end = min(bytes, beg + per);
    auto load = [&](int s, size_t at) {
        uint32_t n = at < end ? (uint32_t)min((size_t)TB, end - at) : 0u;
        for (uint32_t i = tid * 16; i < n; i += 256 * 16)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s32(sm + (size_t)s * TB + i)), "l"(p + at + i) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    size_t nxt = beg;
    for (int s = 0; s < S; ++s) { load(s, nxt); nxt += TB; }
    float m = -1e30f; int q = 0;
    for (size_t cur = beg; cur < end; cur += TB, ++q) {
        int s = q % S; uint32_t n = (uint32_t)min((size_t)TB, end - cur);
        asm volatile("cp.async.wait_group %0;" ::"n"(S - 1) : "memory");
        __syncthreads();
        const float4* t = (const float4*)(sm + (size_t)s * TB);
        for (uint32_t i = tid; i < n / 16; i += 256) { float4 v = t[i]; m = fmaxf(m, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w))); }
        __syncthreads();
        load(s, nxt); nxt += TB;
    }
    if (m == 12345.f) out[0] = m;
}
// one tile per CTA, non-persistent TMA (the first design)
__global__ void __launch_bounds__(256) tma_once(const char* p, size_t bytes, int TB, float* out) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ uint64_t full;
    const int tid = threadIdx.x;
    size_t beg = (size_t)blockIdx.x * TB; if (beg >= bytes) return;
    uint32_t n = (uint32_t)min((size_t)TB, bytes - beg);
    if (tid == 0) { mb_init(&full, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    if (tid == 0) { mb_expect(&full, n); bulk(sm, p + beg, n, &full); }
    mb_wait(&full, 0);
    float m = -1e30f; const float4* t = (const float4*)sm;
    for (uint32_t i = tid; i < n / 16; i += 256) { float4 v = t[i]; m = fmaxf(m, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w))); }
    if (m == 12345.f) out[0] = m;
}
int main() {
    size_t sz = 28672000;
    float* out; cudaMalloc(&out, 4);
    char* flush; size_t fl = 512u << 20; cudaMalloc(&flush, fl);
    char* p; cudaMalloc(&p, sz); cudaMemset(p, 0, sz);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    auto run = [&](const char* name, auto launch) {
        std::vector<float> t;
        for (int it = 0; it < 20; ++it) {
            cudaMemsetAsync(flush, it, fl);
            cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b); t.push_back(ms * 1e3f);
        }
        std::sort(t.begin(), t.end());
        printf("%-40s cold  min %.2f us  med %.2f us  %s\n", name, t[0], t[10], cudaGetErrorString(cudaGetLastError()));
    };
#define ATTR(k, n) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, n)
    ATTR(tma_ring<3>, 100 * 1024); ATTR(tma_ring<4>, 100 * 1024); ATTR(tma_ring<8>, 100 * 1024); ATTR(tma_ring<2>, 200 * 1024);
    ATTR(cpasync_ring<3>, 100 * 1024); ATTR(cpasync_ring<4>, 100 * 1024); ATTR(cpasync_ring<8>, 100 * 1024);
    ATTR(tma_once, 64 * 1024);
    run("tma_ring<3> 24K x296", [&] { tma_ring<3><<<296, 256, 3 * 24576>>>(p, sz, 24576, out); });
    run("tma_ring<4> 24K x296", [&] { tma_ring<4><<<296, 256, 4 * 24576>>>(p, sz, 24576, out); });
    run("tma_ring<8> 12K x296", [&] { tma_ring<8><<<296, 256, 8 * 12288>>>(p, sz, 12288, out); });
    run("tma_ring<8> 8K x296", [&] { tma_ring<8><<<296, 256, 8 * 8192>>>(p, sz, 8192, out); });
    run("tma_ring<8> 4K x592", [&] { tma_ring<8><<<592, 256, 8 * 4096>>>(p, sz, 4096, out); });
    run("tma_ring<2> 96K x148", [&] { tma_ring<2><<<148, 256, 2 * 98304>>>(p, sz, 98304, out); });
    run("tma_ring<4> 24K x148", [&] { tma_ring<4><<<148, 256, 4 * 24576>>>(p, sz, 24576, out); });
    run("cpasync_ring<3> 24K x296", [&] { cpasync_ring<3><<<296, 256, 3 * 24576>>>(p, sz, 24576, out); });
    run("cpasync_ring<4> 24K x296", [&] { cpasync_ring<4><<<296, 256, 4 * 24576>>>(p, sz, 24576, out); });
    run("cpasync_ring<8> 8K x296", [&] { cpasync_ring<8><<<296, 256, 8 * 8192>>>(p, sz, 8192, out); });
    run("cpasync_ring<8> 8K x592", [&] { cpasync_ring<8><<<592, 256, 8 * 8192>>>(p, sz, 8192, out); });
    run("tma_once 24K x1167", [&] { tma_once<<<(unsigned)((sz + 24575) / 24576), 256, 24576>>>(p, sz, 24576, out); });
    run("tma_once 48K x584", [&] { tma_once<<<(unsigned)((sz + 49151) / 49152), 256, 49152>>>(p, sz, 49152, out); });
    run("tma_once 12K x2334", [&] { tma_once<<<(unsigned)((sz + 12287) / 12288), 256, 12288>>>(p, sz, 12288, out); });
    return 0;
}
