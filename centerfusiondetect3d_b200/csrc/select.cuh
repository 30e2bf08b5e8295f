// select.cuh -- exact block-level top-K' selection over 64-bit composites (see common.cuh).
//
// block_select<T, J>(src, Kp, scratch, emit) hands the Kp largest composites of the block's elements to
// emit(slot, comp), slot in [0, Kp), in unspecified order.  Composites are pairwise distinct, so the
// winner set is unique and the result is deterministic whatever the scheduling.
//
//   1. every thread keeps the J largest keys of its own elements (registers);
//   2. each warp peels its ceil(Kp/NW) largest values with redux.max + ballot; the block minimum L of
//      those per-warp bounds is a lower bound of the Kp-th largest key: at least NW*ceil(Kp/NW) >= Kp
//      distinct elements are >= L;
//   3. elements with key >= L ("survivors", typically ~2*Kp) are compacted into shared memory with one
//      warp-aggregated atomic per warp-iteration that has any;
//   4. an MSB-first 8-bit radix select over the survivors (shared-memory histogram, warp-0 suffix scan),
//      starting at the highest bit in which [L, max] can differ, finds the exact Kp-th composite;
//   5. survivors >= that composite are emitted.
// If the survivors overflow the shared list (massive ties, tiny inputs with L == 0) steps 4-5 iterate the
// source instead of the list: slower, same result.
//
// Src contract:  template<class F> __device__ void for_each(F&& f)  calls f(comp) a BLOCK-UNIFORM number
// of times in every thread (warp collectives run inside f); comp == 0 means "no element here".
#pragma once
#include "common.cuh"

namespace cf {

constexpr int SEL_CAP = 2048;

struct SelScratch {
    unsigned long long surv[SEL_CAP];
    unsigned int hist[256];
    unsigned int L, kmax, n_surv, n_out, dstar, above, cnt, pad_;
};

template <int T, int J, class Src, class Emit>
__device__ __forceinline__ void block_select(Src& src, const int Kp, SelScratch& s, Emit&& emit) {
    static_assert(T >= 256 && T % 32 == 0, "block size");
    const int tid = threadIdx.x, lane = tid & 31;
    constexpr int NW = T / 32;

    if (tid == 0) {
        s.L = 0xffffffffu;
        s.kmax = 0u;
        s.n_surv = 0u;
        s.n_out = 0u;
    }
    // 1. thread-local top-J keys
    uint32_t top[J];
#pragma unroll
    for (int j = 0; j < J; ++j) top[j] = 0u;
    src.for_each([&](unsigned long long c) {
        uint32_t k = comp_key(c);
#pragma unroll
        for (int j = 0; j < J; ++j) {
            if (k > top[j]) {
                uint32_t t = top[j];
                top[j] = k;
                k = t;
            }
        }
    });
    __syncthreads();

    // 2. per-warp bound, block minimum
    {
        const int rounds = (Kp + NW - 1) / NW;
        uint32_t first = __reduce_max_sync(FULL, top[0]);
        uint32_t m = 0u;
        if (rounds <= 32 * J) {
            for (int r = 0; r < rounds; ++r) {
                m = __reduce_max_sync(FULL, top[0]);
                const uint32_t bal = __ballot_sync(FULL, top[0] == m);
                if (lane == __ffs(bal) - 1) {
#pragma unroll
                    for (int j = 0; j + 1 < J; ++j) top[j] = top[j + 1];
                    top[J - 1] = 0u;
                }
            }
        }
        if (lane == 0) {
            atomicMin(&s.L, m);
            atomicMax(&s.kmax, first);
        }
    }
    __syncthreads();
    const uint32_t L = s.L, kmax = s.kmax;

    // 3. compact survivors
    src.for_each([&](unsigned long long c) {
        const bool p = (c != 0ull) && (comp_key(c) >= L);
        const uint32_t bal = __ballot_sync(FULL, p);
        if (bal) {
            const int leader = __ffs(bal) - 1;
            uint32_t base = 0u;
            if (lane == leader) base = atomicAdd(&s.n_surv, (unsigned)__popc(bal));
            base = __shfl_sync(FULL, base, leader);
            if (p) {
                const uint32_t slot = base + __popc(bal & lanemask_lt());
                if (slot < (uint32_t)SEL_CAP) s.surv[slot] = c;
            }
        }
    });
    __syncthreads();
    const uint32_t n = s.n_surv;
    const bool listed = n <= (uint32_t)SEL_CAP;

    auto iter = [&](auto&& f) {
        if (listed) {
            for (uint32_t i = tid; i < n; i += T) f(s.surv[i]);
        } else {
            src.for_each([&](unsigned long long c) {
                if (c != 0ull && comp_key(c) >= L) f(c);
            });
        }
    };

    // 4. radix select: exact Kp-th largest composite among the survivors (n >= Kp by construction)
    const unsigned long long cmax = ((unsigned long long)kmax << 32) | 0xffffffffull;
    const unsigned long long cmin = (unsigned long long)L << 32;
    unsigned long long Tc = cmin;
    if (n != (uint32_t)Kp) {
        int hb = 63 - __clzll((long long)(cmax ^ cmin));
        unsigned long long known = (hb == 63) ? 0ull : ~((1ull << (hb + 1)) - 1ull);
        unsigned long long prefix = cmax & known;
        uint32_t want = (uint32_t)Kp;
        for (;;) {
            const int shift = hb >= 7 ? hb - 7 : 0;
            const int nb = hb - shift + 1;
            const uint32_t dmask = (1u << nb) - 1u;
            if (tid < 256) s.hist[tid] = 0u;
            __syncthreads();
            iter([&](unsigned long long c) {
                if ((c & known) == prefix) atomicAdd(&s.hist[(uint32_t)(c >> shift) & dmask], 1u);
            });
            __syncthreads();
            if (tid < 32) {
                uint32_t h[8], sum = 0u;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    h[j] = s.hist[lane * 8 + j];
                    sum += h[j];
                }
                uint32_t suf = sum;  // inclusive suffix sum over lanes >= lane
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const uint32_t t = __shfl_down_sync(FULL, suf, off);
                    if (lane + off < 32) suf += t;
                }
                if (suf >= want && suf - sum < want) {
                    uint32_t acc = suf - sum;
#pragma unroll
                    for (int j = 7; j >= 0; --j) {
                        if (acc < want && acc + h[j] >= want) {
                            s.dstar = (uint32_t)(lane * 8 + j);
                            s.above = acc;
                            s.cnt = h[j];
                        }
                        acc += h[j];
                    }
                }
            }
            __syncthreads();
            const uint32_t d = s.dstar, above = s.above, cnt = s.cnt;
            want -= above;
            prefix |= (unsigned long long)d << shift;
            known |= (unsigned long long)dmask << shift;
            if (cnt == want || shift == 0) {
                Tc = prefix;
                break;
            }
            hb = shift - 1;
        }
    }

    // 5. emit the winners
    iter([&](unsigned long long c) {
        if (c >= Tc) {
            const uint32_t slot = atomicAdd(&s.n_out, 1u);
            emit(slot, c);
        }
    });
    __syncthreads();
}

}  // namespace cf
