// stage.cu -- the fused fusion-stage kernel: ONE launch per call, one thread-block cluster per frame.
//
// Replaces (reference, src/lib/): model/utils.py:112-128 nms, :6-38 topk, :41-72 transposeAndGetFeature;
// utils/pointcloud.py:331-394 getPcFrustumHeatmap with :397-481 cvtPcDepthToHeatmap under its B x K loop;
// model/decode.py:10-174 fusionDecode.
//
// A frame's heat-map is a (C*H) x W row matrix.  The CS CTAs of a cluster (CS = 1..16, chosen by the host so that
// the batch covers the chip in one wave) split its rows; each CTA streams its share from HBM ONCE with 1-D TMA bulk
// copies (cp.async.bulk + mbarrier complete_tx) into shared-memory slots and KEEPS it there: at the reference
// sizes (448x800 input, batch 32: 4 CTAs x 224 000 B) the whole batch of heat-maps sits in the chip's shared
// memory between the two looks the exact selection needs.  Per frame:
//
//   A  bound    every lane takes the maximum of its own float4 items of a chunk; the lane groups of a frame are
//               disjoint, so the K-th largest group maximum is a lower bound of the K-th largest pixel (for the
//               nms'd list a group maximum only counts if that pixel survives the 3x3 suppression).  Group maxima
//               go into a 4096-bin histogram DISTRIBUTED over the cluster's shared memory (bin -> CTA bin % CS,
//               16-bit counters, red.shared::cluster).
//   -- cluster barrier --
//      tau      every CTA scans the distributed histogram top-down (ld.shared::cluster): tau = lower edge of the
//               bin holding the K-th largest group maximum.
//   B  filter   second look, from shared memory: B1 scans (one vote per 512 pixels) and parks the candidate pixels in
//               a CTA-wide work list; B2 tests them side by side, one per thread (threshold, 3x3 suppression) and
//               appends the survivors to the frame's lists (typically 1.1-3 x K entries; one reservation per CTA).
//   -- cluster barrier --
//      select   every CTA ranks the two survivor lists itself (the two jobs on the two halves of the CTA, rank by
//               counting over key buckets): no delivery, no barrier.  Meanwhile the frame's radar depth plane
//               pc_dep[b,0] lands in the shared memory the heat-map no longer needs (one bulk copy).
//      tail     detection k belongs to CTA k % CS: head gathers, frustum geometry / decode epilogue (merge.cuh);
//   D  assoc    one warp per detection: ROI scan of the staged depth plane, warp-shuffle lexicographic (depth, pixel)
//               minimum (assoc.cuh); the K paint records are broadcast to every CTA.
//   -- cluster barrier --
//   E  paint    each CTA paints its band of rows of pc_hm: owner map in shared memory, every pixel written once.
//
// Shares that do not fit (full-resolution maps, single-CTA clusters) run the same code as a ring: the slots are
// recycled in pass A, the last S chunks stay resident, and pass B re-fetches the others (L2 hits) through the slots
// it frees.  The warp that finishes with a slot issues the next bulk copy into it, so there is no producer warp and
// no "empty" barrier.  Inputs the TMA cannot take (W % 4 != 0, unaligned base) are staged by the consuming warp.
#include <math_constants.h>

#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>

#include "assoc.cuh"
#include "geometry.cuh"
#include "kernels.cuh"
#include "paint.cuh"
#include "select.cuh"

namespace cf {

constexpr int ST = 512;          // threads per CTA
constexpr int SNW = ST / 32;     // warps
constexpr int WL_CAP = 640;      // pass-B work list: candidate pixels a CTA parks before it tests them side by side
constexpr int WL_CAP_RING = 4096;  // ... when the share streams through a ring of slots (shared memory to spare)
constexpr int FL_CAP = 256;      // pass-B flagged-group list (resident shares)
constexpr int FL_CAP_STREAM = 2048;  // ... of a streamed share (a CTA of a small cluster holds several times the rows)
constexpr int HOT_CAP = 192;     // hot list: survivors whose head values are prefetched early (a hint: overflow is dropped)
constexpr int HOT_DIV = 2;       // raw list: hot = at or above the bound of the K / HOT_DIV largest group maxima (covers nearly
                                 // all winners at ~1.3 x K entries; the nms'd list has only ~1.3 x K survivors: all are hot)
constexpr int MAX_SLOTS = 64;
constexpr int MAX_CS = 16;
constexpr size_t SMEM_DYN_CAP = 227 * 1024 - 1792;  // leaves room for the kernel's static shared memory

// ---- cluster primitives ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cl_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cl_size() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
// all threads of all CTAs of the cluster; release/acquire at cluster scope (shared::cluster and global memory)
__device__ __forceinline__ void cl_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void cl_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cl_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t cl_map(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ uint32_t cl_atom_add(uint32_t raddr, uint32_t v) {
    uint32_t r;
    asm volatile("atom.shared::cluster.add.u32 %0, [%1], %2;" : "=r"(r) : "r"(raddr), "r"(v) : "memory");
    return r;
}
__device__ __forceinline__ void cl_red_add(uint32_t raddr, uint32_t v) {
    asm volatile("red.shared::cluster.add.u32 [%0], %1;" ::"r"(raddr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t cl_ld(uint32_t raddr) {
    uint32_t v;
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(raddr) : "memory");
    return v;
}
__device__ __forceinline__ void cl_st(uint32_t raddr, uint32_t v) {
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(raddr), "r"(v) : "memory");
}
// generic-proxy accesses to shared memory before this fence are ordered before later async-proxy (TMA) accesses
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// development aid (library built with -DCFD3D_PROFILE, tools/prof_phases.py): per-CTA clock64 stamps at the phase
// boundaries of the last frame a CTA processed.  The product build carries neither the array nor the stamps.
#ifdef CFD3D_PROFILE
constexpr int PROF_PTS = 48;
__device__ long long g_prof[1024 * PROF_PTS];
#define CF_MERGE_PROF(i)                                                                                  \
    do {                                                                                                  \
        if (p.prof && threadIdx.x == 0 && blockIdx.x < 1024) g_prof[blockIdx.x * PROF_PTS + 16 + (i)] = clock64(); \
    } while (0)
#define CF_MERGE_PROF_DEP(i, val)                                                                          \
    do {                                                                                                  \
        if (p.prof && threadIdx.x == 0 && blockIdx.x < 1024) {                                            \
            long long t_;                                                                                 \
            asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_) : "r"(val) : "memory");                      \
            g_prof[blockIdx.x * PROF_PTS + 16 + (i)] = t_;                                                \
        }                                                                                                 \
    } while (0)
#define CF_PROF_SLOT(thr, i, val)                                                                            \
    do {                                                                                                  \
        if (p.prof && threadIdx.x == (thr) && blockIdx.x < 1024) {                                        \
            long long t_;                                                                                 \
            asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_) : "r"(val) : "memory");                      \
            g_prof[blockIdx.x * PROF_PTS + (i)] = t_;                                                     \
        }                                                                                                 \
    } while (0)
#define PROF(i)                                                                              \
    do {                                                                                     \
        if (p.prof && tid == 0 && blockIdx.x < 1024) g_prof[blockIdx.x * PROF_PTS + (i)] = clock64(); \
    } while (0)
#define PROF_VAL(i, v)                                                                         \
    do {                                                                                     \
        if (p.prof && tid == 0 && blockIdx.x < 1024) g_prof[blockIdx.x * PROF_PTS + (i)] = (long long)(v); \
    } while (0)
#define PROF_DEP(i, val)                                                                        \
    do {                                                                                     \
        if (p.prof && tid == 0 && blockIdx.x < 1024) {                                       \
            long long t_;                                                                    \
            asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_) : "r"(val) : "memory");         \
            g_prof[blockIdx.x * PROF_PTS + (i)] = t_;                                        \
        }                                                                                    \
    } while (0)
#else
#define CF_MERGE_PROF(i) do { } while (0)
#define CF_MERGE_PROF_DEP(i, val) do { } while (0)
#define CF_PROF_SLOT(thr, i, val) do { } while (0)
#define PROF(i) do { } while (0)
#define PROF_VAL(i, v) do { } while (0)
#define PROF_DEP(i, val) do { } while (0)
#endif

}  // namespace cf
#include "merge.cuh"
namespace cf {

struct StageParams {
    int prof;
    const float* heat;
    int B, C, H, W, Wp, K;
    int R;             // rows per chunk
    int S;             // shared-memory slots
    int slot_floats;   // floats per slot (stride)
    int lists;         // LIST_RAW | LIST_NMS
    int use_tma;
    int cs_log2;       // log2(cluster size)
    int hist_words;    // 32-bit words of this CTA's part of the histogram (both lists)
    int slot_area;     // bytes of the slot area (also hosts the merge scratch, the depth plane and the owner map)
    int dep_off;       // byte offset of the staged depth plane inside the slot area, -1: read it from global memory
    // byte offsets inside the slot area once the heat-map is done with it: [paint records K | sorted composites 2K | their pixels 2K |
    // selection scratch | gather buffers | frustum parameters | matches | depth plane]; the owner map of the paint
    // takes everything behind the records
    int off_comp, off_ind, off_msm, off_gath, off_det, off_best, off_par, off_own;
    int GB;            // detections per gather batch
    int wl_entries;    // words of the pass-B work list
    int g_log2;        // a lane's group in passes A / B = 2^g_log2 float4 items of a chunk (32 << g_log2 >= R * Wp / 4) ...
    int g_contig;      // ... consecutive ones (1) or every 32nd (0)
    int late_iw;       // ... of which this many issue the copies
    int late_decode;   // > 0: this many decode warps issue the decode job's gathers themselves and run its epilogue beside
                       // the frustum epilogue and the ROI scans (merge_tail, late_dw)
    int hot_passes;    // hot prefetch: 2 = frustum + decode channels, 1 = frustum only, 0 = none
    int gmax_off;      // streamed shares: byte offset (from the histogram) of the per-(chunk, lane) group maxima, then the flagged list
    int fl_cap;        // capacity of the flagged-group list
    uint32_t* ctr;
    unsigned long long* cand;
    MergeParams mp;
    int n_jobs;
    int do_frustum;
    const float* pc_dep;
    const float* pc_vel;  // (B,3,H,W) whose planes 1, 2 hold the velocities (== pc_dep unless they stayed in host memory)
    float max_pc_dist;
    float* pc_hm;
    long long pc_hm_stride;  // floats between the pc_hm of consecutive frames (3*H*W, or 67*H*W inside the combiner input)
    int32_t* assoc;
    PaintRec* recs;
};

// order-preserving key -> the float it encodes; key 0 ("everything") -> -inf
__device__ __forceinline__ float tau_to_float(uint32_t key) { return key ? key2f(key) : -CUDART_INF_F; }

// group maximum -> distributed histogram: one remote reduction per distinct bin per warp
__device__ __forceinline__ void hist_add_cluster(uint32_t hist_saddr, int list_words, int list, uint32_t key,
                                                 int cs_log2) {
    const uint32_t bin = key ? (key >> HIST_SHIFT) : 0xffffffffu;
    const uint32_t peers = __match_any_sync(FULL, bin);
    if (key && (threadIdx.x & 31) == __ffs(peers) - 1) {
        const uint32_t owner = bin & ((1u << cs_log2) - 1u), lb = bin >> cs_log2;
        const uint32_t word = (uint32_t)(list * list_words) + (lb >> 1);
        cl_red_add(cl_map(hist_saddr + word * 4u, owner), (uint32_t)__popc(peers) << (16u * (lb & 1u)));
    }
}
__device__ __forceinline__ uint32_t hist_read_cluster(uint32_t hist_saddr, int list_words, int list, int bin,
                                                      int cs_log2) {
    const uint32_t owner = (uint32_t)bin & ((1u << cs_log2) - 1u), lb = (uint32_t)bin >> cs_log2;
    const uint32_t word = (uint32_t)(list * list_words) + (lb >> 1);
    return (cl_ld(cl_map(hist_saddr + word * 4u, owner)) >> (16u * (lb & 1u))) & 0xffffu;
}

// tau of both lists: lower edge of the highest bin b* with count(bins >= b*) >= K, 0 ("everything") when the frame
// has fewer than K groups.  Scans downwards from the top occupied bin, ST bins per round, both lists side by side.
// hot[l]: the same for the smaller count Kh (hot[l] >= tau[l]; 0xffffffff when there is none) -- survivors at or above it
// are almost surely among the K best, and the filter pass starts the L2 prefetch of their head values right away.
__device__ __forceinline__ void tau_scan_cluster(uint32_t hist_saddr, int list_words, int cs_log2, const uint32_t top_in[2],
                                                 int K, int Kh, bool want0, bool want1, uint32_t* s_scan, uint32_t tau[2],
                                                 uint32_t hot[2]) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = SNW;
    constexpr int STRIDE = NW + 2;  // per list: NW warp totals, [NW] = tau bin found, [NW + 1] = hot bin found
    const bool want[2] = {want0, want1};
    uint32_t top[2], carry[2] = {0u, 0u};
    bool done[2];
#pragma unroll
    for (int l = 0; l < 2; ++l) {
        top[l] = want[l] ? top_in[l] : 0u;
        done[l] = !want[l];
        tau[l] = 0u;
        hot[l] = 0xffffffffu;
    }
    if (tid < 2) s_scan[tid * STRIDE + NW + 1] = 0xffffffffu;  // (ordered before the first write by the round's barrier)
    for (int round = 0; !(done[0] && done[1]); ++round) {
        uint32_t c[2], inc[2];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            const int bin = (int)top[l] - (round * ST + tid);
            c[l] = (!done[l] && bin >= 0) ? hist_read_cluster(hist_saddr, list_words, l, bin, cs_log2) : 0u;
        }
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            inc[l] = c[l];  // inclusive prefix over tid = over descending bins
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t o = __shfl_up_sync(FULL, inc[l], off);
                if (lane >= off) inc[l] += o;
            }
            if (lane == 31) s_scan[l * STRIDE + warp] = inc[l];
            if (tid == 0) s_scan[l * STRIDE + NW] = 0xffffffffu;
        }
        __syncthreads();
        uint32_t total[2];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            // prefix over the warps' totals: every warp scans them itself (lanes 0..NW-1), no extra barrier
            static_assert(NW <= 32, "one lane per warp total");
            const uint32_t wt = lane < NW ? s_scan[l * STRIDE + lane] : 0u;
            uint32_t winc = wt;
#pragma unroll
            for (int off = 1; off < NW; off <<= 1) {
                const uint32_t o = __shfl_up_sync(FULL, winc, off);
                if (lane >= off) winc += o;
            }
            const uint32_t below = carry[l] + __shfl_sync(FULL, winc - wt, warp);
            total[l] = carry[l] + __shfl_sync(FULL, winc, NW - 1);
            inc[l] += below;
            const int bin = (int)top[l] - (round * ST + tid);
            if (!done[l] && inc[l] >= (uint32_t)K && inc[l] - c[l] < (uint32_t)K)
                s_scan[l * STRIDE + NW] = (uint32_t)bin << HIST_SHIFT;  // exactly one thread
            if (!done[l] && inc[l] >= (uint32_t)Kh && inc[l] - c[l] < (uint32_t)Kh)
                s_scan[l * STRIDE + NW + 1] = (uint32_t)bin << HIST_SHIFT;  // exactly one thread, in one round
        }
        __syncthreads();
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            if (done[l]) continue;
            hot[l] = s_scan[l * STRIDE + NW + 1];
            const uint32_t found = s_scan[l * STRIDE + NW];
            if (found != 0xffffffffu) {
                tau[l] = found;
                done[l] = true;
            } else if ((int)top[l] - (round + 1) * ST < 0) {
                done[l] = true;  // histogram exhausted: fewer than K groups, tau stays 0
            } else {
                carry[l] = total[l];
            }
        }
        __syncthreads();
    }
}

// The whole CTA ranks the jobs' lists one after the other (single-job calls, and lists that need the exact block-level
// selection first).  Out of line: on the fused stage's common path this is cold code.
__device__ __noinline__ void select_sequential(const MergeParams& mp, MergeSmem& msm, int b, int nj, uint32_t n0, uint32_t n1,
                                               unsigned long long* xs_comp, int32_t* xs_ind, uint32_t HWu) {
    const int K = mp.K;
    for (int li = 0; li < nj; ++li) {
        unsigned long long* dst = xs_comp + li * K;
        int32_t* dind = xs_ind + li * K;
        merge_select<ST, true>(mp, msm, b, li, (int)threadIdx.x, 0, li ? n1 : n0, [&](int r, unsigned long long c) {
            dst[r] = c;
            dind[r] = (int32_t)(comp_gidx(c) % HWu);
        });
    }
}

// MODE (compile-time: a launch carries only the code of its own plan -- the once-per-frame path is bound by instruction
// supply): 0 = share resident in shared memory, 1 = share streamed through the slot ring once, 2 = chunks staged by the
// consuming warp (inputs the TMA cannot take)
enum { MODE_RESIDENT = 0, MODE_STREAM = 1, MODE_MANUAL = 2 };
// SPEC: 0 = any job mix (read from the parameters); 1 = the fused stage -- raw list -> frustum job, nms'd list -> decode
// job, association and paint -- with the depth plane staged in shared memory; 2 = the same, plane scanned from L2
enum { SPEC_ANY = 0, SPEC_STAGE_STAGED = 1, SPEC_STAGE_L2 = 2 };
template <int MODE, int SPEC>
__global__ void __launch_bounds__(ST, 1) fusion_stage_kernel(const __grid_constant__ StageParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t full[MAX_SLOTS];
    __shared__ uint64_t dep_bar;
    __shared__ volatile uint32_t s_issued[MAX_SLOTS];  // ring mode: bulk copies started into each slot so far
    __shared__ uint32_t s_scan[2 * (SNW + 2)];
    __shared__ uint32_t s_tops[2][MAX_CS];
    __shared__ uint32_t s_mytop[2];
    __shared__ uint32_t s_nwl;  // entries of the pass-B work list
    __shared__ uint32_t s_nhot;  // entries of the hot list (survivors whose head values are prefetched early)
    __shared__ uint32_t s_ctr[2];  // in the cluster's CTA 0: survivors appended to the frame's two lists (a shared-memory
                                   // counter every CTA reaches over DSMEM: a third of the round trip of a global atomic)
    __shared__ float s_cal[2];    // calib[0,0], calib[0,2] of the frame (fetched at frame start)
    __shared__ uint32_t s_nfl;  // pass B, resident shares: flagged groups (a group = the 1/32 of a chunk one lane took in pass A)
    __shared__ uint16_t s_fl[FL_CAP];
    __shared__ uint32_t s_cnt[2], s_base[2];  // pass B2: survivors of this CTA per list, their base in the frame's list

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rank = cl_rank(), cs = cl_size();
    const int cluster_id = (int)(blockIdx.x >> p.cs_log2), n_clusters = (int)(gridDim.x >> p.cs_log2);
    const int C = p.C, H = p.H, W = p.W, Wp = p.Wp, R = p.R, S = p.S, K = p.K;
    const int CH = C * H, W4 = Wp >> 2, g_log2 = p.g_log2;
    const bool g_contig = p.g_contig != 0;
    const size_t HW = (size_t)H * W;
    constexpr bool FUSED = SPEC != SPEC_ANY;  // (not `FULL`: that is the full-warp mask of common.cuh)
    const bool want_raw = FUSED || (p.lists & LIST_RAW) != 0, want_nms = FUSED || (p.lists & LIST_NMS) != 0;
    const bool do_frustum = FUSED || p.do_frustum != 0;
    const int n_jobs = FUSED ? 2 : p.n_jobs;
    const bool dep_staged = SPEC == SPEC_STAGE_STAGED || (SPEC == SPEC_ANY && p.dep_off >= 0);
    constexpr bool use_tma = MODE != MODE_MANUAL;

    float* slots = reinterpret_cast<float*>(smem);
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem + p.slot_area);
    uint32_t* wl = hist + p.hist_words;  // pass-B work list, p.wl_entries words
    const uint32_t hist_saddr = smem_u32(hist);
    const int list_words = p.hist_words >> 1;

    // this CTA's rows of the frame matrix, in chunks of R rows
    const int row_begin = (int)((long long)CH * rank / cs), row_end = (int)((long long)CH * (rank + 1) / cs);
    const int n = (row_end - row_begin + R - 1) / R;   // chunks
    const int LPF = n;                                 // bulk copies per frame: every chunk once
    constexpr bool resident = MODE == MODE_RESIDENT;   // the share stays in shared memory between the two looks (n <= S)
    constexpr bool stream1 = MODE == MODE_STREAM;      // streamed through the slot ring once; the filter pass reads the
                                                       // flagged groups again from global memory (L2)
    // streamed shares: group maximum per (chunk, lane), then the long flagged-group list
    float* s_gm = reinterpret_cast<float*>(smem + p.slot_area + p.gmax_off);
    uint16_t* fl_list = stream1 ? reinterpret_cast<uint16_t*>(s_gm + (size_t)n * 32) : s_fl;

    // The merge parameters (two jobs with their head tables) are indexed at run time (job of this half of the CTA, head
    // channel of this thread).  Out of the kernel's parameter space such a read costs microseconds the first time, so
    // every thread fetches one word of them now -- the latency passes during pass A -- and parks it in shared memory
    // once the heat-map has left the slot area.
    static_assert(sizeof(MergeParams) <= ST * 4 && sizeof(MergeParams) % 4 == 0, "one word per thread");
    uint32_t par_word = 0u;
    if (tid < (int)(sizeof(MergeParams) / 4)) par_word = reinterpret_cast<const uint32_t*>(&p.mp)[tid];
    if (tid < MAX_SLOTS) s_issued[tid] = 0u;
    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(&full[s], 1);
        mbar_init(&dep_bar, 1);
        mbar_fence_init();
    }
    // Frame-start state of the cluster's FIRST frame: histogram part and counters cleared, and the cluster barrier that
    // tells the peers so, all of it shared memory only -- done BEFORE the wait for the kernel ahead of us, so that in a
    // stream of back-to-back calls (programmatic dependent launch, see launch_fusion_stage) it overlaps that kernel's tail
    // instead of standing between the launch and the first bulk copy.
    auto frame_reset = [&]() {
        for (int i = tid; i < p.hist_words; i += ST) hist[i] = 0u;
        if (tid < 2 * MAX_CS) (&s_tops[0][0])[tid] = 0u;
        if (tid < 2) s_mytop[tid] = 0u;
        if (tid == 0) s_nwl = s_nfl = s_nhot = 0u;
        if (tid < 2) s_ctr[tid] = 0u;  // (cleared before this CTA arrives at the frame's first cluster barrier)
    };
    frame_reset();
    __syncthreads();
    cl_sync();  // every CTA of the cluster runs and its histogram part is clear: remote reductions may land
    // programmatic dependent launch: let the next kernel of the stream place its CTAs as ours retire, and do not touch
    // global memory before the kernel ahead of us has completed
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    int it = 0;
    for (int b = cluster_id; b < p.B; b += n_clusters, ++it) {
        const float* frame = p.heat + (size_t)b * CH * W;
        const uint32_t ctr_raw = cl_map(smem_u32(&s_ctr[0]), 0u), ctr_nms = cl_map(smem_u32(&s_ctr[1]), 0u);  // CTA 0's
        unsigned long long* cand_raw = p.cand + (size_t)b * 2 * CAND_CAP;
        unsigned long long* cand_nms = cand_raw + CAND_CAP;

        // sequence number (1-based, over the whole launch) of bulk copy L of this frame on its slot
        auto load_seq = [&](int L) -> uint32_t {
            const int s = L % S;
            const int per_frame = (LPF - 1 - s) / S + 1;
            return (uint32_t)(it * per_frame + L / S) + 1u;
        };
        auto issue_load = [&](int L) {  // one thread: bulk copy L of this frame (L >= n: pass-B re-fetch)
            const int q = L < n ? L : L - n;
            const int s = L % S;
            const int rr = row_begin + q * R, nr = min(R, row_end - rr);
            const uint32_t bytes = (uint32_t)nr * (uint32_t)W * 4u;
            mbar_expect_tx(&full[s], bytes);
            tma_bulk_g2s(slots + (size_t)s * p.slot_floats, frame + (size_t)rr * W, bytes, &full[s]);
            if (!resident) {
                __threadfence_block();
                s_issued[s] = load_seq(L);
            }
        };
        // Wait for bulk copy L.  A parity wait is only meaningful once the copy has been STARTED (a barrier that is
        // still two phases behind reports the awaited parity as complete), and in ring mode the warp that starts it
        // is whichever one frees the slot: wait for the start first, then for the bytes.
        auto wait_load = [&](int L) {
            const int s = L % S;
            const uint32_t seq = load_seq(L);
            if (!resident) {
                for (uint32_t spins = 0; s_issued[s] < seq; ++spins)
                    if (spins > (1u << 26)) __trap();
                __threadfence_block();
            }
            mbar_wait(&full[s], (seq - 1u) & 1u);
        };
        // consuming warp stages a chunk itself (inputs the TMA cannot take): padded row stride, -inf padding
        auto manual_fill = [&](float* tile, int rr, int nr) {
            const int cnt = nr * Wp;
            for (int i = lane; i < cnt; i += 32) {
                const int row = i / Wp, x = i - row * Wp;
                tile[i] = x < W ? __ldg(frame + (size_t)(rr + row) * W + x) : -CUDART_INF_F;
            }
            __syncwarp();
        };
        // a row of the frame matrix: shared memory when it is one of this CTA's resident rows, else global (L2)
        auto row_ptr = [&](int g) -> const float* {
            if (resident && g >= row_begin && g < row_end) return slots + (size_t)(g - row_begin) * Wp;  // slots are contiguous
            return frame + (size_t)g * W;
        };

        PROF(0);
        // ---- frame start: clear this CTA's histogram part and counters, start the first S bulk copies ------------
        if (it > 0) {  // (the first frame's reset ran in the prologue)
            frame_reset();
            fence_async_smem();  // the previous frame's merge scratch / owner map were generic writes to the slot area
            __syncthreads();
            // "this CTA's histogram part is clear": arrive now, wait right before the first remote reduction
            // (the release fence of the arrival then has nothing outstanding to wait for)
            cl_arrive();
        }
        if (use_tma && tid < (LPF < S ? LPF : S)) issue_load(tid);
        if (do_frustum && tid >= ST - 2)  // (the last warp: it stalls on this DRAM read before its first chunk)
            s_cal[tid - (ST - 2)] = __ldg(p.mp.job[0].fh.calib + 12 * (size_t)b + 2 * (tid - (ST - 2)));
        if (do_frustum && tid == 0) {  // the association's depth plane: into L2 while the heat-map streams
            // (NOT the velocity planes: they are read at <= K matched pixels, and prefetching them whole cost a streamed
            // frame 2 % -- 179 KB of DRAM reads beside the 896 KB it streams -- and gained the others nothing)
            const size_t total = HW * 4, per = ((total + cs - 1) / cs + 15) & ~(size_t)15;
            const size_t off = per * rank;
            if (off < total) {
                const size_t nbytes = (min(per, total - off)) & ~(size_t)15;
                const char* pd = reinterpret_cast<const char*>(p.pc_dep + (size_t)b * 3 * HW) + off;
                if (nbytes && ((reinterpret_cast<uintptr_t>(pd) & 15u) == 0))
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(pd), "r"((uint32_t)nbytes) : "memory");
            }
        }
        PROF(15);
        if (it > 0) cl_wait();
        PROF(1);

        // ---- pass A: group maxima -> distributed histogram ------------------------------------------------------
        float gmax[4] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};  // this lane's groups (resident: n <= 64)
        {
            const float inv_w4 = 1.0f / (float)W4;
            uint32_t tkraw = 0u, tknms = 0u;
            const int G = 1 << g_log2;
            for (int q = warp; q < n; q += SNW) {
                const int rr = row_begin + q * R, nr = min(R, row_end - rr);
                float* tile;
                if (use_tma) {
                    wait_load(q);
                    tile = slots + (size_t)(q % S) * p.slot_floats;
                } else {
                    tile = slots + (size_t)warp * p.slot_floats;
                    manual_fill(tile, rr, nr);
                }
                const int n_items = nr * W4;
                const float4* c4p = reinterpret_cast<const float4*>(tile);
                float best = -CUDART_INF_F;
                // a lane's group = G float4 items.  Consecutive ones (g_contig): the filter pass re-reads a flagged group
                // as one contiguous run -- whole sectors from L2, conflict-free out of shared memory -- and a blob
                // flags half as many groups; the visiting order is rotated by the lane index so that the eight lanes
                // of a quarter-warp hit eight different bank groups.  Or every 32nd item: a blob's pixels spread over
                // many groups, which makes the bound below tighter (about half the survivors) -- the choice of large K,
                // where the survivors must still fit their list (launch_fusion_stage).
                const int gbase = g_contig ? lane << g_log2 : lane;
                int bi = gbase;
#pragma unroll 4
                for (int t = 0; t < G; ++t) {
                    const int i = g_contig ? gbase + ((t + lane) & (G - 1)) : lane + 32 * t;
                    if (i < n_items) {
                        const float4 v = c4p[i];
                        const float m = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w));
                        if (m > best) {
                            best = m;
                            bi = i;
                        }
                    }
                }
                uint32_t kraw = 0u, knms = 0u;
                int r = 0, x = -1;
                if (gbase < n_items) {
                    r = (int)(((float)bi + 0.5f) * inv_w4);  // bi / W4, exact for bi < 2^21
                    int cc = bi - r * W4;
                    if (cc < 0) {
                        --r;
                        cc += W4;
                    } else if (cc >= W4) {
                        ++r;
                        cc -= W4;
                    }
                    const float4 v = c4p[bi];
                    const float cv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int j = 3; j >= 0; --j)
                        if (cv[j] == best && 4 * cc + j < W) x = 4 * cc + j;
                    if (x >= 0) kraw = f2key(best);  // (x < 0: all -inf in a padded float4)
                }
                if (want_nms) {
                    // the group maximum backs the nms'd bound only if that pixel survives the 3x3 suppression; a
                    // group whose maximum is suppressed contributes nothing (weaker bound, still a bound)
                    const bool have = x >= 0;
                    if (resident) {  // rows above / below may sit in the neighbouring chunks: wait for their copies
                        if (__any_sync(FULL, have && r == 0 && q > 0)) wait_load(q - 1);
                        if (__any_sync(FULL, have && r == nr - 1 && q < n - 1)) wait_load(q + 1);
                    }
                    if (have) {
                        const int g = rr + r;   // frame-matrix row
                        const int py = g % H;   // row inside its class plane: planes do not see each other
                        const float* rows[3];
                        rows[0] = r > 0 ? tile + (r - 1) * Wp : row_ptr(g - 1);
                        rows[1] = tile + r * Wp;
                        rows[2] = r < nr - 1 ? tile + (r + 1) * Wp : row_ptr(g + 1);
                        float nb = -CUDART_INF_F;
#pragma unroll
                        for (int dy = -1; dy <= 1; ++dy) {
                            const bool yok = py + dy >= 0 && py + dy < H;
#pragma unroll
                            for (int dx = -1; dx <= 1; ++dx) {
                                const bool ok = yok && x + dx >= 0 && x + dx < W;
                                const float t = ok ? rows[dy + 1][x + dx] : -CUDART_INF_F;
                                nb = fmaxf(nb, t);
                            }
                        }
                        knms = (nb > best) ? 0u : kraw;
                    }
                }
                if (want_raw) hist_add_cluster(hist_saddr, list_words, 0, kraw, p.cs_log2);
                if (want_nms) hist_add_cluster(hist_saddr, list_words, 1, knms, p.cs_log2);
                tkraw = max(tkraw, kraw);
                tknms = max(tknms, knms);
                if (stream1) {
                    s_gm[q * 32 + lane] = best;
                } else {
                    const int ci = q / SNW;
                    if (ci == 0) gmax[0] = best;
                    else if (ci == 1) gmax[1] = best;
                    else if (ci == 2) gmax[2] = best;
                    else if (ci == 3) gmax[3] = best;
                }
                // a chunk that will not stay resident frees its slot: this warp starts the copy that reuses it
                if (use_tma && q + S < n) {
                    __syncwarp();
                    if (lane == 0 && q + S < LPF) issue_load(q + S);
                }
            }
            tkraw = __reduce_max_sync(FULL, tkraw);
            tknms = __reduce_max_sync(FULL, tknms);
            if (lane == 0) {
                if (tkraw) atomicMax(&s_mytop[0], tkraw >> HIST_SHIFT);
                if (tknms) atomicMax(&s_mytop[1], tknms >> HIST_SHIFT);
            }
            __syncthreads();
            if (tid < 2 * (int)cs) {  // publish this CTA's top occupied bins to every CTA of the cluster
                const int l = tid / (int)cs, r = tid - l * (int)cs;
                cl_st(cl_map(smem_u32(&s_tops[l][rank]), (uint32_t)r), s_mytop[l]);
            }
        }
        PROF(2);
        cl_sync();  // histogram complete
        PROF(3);

        // ---- tau ---------------------------------------------------------------------------------------------------
        uint32_t top[2] = {0u, 0u}, tau_key[2];
        for (uint32_t r = 0; r < cs; ++r) {
            top[0] = max(top[0], s_tops[0][r]);
            top[1] = max(top[1], s_tops[1][r]);
        }
        uint32_t hot_key[2];
        tau_scan_cluster(hist_saddr, list_words, p.cs_log2, top, K, max(1, K / HOT_DIV), want_raw, want_nms, s_scan, tau_key,
                         hot_key);
        PROF(4);
        const float INF = CUDART_INF_F;
        float tau_raw = want_raw ? tau_to_float(tau_key[0]) : INF;
        float tau_nms = want_nms ? tau_to_float(tau_key[1]) : INF;
        // survivors at or above these are almost surely winners: their head values are prefetched into L2 while the
        // cluster gathers at the barrier behind the filter pass, two phases ahead of the gathers
        const float hot_raw = (want_raw && hot_key[0] != 0xffffffffu) ? tau_to_float(hot_key[0]) : INF;
        const float hot_nms = want_nms ? -INF : INF;  // (every survivor of the nms'd list)
        if (want_raw && tau_key[0] == 0u) {
            // fewer than K groups: no useful bound -> the merge selects from the frame itself (list flagged overflowed)
            if (rank == 0 && tid == 0) atomicAdd(&s_ctr[0], (uint32_t)CAND_CAP + 1u);
            tau_raw = INF;
        }
        if (want_nms && !(tau_nms > 0.0f)) {
            // fewer than K positive peaks: suppressed pixels (worth 0) belong to the top-K too -> same
            if (rank == 0 && tid == 0) atomicAdd(&s_ctr[1], (uint32_t)CAND_CAP + 1u);
            tau_nms = INF;
        }

        // ---- pass B: survivors >= tau, from shared memory ----------------------------------------------------------
        // B1 scans (one vote per 512 pixels) and parks the few candidate pixels in a CTA-wide work list; B2 tests
        // them side by side, one candidate per thread.  (Code executed once per launch is latency bound: a lane
        // working through a blob's pixels on its own would cost microseconds.)
        {
            const float tau_min = fminf(tau_raw, tau_nms);
            const float inv_w4 = 1.0f / (float)W4;
            constexpr int FU = 4;  // float4 items per lane per block
            // candidate (chunk q, pixel px of its tile) -> value, frame-flat pixel, list membership
            auto cand_eval = [&](uint32_t code, float& v, uint32_t& gi, bool& sraw, bool& snms) {
                const int q = (int)(code >> 16), px = (int)(code & 0xffffu);
                const int i = px >> 2, j = px & 3;
                int r = (int)(((float)i + 0.5f) * inv_w4);  // i / W4, exact for i < 2^21
                int cc = i - r * W4;
                if (cc < 0) {
                    --r;
                    cc += W4;
                } else if (cc >= W4) {
                    ++r;
                    cc -= W4;
                }
                const int x = 4 * cc + j;
                sraw = snms = false;
                v = 0.0f;
                gi = 0u;
                if (x >= W) return;  // padded pixel (manual staging)
                const int g = row_begin + q * R + r;
                int py = g;  // row inside its class plane (at most C subtractions instead of a division)
                while (py >= H) py -= H;
                // 3x3 neighbourhood, branch-free: out-of-plane neighbours are replaced by in-plane pixels of the same
                // window (clamped coordinates), which cannot change the maximum
                const float* row0 = row_ptr(g);
                const float* rowm = py > 0 ? row_ptr(g - 1) : row0;
                const float* rowp = py < H - 1 ? row_ptr(g + 1) : row0;
                const int xl = max(x - 1, 0), xr = min(x + 1, W - 1);
                v = row0[x];
                const float nb = fmaxf(fmaxf(fmax3(rowm[xl], rowm[x], rowm[xr]), fmaxf(row0[xl], row0[xr])),
                                       fmax3(rowp[xl], rowp[x], rowp[xr]));
                gi = (uint32_t)g * (uint32_t)W + (uint32_t)x;
                sraw = want_raw && v >= tau_raw;
                snms = want_nms && v >= tau_nms && !(nb > v);
            };
            // park this lane's candidates (bit j of mask: pixel j of float4 item_of(bit)) in the work list from `slot` on;
            // lanes that hold several park one per round, so that a blob's pixels are written side by side
            uint16_t* wl16 = reinterpret_cast<uint16_t*>(wl);
            // (the last HOT_CAP words of the work list hold the hot list)
            const uint32_t wl_cap = resident ? 2u * (uint32_t)(p.wl_entries - HOT_CAP) : (uint32_t)(p.wl_entries - HOT_CAP);
            uint32_t* s_hot = wl + (p.wl_entries - HOT_CAP);
            auto park = [&](uint32_t mask, uint32_t slot, auto&& code_of) {  // code_of(bit): (chunk << 16) | 4 * item of bit's float4
                for (;;) {
                    const bool mine = mask != 0u;
                    const uint32_t bal = __ballot_sync(FULL, mine);
                    if (!bal) break;
                    const uint32_t my = slot + (uint32_t)__popc(bal & lanemask_lt());
                    slot += (uint32_t)__popc(bal);
                    if (!mine) continue;
                    const int bit = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const uint32_t code = (uint32_t)code_of(bit) + (uint32_t)(bit & 3);
                    if (my < wl_cap) {
                        if (resident) wl16[my] = (uint16_t)code;  // q == 0: the pixel index alone, and it fits 16 bits
                        else wl[my] = code;
                    } else {  // work list full (dense survivors): test and append right here, lane by lane
                        float v;
                        uint32_t gi;
                        bool sraw, snms;
                        cand_eval(code, v, gi, sraw, snms);
                        if (sraw) {
                            const uint32_t gs = cl_atom_add(ctr_raw, 1u);
                            if (gs < (uint32_t)CAND_CAP) cand_raw[gs] = make_comp(f2key(v), gi);
                        }
                        if (snms) {
                            const uint32_t gs = cl_atom_add(ctr_nms, 1u);
                            if (gs < (uint32_t)CAND_CAP) cand_nms[gs] = make_comp(f2key(v), gi);
                        }
                    }
                }
            };
            // Fast path: pass A left the maximum of every group (a group = the 1/32 of a chunk one lane took) -- in
            // registers when the share is resident, in shared memory when it streamed through the ring -- so the groups
            // that hold a candidate are known without touching the share again.  They go to a CTA-wide list, and half-warps
            // rescan just those, whichever warp found them: out of shared memory (resident), or out of global memory, where
            // the chunks have just passed through L2 (streamed: a few hundred groups of 16 sectors instead of a second
            // pass over the whole share).
            bool scanned = false;
            if (resident || stream1) {
                // (resident: every chunk's copy was awaited by the warp that took the chunk in pass A; the CTA barrier
                // below orders those observations before the reads of all the other warps)
                PROF(30);
                const int n_ci = (n + SNW - 1) / SNW;
                for (int ci = 0; ci < n_ci; ++ci) {
                    const int q = warp + ci * SNW;
                    float gm = -INF;
                    if (stream1) {
                        if (q < n) gm = s_gm[q * 32 + lane];
                    } else {
                        gm = ci == 0 ? gmax[0] : (ci == 1 ? gmax[1] : (ci == 2 ? gmax[2] : gmax[3]));
                    }
                    const bool flag = q < n && gm >= tau_min;
                    const uint32_t bal = __ballot_sync(FULL, flag);
                    if (bal) {
                        uint32_t base = 0u;
                        if (lane == 0) base = atomicAdd(&s_nfl, (uint32_t)__popc(bal));
                        base = __shfl_sync(FULL, base, 0) + (uint32_t)__popc(bal & lanemask_lt());
                        if (flag && base < (uint32_t)p.fl_cap) fl_list[base] = (uint16_t)((q << 5) | lane);
                    }
                }
                PROF(31);
                __syncthreads();
                const int nfl = (int)s_nfl;
                PROF(29);
                scanned = nfl <= p.fl_cap;  // else (dense frame): the block scan below
                if (scanned) {
                    // every half-warp takes up to 8 groups per round; a lane collects the candidate bits of all of them in
                    // one register first (loads back to back), so that the reservation and the parking happen once
                    const int hw = tid >> 4, hl = tid & 15;
                    const int chunk_items = R * W4, t_max = 1 << g_log2;
                    const float4* c4p = resident ? reinterpret_cast<const float4*>(slots)
                                                 : reinterpret_cast<const float4*>(frame + (size_t)row_begin * W);
                    for (int f0 = 0; f0 < nfl; f0 += 8 * (ST / 16)) {
                        for (int t0 = 0; t0 < t_max; t0 += 16) {
                            uint32_t mask = 0u;
#pragma unroll
                            for (int u = 0; u < 8; ++u) {
                                const int f = f0 + u * (ST / 16) + hw;
                                if (f < nfl) {
                                    const int code = (int)fl_list[f];
                                    const int q = code >> 5, i = g_contig ? ((code & 31) << g_log2) + t0 + hl : (code & 31) + 32 * (t0 + hl);
                                    if (t0 + hl < t_max && i < min(R, row_end - (row_begin + q * R)) * W4) {
                                        const float4 v = c4p[q * chunk_items + i];  // (chunks are contiguous in both memories)
                                        const uint32_t m4 = (v.x >= tau_min ? 1u : 0u) | (v.y >= tau_min ? 2u : 0u) |
                                                            (v.z >= tau_min ? 4u : 0u) | (v.w >= tau_min ? 8u : 0u);
                                        mask |= m4 << (4 * u);
                                    }
                                }
                            }
                            const uint32_t total = __reduce_add_sync(FULL, (uint32_t)__popc(mask));
                            if (!total) continue;
                            uint32_t slot = 0u;
                            if (lane == 0) slot = atomicAdd(&s_nwl, total);
                            slot = __shfl_sync(FULL, slot, 0);
                            park(mask, slot, [&](int bit) {
                                const int code = (int)fl_list[f0 + (bit >> 2) * (ST / 16) + hw];
                                const int q = code >> 5, i = g_contig ? ((code & 31) << g_log2) + t0 + hl : (code & 31) + 32 * (t0 + hl);
                                // resident: one contiguous tile, the pixel index alone (16 bits); streamed: (chunk, pixel)
                                return resident ? (q * chunk_items + i) * 4 : ((q << 16) | (i * 4));
                            });
                        }
                    }
                }
            }
            // Block scan (dense frames whose flagged groups overflow the list, and inputs the TMA cannot take).  Visits:
            // resident -> ONE visit, the whole share as one contiguous tile, 128-item blocks interleaved over the warps;
            // streamed / manual -> one visit per chunk per warp, straight from global memory / staged by the warp.
            const int n_vis = scanned ? 0 : (resident ? 1 : n);
            for (int vb = resident ? 0 : warp; vb < n_vis; vb += SNW) {
                int q = 0, nr = row_end - row_begin;
                const float* tile = slots;
                if (resident) {
                    for (int c = lane; c < n; c += 32) mbar_wait(&full[c], (uint32_t)it & 1u);  // every copy, seen by this warp
                    __syncwarp();
                } else {
                    q = vb;
                    const int rr = row_begin + q * R;
                    nr = min(R, row_end - rr);
                    if (use_tma) {
                        tile = frame + (size_t)rr * W;  // (streamed share, flagged-group list overflowed: the chunk from L2)
                    } else {
                        float* mt = slots + (size_t)warp * p.slot_floats;
                        manual_fill(mt, rr, nr);
                        tile = mt;
                    }
                }
                const int n_items = nr * W4;
                const float4* c4p = reinterpret_cast<const float4*>(tile);
                const int i_step = resident ? SNW * 32 * FU : 32 * FU;
                for (int item0 = resident ? warp * (32 * FU) : 0; item0 < n_items; item0 += i_step) {
                    float4 q4[FU];
                    float m = -INF;
#pragma unroll
                    for (int u = 0; u < FU; ++u) {
                        const int i = item0 + u * 32 + lane;
                        q4[u] = i < n_items ? c4p[i] : make_float4(-INF, -INF, -INF, -INF);
                        m = fmaxf(m, fmaxf(fmaxf(q4[u].x, q4[u].y), fmaxf(q4[u].z, q4[u].w)));
                    }
                    // almost every block stops here: none of its 512 pixels reaches either threshold
                    if (!__any_sync(FULL, m >= tau_min)) continue;
                    uint32_t mask = 0u;  // bit u*4+j: pixel j of this lane's u-th float4 is a candidate
#pragma unroll
                    for (int u = 0; u < FU; ++u) {
                        if (q4[u].x >= tau_min) mask |= 1u << (4 * u);
                        if (q4[u].y >= tau_min) mask |= 2u << (4 * u);
                        if (q4[u].z >= tau_min) mask |= 4u << (4 * u);
                        if (q4[u].w >= tau_min) mask |= 8u << (4 * u);
                    }
                    // park the candidates: one reservation per block, then one ballot per (float4, pixel) position,
                    // so that a blob's pixels, which cluster in a few lanes, are written side by side
                    const uint32_t total = __reduce_add_sync(FULL, (uint32_t)__popc(mask));
                    uint32_t slot = 0u;
                    if (lane == 0) slot = atomicAdd(&s_nwl, total);
                    slot = __shfl_sync(FULL, slot, 0);
                    park(mask, slot, [&](int bit) { return (q << 16) | ((item0 + (bit >> 2) * 32 + lane) * 4); });
                }
            }
            PROF(12);
            __syncthreads();
            PROF(13);
            // B2: candidates side by side, one per thread and trip.  Survivors are counted per CTA (shared-memory atomics), the frame's lists
            // see ONE reservation per (CTA, list): same-address global atomics serialise at a few per microsecond.
            const uint32_t n_wl = min(s_nwl, wl_cap);
            PROF_VAL(27, s_nfl);
            PROF_VAL(28, s_nwl);
            // (up to BU candidates per thread in one trip: a CTA that parked more than 512 pays the counting, the two
            // barriers and the round trip of the reservation once -- such a CTA sets the pace of its whole cluster)
            constexpr int BU = 3;
            for (uint32_t c0 = 0; c0 < n_wl; c0 += ST * BU) {
                unsigned long long comp[BU];
                uint32_t off[BU], flags = 0u, tot0 = 0u, tot1 = 0u;  // off: position among the warp's survivors (raw | nms << 16)
                if (tid < 2) s_cnt[tid] = 0u;
                __syncthreads();
#pragma unroll
                for (int u = 0; u < BU; ++u) {
                    const uint32_t c = c0 + (uint32_t)(u * ST + tid);
                    float v = 0.0f;
                    uint32_t gi = 0u;
                    bool sraw = false, snms = false;
                    if (c0 + (uint32_t)(u * ST) < n_wl) {  // (uniform: trips beyond the list cost nothing)
                        if (c < n_wl) cand_eval(resident ? (uint32_t)wl16[c] : wl[c], v, gi, sraw, snms);
                        if (((sraw && v >= hot_raw) || (snms && v >= hot_nms)) && gi < 0x40000000u) {  // (rare: about K / HOT_DIV per list and frame)
                            const uint32_t code = gi | ((sraw && v >= hot_raw) ? 0x40000000u : 0u) |
                                                  ((snms && v >= hot_nms) ? 0x80000000u : 0u);
                            const uint32_t hs = atomicAdd(&s_nhot, 1u);
                            if (hs < (uint32_t)HOT_CAP) s_hot[hs] = code;
                        }
                        const uint32_t bal0 = __ballot_sync(FULL, sraw), bal1 = __ballot_sync(FULL, snms);
                        off[u] = (tot0 + (uint32_t)__popc(bal0 & lanemask_lt())) |
                                 ((tot1 + (uint32_t)__popc(bal1 & lanemask_lt())) << 16);
                        tot0 += (uint32_t)__popc(bal0);
                        tot1 += (uint32_t)__popc(bal1);
                    } else {
                        off[u] = 0u;
                    }
                    comp[u] = make_comp(f2key(v), gi);
                    flags |= (sraw ? 1u : 0u) << (2 * u) | (snms ? 2u : 0u) << (2 * u);
                }
                uint32_t lb0 = 0u, lb1 = 0u;
                if (tot0 && lane == 0) lb0 = atomicAdd(&s_cnt[0], tot0);
                if (tot1 && lane == 0) lb1 = atomicAdd(&s_cnt[1], tot1);
                lb0 = __shfl_sync(FULL, lb0, 0);
                lb1 = __shfl_sync(FULL, lb1, 0);
                __syncthreads();
                if (tid < 2 && s_cnt[tid]) s_base[tid] = cl_atom_add(tid ? ctr_nms : ctr_raw, s_cnt[tid]);
                __syncthreads();
                lb0 += s_base[0];
                lb1 += s_base[1];
#pragma unroll
                for (int u = 0; u < BU; ++u) {
                    const uint32_t s0 = lb0 + (off[u] & 0xffffu), s1 = lb1 + (off[u] >> 16);
                    if ((flags >> (2 * u) & 1u) && s0 < (uint32_t)CAND_CAP) cand_raw[s0] = comp[u];
                    if ((flags >> (2 * u) & 2u) && s1 < (uint32_t)CAND_CAP) cand_nms[s1] = comp[u];
                }
            }
            PROF(14);
            fence_async_smem();
            __syncthreads();
            if (tid == 0) s_nfl = 0u;  // becomes the association's work counter
            // the slot area is free from here on: the merge parameters move in (visible after the cluster barrier)
            if (tid < (int)(sizeof(MergeParams) / 4)) reinterpret_cast<uint32_t*>(smem + p.off_par)[tid] = par_word;
            // Stage the radar depth plane for the association.
            if (do_frustum && dep_staged && tid == 0) {
                const uint32_t bytes = (uint32_t)(HW * 4);
                mbar_expect_tx(&dep_bar, bytes);
                tma_bulk_g2s(smem + p.dep_off, p.pc_dep + (size_t)b * 3 * HW, bytes, &dep_bar);
            }
        }
        PROF(5);
        // survivor lists complete -- and while the barrier gathers the cluster, the head values of this CTA's hot
        // survivors start their way from DRAM into L2: the gathers that want them run two phases from now (140 k
        // scattered sectors per batch, DRAM random-access bound when every cluster asks for them at the same moment)
        __syncthreads();  // (the merge parameters just parked in shared memory, by other threads)
        cl_arrive();
        {
            const MergeParams& mpp = *reinterpret_cast<const MergeParams*>(smem + p.off_par);
            const uint32_t* s_hot = wl + (p.wl_entries - HOT_CAP);
            const int n_hot = (int)min(s_nhot, (uint32_t)HOT_CAP);
            const uint32_t HWu = (uint32_t)HW;
            // the frustum job's channels first: the association waits for them, the decode outputs are due at the frame's end
#pragma unroll
            for (int pass = 0; pass < p.hot_passes; ++pass) {
                const int epi = pass ? EPI_DECODE : EPI_FRUSTUM;
                const int NS = pass ? (int)DS_N : (int)FS_N;
                for (int li = FUSED ? pass : 0; li < (FUSED ? pass + 1 : n_jobs); ++li) {
                    const MergeJob& jb = mpp.job[li];
                    if (!FUSED && jb.epilogue != epi) continue;
                    const uint32_t flag = (FUSED ? li != 0 : mpp.list[li] != 0) ? 0x80000000u : 0x40000000u;
                    for (int t = tid; t < n_hot * NS; t += ST) {
                        const int e = t / NS, slot = t - e * NS;
                        const uint32_t code = s_hot[e];
                        const float* base = jb.gt.base[slot];
                        if ((code & flag) && base) {
                            const uint32_t ind = (code & 0x3fffffffu) % HWu;
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(base + ((size_t)b * jb.gt.nch[slot] + jb.gt.ch[slot]) * HWu + ind));
                        }
                    }
                }
            }
        }
        cl_wait();
        PROF(6);

        // Several frames per cluster, shares resident: this CTA's share of the NEXT frame's heat-map starts its way from HBM
        // into L2 now, under the once-per-frame phases that leave the memory system idle (one bulk prefetch; a round's
        // shares together are a fraction of the L2).  Pass A of the next frame then streams from L2.
        if (resident && b + n_clusters < p.B && tid == 64) {
            const char* nf = reinterpret_cast<const char*>(p.heat + (size_t)(b + n_clusters) * CH * W + (size_t)row_begin * W);
            const uint32_t nbytes = (uint32_t)((size_t)(row_end - row_begin) * W * 4) & ~15u;
            if (nbytes && ((reinterpret_cast<uintptr_t>(nf) & 15u) == 0))
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nf), "r"(nbytes) : "memory");
        }

        // ---- select: every CTA ranks the jobs' survivor lists itself; the sorted top-K stays in its own shared memory ----
        // (ranking twice the entries costs a CTA less than the cluster barrier and the remote deliveries a split
        // ranking needs; the two jobs run side by side on the two halves of the CTA)
        unsigned long long* xs_comp = reinterpret_cast<unsigned long long*>(smem + p.off_comp);  // [job][K]
        int32_t* xs_ind = reinterpret_cast<int32_t*>(smem + p.off_ind);  // [job][K]: pixel of the ranked composite
        const MergeParams& mp = *reinterpret_cast<const MergeParams*>(smem + p.off_par);
        {
            MergeSmem* msm = reinterpret_cast<MergeSmem*>(smem + p.off_msm);  // one scratch per job
            const int nj = n_jobs;
            const uint32_t n0 = cl_ld(mp.list[0] ? ctr_nms : ctr_raw);  // survivors of the jobs' lists
            const uint32_t n1 = nj == 2 ? cl_ld(mp.list[1] ? ctr_nms : ctr_raw) : 0u;
            if (nj == 2 && n0 <= (uint32_t)MERGE_WIN && n1 <= (uint32_t)MERGE_WIN) {
                const int li = tid < ST / 2 ? 0 : 1;  // job of this half of the CTA (named barrier 1 + li)
                unsigned long long* dst = xs_comp + li * K;
                int32_t* dind = xs_ind + li * K;
                const uint32_t HWu = (uint32_t)HW;
                merge_select<ST / 2, false>(mp, msm[li], b, li, tid - li * (ST / 2), 1 + li, li ? n1 : n0,
                                            [&](int r, unsigned long long c) {
                                                dst[r] = c;
                                                dind[r] = (int32_t)(comp_gidx(c) % HWu);
                                            });
            } else {  // one job, or a list that needs the exact block-level selection first: the whole CTA, job by job
                select_sequential(mp, msm[0], b, nj, n0, n1, xs_comp, xs_ind, (uint32_t)HW);  // (out of line: cold)
            }
            __syncthreads();
        }
        PROF(7);
        // (no tail reads the peers' shared memory; without the frustum tail there is no later cluster barrier either, and
        // CTA 0, which holds the lists' counters, must not run ahead into its next frame or leave)
        if (!do_frustum) cl_sync();
        PROF(8);

        // ---- tail: detection k belongs to CTA k % CS: gathers, frustum geometry / decode outputs ----------------------
        const int n_local = (K - (int)rank + (int)cs - 1) / (int)cs;
        DetParam* sdet = reinterpret_cast<DetParam*>(smem + p.off_det);
        // (one batch and a frustum job: the decode epilogue waits until the association's finishing threads leave the upper
        // half of the CTA idle -- its 29 channels per detection have landed long before)
        const bool defer_decode = do_frustum && n_jobs == 2 && n_local <= p.GB && n_local <= ST / 2;
        merge_tail<ST, FUSED>(mp, n_jobs, b, (int)rank, (int)cs, n_local, xs_comp, xs_ind, reinterpret_cast<float*>(smem + p.off_gath),
                       p.GB, sdet, do_frustum ? reinterpret_cast<short4*>(smem + p.off_best) : nullptr,
                       do_frustum ? s_cal : nullptr, defer_decode, (FUSED && defer_decode) ? p.late_decode : 0);
        PROF(24);
        if (!do_frustum) continue;

        // ---- association: one warp per detection, ROI scan of the staged depth plane ---------------------------------
        PaintRec* srecs = reinterpret_cast<PaintRec*>(smem);  // all K records, in detection order, in every CTA
        {
            unsigned long long* sbest = reinterpret_cast<unsigned long long*>(smem + p.off_best);
            const float* dep = p.pc_dep + (size_t)b * 3 * HW;
            const float* dep0 = dep;
            if (dep_staged) {
                mbar_wait(&dep_bar, (uint32_t)it & 1u);
                dep0 = reinterpret_cast<const float*>(smem + p.dep_off);
            }
            PROF(32);
            // (the ROI bounds of detection kl wait in sbest[kl], left there by the epilogue's idle warps: this loop runs
            // once per detection on one warp and is priced by its instruction count)
            // (deferred decode epilogue: the last DW warps run it now, beside the scans of the others -- its copies landed
            // before the tail's closing barrier)
            const bool late_d = FUSED && defer_decode && p.late_decode;  // (FUSED: job 1 is the decode job)
            const int DW = late_d ? p.late_decode : 3;
            const int nsw = defer_decode ? SNW - DW : SNW;  // scanning warps
            // The two velocity values of a match are LOADED by the warp that found it (lanes 0, 1) and parked in the
            // detection's parameters one scan later (the box corners there are dead by now): the finishing threads below
            // then wait for no memory at all.
            float vpend = 0.0f;
            int klpend = -1;
            auto park_vel = [&]() {
                if (klpend >= 0 && lane < 2) reinterpret_cast<float*>(&sdet[klpend])[lane] = vpend;
            };
            auto load_vel = [&](int kl, unsigned long long best) {
                park_vel();
                klpend = kl;
                vpend = 0.0f;
                if (lane < 2 && best != ~0ull)
                    vpend = __ldg(p.pc_vel + (size_t)b * 3 * HW + (size_t)(lane + 1) * HW + (uint32_t)best);
            };
            if (warp >= nsw) {
                if (late_d) {  // the decode job's copies: issued, awaited and consumed by these warps alone
                    const int iw = min(DW, max(1, p.late_iw));  // (issuing warps: few, so that the SM's queue of scattered
                                                                 // requests leaves room for everybody else's accesses)
                    if (warp - nsw < iw)
                        gather_issue<DS_N>(mp.job[1].gt, b, HW, xs_ind + K, (int)rank, (int)cs, 0, n_local,
                                           reinterpret_cast<float*>(smem + p.off_gath) + (size_t)p.GB * GATH_STRIDE,
                                           tid - (ST - DW * 32), iw * 32);
                    cp_async_commit();
                    cp_async_wait<0>();
                    asm volatile("bar.sync 3, %0;" ::"r"(DW * 32) : "memory");
                }
                CF_PROF_SLOT(ST - DW * 32, 35, 0);
                merge_tail_decode<FUSED>(mp, n_jobs, b, (int)rank, (int)cs, 0, n_local, xs_comp,
                                  reinterpret_cast<const float*>(smem + p.off_gath), p.GB, tid - (ST - DW * 32), DW * 32);
                CF_PROF_SLOT(ST - DW * 32, 36, 0);
            } else if (dep_staged) {
                for (int kl = warp;; kl = nsw) {
                    if (kl >= nsw) {  // first detection by warp index, the rest from the shared counter (big ROIs wait less)
                        if (lane == 0) kl = nsw + (int)atomicAdd(&s_nfl, 1u);
                        kl = __shfl_sync(FULL, kl, 0);
                    }
                    if (kl >= n_local) break;
                    const short4 roi = *reinterpret_cast<const short4*>(&sbest[kl]);
                    const float2 dt = *reinterpret_cast<const float2*>(&sdet[kl].depth);
                    __syncwarp();
                    const unsigned long long best =
                        warp_min_key(roi_scan_box<8, true>(dep0, W, roi.x, roi.y, roi.z, roi.w, dt.x, dt.y, lane, 32));
                    if (lane == 0) sbest[kl] = best;
                    load_vel(kl, best);
                    if (kl == 0) PROF_DEP(33, (uint32_t)best);
                }
                park_vel();
            } else {
                // A plane too big to stage is read from L2: float4 row loads (fewer, wider requests; the staged plane
                // is bound by its per-pixel tests, and the narrow ROIs of the reference sizes fill a warp better pixel
                // by pixel), and the warps take detections from a shared counter as they become free (at K = 500 the
                // ROI sizes of a CTA's 60-odd detections differ widely).
                const bool quads = (W & 3) == 0 && (reinterpret_cast<uintptr_t>(dep0) & 15u) == 0;
                for (;;) {
                    int kl = 0;
                    if (lane == 0) kl = (int)atomicAdd(&s_nfl, 1u);  // (pass B's counter, reset when it was done with it)
                    kl = __shfl_sync(FULL, kl, 0);
                    if (kl >= n_local) break;
                    const short4 roi = *reinterpret_cast<const short4*>(&sbest[kl]);
                    const float2 dt = *reinterpret_cast<const float2*>(&sdet[kl].depth);
                    __syncwarp();
                    const unsigned long long best = warp_min_key(
                        quads ? roi_scan_box4<8>(dep0, W, roi.x, roi.y, roi.z, roi.w, dt.x, dt.y, lane)
                              : roi_scan_box<24, true>(dep0, W, roi.x, roi.y, roi.z, roi.w, dt.x, dt.y, lane, 32));
                    if (lane == 0) sbest[kl] = best;
                    load_vel(kl, best);
                }
                park_vel();
            }
            PROF(25);
            __syncthreads();
            PROF(26);
            // paint records: the velocity reads of all matches go out together; every CTA of the cluster gets a copy
            // of every record in its shared memory (st.shared::cluster), in detection order
            for (int kl = tid; kl < n_local; kl += ST) {
                const int k = (int)rank + kl * (int)cs;
                const size_t o = (size_t)b * K + k;
                const PaintRec r = assoc_record(sbest[kl], sdet[kl].x1, sdet[kl].y1, sdet[kl], p.max_pc_dist, p.assoc ? p.assoc + o : nullptr);
                const uint32_t* rw = reinterpret_cast<const uint32_t*>(&r);
                PROF_DEP(34, rw[0] ^ rw[1] ^ rw[2] ^ rw[3] ^ rw[4]);
                const uint32_t dst = smem_u32(srecs + k);
                for (uint32_t c = 0; c < cs; ++c) {
                    const uint32_t ra = cl_map(dst, c);
#pragma unroll
                    for (int w = 0; w < 8; ++w) cl_st(ra + 4u * w, rw[w]);
                }
            }
        }
        PROF(9);
        // ---- paint: this CTA's band of rows of pc_hm ----------------------------------------------------------------
        {
            const int y0 = (int)((long long)H * rank / cs), y1 = (int)((long long)H * (rank + 1) / cs);
            const size_t own_off = (size_t)p.off_own;
            const int max_rows = max(1, (int)(((size_t)p.slot_area - own_off) / ((size_t)W * sizeof(int))));
            int* sown = reinterpret_cast<int*>(smem + own_off);
            // paint records complete in every CTA -- while the barrier gathers the cluster, the owner map of the band's
            // first piece is cleared (this CTA is done with the scratch it overlays: the barrier below starts with a
            // CTA-wide one)
            __syncthreads();
            cl_arrive();
            const int npx0 = min(max_rows, y1 - y0) * W;
            for (int i = tid; i < npx0; i += ST) sown[i] = -1;
            cl_wait();
            PROF(10);
            paint_band<ST, true, true>(srecs, K, H, W, y0, y1, p.pc_hm + (size_t)b * (size_t)p.pc_hm_stride, nullptr, sown,
                                       max_rows);
        }
        PROF(11);
    }
}

// -----------------------------------------------------------------------------------------------------
// host side: plan + launch
// -----------------------------------------------------------------------------------------------------
// Development switches, read from the environment ONCE per process (never on the launch path).
struct StageEnv {
    bool verbose, no_pdl;
    int prof, force_cs;
    long long chunk_bytes;
    int late_min, late_dw, late_iw;
    StageEnv() {
        verbose = getenv("CFD3D_VERBOSE") != nullptr;
        no_pdl = getenv("CFD3D_NO_PDL") != nullptr;
        prof = getenv("CFD3D_PROF") ? atoi(getenv("CFD3D_PROF")) : 0;
        force_cs = getenv("CFD3D_FORCE_CS") ? atoi(getenv("CFD3D_FORCE_CS")) : 0;  // plan exploration (tools/plan_sweep.sh)
        chunk_bytes = 8192;
        late_min = getenv("CFD3D_LATE_MIN") ? atoi(getenv("CFD3D_LATE_MIN")) : 16;  // detections per CTA (development sweep: 25 per CTA, cfg2, 33.9 -> 32.9 us)
        // (measured on 256 single-CTA frames: 4 decode warps of which 1 issues -> 151.7 us; 3 / 3 -> 153.6; 5 / 5 -> 158.9;
        // without the late decode 162.4)
        late_dw = getenv("CFD3D_LATE_DW") ? atoi(getenv("CFD3D_LATE_DW")) : 4;
        late_iw = getenv("CFD3D_LATE_IW") ? atoi(getenv("CFD3D_LATE_IW")) : 1;
        if (const char* e = getenv("CFD3D_CHUNK_BYTES")) {
            if (atoll(e) >= 256) chunk_bytes = atoll(e);
        }
    }
};
static const StageEnv& stage_env() {
    static const StageEnv e;
    return e;
}

static size_t align_up_sz(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct StagePlan {
    int cs, cs_log2, R, S, slot_floats, hist_words, slot_area, dep_off, n_clusters;
    int off_comp, off_ind, off_msm, off_gath, off_det, off_best, off_par, off_own, GB, wl_entries, gmax_off, fl_cap;
    bool resident;
    int mode;  // MODE_*
    size_t smem;
};

// geometry of one cluster size; false when it cannot run
static bool plan_for_cs(int cs, int C, int H, int W, int K, bool use_tma, bool stage_dep, StagePlan* pl) {
    const long long CH = (long long)C * H;
    if (cs > CH || HIST_BINS % (2 * cs) != 0) return false;
    const int Wp = (W + 3) & ~3;
    int cs_log2 = 0;
    while ((1 << cs_log2) < cs) ++cs_log2;
    const long long share = (CH + cs - 1) / cs;  // rows of the largest share
    const long long chunk_bytes = stage_env().chunk_bytes;
    long long R = chunk_bytes / ((long long)Wp * 4);
    if (R < 1) R = 1;
    if (R > share) R = share;
    // 16-bit histogram counters: a frame must not have more than 65535 lane groups
    while (((share + R - 1) / R) * 32 * cs > 65535) ++R;
    const int hist_words = 2 * (HIST_BINS / cs) / 2;
    size_t fixed = (size_t)hist_words * 4 + (size_t)WL_CAP * sizeof(uint32_t);
    // A share that fits the shared memory as a whole but not as a whole number of default-sized chunks (full-resolution
    // maps on clusters of 16: 141 rows of 1600 bytes) stays resident with smaller chunks: a resident share is read from
    // HBM once and filtered out of shared memory, a streamed one is fetched twice.
    if (use_tma && (size_t)share * Wp * 4 + fixed <= SMEM_DYN_CAP) {
        for (long long r = R; r >= 1; --r) {
            const long long nn = (share + r - 1) / r;
            if (nn > MAX_SLOTS || nn * 32 * cs > 65535) break;
            if ((size_t)nn * align_up_sz((size_t)r * Wp * 4, 16) + fixed <= SMEM_DYN_CAP) {
                R = r;
                break;
            }
        }
    }
    {  // a streamed share keeps one group maximum per (chunk, lane) in shared memory, and the flagged-group list packs
       // the chunk index into 11 bits: bound the number of chunks (bigger chunks: a weaker bound, still a bound)
        const bool res = use_tma && (share + R - 1) / R <= MAX_SLOTS &&
                         (size_t)((share + R - 1) / R) * align_up_sz((size_t)R * Wp * 4, 16) + fixed <= SMEM_DYN_CAP;
        if (use_tma && !res)
            while ((share + R - 1) / R > 256) ++R;
    }
    const size_t slot_bytes = align_up_sz((size_t)R * Wp * 4, 16);
    if (fixed + slot_bytes > SMEM_DYN_CAP) return false;
    if (R * Wp > 65535) return false;  // (work-list codes keep the pixel of a chunk in 16 bits)
    const long long n = (share + R - 1) / R;
    long long S = (long long)((SMEM_DYN_CAP - fixed) / slot_bytes);
    if (S > MAX_SLOTS) S = MAX_SLOTS;
    int wl_entries = WL_CAP;
    int gmax_off = 0, fl_cap = FL_CAP;
    if (S < n || !use_tma) {
        // the share does not stay resident: its CTA sees more candidates (bigger share) and a ring of a few slots less
        // streams just as well -> room for a long work list, and (bulk copies) for the group maxima of every chunk plus a
        // long flagged-group list: the filter pass then re-reads only the flagged groups, from L2, instead of streaming
        // the share a second time
        size_t extra = 0;
        if (use_tma) {
            gmax_off = (int)((size_t)hist_words * 4 + (size_t)WL_CAP_RING * sizeof(uint32_t));
            fl_cap = FL_CAP_STREAM;
            extra = (size_t)n * 32 * sizeof(float) + (size_t)FL_CAP_STREAM * sizeof(uint16_t);
        }
        const size_t fixed_long = (size_t)hist_words * 4 + (size_t)WL_CAP_RING * sizeof(uint32_t) + extra;
        const long long S_long = fixed_long + slot_bytes <= SMEM_DYN_CAP ? (long long)((SMEM_DYN_CAP - fixed_long) / slot_bytes) : 0;
        if (S_long >= (use_tma ? 4 : SNW)) {
            wl_entries = WL_CAP_RING;
            fixed = fixed_long;
            S = S_long < MAX_SLOTS ? S_long : MAX_SLOTS;
        } else if (use_tma) {
            return false;  // (rows too wide for four slots beside the lists: another cluster size, or no plan at all)
        }
    }
    if (use_tma) {
        if (S > n) S = n;
    } else {
        S = SNW;  // every warp stages its own chunk
        if ((size_t)S * slot_bytes + fixed > SMEM_DYN_CAP) return false;
    }
    if (S < 1) return false;
    // layout of the slot area once the heat-map is done with it (StageParams).  The gather buffers take GB detections at
    // a time: as many as the CTA owns when they fit, else the tail runs in batches
    const int n_local_max = (K + cs - 1) / cs;
    int GB = n_local_max < 256 ? n_local_max : 256;
    size_t off_comp, off_ind, off_msm, off_gath, off_det, off_best, off_par, tail_end;
    for (;;) {
        off_comp = align_up_sz((size_t)K * sizeof(PaintRec), 128);
        off_ind = align_up_sz(off_comp + 2 * (size_t)K * 8, 128);
        off_msm = align_up_sz(off_ind + 2 * (size_t)K * 4, 128);
        off_gath = align_up_sz(off_msm + 2 * sizeof(MergeSmem), 128);
        off_det = align_up_sz(off_gath + 2 * (size_t)GB * GATH_STRIDE * 4, 128);
        off_best = align_up_sz(off_det + (size_t)n_local_max * sizeof(DetParam), 128);
        off_par = align_up_sz(off_best + (size_t)n_local_max * 8, 128);
        tail_end = align_up_sz(off_par + sizeof(MergeParams), 128);
        if (tail_end + fixed <= SMEM_DYN_CAP || GB <= 16) break;
        GB = (GB + 1) / 2;
    }
    const size_t off_own = align_up_sz((size_t)K * sizeof(PaintRec), 16);
    size_t area = (size_t)S * slot_bytes;
    if (area < tail_end) area = tail_end;
    if (area < off_own + (size_t)W * sizeof(int)) area = off_own + (size_t)W * sizeof(int);  // >= 1 row of owner map
    int dep_off = -1;
    if (stage_dep) {
        const size_t dep_bytes = (size_t)H * W * 4;
        if (dep_bytes < (1u << 20) && tail_end + dep_bytes + fixed <= SMEM_DYN_CAP) {
            dep_off = (int)tail_end;
            if (area < tail_end + dep_bytes) area = tail_end + dep_bytes;
        }
    }
    area = align_up_sz(area, 128);
    if (area + fixed > SMEM_DYN_CAP) return false;
    pl->cs = cs;
    pl->cs_log2 = cs_log2;
    pl->R = (int)R;
    pl->S = (int)S;
    pl->slot_floats = (int)(slot_bytes / 4);
    pl->hist_words = hist_words;
    pl->slot_area = (int)area;
    pl->dep_off = dep_off;
    pl->off_comp = (int)off_comp;
    pl->off_ind = (int)off_ind;
    pl->off_msm = (int)off_msm;
    pl->off_gath = (int)off_gath;
    pl->off_det = (int)off_det;
    pl->off_best = (int)off_best;
    pl->off_par = (int)off_par;
    pl->off_own = (int)off_own;
    pl->GB = GB;
    pl->wl_entries = wl_entries;
    pl->gmax_off = gmax_off;
    pl->fl_cap = fl_cap;
    pl->resident = use_tma && S >= n;
    pl->mode = !use_tma ? MODE_MANUAL : (pl->resident ? MODE_RESIDENT : MODE_STREAM);
    pl->smem = area + fixed;
    return true;
}

static cudaError_t launch_cfg(const StagePlan& pl, int grid, cudaStream_t st, cudaLaunchConfig_t* cfg,
                              cudaLaunchAttribute* attr) {
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)pl.cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg->gridDim = dim3((unsigned)grid);
    cfg->blockDim = dim3(ST);
    cfg->dynamicSmemBytes = pl.smem;
    cfg->stream = st;
    cfg->attrs = attr;
    cfg->numAttrs = 1;
    return cudaSuccess;
}

// test hook (cf_debug_set_plan): pin the cluster size / cap the co-resident clusters; part of the plan-cache key
static std::atomic<int> g_force_cs{0}, g_force_cap{0};
void debug_set_plan(int force_cs, int max_clusters) {
    g_force_cs.store(force_cs > 0 ? force_cs : 0);
    g_force_cap.store(max_clusters > 0 ? max_clusters : 0);
}
static thread_local cf_plan_info_t g_last_plan = {0, 0, 0, 0, 0, 0, 0, 0};
void last_plan(cf_plan_info_t* out) { *out = g_last_plan; }

typedef void (*StageKernel)(const StageParams);
static StageKernel stage_kernel_of(int mode, int spec) {
    static const StageKernel table[3][3] = {
        {fusion_stage_kernel<MODE_RESIDENT, SPEC_ANY>, fusion_stage_kernel<MODE_RESIDENT, SPEC_STAGE_STAGED>,
         fusion_stage_kernel<MODE_RESIDENT, SPEC_STAGE_L2>},
        {fusion_stage_kernel<MODE_STREAM, SPEC_ANY>, fusion_stage_kernel<MODE_STREAM, SPEC_STAGE_STAGED>,
         fusion_stage_kernel<MODE_STREAM, SPEC_STAGE_L2>},
        {fusion_stage_kernel<MODE_MANUAL, SPEC_ANY>, fusion_stage_kernel<MODE_MANUAL, SPEC_STAGE_STAGED>,
         fusion_stage_kernel<MODE_MANUAL, SPEC_STAGE_L2>}};
    return table[mode][spec];
}

// The kernel's function attributes are process-wide state of the CUDA runtime: they are set ONCE per device, to the
// largest values any plan uses, so that concurrent host threads with different shapes never race on them.
static cudaError_t device_setup(int dev) {
    static std::mutex mu;
    static std::map<int, cudaError_t> done;
    std::lock_guard<std::mutex> g(mu);
    auto f = done.find(dev);
    if (f != done.end()) return f->second;
    cudaError_t e = cudaSuccess;
    for (int m = 0; m < 9 && e == cudaSuccess; ++m) {
        const StageKernel k = stage_kernel_of(m / 3, m % 3);
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_DYN_CAP);
        // clusters of 16 are an opt-in; a device that refuses just never gets such a plan (occupancy query says 0)
        if (e == cudaSuccess && cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess)
            cudaGetLastError();
    }
    done[dev] = e;
    return e;
}

// co-resident clusters of this size with this much shared memory
static int max_clusters(const StagePlan& pl) {
    int n = 0;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    memset(&cfg, 0, sizeof(cfg));
    launch_cfg(pl, pl.cs, nullptr, &cfg, attr);
    if (cudaOccupancyMaxActiveClusters(&n, stage_kernel_of(pl.mode, SPEC_ANY), &cfg) != cudaSuccess) n = 0;
    cudaGetLastError();  // a refused cluster size is not an error of the call
    return n;
}

// cluster size: cheapest sum of rounds over the batch, then the most CTAs per frame.  Cached per
// (device, shape, flags): a steady-state call does one map lookup under a mutex and no CUDA runtime query.
static bool choose_plan(int dev, int B, int C, int H, int W, int K, bool use_tma, bool stage_dep, StagePlan* out) {
    typedef std::tuple<int, int, int, int, int, int, int, int, int> Key;
    static std::mutex mu;
    static std::map<Key, std::pair<bool, StagePlan>> cache;
    const int force_cs = g_force_cs.load() ? g_force_cs.load() : stage_env().force_cs, force_cap = g_force_cap.load();
    const Key key(dev, B, C, H, W, K, (use_tma ? 1 : 0) | (stage_dep ? 2 : 0), force_cs, force_cap);
    {
        std::lock_guard<std::mutex> g(mu);
        auto f = cache.find(key);
        if (f != cache.end()) {
            *out = f->second.second;
            return f->second.first;
        }
    }
    StagePlan best;
    memset(&best, 0, sizeof(best));
    long long best_rounds = -1, best_num = 0, best_den = 1;
    for (int cs = MAX_CS; cs >= 1; cs >>= 1) {
        if (force_cs && cs != force_cs) continue;
        StagePlan pl;
        if (!plan_for_cs(cs, C, H, W, K, use_tma, stage_dep, &pl)) continue;
        int cap = max_clusters(pl);
        if (force_cap && force_cap < cap) cap = force_cap;  // tests: force several frames per cluster
        if (stage_env().verbose) fprintf(stderr, "[cfd3d]   cs=%d: %d co-resident clusters (smem %zu)\n", cs, cap, pl.smem);
        if (cap < 1) continue;
        pl.n_clusters = B < cap ? B : cap;
        // Cost of a plan = rounds x the weight of one round.  A round over a resident share weighs 1; a round that streams
        // its share twice weighs 1 + 2 / cs (measured on B200, tools/plan_sweep.sh: 3 for single-CTA clusters, whose CTA
        // also runs the whole tail of its frame alone, 2 for pairs, 1.25 for clusters of 8).  Ties go to the plan with
        // fewer rounds, then to the larger cluster.
        const long long rounds = ((long long)B + pl.n_clusters - 1) / pl.n_clusters;
        const long long num = rounds * (cs + (pl.resident ? 0 : 2)), den = cs;
        const bool better = best_rounds < 0 || num * best_den < best_num * den ||
                            (num * best_den == best_num * den && rounds < best_rounds);
        if (better) {
            best = pl;
            best_rounds = rounds;
            best_num = num;
            best_den = den;
        }
    }
    std::lock_guard<std::mutex> g(mu);
    cache[key] = std::make_pair(best_rounds >= 0, best);
    *out = best;
    return best_rounds >= 0;
}

cudaError_t launch_fusion_stage(const float* heat, int B, int C, int H, int W, int K, const Workspace& ws,
                                const MergeJob* raw_job, const MergeJob* nms_job, const float* pc_dep,
                                const float* pc_vel, float max_pc_dist, float* pc_hm, long long pc_hm_stride,
                                int32_t* assoc, cudaStream_t st, int* launches) {
    const bool do_frustum = raw_job && raw_job->epilogue == EPI_FRUSTUM && pc_dep && pc_hm;
    const bool use_tma = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0);
    const bool stage_dep = do_frustum && (((size_t)H * W) % 4 == 0) && ((reinterpret_cast<uintptr_t>(pc_dep) & 15u) == 0);
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    e = device_setup(dev);
    if (e != cudaSuccess) return e;
    StagePlan best;
    if (!choose_plan(dev, B, C, H, W, K, use_tma, stage_dep, &best)) return cudaErrorInvalidConfiguration;

    StageParams p;
    memset(&p, 0, sizeof(p));
    p.heat = heat;
    p.B = B;
    p.C = C;
    p.H = H;
    p.W = W;
    p.Wp = (W + 3) & ~3;
    p.K = K;
    p.R = best.R;
    p.S = best.S;
    p.slot_floats = best.slot_floats;
    p.lists = (raw_job ? LIST_RAW : 0) | (nms_job ? LIST_NMS : 0);
    p.use_tma = use_tma ? 1 : 0;
    p.cs_log2 = best.cs_log2;
    p.hist_words = best.hist_words;
    p.slot_area = best.slot_area;
    p.dep_off = best.dep_off;
    p.off_comp = best.off_comp;
    p.off_ind = best.off_ind;
    p.off_msm = best.off_msm;
    p.off_gath = best.off_gath;
    p.off_det = best.off_det;
    p.off_best = best.off_best;
    p.off_par = best.off_par;
    p.off_own = best.off_own;
    p.GB = best.GB;
    p.wl_entries = best.wl_entries;
    p.g_log2 = 0;
    while ((32ll << p.g_log2) < (long long)p.R * (p.Wp / 4)) ++p.g_log2;
    // consecutive groups roughly double the survivors of the raw list (measured: 2.8 -> 4.7 x K at the reference size,
    // 2.3 -> 4.5 x K at 225 x 400, K = 500); a list that overflows CAND_CAP falls back to the exact selection from the
    // frame, which is milliseconds
    p.g_contig = 8ll * K <= CAND_CAP ? 1 : 0;
    // An SM moves ~0.15 scattered sectors per clock out of DRAM whatever asks for them -- loads, cp.async or prefetches: a
    // cluster's 4400 gathers per frame are 2 us on four SMs, where prefetching the hot survivors' sectors under the
    // barrier before hides most of them, and 15 us on one, where the prefetches only move the same wait into that barrier
    // (measured: 4.8 + 12.1 us with, 1.2 + 15.5 us without; 256 frames 163.8 -> 161.6 us).
    p.hot_passes = best.cs == 1 ? 0 : 2;
    p.late_iw = stage_env().late_iw;
    p.late_decode = (K + best.cs - 1) / best.cs >= stage_env().late_min ? stage_env().late_dw : 0;
    // (with the decode job's copies off the association's path, prefetching THEIR sectors only delays the frustum job's
    // in the same queue: frustum channels alone for clusters of 2 and 4 -- cfg2 32.8 -> 31.9 us, cfg4 89.7 -> 88.9; a
    // cluster of 8 on a full-resolution map still gains from both, 72.9 vs 73.9)
    if (p.late_decode && best.cs > 1 && best.cs <= 4) p.hot_passes = 1;
    p.gmax_off = best.gmax_off;
    p.fl_cap = best.fl_cap;
    p.ctr = ws.ctr;
    p.cand = ws.cand;
    p.mp.heat = heat;
    p.mp.C = C;
    p.mp.H = H;
    p.mp.W = W;
    p.mp.K = K;
    p.mp.ctr = ws.ctr;
    p.mp.cand = ws.cand;
    p.mp.det = ws.det;
    int n = 0;
    if (raw_job) {
        p.mp.list[n] = 0;
        p.mp.job[n] = *raw_job;
        ++n;
    }
    if (nms_job) {
        p.mp.list[n] = 1;
        p.mp.job[n] = *nms_job;
        ++n;
    }
    if (n == 1) {
        p.mp.list[1] = p.mp.list[0];
        p.mp.job[1] = p.mp.job[0];
    }
    p.n_jobs = n;
    p.do_frustum = do_frustum ? 1 : 0;
    p.pc_dep = pc_dep;
    p.pc_vel = pc_vel ? pc_vel : pc_dep;
    p.max_pc_dist = max_pc_dist;
    p.pc_hm = pc_hm;
    p.pc_hm_stride = pc_hm_stride > 0 ? pc_hm_stride : 3ll * H * W;
    p.assoc = assoc;
    p.recs = ws.recs;
    p.prof = stage_env().prof;
    p.mp.prof = p.prof;

    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[2];
    memset(&cfg, 0, sizeof(cfg));
    launch_cfg(best, best.n_clusters * best.cs, st, &cfg, attr);
    if (!stage_env().no_pdl) {
        // programmatic dependent launch: this grid's CTAs may be placed (and run their prologue) while the kernel before
        // it in the stream drains; they touch global memory only after griddepcontrol.wait.  Back-to-back stage calls
        // (batches streaming through a graph) hide the grid launch latency that way.
        attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.numAttrs = 2;
    }
    if (stage_env().verbose)
        fprintf(stderr, "[cfd3d] stage plan: B=%d cs=%d clusters=%d R=%d S=%d slot=%dB area=%d dep_off=%d smem=%zu tma=%d\n", B,
                best.cs, best.n_clusters, best.R, best.S, best.slot_floats * 4, best.slot_area, best.dep_off, best.smem,
                (int)use_tma);
    g_last_plan.cluster_size = best.cs;
    g_last_plan.n_clusters = best.n_clusters;
    g_last_plan.rounds = (B + best.n_clusters - 1) / best.n_clusters;
    g_last_plan.resident = best.resident ? 1 : 0;
    g_last_plan.depth_plane_staged = best.dep_off >= 0 ? 1 : 0;
    g_last_plan.chunk_rows = best.R;
    g_last_plan.slots = best.S;
    g_last_plan.smem_bytes = (int)best.smem;
    // the fused stage (both lists, frustum + decode jobs, association) runs the kernel specialised for it
    const int spec = (do_frustum && raw_job && nms_job && nms_job->epilogue == EPI_DECODE)
                         ? (best.dep_off >= 0 ? SPEC_STAGE_STAGED : SPEC_STAGE_L2)
                         : SPEC_ANY;
    e = cudaLaunchKernelEx(&cfg, stage_kernel_of(best.mode, spec), p);
    if (e == cudaSuccess) ++*launches;
    return e;
}

#ifdef CFD3D_PROFILE
int debug_read_prof(long long* out, int n_cta) {
    if (n_cta > 1024) n_cta = 1024;
    return (int)cudaMemcpyFromSymbol(out, g_prof, sizeof(long long) * PROF_PTS * n_cta);
}
#endif

}  // namespace cf
