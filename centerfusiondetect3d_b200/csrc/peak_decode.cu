// peak_decode.cu -- heat-map peak decode: frame-wide top-K of the raw and of the 3x3-nms'd heat-map in two
// streaming passes plus a small merge, and the stand-alone nms / gather kernels.
//
// Replaces (reference, src/lib/): model/utils.py:112-128 nms, :6-38 topk, :41-72 getFeature /
// transposeAndGetFeature; the gathers + box/alpha/threshold block of utils/pointcloud.py:347-380; and
// model/decode.py:10-174 fusionDecode.
//
// The reference's two-level topk (per class, then across classes) is the frame-wide top-K of C*H*W values.
// A batch is viewed as a (B*C*H) x W row matrix.
//
//   pass A  peak_bound   persistent: 2 CTAs per SM, each owning a contiguous range of rows, streamed through a
//                        3-stage shared-memory ring of R-row tiles (+1 halo row each side) by 1-D TMA bulk copies
//                        (cp.async.bulk + mbarrier complete_tx).  Per tile every thread takes the maximum of
//                        its own float4 items; the thread groups of a frame are disjoint, so the K-th largest
//                        group maximum is a lower bound of the K-th largest pixel.  For the nms'd list a group
//                        maximum counts only if that pixel survives the 3x3 suppression (checked in the tile).
//                        Group maxima go into a 4096-bin shared-memory histogram, flushed to the frame's global
//                        histogram once per (CTA, frame); the last CTA to finish a frame scans it and publishes
//                        tau = lower edge of the bin holding the K-th largest group maximum.
//                        This is the only HBM read of the heat-map.
//   pass B  peak_filter  re-reads the heat-map (L2 hits at the reference's batch sizes) with plain float4 loads
//                        and appends every pixel >= tau (and, for the nms'd list, surviving the 3x3 test) to
//                        the frame's survivor list: typically 1.1-3 x K entries, CAND_CAP slots.
//   merge                one CTA per (frame, list): exact top-K of the survivors under the total order
//                        (score desc, class asc, index asc), sorted, + the frustum / decode / topk epilogue.
//                        If the survivors overflowed (massive ties) it selects from the frame itself.
#include <math_constants.h>

#include <cstdio>
#include <cstdlib>

#include "geometry.cuh"
#include "kernels.cuh"
#include "merge.cuh"
#include "select.cuh"

namespace cf {

// -----------------------------------------------------------------------------------------------------
// plan + workspace
// -----------------------------------------------------------------------------------------------------
static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

constexpr int PT_THREADS = 256;
constexpr int PB_STAGES = 5;       // TMA ring depth of the bound pass
constexpr int PB_CONSUMERS = PB_STAGES;  // consumer warp w owns stage w: a warp never waits two phases ahead
constexpr int PB_THREADS = (PB_CONSUMERS + 1) * 32;  // + the producer warp
constexpr size_t SMEM_CAP = 227 * 1024 - 1024;  // dynamic budget: leaves room for the kernels' static shared memory
constexpr size_t PB_HIST_BYTES = 2 * HIST_BINS * sizeof(uint32_t);

static size_t bound_smem(long long R, int Wp) {
    return PB_STAGES * align_up((size_t)R * Wp * sizeof(float), 128) + PB_HIST_BYTES;
}

bool make_plan(int C, int H, int W, int K, Plan* p) {
    (void)K;
    const long long CH = (long long)C * H;
    const int Wp = (W + 3) & ~3;
    if (bound_smem(1, Wp) > SMEM_CAP) return false;  // one row per stage must fit
    int tile_floats = 4096;  // 16 KB stages: 5 stages + the histograms leave room for 2 CTAs per SM
    const char* e = getenv("CFD3D_TILE_FLOATS");
    if (e && atoi(e) >= 256) tile_floats = atoi(e);
    long long R = tile_floats / Wp;
    if (R < 1) R = 1;
    if (R > CH) R = CH;
    while (R > 1 && bound_smem(R, Wp) > SMEM_CAP) --R;
    p->T = PT_THREADS;
    p->R = (int)R;
    p->NT = (int)((CH + R - 1) / R);
    p->Wp = Wp;
    p->smem = bound_smem(R, Wp);
    return true;
}

size_t workspace_layout(int B, int C, int H, int W, int K, const Plan& plan, void* base, Workspace* out) {
    (void)C;
    (void)H;
    (void)W;
    (void)plan;
    size_t off = 0;
    char* b = (char*)base;
    const size_t hist = align_up((size_t)B * 2 * HIST_BINS * sizeof(uint32_t), 256);
    const size_t ctr = align_up((size_t)B * 2 * CTR_STRIDE * sizeof(uint32_t), 256);
    const size_t tau = align_up((size_t)B * 2 * sizeof(uint32_t), 256);
    const size_t cand = align_up((size_t)B * 2 * CAND_CAP * sizeof(unsigned long long), 256);
    const size_t det = align_up((size_t)B * K * sizeof(DetParam), 256);
    const size_t rec = align_up((size_t)B * K * sizeof(PaintRec), 256);
    if (out) {
        out->zero_base = b;
        out->zero_bytes = hist + ctr;
        out->hist = (uint32_t*)(b + off);
    }
    off += hist;
    if (out) out->ctr = (uint32_t*)(b + off);
    off += ctr;
    if (out) out->tau = (uint32_t*)(b + off);
    off += tau;
    if (out) out->cand = (unsigned long long*)(b + off);
    off += cand;
    if (out) out->det = (DetParam*)(b + off);
    off += det;
    if (out) out->recs = (PaintRec*)(b + off);
    off += rec;
    return off;
}

struct PeakParams {
    const float* heat;
    int B, C, H, W, Wp, K, R, lists, use_tma, dbg;
    uint32_t* hist;
    uint32_t* ctr;
    uint32_t* tau;
    unsigned long long* cand;
};

// tau = lower edge of the highest bin b* with count(bins >= b*) >= K; 0 when the frame has fewer than K groups.
// Scans the frame's histogram downwards from its top occupied bin, T bins per round (the K-th largest group
// maximum sits a few dozen bins below the top on real heat-maps: one round).  Whole CTA; returns the key.
template <int T>
__device__ __forceinline__ uint32_t tau_scan(const uint32_t* hist, uint32_t top, int K, uint32_t* s_scan) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = T / 32;
    uint32_t carry = 0u;
    for (int round = 0;; ++round) {
        const int bin = (int)top - (round * T + tid);
        const uint32_t c = bin >= 0 ? __ldg(hist + bin) : 0u;
        uint32_t inc = c;  // inclusive prefix over tid = over descending bins
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_up_sync(FULL, inc, off);
            if (lane >= off) inc += o;
        }
        if (lane == 31) s_scan[warp] = inc;
        if (tid == 0) s_scan[NW] = 0xffffffffu;
        __syncthreads();
        uint32_t below = carry, total = carry;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            const uint32_t t = s_scan[w];
            if (w < warp) below += t;
            total += t;
        }
        inc += below;
        if (inc >= (uint32_t)K && inc - c < (uint32_t)K) s_scan[NW] = (uint32_t)bin << HIST_SHIFT;  // one thread
        __syncthreads();
        const uint32_t found = s_scan[NW];
        __syncthreads();
        if (found != 0xffffffffu) return found;
        if ((int)top - (round + 1) * T < 0) return 0u;  // histogram exhausted: fewer than K groups
        carry = total;
    }
}

// -----------------------------------------------------------------------------------------------------
// pass A  peak_bound: group maxima -> histogram -> tau
// -----------------------------------------------------------------------------------------------------
// one shared-memory atomic per distinct bin per warp
__device__ __forceinline__ void hist_add(uint32_t* hist, uint32_t key) {
    const uint32_t bin = key ? (key >> HIST_SHIFT) : 0xffffffffu;
    const uint32_t peers = __match_any_sync(FULL, bin);
    if (key && (threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(hist + bin, (uint32_t)__popc(peers));
}

// order-preserving key -> the float it encodes; key 0 ("everything") -> -inf
__device__ __forceinline__ float tau_key_to_float(uint32_t key) { return key ? key2f(key) : -CUDART_INF_F; }

// One chunk in shared memory, processed by ONE warp: rows [rr, rr+n) of a frame's (C*H) x W matrix, row stride
// Wp, no halo.  Every lane takes the maximum of its own float4 items (group = the lane's items of this chunk);
// returns it as keys: kraw for the raw list, knms for the nms'd list (0 = no contribution).
__device__ __forceinline__ void chunk_group_max(const float* tile, const float* frame, int W, int Wp, int H, int rr,
                                                int n, float inv_w4, bool want_nms, uint32_t& kraw, uint32_t& knms) {
    const int lane = threadIdx.x & 31;
    const int W4 = Wp >> 2;
    const int n_items = n * W4;
    const float4* c4p = reinterpret_cast<const float4*>(tile);
    float best = -CUDART_INF_F;
    int bi = lane;
#pragma unroll 8
    for (int i = lane; i < n_items; i += 32) {
        const float4 v = c4p[i];
        const float m = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w));
        if (m > best) {
            best = m;
            bi = i;
        }
    }
    kraw = 0u;
    knms = 0u;
    if (lane >= n_items) return;
    int r = (int)(((float)bi + 0.5f) * inv_w4);  // bi / W4, exact for bi < 2^21
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
    int x = -1;
#pragma unroll
    for (int j = 3; j >= 0; --j)
        if (cv[j] == best && 4 * cc + j < W) x = 4 * cc + j;
    if (x < 0) return;  // (all -inf in a padded float4)
    kraw = f2key(best);
    if (!want_nms) return;
    // the group maximum backs the nms'd bound only if that pixel survives the 3x3 suppression; a group whose
    // maximum is suppressed contributes nothing (weaker bound, still a bound).  Rows above / below come from the
    // chunk, or from global memory (L2) when the pixel sits on the chunk's first / last row.
    const int g = rr + r;    // frame-matrix row
    const int py = g % H;    // row inside its class plane: planes do not see each other
    const float* rows[3];
    rows[0] = r > 0 ? tile + (r - 1) * Wp : frame + (size_t)(g - 1) * W;
    rows[1] = tile + r * Wp;
    rows[2] = r < n - 1 ? tile + (r + 1) * Wp : frame + (size_t)(g + 1) * W;
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

// grid = (CTAs per frame, B).  Warps 0..PB_CONSUMERS-1 consume chunks (chunk q -> warp q % PB_CONSUMERS, stage
// q % PB_STAGES); warp PB_CONSUMERS is the TMA producer.  full[s]: tx-count barrier the bulk copy completes;
// empty[s]: the consuming warp's release.  No CTA-wide barrier between the prologue and the final flush.
__global__ void __launch_bounds__(PB_THREADS, 2) peak_bound_kernel(const PeakParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint64_t full[PB_STAGES], empty[PB_STAGES];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int CH = p.C * p.H, R = p.R, W = p.W, Wp = p.Wp;
    const bool want_raw = (p.lists & LIST_RAW) != 0, want_nms = (p.lists & LIST_NMS) != 0;
    const size_t stage_floats = (((size_t)R * Wp * sizeof(float) + 127) & ~(size_t)127) / sizeof(float);
    float* stages = reinterpret_cast<float*>(smem_raw);
    uint32_t* sh_raw = reinterpret_cast<uint32_t*>(stages + PB_STAGES * stage_floats);
    uint32_t* sh_nms = sh_raw + HIST_BINS;
    const int b = blockIdx.y;
    const int row_begin = (int)((long long)CH * blockIdx.x / gridDim.x);
    const int row_end = (int)((long long)CH * (blockIdx.x + 1) / gridDim.x);
    const int n_chunks = (row_end - row_begin + R - 1) / R;
    const float* frame = p.heat + (size_t)b * CH * W;

    for (int i = tid; i < 2 * HIST_BINS; i += PB_THREADS) sh_raw[i] = 0u;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < PB_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == PB_CONSUMERS) {
        if (p.use_tma && lane == 0) {
            for (int q = 0; q < n_chunks; ++q) {
                const int s = q % PB_STAGES;
                if (q >= PB_STAGES) mbar_wait(&empty[s], (uint32_t)(q / PB_STAGES - 1) & 1u);
                const int rr = row_begin + q * R, n = min(R, row_end - rr);
                const uint32_t bytes = (uint32_t)n * (uint32_t)W * 4u;
                mbar_expect_tx(&full[s], bytes);
                tma_bulk_g2s(stages + s * stage_floats, frame + (size_t)rr * W, bytes, &full[s]);
            }
        }
    } else {
        const float inv_w4 = 1.0f / (float)(Wp >> 2);
        for (int q = warp; q < n_chunks; q += PB_CONSUMERS) {
            const int s = q % PB_STAGES;
            const int rr = row_begin + q * R, n = min(R, row_end - rr);
            float* tile = stages + s * stage_floats;
            if (p.use_tma) {
                mbar_wait(&full[s], (uint32_t)(q / PB_STAGES) & 1u);
            } else {
                // fallback (W % 4 != 0 or unaligned base): the consuming warp fills its own stage (padded row stride)
                const int cnt = n * Wp;
                for (int i = lane; i < cnt; i += 32) {
                    const int row = i / Wp, x = i - row * Wp;
                    tile[i] = x < W ? __ldg(frame + (size_t)(rr + row) * W + x) : -CUDART_INF_F;
                }
                __syncwarp();
            }
            uint32_t kraw = 0u, knms = 0u;
            if (!(p.dbg & 4)) chunk_group_max(tile, frame, W, Wp, p.H, rr, n, inv_w4, want_nms, kraw, knms);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);  // stage free: the producer may refill it
            if (!(p.dbg & 1)) {
                if (want_raw) hist_add(sh_raw, kraw);
                if (want_nms) hist_add(sh_nms, knms);
            }
        }
    }
    __syncthreads();
    if (p.dbg & 2) return;

    // flush this CTA's histogram into the frame's global histogram (fire-and-forget REDs) and raise the frame's
    // top occupied bin; the filter pass derives tau from both (kernel boundary = visibility)
    uint32_t* gh = p.hist + (size_t)b * 2 * HIST_BINS;
#pragma unroll
    for (int l = 0; l < 2; ++l) {
        uint32_t top = 0u;
        for (int i = tid; i < HIST_BINS; i += PB_THREADS) {
            const uint32_t c = sh_raw[l * HIST_BINS + i];
            if (c) {
                atomicAdd(gh + l * HIST_BINS + i, c);
                top = (uint32_t)i;  // i ascends
            }
        }
        const uint32_t wt = __reduce_max_sync(FULL, top);
        if (lane == 0 && wt) atomicMax(p.ctr + ((size_t)b * 2 + l) * CTR_STRIDE + 2, wt);
    }
}

// -----------------------------------------------------------------------------------------------------
// pass B  peak_filter: survivors (value >= tau) -> per-frame candidate list
// -----------------------------------------------------------------------------------------------------
// 3x3 suppression test of one pixel straight from global memory (rare path: only pixels above tau)
__device__ __forceinline__ bool is_peak(const float* frame, int H, int W, uint32_t gi, float v) {
    const int g = (int)(gi / (uint32_t)W), x = (int)(gi - (uint32_t)g * (uint32_t)W);
    const int py = g % H;
    bool keep = true;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
        if (py + dy < 0 || py + dy >= H) continue;
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) {
            if (x + dx < 0 || x + dx >= W) continue;
            keep = keep && !(__ldg(frame + (size_t)(g + dy) * W + x + dx) > v);
        }
    }
    return keep;
}

constexpr int FT = 256;  // threads
constexpr int FI = 8;    // vector items per thread per batch
constexpr int FWCAP = 128;  // nms'd candidates a warp stages per batch for the side-by-side 3x3 tests

// warp-aggregated append of at most one survivor per lane
__device__ __forceinline__ void append_survivors(bool pred, unsigned long long comp, uint32_t* ctr,
                                                 unsigned long long* cand) {
    const uint32_t bal = __ballot_sync(FULL, pred);
    if (!bal) return;
    const int lane = threadIdx.x & 31, leader = __ffs(bal) - 1;
    uint32_t base = 0u;
    if (lane == leader) base = atomicAdd(ctr, (uint32_t)__popc(bal));
    base = __shfl_sync(FULL, base, leader);
    if (pred) {
        const uint32_t slot = base + __popc(bal & lanemask_lt());
        if (slot < (uint32_t)CAND_CAP) cand[slot] = comp;  // the counter keeps counting: overflow is detectable
    }
}

// warp-aggregated append of each lane's survivors (bit k*VEC+j of `mask` = pixel j of item k) to a frame list:
// one atomic per warp; the counter keeps counting past CAND_CAP so that overflow is detectable
template <int VEC>
__device__ __forceinline__ void warp_append(uint32_t mask, const float* frame, uint32_t item0, uint32_t* ctr,
                                            unsigned long long* cand) {
    const int lane = threadIdx.x & 31;
    const uint32_t cnt = (uint32_t)__popc(mask);
    uint32_t inc = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_up_sync(FULL, inc, off);
        if (lane >= off) inc += o;
    }
    const uint32_t total = __shfl_sync(FULL, inc, 31);
    if (!total) return;
    uint32_t base = 0u;
    if (lane == 31) base = atomicAdd(ctr, total);
    base = __shfl_sync(FULL, base, 31);
    uint32_t slot = base + inc - cnt;
    while (mask) {
        const int bit = __ffs(mask) - 1;
        mask &= mask - 1;
        const uint32_t gi = (item0 + (uint32_t)(bit / VEC) * FT) * VEC + (uint32_t)(bit % VEC);
        if (slot < (uint32_t)CAND_CAP) cand[slot] = make_comp(f2key(__ldg(frame + gi)), gi);
        ++slot;
    }
}

// tau of both lists in one go (see tau_scan): the two scans share their barriers.  tau[l] = 0 -> "everything".
template <int T>
__device__ __forceinline__ void tau_scan2(const uint32_t* gh, const uint32_t* ctr, int K, bool want0, bool want1,
                                          uint32_t* s_scan /* 2*(T/32+1) */, uint32_t tau[2]) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = T / 32;
    const bool want[2] = {want0, want1};
    uint32_t top[2], carry[2] = {0u, 0u};
    bool done[2];
#pragma unroll
    for (int l = 0; l < 2; ++l) {
        top[l] = want[l] ? __ldg(ctr + l * CTR_STRIDE + 2) : 0u;
        done[l] = !want[l];
        tau[l] = 0u;
    }
    for (int round = 0; !(done[0] && done[1]); ++round) {
        uint32_t c[2], inc[2];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            const int bin = (int)top[l] - (round * T + tid);
            c[l] = (!done[l] && bin >= 0) ? __ldg(gh + l * HIST_BINS + bin) : 0u;
        }
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            inc[l] = c[l];  // inclusive prefix over tid = over descending bins
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t o = __shfl_up_sync(FULL, inc[l], off);
                if (lane >= off) inc[l] += o;
            }
            if (lane == 31) s_scan[l * (NW + 1) + warp] = inc[l];
            if (tid == 0) s_scan[l * (NW + 1) + NW] = 0xffffffffu;
        }
        __syncthreads();
        uint32_t total[2];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            uint32_t below = carry[l];
            total[l] = carry[l];
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                const uint32_t t = s_scan[l * (NW + 1) + w];
                if (w < warp) below += t;
                total[l] += t;
            }
            inc[l] += below;
            const int bin = (int)top[l] - (round * T + tid);
            if (!done[l] && inc[l] >= (uint32_t)K && inc[l] - c[l] < (uint32_t)K)
                s_scan[l * (NW + 1) + NW] = (uint32_t)bin << HIST_SHIFT;  // exactly one thread
        }
        __syncthreads();
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            if (done[l]) continue;
            const uint32_t found = s_scan[l * (NW + 1) + NW];
            if (found != 0xffffffffu) {
                tau[l] = found;
                done[l] = true;
            } else if ((int)top[l] - (round + 1) * T < 0) {
                done[l] = true;  // histogram exhausted: fewer than K groups, tau stays 0
            } else {
                carry[l] = total[l];
            }
        }
        __syncthreads();
    }
}

// grid = (CTAs per frame, B).  VEC = 4: the frame is read as float4 (frame size and base 16-byte aligned);
// VEC = 1: scalar fallback.  The last CTA of a frame to finish runs the frame's merge jobs.
template <int VEC>
__global__ void __launch_bounds__(FT) peak_filter_kernel(const PeakParams p, const MergeParams mp, int n_jobs) {
    __shared__ uint32_t s_scan[2 * (FT / 32 + 1)];
    __shared__ int s_last;
    __shared__ uint32_t s_wbuf[FT / 32][FWCAP];
    __shared__ MergeSmem msm;
    const int tid = threadIdx.x;
    const int b = blockIdx.y;
    const bool want_raw = (p.lists & LIST_RAW) != 0, want_nms = (p.lists & LIST_NMS) != 0;
    const uint32_t n_px = (uint32_t)p.C * (uint32_t)p.H * (uint32_t)p.W;
    const uint32_t n_items = n_px / VEC;
    const float* frame = p.heat + (size_t)b * n_px;
    uint32_t* ctr_raw = p.ctr + ((size_t)b * 2) * CTR_STRIDE;
    uint32_t* ctr_nms = ctr_raw + CTR_STRIDE;
    const float INF = CUDART_INF_F;
    const bool rowvec = (p.W % 4) == 0;
    // this CTA's contiguous share of the frame, in batches of FT*FI items
    const uint32_t it_begin = (uint32_t)((unsigned long long)n_items * blockIdx.x / gridDim.x);
    const uint32_t it_end = (uint32_t)((unsigned long long)n_items * (blockIdx.x + 1) / gridDim.x);

    auto load_batch = [&](uint32_t item0, float (&v)[FI][VEC]) {
#pragma unroll
        for (int k = 0; k < FI; ++k) {
            const uint32_t i = item0 + (uint32_t)(k * FT);
            if (VEC == 4) {
                float4 q = make_float4(-INF, -INF, -INF, -INF);
                if (i < it_end) q = __ldg(reinterpret_cast<const float4*>(frame) + i);
                v[k][0] = q.x;
                v[k][VEC > 1 ? 1 : 0] = q.y;
                v[k][VEC > 2 ? 2 : 0] = q.z;
                v[k][VEC > 3 ? 3 : 0] = q.w;
            } else {
                v[k][0] = i < it_end ? __ldg(frame + i) : -INF;
            }
        }
    };

    // first batch in flight while tau is derived from the frame's histogram
    float v[FI][VEC];
    load_batch(it_begin + (uint32_t)tid, v);
    uint32_t tau_key[2];
    tau_scan2<FT>(p.hist + (size_t)b * 2 * HIST_BINS, ctr_raw, p.K, want_raw, want_nms, s_scan, tau_key);
    const float tau_raw = want_raw ? tau_key_to_float(tau_key[0]) : INF;
    float tau_nms = want_nms ? tau_key_to_float(tau_key[1]) : INF;
    if (want_nms && !(tau_nms > 0.0f)) {
        // fewer than K positive peaks: suppressed pixels (worth 0) belong to the top-K too -> the merge selects
        // from the frame itself; flag the list as overflowed once
        if (blockIdx.x == 0 && tid == 0) atomicAdd(ctr_nms, (uint32_t)CAND_CAP + 1u);
        tau_nms = INF;
    }

    for (uint32_t base = it_begin; base < it_end; base += FT * FI) {
        const uint32_t item0 = base + (uint32_t)tid;
        if (base != it_begin) load_batch(item0, v);
        uint32_t mraw = 0u, mcan = 0u;  // per-lane pixel masks: raw survivors, nms'd candidates awaiting the 3x3 test
#pragma unroll
        for (int k = 0; k < FI; ++k) {
            const bool in = item0 + (uint32_t)(k * FT) < it_end;  // out-of-range items must not pass tau = -inf
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                if (in && v[k][j] >= tau_raw) mraw |= 1u << (k * VEC + j);
                // a pixel with a larger left / right neighbour is suppressed: checked in registers when a float4
                // lies inside one map row
                bool can = in && v[k][j] >= tau_nms;
                if (VEC == 4 && rowvec) {
                    if (j > 0) can = can && !(v[k][j > 0 ? j - 1 : 0] > v[k][j]);
                    if (j < VEC - 1) can = can && !(v[k][j < VEC - 1 ? j + 1 : 0] > v[k][j]);
                }
                if (can) mcan |= 1u << (k * VEC + j);
            }
        }
        // almost every warp skips the rest: none of its 1 K pixels reaches either threshold
        if (!__any_sync(FULL, (mraw | mcan) != 0u)) continue;
        if (want_raw) warp_append<VEC>(mraw, frame, item0, ctr_raw, p.cand + (size_t)b * 2 * CAND_CAP);
        if (!want_nms) continue;
        // nms'd candidates: 3x3 suppression test straight from global memory (L2).  The warp's candidates are
        // compacted into shared memory and spread over its lanes, so the tests run side by side.
        unsigned long long* cand_nms = p.cand + ((size_t)b * 2 + 1) * CAND_CAP;
        const int lane = tid & 31, warp = tid >> 5;
        const uint32_t cnt = (uint32_t)__popc(mcan);
        uint32_t inc = cnt;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_up_sync(FULL, inc, off);
            if (lane >= off) inc += o;
        }
        const uint32_t total = __shfl_sync(FULL, inc, 31);
        if (total == 0u) continue;
        if (total <= (uint32_t)FWCAP) {
            uint32_t off = inc - cnt;
            for (uint32_t m = mcan; m;) {
                const int bit = __ffs(m) - 1;
                m &= m - 1;
                s_wbuf[warp][off++] = (item0 + (uint32_t)(bit / VEC) * FT) * VEC + (uint32_t)(bit % VEC);
            }
            __syncwarp();
            for (uint32_t e0 = 0; e0 < total; e0 += 32) {
                const uint32_t e = e0 + (uint32_t)lane;
                const bool have = e < total;
                const uint32_t gi = have ? s_wbuf[warp][e] : 0u;
                const float x = __ldg(frame + gi);
                const bool pk = have && is_peak(frame, p.H, p.W, gi, x);
                append_survivors(pk, make_comp(f2key(x), gi), ctr_nms, cand_nms);
            }
            __syncwarp();
        } else {  // (a warp full of candidates: plateaus) one lane at a time
            uint32_t mnms = 0u;
            for (uint32_t m = mcan; m;) {
                const int bit = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t gi = (item0 + (uint32_t)(bit / VEC) * FT) * VEC + (uint32_t)(bit % VEC);
                if (is_peak(frame, p.H, p.W, gi, __ldg(frame + gi))) mnms |= 1u << bit;
            }
            warp_append<VEC>(mnms, frame, item0, ctr_nms, cand_nms);
        }
    }

    // ---- merge tail: the last CTA of the frame to get here owns the complete survivor lists ----------------------
    if (n_jobs == 0) return;
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        s_last = atomicAdd(ctr_raw + 1, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int li = 0; li < n_jobs; ++li) merge_frame<FT>(mp, msm, b, li);
}

cudaError_t launch_peak_clear(const Workspace& ws, cudaStream_t st) {
    return cudaMemsetAsync(ws.zero_base, 0, ws.zero_bytes, st);
}

static void fill_params(PeakParams& pp, const float* heat, int B, int C, int H, int W, int K, int lists,
                        const Plan& plan, const Workspace& ws) {
    pp.heat = heat;
    pp.B = B;
    pp.C = C;
    pp.H = H;
    pp.W = W;
    pp.Wp = plan.Wp;
    pp.K = K;
    pp.R = plan.R;
    pp.lists = lists;
    pp.use_tma = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0);
    { const char* e = getenv("CFD3D_DBG"); pp.dbg = e ? atoi(e) : 0; }
    pp.hist = ws.hist;
    pp.ctr = ws.ctr;
    pp.tau = ws.tau;
    pp.cand = ws.cand;
}

cudaError_t launch_peak_bound(const float* heat, int B, int C, int H, int W, int K, int lists, const Plan& plan,
                              const Workspace& ws, cudaStream_t st, int* launches) {
    PeakParams pp;
    fill_params(pp, heat, B, C, H, W, K, lists, plan, ws);
    cudaError_t e = cudaFuncSetAttribute(peak_bound_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)plan.smem);
    if (e != cudaSuccess) return e;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // CTAs per frame: fill the chip's 2-per-SM slots in one wave, at least one full chunk per CTA
    const long long CH = (long long)C * H;
    long long cpf = (2ll * sms) / B;
    const long long max_cpf = CH / plan.R > 0 ? CH / plan.R : 1;
    if (cpf > max_cpf) cpf = max_cpf;
    if (cpf < 1) cpf = 1;
    peak_bound_kernel<<<dim3((unsigned)cpf, B), PB_THREADS, plan.smem, st>>>(pp);
    e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

cudaError_t launch_peak_filter(const float* heat, int B, int C, int H, int W, int K, int lists, const Plan& plan,
                               const Workspace& ws, const MergeJob* raw_job, const MergeJob* nms_job,
                               cudaStream_t st, int* launches) {
    PeakParams pp;
    fill_params(pp, heat, B, C, H, W, K, lists, plan, ws);
    MergeParams mp;
    mp.heat = heat;
    mp.C = C;
    mp.H = H;
    mp.W = W;
    mp.K = K;
    mp.ctr = ws.ctr;
    mp.cand = ws.cand;
    mp.det = ws.det;
    int n = 0;
    if (raw_job) {
        mp.list[n] = 0;
        mp.job[n] = *raw_job;
        ++n;
    }
    if (nms_job) {
        mp.list[n] = 1;
        mp.job[n] = *nms_job;
        ++n;
    }
    if (n < 2) {
        mp.list[1] = mp.list[0];
        mp.job[1] = mp.job[0];
    }
    const long long n_px = (long long)C * H * W;
    const bool vec = (n_px % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0);
    const long long n_items = vec ? n_px / 4 : n_px;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // CTAs per frame: two per SM over the batch, at least one batch of items each
    long long cpf = (2ll * sms) / B;
    const long long max_cpf = (n_items + FT * FI - 1) / (FT * FI);
    if (cpf > max_cpf) cpf = max_cpf;
    if (cpf < 1) cpf = 1;
    if (vec)
        peak_filter_kernel<4><<<dim3((unsigned)cpf, B), FT, 0, st>>>(pp, mp, n);
    else
        peak_filter_kernel<1><<<dim3((unsigned)cpf, B), FT, 0, st>>>(pp, mp, n);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// merge as its own kernel: one CTA of 1024 threads per (frame, job) -- the two jobs of a frame run side by side
// -----------------------------------------------------------------------------------------------------
constexpr int MT = 1024;
__global__ void __launch_bounds__(MT) merge_kernel(const MergeParams p) {
    __shared__ MergeSmem msm;
    merge_frame<MT>(p, msm, blockIdx.x, blockIdx.y);
}

static int fill_merge(MergeParams& mp, const float* heat, int C, int H, int W, int K, const Workspace& ws,
                      const MergeJob* raw_job, const MergeJob* nms_job) {
    mp.heat = heat;
    mp.C = C;
    mp.H = H;
    mp.W = W;
    mp.K = K;
    mp.ctr = ws.ctr;
    mp.cand = ws.cand;
    mp.det = ws.det;
    int n = 0;
    if (raw_job) {
        mp.list[n] = 0;
        mp.job[n] = *raw_job;
        ++n;
    }
    if (nms_job) {
        mp.list[n] = 1;
        mp.job[n] = *nms_job;
        ++n;
    }
    if (n == 1) {
        mp.list[1] = mp.list[0];
        mp.job[1] = mp.job[0];
    }
    return n;
}

cudaError_t launch_merge(const float* heat, int B, int C, int H, int W, int K, const Plan& plan,
                         const Workspace& ws, const MergeJob* raw_job, const MergeJob* nms_job, cudaStream_t st,
                         int* launches) {
    (void)plan;
    MergeParams mp;
    const int n = fill_merge(mp, heat, C, H, W, K, ws, raw_job, nms_job);
    if (n == 0) return cudaSuccess;
    merge_kernel<<<dim3(B, n), MT, 0, st>>>(mp);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// stand-alone nms (materialised heat-map) and NCHW gather
// -----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nms_kernel(const float* __restrict__ heat, int H, int W, float* __restrict__ out,
                                                  long long total) {
    const long long HW = (long long)H * W;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256ll) {
        const long long pc = i / HW;
        const int ind = (int)(i - pc * HW);
        const int y = ind / W, x = ind - y * W;
        const float* pl = heat + pc * HW;
        const float v = __ldg(pl + ind);
        float m = -CUDART_INF_F;
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy) {
            const int yy = y + dy;
            if (yy < 0 || yy >= H) continue;
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                const int xx = x + dx;
                if (xx < 0 || xx >= W) continue;
                m = fmaxf(m, __ldg(pl + (size_t)yy * W + xx));
            }
        }
        out[i] = __fmul_rn(v, (m == v) ? 1.0f : 0.0f);
    }
}

cudaError_t launch_nms(const float* heat, int B, int C, int H, int W, float* out, cudaStream_t st, int* launches) {
    const long long total = (long long)B * C * H * W;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    nms_kernel<<<(int)blocks, 256, 0, st>>>(heat, H, W, out, total);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

__global__ void __launch_bounds__(256) gather_kernel(const float* __restrict__ feat, int ch, long long HW,
                                                     const int64_t* __restrict__ inds, int M, float* __restrict__ out,
                                                     long long total) {
    for (long long t = blockIdx.x * 256ll + threadIdx.x; t < total; t += (long long)gridDim.x * 256ll) {
        const int c = (int)(t % ch);
        const long long bm = t / ch;
        const long long b = bm / M;
        out[t] = __ldg(feat + ((size_t)b * ch + c) * HW + inds[bm]);
    }
}

cudaError_t launch_gather(const float* feat, int B, int ch, int64_t HW, const int64_t* inds, int M, float* out,
                          cudaStream_t st, int* launches) {
    const long long total = (long long)B * M * ch;
    if (total == 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    gather_kernel<<<(int)blocks, 256, 0, st>>>(feat, ch, (long long)HW, inds, M, out, total);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

}  // namespace cf


