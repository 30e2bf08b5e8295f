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
constexpr int FI = 8;    // vector items per thread

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

// VEC = 4: the frame is read as float4 (frame size and base 16-byte aligned); VEC = 1: scalar fallback
template <int VEC>
__global__ void __launch_bounds__(FT) peak_filter_kernel(const PeakParams p) {
    __shared__ uint32_t s_scan[FT / 32 + 1];
    const int tid = threadIdx.x;
    const int b = blockIdx.y;
    const bool want_raw = (p.lists & LIST_RAW) != 0, want_nms = (p.lists & LIST_NMS) != 0;
    const uint32_t n_px = (uint32_t)p.C * (uint32_t)p.H * (uint32_t)p.W;
    const uint32_t n_items = n_px / VEC;
    const float* frame = p.heat + (size_t)b * n_px;
    uint32_t* ctr_raw = p.ctr + ((size_t)b * 2) * CTR_STRIDE;
    uint32_t* ctr_nms = ctr_raw + CTR_STRIDE;
    const float INF = CUDART_INF_F;

    // this CTA's pixels first: the loads fly while tau is derived from the frame's histogram
    const uint32_t item0 = blockIdx.x * (uint32_t)(FT * FI) + (uint32_t)tid;
    float v[FI][VEC];
#pragma unroll
    for (int k = 0; k < FI; ++k) {
        const uint32_t i = item0 + (uint32_t)(k * FT);
        if (VEC == 4) {
            float4 q = make_float4(-INF, -INF, -INF, -INF);
            if (i < n_items) q = __ldg(reinterpret_cast<const float4*>(frame) + i);
            v[k][0] = q.x;
            v[k][VEC > 1 ? 1 : 0] = q.y;
            v[k][VEC > 2 ? 2 : 0] = q.z;
            v[k][VEC > 3 ? 3 : 0] = q.w;
        } else {
            v[k][0] = i < n_items ? __ldg(frame + i) : -INF;
        }
    }

    const uint32_t* gh = p.hist + (size_t)b * 2 * HIST_BINS;
    float tau_raw = INF, tau_nms = INF;
    if (want_raw) tau_raw = tau_key_to_float(tau_scan<FT>(gh, ctr_raw[2], p.K, s_scan));
    if (want_nms) {
        tau_nms = tau_key_to_float(tau_scan<FT>(gh + HIST_BINS, ctr_nms[2], p.K, s_scan));
        if (!(tau_nms > 0.0f)) {
            // fewer than K positive peaks: suppressed pixels (worth 0) belong to the top-K too -> the merge
            // selects from the frame itself; flag the list as overflowed once
            if (blockIdx.x == 0 && tid == 0) atomicAdd(ctr_nms, (uint32_t)CAND_CAP + 1u);
            tau_nms = INF;
        }
    }
    const float tau_min = fminf(tau_raw, tau_nms);

    uint32_t mraw = 0u, mcan = 0u;  // per-lane pixel masks: raw survivors, nms'd candidates awaiting the 3x3 test
#pragma unroll
    for (int k = 0; k < FI; ++k) {
        const bool in = item0 + (uint32_t)(k * FT) < n_items;  // out-of-range items must not pass tau = -inf
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (in && v[k][j] >= tau_raw) mraw |= 1u << (k * VEC + j);
            if (in && v[k][j] >= tau_nms) mcan |= 1u << (k * VEC + j);
        }
    }
    // almost every warp ends here: none of its 1 K pixels reaches either threshold
    if (!__any_sync(FULL, (mraw | mcan) != 0u)) return;
    (void)tau_min;

    uint32_t mnms = 0u;
    for (uint32_t m = mcan; m;) {  // rare: 3x3 suppression test straight from global memory
        const int bit = __ffs(m) - 1;
        m &= m - 1;
        const uint32_t gi = (item0 + (uint32_t)(bit / VEC) * FT) * VEC + (uint32_t)(bit % VEC);
        if (is_peak(frame, p.H, p.W, gi, __ldg(frame + gi))) mnms |= 1u << bit;
    }
    if (want_raw) warp_append<VEC>(mraw, frame, item0, ctr_raw, p.cand + (size_t)b * 2 * CAND_CAP);
    if (want_nms) warp_append<VEC>(mnms, frame, item0, ctr_nms, p.cand + ((size_t)b * 2 + 1) * CAND_CAP);
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
                               const Workspace& ws, cudaStream_t st, int* launches) {
    PeakParams pp;
    fill_params(pp, heat, B, C, H, W, K, lists, plan, ws);
    const long long n_px = (long long)C * H * W;
    const bool vec = (n_px % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0);
    const long long n_items = vec ? n_px / 4 : n_px;
    const unsigned grid = (unsigned)((n_items + FT * FI - 1) / (FT * FI));
    if (vec)
        peak_filter_kernel<4><<<dim3(grid, B), FT, 0, st>>>(pp);
    else
        peak_filter_kernel<1><<<dim3(grid, B), FT, 0, st>>>(pp);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// merge: frame-wide sorted top-K of the survivors + epilogue
// -----------------------------------------------------------------------------------------------------
constexpr int MT = 1024;

struct ListSrc {
    const unsigned long long* cand;
    int n_in, n_round;
    template <class F>
    __device__ __forceinline__ void for_each(F&& f) const {
        for (int i = threadIdx.x; i < n_round; i += MT) f(i < n_in ? __ldcg(cand + i) : 0ull);
    }
};

// overflow path: the frame itself as the source (slow, exact; only reached under massive ties)
struct FrameSrc {
    const float* frame;
    int CH, H, W, n_round;
    bool nms;
    template <class F>
    __device__ __forceinline__ void for_each(F&& f) const {
        const int n_in = CH * W;
        for (int i = threadIdx.x; i < n_round; i += MT) {
            unsigned long long c = 0ull;
            if (i < n_in) {
                const int g = i / W, x = i - g * W;
                const float v = __ldg(frame + i);
                uint32_t key = f2key(v);
                if (nms) {
                    const int py = g % H;
                    float m = -CUDART_INF_F;
                    for (int dy = -1; dy <= 1; ++dy) {
                        if (py + dy < 0 || py + dy >= H) continue;
                        for (int dx = -1; dx <= 1; ++dx) {
                            if (x + dx < 0 || x + dx >= W) continue;
                            m = fmaxf(m, __ldg(frame + (size_t)(g + dy) * W + x + dx));
                        }
                    }
                    if (!(m == v)) key = KEY_ZERO;
                }
                c = make_comp(key, (uint32_t)i);
            }
            f(c);
        }
    }
};

struct MergeParams {
    const float* heat;
    int C, H, W, K;
    const uint32_t* ctr;
    const unsigned long long* cand;
    int list[2];  // 0 = raw, 1 = nms'd
    MergeJob job[2];
    DetParam* det;
};

__device__ __forceinline__ void epilogue_frustum(const MergeParams& p, const MergeJob& jb, int b, int k, int ind,
                                                 int x, int y) {
    const size_t HW = (size_t)p.H * p.W;
    const FrustumHeads& fh = jb.fh;
    // reference src/lib/utils/pointcloud.py:348-367
    const float xf = __fadd_rn((float)x, 0.5f), yf = __fadd_rn((float)y, 0.5f);
    float w = __ldg(fh.wh + ((size_t)b * 2 + 0) * HW + ind);
    float h = __ldg(fh.wh + ((size_t)b * 2 + 1) * HW + ind);
    if (w < 0.0f) w = 0.0f;
    if (h < 0.0f) h = 0.0f;
    const float hw2 = __fdiv_rn(w, 2.0f), hh2 = __fdiv_rn(h, 2.0f);
    DetParam d;
    d.x1 = __fsub_rn(xf, hw2);
    d.y1 = __fsub_rn(yf, hh2);
    d.x2 = __fadd_rn(xf, hw2);
    d.y2 = __fadd_rn(yf, hh2);
    float r[8], dm[3];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = __ldg(fh.rot + ((size_t)b * 8 + j) * HW + ind);
#pragma unroll
    for (int j = 0; j < 3; ++j) dm[j] = __ldg(fh.dim + ((size_t)b * 3 + j) * HW + ind);
    d.depth = __ldg(fh.depth + (size_t)b * HW + ind);
    const float alpha = get_alpha(r);                                   // :372
    const float cx = __fdiv_rn(__fadd_rn(d.x1, d.x2), 2.0f);            // :373-379
    d.thr = dist_thresh_one(fh.calib + 12 * (size_t)b, cx, dm, alpha);  // :380
    d.ind = ind;
    d.pad_ = 0;
    const size_t o = (size_t)b * p.K + k;
    p.det[o] = d;
    if (jb.fo.thr) jb.fo.thr[o] = d.thr;
    if (jb.fo.inds) jb.fo.inds[o] = ind;
}

__device__ __forceinline__ void epilogue_decode(const MergeParams& p, const MergeJob& jb, int b, int k, float score,
                                                int cls, int ind, int x, int y) {
    // reference src/lib/model/decode.py:38-41, :80-85, :132-163
    const size_t HW = (size_t)p.H * p.W;
    const cf_decode_heads_t& hd = jb.dec.heads;
    const cf_decode_out_t& o = jb.dec.out;
    const size_t ok = (size_t)b * p.K + k;
    if (hd.uncertainty) {
        const float u = __ldg(hd.uncertainty + (size_t)b * HW + ind);
        score = __fmul_rn(score, expf(-expf(u)));
    }
    o.scores[ok] = score;
    o.class_ids[ok] = (float)cls;
    const float xn = __fdiv_rn((float)x, (float)p.W), yn = __fdiv_rn((float)y, (float)p.H);
    o.centers[2 * ok + 0] = xn;
    o.centers[2 * ok + 1] = yn;
    float xs = __fmul_rn(xn, (float)jb.dec.out_w), ys = __fmul_rn(yn, (float)jb.dec.out_h);
    if (hd.reg) {
        xs = __fadd_rn(xs, __ldg(hd.reg + ((size_t)b * 2 + 0) * HW + ind));
        ys = __fadd_rn(ys, __ldg(hd.reg + ((size_t)b * 2 + 1) * HW + ind));
    } else {
        xs = __fadd_rn(xs, 0.5f);
        ys = __fadd_rn(ys, 0.5f);
    }
    if (hd.wh) {
        float w = __ldg(hd.wh + ((size_t)b * 2 + 0) * HW + ind);
        float h = __ldg(hd.wh + ((size_t)b * 2 + 1) * HW + ind);
        if (w < 0.0f) w = 0.0f;
        if (h < 0.0f) h = 0.0f;
        if (jb.dec.norm2d) {
            w = __fmul_rn(w, (float)jb.dec.out_w);
            h = __fmul_rn(h, (float)jb.dec.out_h);
        }
        const float hw2 = __fdiv_rn(w, 2.0f), hh2 = __fdiv_rn(h, 2.0f);
        float4 bb;
        bb.x = __fsub_rn(xs, hw2);
        bb.y = __fsub_rn(ys, hh2);
        bb.z = __fadd_rn(xs, hw2);
        bb.w = __fadd_rn(ys, hh2);
        o.bboxes[4 * ok + 0] = bb.x;
        o.bboxes[4 * ok + 1] = bb.y;
        o.bboxes[4 * ok + 2] = bb.z;
        o.bboxes[4 * ok + 3] = bb.w;
    }
    if (hd.amodal) {  // :165 de-normalise
        float ax = __ldg(hd.amodal + ((size_t)b * 2 + 0) * HW + ind);
        float ay = __ldg(hd.amodal + ((size_t)b * 2 + 1) * HW + ind);
        if (jb.dec.norm2d) {
            ax = __fmul_rn(ax, (float)jb.dec.out_w);
            ay = __fmul_rn(ay, (float)jb.dec.out_h);
        }
        o.amodal[2 * ok + 0] = ax;
        o.amodal[2 * ok + 1] = ay;
    }
    if (hd.depth) o.depth[ok] = __ldg(hd.depth + (size_t)b * HW + ind);
    if (o.inds) o.inds[ok] = ind;
}

// plain channel copies of the decode heads, spread over the whole block: out[b,k,ch] = head[b,ch,ind_k]
__device__ __forceinline__ void copy_head(const float* head, float* out, int nch, int b, int K, size_t HW,
                                          const int* sind) {
    if (!head) return;
    for (int t = threadIdx.x; t < K * nch; t += MT) {
        const int k = t / nch, ch = t - k * nch;
        out[((size_t)b * K + k) * nch + ch] = __ldg(head + ((size_t)b * nch + ch) * HW + sind[k]);
    }
}

__global__ void __launch_bounds__(MT) merge_kernel(const MergeParams p) {
    __shared__ SelScratch sc;
    __shared__ unsigned long long win[MT];
    __shared__ unsigned long long sorted[CF_MAX_K];
    __shared__ int sind[CF_MAX_K];
    const int tid = threadIdx.x;
    const int b = blockIdx.x, li = blockIdx.y;
    const MergeJob& jb = p.job[li];
    const int list = p.list[li];
    const int K = p.K;
    const int HWi = p.H * p.W;
    const uint32_t n = __ldcg(p.ctr + ((size_t)b * 2 + list) * CTR_STRIDE);  // survivors appended (>= K by construction)
    const unsigned long long* cand = p.cand + ((size_t)b * 2 + list) * CAND_CAP;

    int nw;  // entries of win[] to rank
    if (n <= (uint32_t)MT) {
        win[tid] = tid < (int)n ? __ldcg(cand + tid) : 0ull;
        nw = (int)n;
        __syncthreads();
    } else {
        if (n <= (uint32_t)CAND_CAP) {
            ListSrc src;
            src.cand = cand;
            src.n_in = (int)n;
            src.n_round = (src.n_in + MT - 1) / MT * MT;
            block_select<MT, 1>(src, K, sc, [&](uint32_t slot, unsigned long long c) { win[slot] = c; });
        } else {
            FrameSrc src;
            src.frame = p.heat + (size_t)b * p.C * HWi;
            src.CH = p.C * p.H;
            src.H = p.H;
            src.W = p.W;
            src.n_round = (p.C * HWi + MT - 1) / MT * MT;
            src.nms = list == 1;
            block_select<MT, 1>(src, K, sc, [&](uint32_t slot, unsigned long long c) { win[slot] = c; });
        }
        nw = K;
        __syncthreads();
    }

    // rank by counting (composites are distinct: ranks are a permutation); the K best land in sorted[]
    if (tid < nw) {
        const unsigned long long me = win[tid];
        int rank = 0;
        for (int j = 0; j < nw; ++j) rank += (win[j] > me) ? 1 : 0;
        if (rank < K) sorted[rank] = me;
    }
    __syncthreads();

    if (tid < K) {
        const unsigned long long c = sorted[tid];
        const uint32_t g = comp_gidx(c), key = comp_key(c);
        const int cls = (int)(g / (uint32_t)HWi);
        const int ind = (int)(g - (uint32_t)cls * (uint32_t)HWi);
        const int y = ind / p.W, x = ind - y * p.W;
        sind[tid] = ind;
        float score = key2f(key);
        if (key == KEY_ZERO) score = __fmul_rn(__ldg(p.heat + ((size_t)b * p.C + cls) * HWi + ind), 0.0f);
        const size_t o = (size_t)b * K + tid;
        if (jb.epilogue == EPI_TOPK) {
            if (jb.topk.scores) jb.topk.scores[o] = score;
            if (jb.topk.inds) jb.topk.inds[o] = ind;
            if (jb.topk.classes) jb.topk.classes[o] = cls;
            if (jb.topk.ys) jb.topk.ys[o] = y;
            if (jb.topk.xs) jb.topk.xs[o] = x;
        } else if (jb.epilogue == EPI_FRUSTUM) {
            epilogue_frustum(p, jb, b, tid, ind, x, y);
        } else if (jb.epilogue == EPI_DECODE) {
            epilogue_decode(p, jb, b, tid, score, cls, ind, x, y);
        }
    }
    if (jb.epilogue == EPI_DECODE) {
        __syncthreads();
        const size_t HW = (size_t)HWi;
        copy_head(jb.dec.heads.rot, jb.dec.out.rot, 8, b, K, HW, sind);
        copy_head(jb.dec.heads.dim, jb.dec.out.dim, 3, b, K, HW, sind);
        copy_head(jb.dec.heads.nuscenes_att, jb.dec.out.nuscenes_att, 8, b, K, HW, sind);
        copy_head(jb.dec.heads.velocity, jb.dec.out.velocity, 3, b, K, HW, sind);
    }
}

cudaError_t launch_merge(const float* heat, int B, int C, int H, int W, int K, const Plan& plan,
                         const Workspace& ws, const MergeJob* raw_job, const MergeJob* nms_job, cudaStream_t st,
                         int* launches) {
    (void)plan;
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
    if (n == 0) return cudaSuccess;
    if (n == 1) {
        mp.list[1] = mp.list[0];
        mp.job[1] = mp.job[0];
    }
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


