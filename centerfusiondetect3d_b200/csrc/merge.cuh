// merge.cuh -- per-frame merge of the peak-decode survivors: exact sorted top-K under the total order
// (score desc, class asc, index asc) + the topk / frustum / decode epilogue.  Device code of the fused stage kernel
// (stage.cu), spread over the CTAs of a frame's cluster:
//
//   select    every CTA ranks the jobs' survivor lists itself (rank by counting over key buckets), the two jobs side by
//             side on the two halves of the CTA: the sorted top-K lands in its own shared memory (xs_comp[job][rank])
//             and the tail follows without a cluster barrier;
//   tail      detection k belongs to CTA k % CS from here on: that CTA gathers the head values of ITS detections (all
//             loads of both jobs issued back to back: one DRAM round trip; one SM sustains only a few hundred
//             outstanding scattered loads, so the cluster shares them), runs the frustum geometry / decode epilogue
//             from shared memory and keeps the frustum parameters local for the association.
//
// Replaces (reference, src/lib/): model/utils.py:6-38 topk (second level), :41-72 transposeAndGetFeature; the
// gathers + box / alpha / threshold block of utils/pointcloud.py:347-380; model/decode.py:47-172.
#pragma once
#include <math_constants.h>

#include "assoc.cuh"
#include "geometry.cuh"
#include "kernels.cuh"
#include "post.cuh"
#include "select.cuh"

#ifndef CF_MERGE_PROF
#define CF_MERGE_PROF(i) do { } while (0)
#define CF_MERGE_PROF_DEP(i, v) do { } while (0)
#define CF_PROF_SLOT(thr, i, v) do { } while (0)
#endif

namespace cf {

constexpr int MERGE_WIN = 1024;  // survivors a HALF of the CTA ranks directly (the whole CTA: a full list, CAND_CAP)

// overflow path: the frame itself as the source (slow, exact; only reached under massive ties)
template <int T>
struct FrameSrc {
    const float* frame;
    int CH, H, W, n_round;
    bool nms;
    template <class F>
    __device__ __forceinline__ void for_each(F&& f) const {
        const int n_in = CH * W;
        for (int i = threadIdx.x; i < n_round; i += T) {
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

// shared memory of the selection
struct MergeSmem {
    SelScratch sc;
    unsigned long long win[MERGE_WIN];
};

// a list that overflowed its CAND_CAP slots (dense ties), or was flagged so (no useful bound): exact block-level
// selection of the K best into win[] from the frame itself.  Slow, exact, rare -> never inlined.
template <int T>
__device__ __noinline__ void select_overflow(const MergeParams& p, MergeSmem& sm, int b, int list) {
    const int K = p.K, HWi = p.H * p.W;
    FrameSrc<T> src;
    src.frame = p.heat + (size_t)b * p.C * HWi;
    src.CH = p.C * p.H;
    src.H = p.H;
    src.W = p.W;
    src.n_round = (p.C * HWi + T - 1) / T * T;
    src.nms = list == 1;
    block_select<T, 1>(src, K, sm.sc, [&](uint32_t slot, unsigned long long c) { sm.win[slot] = c; });
    __syncthreads();
}

// barrier of a thread group: the whole CTA (BAR == 0) or one of its halves (named barrier BAR, T threads)
template <int T, bool WHOLE>
__device__ __forceinline__ void group_sync(int bar) {
    if (WHOLE) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(T) : "memory");
}

// The sorted top-K of job li by the thread group (tid in [0, T); named barrier `bar`, or the whole CTA): emit(rank, composite) for every element
// of the K best.  Composites are distinct, so ranks are a permutation.  n = survivors appended to the job's list (>= K by
// construction; the caller read the cluster's counter).  A half of the CTA takes lists up to MERGE_WIN entries, the whole
// CTA any list that fits its slots; one that overflowed them (dense ties) or was flagged overflowed needs the exact
// block-level selection from the frame first, which only the whole CTA can run (WHOLE).
template <int T, bool WHOLE, class Emit>
__device__ __forceinline__ void merge_select(const MergeParams& p, MergeSmem& sm, int b, int li, int tid, int bar, uint32_t n,
                                             Emit&& emit) {
    const int list = p.list[li];
    const int K = p.K;
    // rank window: a half of the CTA ranks up to MERGE_WIN survivors, the whole CTA a full list (K = 500 on a
    // 225 x 400 map leaves ~1200: ranking them directly costs a third of the exact selection that would cut them to K)
    constexpr int WIN = WHOLE ? CAND_CAP : MERGE_WIN;
    constexpr int U = WIN / T;  // list entries per thread
    static_assert(WIN % T == 0 && WIN <= CAND_CAP && WIN <= SEL_CAP && 256 * 4 <= MERGE_WIN * 8, "rank window");
    CF_MERGE_PROF(0);
    const unsigned long long* cand = p.cand + ((size_t)b * 2 + list) * CAND_CAP;
    int nw;  // entries of win[] to rank
    unsigned long long me[U];
    if (n <= (uint32_t)WIN) {
        nw = (int)n;
#pragma unroll
        for (int u = 0; u < U; ++u) me[u] = tid + u * T < nw ? __ldcg(cand + tid + u * T) : 0ull;
    } else {
        if (WHOLE) select_overflow<T>(p, sm, b, list);  // rare, out of line: keeps the common path's code small
        nw = K;
#pragma unroll
        for (int u = 0; u < U; ++u) me[u] = tid + u * T < nw ? sm.win[tid + u * T] : 0ull;
        group_sync<T, WHOLE>(bar);  // win[] is read; surv[] (the selection's scratch) is free again
    }
    CF_MERGE_PROF(1);
    // Rank = number of larger composites.  Counting over the whole list costs n^2 compares (microseconds at n ~ 500),
    // so the list is first bucketed by key (256 linear buckets over [min key, max key], a counting sort in shared
    // memory): rank = entries of the higher buckets + larger entries of the own bucket.
    unsigned long long* sorted = sm.sc.surv;
    uint32_t* start = reinterpret_cast<uint32_t*>(sm.win);  // (win[] is free: its K entries sit in me[] by now)
    const int lane = tid & 31;
    if (tid < 256) sm.sc.hist[tid] = 0u;
    if (tid == 0) {
        sm.sc.L = 0xffffffffu;
        sm.sc.kmax = 0u;
    }
    uint32_t kmn = 0xffffffffu, kmx = 0u;
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (tid + u * T < nw) {
            kmn = min(kmn, comp_key(me[u]));
            kmx = max(kmx, comp_key(me[u]));
        }
    }
    kmn = __reduce_min_sync(FULL, kmn);
    kmx = __reduce_max_sync(FULL, kmx);
    group_sync<T, WHOLE>(bar);
    if (lane == 0 && kmn <= kmx) {
        atomicMin(&sm.sc.L, kmn);
        atomicMax(&sm.sc.kmax, kmx);
    }
    group_sync<T, WHOLE>(bar);
    const uint32_t kmin = sm.sc.L, range = sm.sc.kmax - kmin;
    const int shift = range < 256u ? 0 : 24 - __clz(range);  // (range >> shift) < 256
    uint32_t bk[U], pos[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        bk[u] = (comp_key(me[u]) - kmin) >> shift;
        if (tid + u * T < nw) pos[u] = atomicAdd(&sm.sc.hist[bk[u]], 1u);
    }
    group_sync<T, WHOLE>(bar);
    if (tid < 32) {  // start[b] = entries of the buckets above b
        uint32_t h[8], sum = 0u;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            h[j] = sm.sc.hist[lane * 8 + j];
            sum += h[j];
        }
        uint32_t suf = sum;  // inclusive suffix sum over lanes >= lane
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_down_sync(FULL, suf, off);
            if (lane + off < 32) suf += o;
        }
        uint32_t acc = suf - sum;
#pragma unroll
        for (int j = 7; j >= 0; --j) {
            start[lane * 8 + j] = acc;
            acc += h[j];
        }
    }
    group_sync<T, WHOLE>(bar);
#pragma unroll
    for (int u = 0; u < U; ++u)
        if (tid + u * T < nw) sorted[start[bk[u]] + pos[u]] = me[u];
    group_sync<T, WHOLE>(bar);
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (tid + u * T >= nw) continue;
        const int s0 = (int)start[bk[u]], s1 = s0 + (int)sm.sc.hist[bk[u]];
        if (s0 >= K) continue;  // K or more elements in higher buckets: not a winner
        int rank = s0;
        for (int jx = s0; jx < s1; ++jx) rank += (sorted[jx] > me[u]) ? 1 : 0;
        if (rank < K) emit(rank, me[u]);
    }
    group_sync<T, WHOLE>(bar);  // the scratch is reused by the next job of this group
    CF_MERGE_PROF(2);
}

// ---- asynchronous global -> shared copies of single values (cp.async, SASS LDGSTS): no register is held while the
// value is on its way, and a thread can wait for an older group of copies while a younger one is still in flight
__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {  // all but the N youngest groups of this thread have landed
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// detections kl0 .. kl0+nb of this CTA (detection k = first + kl * stride): the head values of a job's N-channel table
// gt on their way into gath[(kl - kl0) * GATH_STRIDE + slot].  inds = the job's peak pixels inds[rank].
// Executed by the thread group [0, T) given as (tid, T); the caller commits the group and waits for it.
template <int N>
__device__ __forceinline__ void gather_issue(const GatherTab& gt, int b, size_t HW, const int32_t* inds, int first,
                                             int stride, int kl0, int nb, float* gath, int tid, int T) {
    const int total = nb * N;
    for (int t = tid; t < total; t += T) {
        const int kl = t / N, s = t - kl * N;
        const float* base = gt.base[s];
        float* dst = gath + kl * GATH_STRIDE + s;
        if (base) cp_async4(dst, base + ((size_t)b * gt.nch[s] + gt.ch[s]) * HW + (uint32_t)inds[first + (kl0 + kl) * stride]);
        else *dst = 0.0f;
    }
}

// The trigonometry of one detection on its 4 lanes (whole warps call this together): lane 0 atan2(rot2, rot3), lane 1
// atan2(rot6, rot7), lane 2 the ray angle, then cos / sin of the yaw on lanes 0 / 1; returns the distance threshold on
// every lane of the quad.
__device__ __forceinline__ float frustum_thr_quad(float ay, float ax, float rot1, float rot5, float dim_w, float dim_l,
                                               int tid) {
    const int j = tid & 3;
    const float at = j < 3 ? f32_atan2(ay, ax) : 0.0f;
    const int l0 = (tid & 31) & ~3;
    const float at1 = __shfl_sync(FULL, at, l0), at2 = __shfl_sync(FULL, at, l0 + 1),
                ray = __shfl_sync(FULL, at, l0 + 2);
    const float yaw = yaw_wrap(alpha_blend(rot1, rot5, at1, at2), ray);
    const float cs_ = j == 0 ? (float)cos((double)yaw) : (j == 1 ? (float)sin((double)yaw) : 0.0f);
    const float c = __shfl_sync(FULL, cs_, l0), s = __shfl_sync(FULL, cs_, l0 + 1);
    return thr_from_cs(c, s, dim_w, dim_l);
}

// frustum geometry of this CTA's detections kl0 .. kl0+nb from the gathered heads: 4 lanes per detection, the three
// atan2 and then sin / cos evaluated side by side (fp64, rounded once: the parity contract of geometry.cuh).  The
// frustum parameters stay in shared memory (sdet[kl]) for the association.
// reference src/lib/utils/pointcloud.py:348-380
// Executed by the thread group [0, T) given as (tid, T): whole warps.
__device__ __forceinline__ void epilogue_frustum(const MergeParams& p, const MergeJob& jb, int b,
                                                 const unsigned long long* comps, const int32_t* inds, int first, int stride,
                                                 int kl0, int nb, const float* gath, DetParam* sdet, short4* sroi,
                                                 const float* cal, int tid, int T) {
    const int K = p.K, W = p.W, HWi = p.H * p.W;
    const int j = tid & 3;
    // cal = {calib[0,0], calib[0,2]} of the frame where the caller fetched them ahead of time (the row is cold in DRAM)
    const float* calib = jb.fh.calib + 12 * (size_t)b;
    const float cal0 = cal ? cal[0] : __ldg(calib + 0), cal2 = cal ? cal[1] : __ldg(calib + 2);
    // box of detection q of this batch (every thread that needs it computes it: a handful of instructions)
    auto box_of = [&](int q, int& ind, DetParam& d) {
        const int k = first + (kl0 + q) * stride;
        const float* g = gath + q * GATH_STRIDE;
        ind = inds[k];
        const int y = ind / W, x = ind - y * W;
        const float xf = __fadd_rn((float)x, 0.5f), yf = __fadd_rn((float)y, 0.5f);
        float w = g[FS_WH], h = g[FS_WH + 1];
        if (w < 0.0f) w = 0.0f;
        if (h < 0.0f) h = 0.0f;
        const float hw2 = __fdiv_rn(w, 2.0f), hh2 = __fdiv_rn(h, 2.0f);
        d.x1 = __fsub_rn(xf, hw2);
        d.y1 = __fsub_rn(yf, hh2);
        d.x2 = __fadd_rn(xf, hw2);
        d.y2 = __fadd_rn(yf, hh2);
    };
    // The paint window of a detection depends on its box alone.  When the quads of the trigonometry leave the upper
    // half of the thread group idle, those warps work the windows out meanwhile (one thread per detection) instead of
    // the association's single finishing thread per detection later.
    const bool side = (T % 64 == 0) && nb * 4 <= T / 2 && nb <= T / 2;
    if (side && tid >= T / 2) {
        const int q = tid - T / 2;
        if (q < nb) {
            int ind;
            DetParam d;
            box_of(q, ind, d);
            det_set_window(d, p.H, W);
            *reinterpret_cast<short4*>(&sdet[kl0 + q].ra) = make_short4(d.ra, d.rb, d.ca, d.cb);
            if (sroi) {  // and the ROI of the association's scan (ya, yb, xa, xb)
                int ya, yb, xa, xb;
                roi_bounds(p.H, W, d.x1, d.y1, d.x2, d.y2, ya, yb, xa, xb);
                sroi[kl0 + q] = make_short4((short)ya, (short)yb, (short)xa, (short)xb);
            }
        }
        return;
    }
    for (int q0 = 0; q0 < nb; q0 += T / 4) {
        const int q = q0 + (tid >> 2);
        const bool live = q < nb;
        const int ql = live ? q : 0;
        const int kl = kl0 + ql, k = first + kl * stride;
        const float* g = gath + ql * GATH_STRIDE;
        int ind;
        DetParam d;
        box_of(ql, ind, d);
        const float cx = __fdiv_rn(__fadd_rn(d.x1, d.x2), 2.0f);  // :373-379
        const float* r = g + FS_ROT;
        // lane 0: atan2(rot2, rot3), lane 1: atan2(rot6, rot7), lane 2: the ray angle
        const float ay = j == 0 ? r[2] : (j == 1 ? r[6] : __fsub_rn(cx, cal2));
        const float ax = j == 0 ? r[3] : (j == 1 ? r[7] : cal0);
        CF_MERGE_PROF_DEP(5, __float_as_int(ay) ^ __float_as_int(ax));
        const float thr = frustum_thr_quad(ay, ax, r[1], r[5], g[FS_DIM + 1], g[FS_DIM + 2], tid);  // :372, :230-235, :380
        CF_MERGE_PROF(6);
        if (live && j == 0) {
            d.depth = g[FS_DEPTH];
            d.thr = thr;
            *reinterpret_cast<float4*>(&sdet[kl].x1) = make_float4(d.x1, d.y1, d.x2, d.y2);
            *reinterpret_cast<float2*>(&sdet[kl].depth) = make_float2(d.depth, d.thr);
            if (!side) {
                det_set_window(d, p.H, W);
                *reinterpret_cast<short4*>(&sdet[kl].ra) = make_short4(d.ra, d.rb, d.ca, d.cb);
                if (sroi) {
                    int ya, yb, xa, xb;
                    roi_bounds(p.H, W, d.x1, d.y1, d.x2, d.y2, ya, yb, xa, xb);
                    sroi[kl] = make_short4((short)ya, (short)yb, (short)xa, (short)xb);
                }
            }
            const size_t o = (size_t)b * K + k;
            if (jb.fo.thr) jb.fo.thr[o] = d.thr;
            if (jb.fo.inds) jb.fo.inds[o] = ind;
        }
    }
}

// score of a sorted composite: the key decodes to the heat value; the nms'd list carries suppressed pixels as key 0
// (value +-0 * heat, i.e. the signed zero the reference's heat * keep produces)
__device__ __forceinline__ float comp_score(const MergeParams& p, int b, unsigned long long c, int cls, int ind) {
    const uint32_t key = comp_key(c);
    float score = key2f(key);
    if (key == KEY_ZERO) score = __fmul_rn(__ldg(p.heat + ((size_t)b * p.C + cls) * (p.H * p.W) + ind), 0.0f);
    return score;
}

// the 2-D box of a decoded detection (decode.py:132-150) out of its gathered channels
__device__ __forceinline__ float4 decode_bbox(const DecodeArgs& da, const float* g, float xn, float yn) {
    float xs = __fmul_rn(xn, (float)da.out_w), ys = __fmul_rn(yn, (float)da.out_h);
    if (da.heads.reg) {
        xs = __fadd_rn(xs, g[DS_REG]);
        ys = __fadd_rn(ys, g[DS_REG + 1]);
    } else {
        xs = __fadd_rn(xs, 0.5f);
        ys = __fadd_rn(ys, 0.5f);
    }
    float w = g[DS_WH], h = g[DS_WH + 1];
    if (w < 0.0f) w = 0.0f;
    if (h < 0.0f) h = 0.0f;
    if (da.norm2d) {
        w = __fmul_rn(w, (float)da.out_w);
        h = __fmul_rn(h, (float)da.out_h);
    }
    const float hw2 = __fdiv_rn(w, 2.0f), hh2 = __fdiv_rn(h, 2.0f);
    return make_float4(__fsub_rn(xs, hw2), __fsub_rn(ys, hh2), __fadd_rn(xs, hw2), __fadd_rn(ys, hh2));
}
// its amodal offset (decode.py:165 de-normalise)
__device__ __forceinline__ float2 decode_amodal(const DecodeArgs& da, const float* g) {
    float ax = g[DS_AMODAL], ay = g[DS_AMODAL + 1];
    if (da.norm2d) {
        ax = __fmul_rn(ax, (float)da.out_w);
        ay = __fmul_rn(ay, (float)da.out_h);
    }
    return make_float2(ax, ay);
}

// postProcess of one decoded detection from its gathered channels (lane 3 of the detection's quad; out of line: the
// epilogues' common path carries a call, not the code)
static __device__ __noinline__ void epilogue_post(const DecodeArgs& da, const float* g, float xn, float yn, int b, size_t ok) {
    const cf_decode_heads_t& hd = da.heads;
    const cf_decode_out_t& o = da.out;
    PostDet d;
    d.has = (hd.wh ? PD_BBOX : 0) | (hd.rot ? PD_ROT : 0) | (hd.depth ? PD_DEPTH : 0) | (hd.dim ? PD_DIM : 0) |
            (hd.amodal ? PD_AMODAL : 0) | (hd.velocity ? PD_VEL : 0);
    d.cx = xn;
    d.cy = yn;
    const float4 bb = hd.wh ? decode_bbox(da, g, xn, yn) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    d.bb[0] = bb.x, d.bb[1] = bb.y, d.bb[2] = bb.z, d.bb[3] = bb.w;
    const float2 am = hd.amodal ? decode_amodal(da, g) : make_float2(0.0f, 0.0f);
    d.amodal[0] = am.x, d.amodal[1] = am.y;
#pragma unroll
    for (int i = 0; i < 8; ++i) d.rot[i] = g[DS_ROT + i];
#pragma unroll
    for (int i = 0; i < 3; ++i) d.dim[i] = g[DS_DIM + i];
#pragma unroll
    for (int i = 0; i < 3; ++i) d.vel[i] = g[DS_VEL + i];
    d.depth = g[DS_DEPTH];
    post_detection(da.post, da.post.calib + 12 * (size_t)b, (long long)ok, d);
    if (hd.velocity) {  // (the plain copy of this head is skipped: see merge_tail_decode)
#pragma unroll
        for (int i = 0; i < 3; ++i) o.velocity[3 * ok + i] = d.vel[i];
    }
}

// reference src/lib/model/decode.py:38-41, :80-85, :132-165
// Four lanes per detection, each with its own part of the outputs (this code runs once per frame on a few threads: its
// cost is the length of the longest thread's instruction sequence): lane 0 score / class / centre / index, lane 1 the
// 2-D box, lane 2 amodal offset and depth, lane 3 -- when the caller asked for it -- the detection's postProcess
// (post.cuh; row f1) from the same gathered channels: class id + 1 and the re-oriented velocity land where the
// reference updates them in place, the rest in the post outputs.
__device__ __forceinline__ void epilogue_decode(const MergeParams& p, const MergeJob& jb, int b,
                                                const unsigned long long* comps, int first, int stride, int kl0, int nb,
                                                const float* gath, int tid, int T) {
    const int K = p.K, HWi = p.H * p.W;
    const cf_decode_heads_t& hd = jb.dec.heads;
    const cf_decode_out_t& o = jb.dec.out;
    const bool post = jb.dec.post.enabled != 0;
    const int j = tid & 3;
    for (int q = tid >> 2; q < nb; q += T >> 2) {
        const int k = first + (kl0 + q) * stride;
        const unsigned long long c = comps[k];
        const uint32_t gi = comp_gidx(c);
        const int cls = (int)(gi / (uint32_t)HWi);
        const int ind = (int)(gi - (uint32_t)cls * (uint32_t)HWi);
        const int y = ind / p.W, x = ind - y * p.W;
        const float* g = gath + q * GATH_STRIDE;
        const size_t ok = (size_t)b * K + k;
        const float xn = __fdiv_rn((float)x, (float)p.W), yn = __fdiv_rn((float)y, (float)p.H);
        if (j == 0) {
            float score = comp_score(p, b, c, cls, ind);
            if (hd.uncertainty) score = __fmul_rn(score, expf(-expf(g[DS_UNC])));
            o.scores[ok] = score;
            o.class_ids[ok] = post ? __fadd_rn((float)cls, 1.0f) : (float)cls;  // (postProcess.py:34)
            o.centers[2 * ok + 0] = xn;
            o.centers[2 * ok + 1] = yn;
            if (o.inds) o.inds[ok] = ind;
        } else if (j == 1) {
            if (hd.wh) *reinterpret_cast<float4*>(o.bboxes + 4 * ok) = decode_bbox(jb.dec, g, xn, yn);
        } else if (j == 2) {
            if (hd.amodal) {
                const float2 a = decode_amodal(jb.dec, g);
                o.amodal[2 * ok + 0] = a.x;
                o.amodal[2 * ok + 1] = a.y;
            }
            if (hd.depth) o.depth[ok] = g[DS_DEPTH];
        } else if (post) {
            epilogue_post(jb.dec, g, xn, yn, b, ok);
        }
    }
}

// plain channel copies of the decode heads out of the gather buffer: out[b,k,ch] = gath[q][slot0 + ch]
__device__ __forceinline__ void copy_head(const float* head, float* out, int nch, int slot0, int b, int K, int first,
                                          int stride, int kl0, int nb, const float* gath, int tid, int T) {
    if (!head) return;
    for (int t = tid; t < nb * nch; t += T) {
        const int q = t / nch, ch = t - q * nch;
        const int k = first + (kl0 + q) * stride;
        out[((size_t)b * K + k) * nch + ch] = gath[q * GATH_STRIDE + slot0 + ch];
    }
}

// epilogue of a TOPK job for detections kl0 .. kl0+nb (thread group (tid, T))
__device__ __forceinline__ void epilogue_topk(const MergeParams& p, const MergeJob& jb, int b, const unsigned long long* comps,
                                              int first, int stride, int kl0, int nb, int tid, int T) {
    const int K = p.K, HWi = p.H * p.W;
    for (int q = tid; q < nb; q += T) {
        const int k = first + (kl0 + q) * stride;
        const unsigned long long c = comps[k];
        const uint32_t gi = comp_gidx(c);
        const int cls = (int)(gi / (uint32_t)HWi);
        const int ind = (int)(gi - (uint32_t)cls * (uint32_t)HWi);
        const int y = ind / p.W, x = ind - y * p.W;
        const size_t o = (size_t)b * K + k;
        if (jb.topk.scores) jb.topk.scores[o] = comp_score(p, b, c, cls, ind);
        if (jb.topk.inds) jb.topk.inds[o] = ind;
        if (jb.topk.classes) jb.topk.classes[o] = cls;
        if (jb.topk.ys) jb.topk.ys[o] = y;
        if (jb.topk.xs) jb.topk.xs[o] = x;
    }
}

// epilogues of the DECODE jobs for detections kl0 .. kl0+nb out of their gather buffers (thread group (tid, T)); the
// caller has waited for the jobs' copies and passed a barrier
template <bool FUSED>  // FUSED: the fused stage -- job 0 is the frustum job, job 1 the decode job (compile time)
__device__ __forceinline__ void merge_tail_decode(const MergeParams& p, int n_jobs, int b, int first, int stride, int kl0,
                                                  int nb, const unsigned long long* xs_comp, const float* gath, int GB,
                                                  int tid, int T) {
    const int K = p.K;
    for (int li = FUSED ? 1 : 0; li < (FUSED ? 2 : n_jobs); ++li) {
        const MergeJob& jb = p.job[li];
        if (!FUSED && jb.epilogue != EPI_DECODE) continue;
        const unsigned long long* comps = xs_comp + li * K;
        const float* g = gath + (size_t)li * GB * GATH_STRIDE;
        epilogue_decode(p, jb, b, comps, first, stride, kl0, nb, g, tid, T);
        copy_head(jb.dec.heads.rot, jb.dec.out.rot, 8, DS_ROT, b, K, first, stride, kl0, nb, g, tid, T);
        copy_head(jb.dec.heads.dim, jb.dec.out.dim, 3, DS_DIM, b, K, first, stride, kl0, nb, g, tid, T);
        copy_head(jb.dec.heads.nuscenes_att, jb.dec.out.nuscenes_att, 8, DS_ATT, b, K, first, stride, kl0, nb, g, tid, T);
        if (!jb.dec.post.enabled)  // (with postProcess fused the velocity is written by the detection's post lane)
            copy_head(jb.dec.heads.velocity, jb.dec.out.velocity, 3, DS_VEL, b, K, first, stride, kl0, nb, g, tid, T);
    }
}

// The tail of the jobs for this CTA's detections (k = first + kl * stride, kl < n_local), in batches of GB detections.
// Every head value travels as an asynchronous 4-byte copy straight into the gather buffers (gath holds
// n_jobs * GB * GATH_STRIDE floats): the frustum job's copies go out first and form their own group, because the
// association waits for them, the decode job's 29 channels per detection follow -- 140 k scattered sectors per batch hit
// the DRAM at the same moment from every cluster, and whoever is served last sets the pace of the phase.  The frustum
// epilogue starts as soon as ITS group has landed.  defer_decode (one batch only): the decode epilogue is left to the
// caller (merge_tail_decode on a few warps beside the association's ROI scans); the closing barrier here comes after
// every thread has waited for its decode copies.  late_d (with defer_decode; CTAs that own many detections): the decode
// job's copies are not even issued here -- an SM moves ~0.15 scattered sectors per clock, and a thread that issues
// copies into a full queue stalls: the caller's late_dw decode warps leave as soon as the frustum job's copies have
// landed, issue the decode job's, wait for them and run the decode epilogue, beside the frustum epilogue and the ROI
// scans of the other warps.
template <int T, bool FUSED>
__device__ __forceinline__ void merge_tail(const MergeParams& p, int n_jobs, int b, int first, int stride, int n_local,
                                           const unsigned long long* xs_comp, const int32_t* xs_ind, float* gath, int GB,
                                           DetParam* sdet, short4* sroi, const float* cal, bool defer_decode, int late_dw = 0) {
    const bool late_d = late_dw > 0;  // (number of decode warps)
    const int K = p.K, HWi = p.H * p.W;
    const int tid = (int)threadIdx.x;
    for (int kl0 = 0; kl0 < n_local; kl0 += GB) {
        const int nb = min(GB, n_local - kl0);
        if (FUSED) {  // group 0: the frustum job (job 0), group 1: the decode job (job 1)
            gather_issue<FS_N>(p.job[0].gt, b, (size_t)HWi, xs_ind, first, stride, kl0, nb, gath, tid, T);
            cp_async_commit();
            if (!late_d)
                gather_issue<DS_N>(p.job[1].gt, b, (size_t)HWi, xs_ind + K, first, stride, kl0, nb,
                                   gath + (size_t)GB * GATH_STRIDE, tid, T);
            cp_async_commit();
        } else {
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {  // group 0: the frustum job, group 1: the others
                for (int li = 0; li < n_jobs; ++li) {
                    const MergeJob& jb = p.job[li];
                    const bool isf = jb.epilogue == EPI_FRUSTUM;
                    if ((pass == 0) != isf) continue;
                    float* g = gath + (size_t)li * GB * GATH_STRIDE;
                    if (isf)
                        gather_issue<FS_N>(jb.gt, b, (size_t)HWi, xs_ind + li * K, first, stride, kl0, nb, g, tid, T);
                    else if (jb.epilogue == EPI_DECODE)
                        gather_issue<DS_N>(jb.gt, b, (size_t)HWi, xs_ind + li * K, first, stride, kl0, nb, g, tid, T);
                }
                cp_async_commit();
            }
        }
        cp_async_wait<1>();
        __syncthreads();
        CF_MERGE_PROF(3);
        if (FUSED && late_d) {
            // the decode warps leave now (the caller has them issue the decode job's copies); the others run the frustum
            // epilogue and meet at a barrier of their own (one batch: defer_decode)
            const int Td = T - late_dw * 32;
            if (tid >= Td) return;
            epilogue_frustum(p, p.job[0], b, xs_comp, xs_ind, first, stride, kl0, nb, gath, sdet, sroi, cal, tid, Td);
            CF_MERGE_PROF(7);
            asm volatile("bar.sync 4, %0;" ::"r"(Td) : "memory");
            CF_MERGE_PROF(4);
            return;
        }
        if (FUSED) {
            epilogue_frustum(p, p.job[0], b, xs_comp, xs_ind, first, stride, kl0, nb, gath, sdet, sroi, cal, tid, T);
        } else {
            for (int li = 0; li < n_jobs; ++li) {
                const MergeJob& jb = p.job[li];
                const unsigned long long* comps = xs_comp + li * K;
                if (jb.epilogue == EPI_TOPK)
                    epilogue_topk(p, jb, b, comps, first, stride, kl0, nb, tid, T);
                else if (jb.epilogue == EPI_FRUSTUM)
                    epilogue_frustum(p, jb, b, comps, xs_ind + li * K, first, stride, kl0, nb,
                                     gath + (size_t)li * GB * GATH_STRIDE, sdet, sroi, cal, tid, T);
            }
        }
        CF_MERGE_PROF(7);
        if (!late_d) cp_async_wait<0>();  // (this thread's decode copies: prefetched into L2 long ago, issued a phase ago)
        if (!defer_decode) {
            __syncthreads();
            merge_tail_decode<FUSED>(p, n_jobs, b, first, stride, kl0, nb, xs_comp, gath, GB, tid, T);
        }
        __syncthreads();
        CF_MERGE_PROF(4);
    }
}

}  // namespace cf
