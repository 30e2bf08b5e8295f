// merge.cuh -- per-frame merge of the peak-decode survivors: exact sorted top-K under the total order
// (score desc, class asc, index asc) + the topk / frustum / decode epilogue.  Device code, executed by ONE CTA of
// T threads per (frame, list): the tail of the filter pass (last CTA of a frame to finish).
//
// Replaces (reference, src/lib/): model/utils.py:6-38 topk (second level), :41-72 transposeAndGetFeature; the
// gathers + box / alpha / threshold block of utils/pointcloud.py:347-380; model/decode.py:47-172.
#pragma once
#include <math_constants.h>

#include "geometry.cuh"
#include "kernels.cuh"
#include "select.cuh"

namespace cf {

constexpr int MERGE_WIN = 1024;  // survivors ranked directly; more go through block_select first


template <int T>
struct ListSrc {
    const unsigned long long* cand;
    int n_in, n_round;
    template <class F>
    __device__ __forceinline__ void for_each(F&& f) const {
        for (int i = threadIdx.x; i < n_round; i += T) f(i < n_in ? __ldcg(cand + i) : 0ull);
    }
};

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
template <int T>
__device__ __forceinline__ void copy_head(const float* head, float* out, int nch, int b, int K, size_t HW,
                                          const int* sind) {
    if (!head) return;
    for (int t = threadIdx.x; t < K * nch; t += T) {
        const int k = t / nch, ch = t - k * nch;
        out[((size_t)b * K + k) * nch + ch] = __ldg(head + ((size_t)b * nch + ch) * HW + sind[k]);
    }
}

// shared memory of merge_frame (static, one instance per CTA)
struct MergeSmem {
    SelScratch sc;
    unsigned long long win[MERGE_WIN];
    unsigned long long sorted[CF_MAX_K];
    int sind[CF_MAX_K];
};

template <int T>
__device__ __forceinline__ void merge_frame(const MergeParams& p, MergeSmem& sm, int b, int li) {
    const int tid = threadIdx.x;
    const MergeJob& jb = p.job[li];
    const int list = p.list[li];
    const int K = p.K;
    const int HWi = p.H * p.W;
    const uint32_t n = __ldcg(p.ctr + ((size_t)b * 2 + list) * CTR_STRIDE);  // survivors appended (>= K by construction)
    const unsigned long long* cand = p.cand + ((size_t)b * 2 + list) * CAND_CAP;

    int nw;  // entries of win[] to rank
    if (n <= (uint32_t)MERGE_WIN) {
        for (int t = tid; t < MERGE_WIN; t += T) sm.win[t] = t < (int)n ? __ldcg(cand + t) : 0ull;
        nw = (int)n;
        __syncthreads();
    } else {
        if (n <= (uint32_t)CAND_CAP) {
            ListSrc<T> src;
            src.cand = cand;
            src.n_in = (int)n;
            src.n_round = (src.n_in + T - 1) / T * T;
            block_select<T, 1>(src, K, sm.sc, [&](uint32_t slot, unsigned long long c) { sm.win[slot] = c; });
        } else {
            FrameSrc<T> src;
            src.frame = p.heat + (size_t)b * p.C * HWi;
            src.CH = p.C * p.H;
            src.H = p.H;
            src.W = p.W;
            src.n_round = (p.C * HWi + T - 1) / T * T;
            src.nms = list == 1;
            block_select<T, 1>(src, K, sm.sc, [&](uint32_t slot, unsigned long long c) { sm.win[slot] = c; });
        }
        nw = K;
        __syncthreads();
    }

    // rank by counting (composites are distinct: ranks are a permutation); the K best land in sorted[]
    for (int t = tid; t < nw; t += T) {
        const unsigned long long me = sm.win[t];
        int rank = 0;
        for (int j = 0; j < nw; ++j) rank += (sm.win[j] > me) ? 1 : 0;
        if (rank < K) sm.sorted[rank] = me;
    }
    __syncthreads();

    for (int k = tid; k < K; k += T) {
        const unsigned long long c = sm.sorted[k];
        const uint32_t g = comp_gidx(c), key = comp_key(c);
        const int cls = (int)(g / (uint32_t)HWi);
        const int ind = (int)(g - (uint32_t)cls * (uint32_t)HWi);
        const int y = ind / p.W, x = ind - y * p.W;
        sm.sind[k] = ind;
        float score = key2f(key);
        if (key == KEY_ZERO) score = __fmul_rn(__ldg(p.heat + ((size_t)b * p.C + cls) * HWi + ind), 0.0f);
        const size_t o = (size_t)b * K + k;
        if (jb.epilogue == EPI_TOPK) {
            if (jb.topk.scores) jb.topk.scores[o] = score;
            if (jb.topk.inds) jb.topk.inds[o] = ind;
            if (jb.topk.classes) jb.topk.classes[o] = cls;
            if (jb.topk.ys) jb.topk.ys[o] = y;
            if (jb.topk.xs) jb.topk.xs[o] = x;
        } else if (jb.epilogue == EPI_FRUSTUM) {
            epilogue_frustum(p, jb, b, k, ind, x, y);
        } else if (jb.epilogue == EPI_DECODE) {
            epilogue_decode(p, jb, b, k, score, cls, ind, x, y);
        }
    }
    if (jb.epilogue == EPI_DECODE) {
        __syncthreads();
        const size_t HW = (size_t)HWi;
        copy_head<T>(jb.dec.heads.rot, jb.dec.out.rot, 8, b, K, HW, sm.sind);
        copy_head<T>(jb.dec.heads.dim, jb.dec.out.dim, 3, b, K, HW, sm.sind);
        copy_head<T>(jb.dec.heads.nuscenes_att, jb.dec.out.nuscenes_att, 8, b, K, HW, sm.sind);
        copy_head<T>(jb.dec.heads.velocity, jb.dec.out.velocity, 3, b, K, HW, sm.sind);
    }
    __syncthreads();  // sm is reused by the next job of this CTA
}

}  // namespace cf
