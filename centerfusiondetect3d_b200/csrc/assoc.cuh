// assoc.cuh -- per-ROI nearest-depth radar association of ONE detection by ONE warp, and the paint window.
//
// Replaces (reference, src/lib/utils/pointcloud.py): :424-461 ROI + threshold test + first-minimum search and
// :468-481 the 0.3-scaled box of cvtPcDepthToHeatmap.
#pragma once
#include "common.cuh"
#include "kernels.cuh"

namespace cf {

__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long o = __shfl_xor_sync(FULL, v, off);
        v = o < v ? o : v;
    }
    return v;
}

// ROI + threshold test + first-minimum search (pointcloud.py:424-461) by NT cooperating threads (`t` of NT, NT a
// multiple of 32 or the whole block); the caller reduces the returned key with a minimum.  `dep` = the frame's depth
// plane pc_dep[b,0] in global OR shared memory (generic loads: the fused stage kernel stages it).  Key = (depth bits,
// flat pixel): depths that pass are positive, so the bit pattern orders like the value and the minimum key is
// the first minimum in row-major ROI order.  Lanes run over columns, rows are unrolled: independent loads, no
// division.
// U = row loads in flight per thread: 8 hides shared-memory latency, global (L2) reads want more.
// Three pieces: roi_bounds (the box -> slices), roi_scan_box (the scan of given slices), roi_scan (both).

// ROI of a box: rows [ya,yb) x cols [xa,xb) (python slices of floor / ceil + 1, pointcloud.py:424-433)
__device__ __forceinline__ void roi_bounds(int H, int W, float x1, float y1, float x2, float y2, int& ya, int& yb, int& xa,
                                           int& xb) {
    pyslice((long long)floorf(y1), (long long)ceilf(y2) + 1, H, ya, yb);
    pyslice((long long)floorf(x1), (long long)ceilf(x2) + 1, W, xa, xb);
}

// the scan proper over a given ROI.  W32: the caller is one warp (t < 32, NT == 32) -- the thread -> (row phase, column)
// map then avoids the two integer divisions (this code runs once per detection on a single warp: it is priced by its
// instruction count, and a division is ~40 of them).
template <int U, bool W32>
__device__ __forceinline__ unsigned long long roi_scan_box(const float* dep, int W, int ya, int yb, int xa, int xb,
                                                           float depth, float thr, int t, int NT) {
    const float hi = __fadd_rn(depth, thr);
    float lo = __fsub_rn(depth, thr);
    if (!(lo > 0.0f)) lo = 0.0f;  // python max(0, t)
    unsigned long long best = ~0ull;
    if (ya >= yb || xa >= xb) return best;
    const int w = xb - xa;
    // thread t -> (row phase, column): columns fastest so that a warp reads contiguous runs
    const int cols = w < NT ? w : NT;  // threads along x
    int rows_par, ty;                  // rows processed side by side, this thread's row phase
    if (W32) {
        // floor(32 / cols) and floor(t / cols) for cols in [1,32], t in [0,32): the +0.5 keeps the approximate
        // quotient on the right side of every integer (no multiple of cols lies in (32, 32.5] or (t, t + 0.5])
        const float inv = __fdividef(1.0f, (float)cols);
        rows_par = (int)(32.5f * inv);
        ty = (int)(((float)t + 0.5f) * inv);
    } else {
        rows_par = NT / cols;
        ty = t / cols;
    }
    const int tx = t - ty * cols;
    if (ty >= rows_par) return best;
    for (int xx = xa + tx; xx < xb; xx += cols) {
        for (int y0 = ya + ty; y0 < yb; y0 += U * rows_par) {
            float v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int yy = y0 + u * rows_par;
                const int yc = yy < yb ? yy : ya;  // clamped: the load is unconditional, 0 fails the test below
                const float t_ = dep[(size_t)yc * W + xx];
                v[u] = yy < yb ? t_ : 0.0f;
            }
            // (lo >= 0, so `> lo` already excludes the empty pixels, 0 and -0, and NaN; almost every pixel fails the
            // window: the common path is two compares per pixel, the key is built behind a branch)
            uint32_t hit = 0u;
#pragma unroll
            for (int u = 0; u < U; ++u) hit |= (v[u] > lo && v[u] < hi) ? (1u << u) : 0u;
            if (hit) {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (hit >> u & 1u) {
                        const int yy = y0 + u * rows_par;
                        const unsigned long long key = ((unsigned long long)__float_as_uint(v[u]) << 32) |
                                                       (unsigned long long)(uint32_t)(yy * W + xx);
                        best = key < best ? key : best;
                    }
                }
            }
        }
    }
    return best;
}

// One warp, four pixels per load (float4; W % 4 == 0 and a 16-byte aligned plane required).  The (row, column quad)
// pairs of the ROI are dealt to the lanes as ONE flat sequence, U loads per lane in flight: every lane is busy whatever
// the ROI's width (lanes over columns leave 14 of 32 idle on an 18-quad box), so a scan out of L2 -- which is bound by
// the number of dependent load batches, one L2 latency each -- needs half the batches.  The big boxes set the pace of
// the association (the cluster waits for the slowest warp of its slowest CTA).  Columns of the first / last quad outside
// [xa, xb) are masked; the minimum key is the same as roi_scan_box's whatever the visiting order.
template <int U>
__device__ __forceinline__ unsigned long long roi_scan_box4(const float* dep, int W, int ya, int yb, int xa, int xb,
                                                            float depth, float thr, int lane) {
    const float hi = __fadd_rn(depth, thr);
    float lo = __fsub_rn(depth, thr);
    if (!(lo > 0.0f)) lo = 0.0f;  // python max(0, t)
    unsigned long long best = ~0ull;
    if (ya >= yb || xa >= xb) return best;
    const int qa = xa >> 2, qb = (xb + 3) >> 2;  // column quads [qa, qb)
    const int nq = qb - qa, total = nq * (yb - ya);
    const float inv = 1.0f / (float)nq;
    for (int j0 = 0; j0 < total; j0 += 32 * U) {
        float4 v[U];
        int px[U], x0[U];  // flat pixel index / column of the quad's first pixel; px = -1: no item
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int j = j0 + u * 32 + lane;
            const int jj = j < total ? j : 0;  // (the load is unconditional)
            int r = (int)(((float)jj + 0.5f) * inv);  // jj / nq; exact after the fix-up below
            int q = jj - r * nq;
            if (q < 0) {
                --r;
                q += nq;
            } else if (q >= nq) {
                ++r;
                q -= nq;
            }
            x0[u] = 4 * (qa + q);
            const int p0 = (ya + r) * W + x0[u];
            v[u] = *reinterpret_cast<const float4*>(dep + p0);
            px[u] = j < total ? p0 : -1;
        }
        // (lo >= 0, so `> lo` already excludes the empty pixels and NaN; almost every pixel fails the window: the common
        // path is two compares per pixel, column mask and key are worked out behind a branch)
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const float c4[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
            uint32_t hit = 0u;
#pragma unroll
            for (int c = 0; c < 4; ++c) hit |= (c4[c] > lo && c4[c] < hi) ? (1u << c) : 0u;
            if (hit && px[u] >= 0) {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    if ((hit >> c & 1u) && x0[u] + c >= xa && x0[u] + c < xb) {
                        const unsigned long long key = ((unsigned long long)__float_as_uint(c4[c]) << 32) |
                                                       (unsigned long long)(uint32_t)(px[u] + c);
                        best = key < best ? key : best;
                    }
                }
            }
        }
    }
    return best;
}

template <int U = 8>
__device__ __forceinline__ unsigned long long roi_scan(const float* dep, int H, int W, float x1, float y1,
                                                       float x2, float y2, float depth, float thr, int t, int NT) {
    int ya, yb, xa, xb;
    roi_bounds(H, W, x1, y1, x2, y2, ya, yb, xa, xb);
    return roi_scan_box<U, false>(dep, W, ya, yb, xa, xb, depth, thr, t, NT);
}

// warp minimum of roi_scan keys with two redux rounds instead of five 64-bit shuffle rounds: smallest depth word first,
// then the smallest pixel among the lanes that hold it
__device__ __forceinline__ unsigned long long warp_min_key(unsigned long long v) {
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    const uint32_t mh = __reduce_min_sync(FULL, hi);
    const uint32_t ml = __reduce_min_sync(FULL, hi == mh ? lo : 0xffffffffu);
    return ((unsigned long long)mh << 32) | ml;
}

// paint window of cvtPcDepthToHeatmap (pointcloud.py:468-481): rows h_min:h_max+1, cols w_min:w_max+1+1
__device__ __forceinline__ void paint_window(float x1, float y1, float x2, float y2, int H, int W, int& ra, int& rb,
                                             int& ca, int& cb) {
    const float cx = __fdiv_rn(__fadd_rn(x1, x2), 2.0f), cy = __fdiv_rn(__fadd_rn(y1, y2), 2.0f);
    const float hw_ = __fdiv_rn(__fmul_rn(0.3f, __fsub_rn(x2, x1)), 2.0f);
    const float hh_ = __fdiv_rn(__fmul_rn(0.3f, __fsub_rn(y2, y1)), 2.0f);
    const long long wmin = (long long)__fsub_rn(cx, hw_), wmax = (long long)__fadd_rn(cx, hw_);  // int(): toward 0
    const long long hmin = (long long)__fsub_rn(cy, hh_), hmax = (long long)__fadd_rn(cy, hh_);
    pyslice(hmin, hmax + 1, H, ra, rb);
    pyslice(wmin, wmax + 2, W, ca, cb);
}

// the window into a detection's parameters (any thread, ahead of the association: it depends on the box alone)
__device__ __forceinline__ void det_set_window(DetParam& d, int H, int W) {
    int ra, rb, ca, cb;
    paint_window(d.x1, d.y1, d.x2, d.y2, H, W, ra, rb, ca, cb);
    const bool ok = ra < rb && ca < cb;
    d.ra = ok ? (int16_t)ra : 0;
    d.rb = ok ? (int16_t)rb : 0;
    d.ca = ok ? (int16_t)ca : 0;
    d.cb = ok ? (int16_t)cb : 0;
}

// The matched pixel of one detection -> paint record + association index (one thread).  d carries its paint window.  best = the reduced roi_scan
// key (~0 = no radar pixel in range), dep = the frame's pc_dep (3,H,W) in global memory (velocity channels).
__device__ __forceinline__ PaintRec assoc_record(unsigned long long best, float v1, float v2, const DetParam& d,
                                                 float max_pc_dist, int32_t* assoc_out) {
    PaintRec r;
    r.ra = r.rb = r.ca = r.cb = 0;
    r.v0 = r.v1 = r.v2 = 0.0f;
    r.pad_ = r.pad2_ = 0;
    int32_t a = -1;
    if (best != ~0ull) {
        a = (int32_t)(uint32_t)best;
        r.v0 = __fdiv_rn(__uint_as_float((uint32_t)(best >> 32)), max_pc_dist);  // :462-466
        r.v1 = v1;
        r.v2 = v2;
        r.ra = d.ra;
        r.rb = d.rb;
        r.ca = d.ca;
        r.cb = d.cb;
    }
    if (assoc_out) *assoc_out = a;
    return r;
}
// ... with the velocity reads (dep = the frame's pc_dep (3,H,W), planes 1 and 2)
__device__ __forceinline__ PaintRec assoc_finalize(unsigned long long best, const float* dep, int H, int W,
                                                   const DetParam& d, float max_pc_dist, int32_t* assoc_out) {
    const size_t HW = (size_t)H * W;
    float v1 = 0.0f, v2 = 0.0f;
    if (best != ~0ull) {
        const uint32_t mp = (uint32_t)best;
        v1 = dep[HW + mp];
        v2 = dep[2 * HW + mp];
    }
    return assoc_record(best, v1, v2, d, max_pc_dist, assoc_out);
}

}  // namespace cf
