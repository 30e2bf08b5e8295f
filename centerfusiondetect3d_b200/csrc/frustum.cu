// frustum.cu -- per-ROI nearest-depth radar association and the ordered, atomic-free-in-HBM paint of the
// pc_hm / pc_dep maps; plus the small stand-alone geometry kernels.
//
// Replaces (reference, src/lib/): utils/pointcloud.py:383-392 (the B x K Python loop) and :397-481
// cvtPcDepthToHeatmap; :195-211 get_alpha; :299-328 getDistanceThresh; the pillar loop of
// dataset/generic_dataset.py:798-826 with dataset/datasets/nuscenes.py:234-263 drawPcHeat.
//
// The reference paints detection after detection, later ones overwriting earlier ones.  Equivalent
// parallel form used here: association per (frame, detection) is independent (one warp each, warp-shuffle
// lexicographic (depth, pixel) minimum over the ROI); the paint is "every pixel takes the covering record
// with the LARGEST index", resolved in shared memory (atomicMax on a per-band owner map), after which
// every output pixel is written to HBM exactly once, coalesced, zeros included.
#include <math_constants.h>

#include <cstring>
#include <map>
#include <mutex>

#include "assoc.cuh"
#include "geometry.cuh"
#include "kernels.cuh"
#include "paint.cuh"

namespace cf {

// -----------------------------------------------------------------------------------------------------
// ordered rectangle paint: out (B,3,H,W) written completely; largest record index wins per pixel
//
// Records in global memory: thousands of radar pillars (drawPcHeat under the pillar loop, nuscenes.py:234-263 /
// generic_dataset.py:798-826) or a few dozen boxes (row f4).  Pillars overlap an order of magnitude deep (2000 pillars
// cover 87 % of a 112x200 map eleven times over), so visiting every pixel of every rectangle is the wrong amount of work.
// Here a CTA owns a strip of 32 columns x 128 rows and each of its four WARPS 8 of those columns over all the rows: lane
// = (group of 32 rows, column), with a 32-bit mask of the lane's rows that already have an owner.  (Pillars are narrow and
// tall: a warp that is narrow and tall meets a third of the rectangles a 32 x 32 tile does, and no tile row is hotter
// than another.)  The CTA first stages the frame's rectangles in shared memory -- ONE pass for up to 2048 records, kept
// in order, only those that touch the strip (a fifth) -- then every warp walks them from the LARGEST index down, 32 at a
// time: each lane tests one record against the warp's columns, a ballot keeps the few that touch them, and for those,
// two per trip, every lane takes `rows of the rectangle inside its group AND NOT filled` -- three shuffles and a handful
// of bit operations per (record, warp) whatever the rectangle's height.  Every pixel gets its owner exactly once; no
// atomics; a warp whose pixels are all owned stops early.  The kernel is bound by the length of a warp's dependent
// instruction chain (~8 cycles per instruction with 3-4 warps per scheduler), not by memory: tools/prof_paint.py.
// (Measured alternatives, all slower on the dense-radar workload: a thread or a warp per rectangle with atomicMax on an
// owner map, 32 x 32 tiles, their records split over four warps with a priority merge, and column-binned lists with a
// sort -- they lose to lane divergence on the few very wide near pillars.)
// -----------------------------------------------------------------------------------------------------
#ifdef CFD3D_PROFILE
__device__ long long g_paint_prof[4096 * 16];
#define PP(i) do { if (lane == 0 && wid < 4096) g_paint_prof[wid * 16 + (i)] = clock64(); } while (0)
#define PPV(i, v) do { if (lane == 0 && wid < 4096) g_paint_prof[wid * 16 + (i)] = (v); } while (0)
#else
#define PP(i) do { } while (0)
#define PPV(i, v) do { } while (0)
#endif
constexpr int PT = 128;              // threads per CTA: a strip of 32 columns x up to 128 rows
constexpr int PAINT_LU = 16;         // records a thread examines per pass (2048 per pass: a frame's pillars in one)
constexpr int PAINT_REC_CAP = PAINT_LU * PT;  // records examined per pass; those that touch the CTA's strip are staged
                                              // in shared memory (10 bytes each; the pair lists are dynamic shared memory:
                                              // 53 KB per CTA together, four CTAs per SM)
constexpr int PAINT_ROWS = 128;      // rows per CTA: 4 lane groups of 32 rows

// rects (optional): the records' rectangles once more as a dense array of 8-byte entries (same indexing as recs).  Every
// strip of a frame examines ALL the frame's rectangles: out of 32-byte records that is a sector per rectangle and CTA
// (7 x 64 x 2000 sectors at the dense-radar workload: the L2's request rate sets the pace), out of the dense array a quarter.
__global__ void __launch_bounds__(PT) paint_kernel(const PaintRec* __restrict__ recs, const short4* __restrict__ rects,
                                                   int n_max, int rec_stride,
                                                   const int32_t* __restrict__ counts, int H, int W, int tiles_x,
                                                   float* __restrict__ out, int32_t* __restrict__ owner_out) {
    __shared__ short4 srect[PAINT_REC_CAP];
    __shared__ uint16_t sidx[PAINT_REC_CAP];  // record index of a staged rectangle, relative to the pass's first
    static_assert(PAINT_REC_CAP <= 65536, "16-bit pass-relative indices");
    __shared__ int swsum[PT / 32];
    // per warp and lane: the (record, newly owned rows) pairs of the lane's 32 pixels, at most 32 (a pair owns >= 1 row)
    extern __shared__ int2 spair_dyn[];
    int2 (*spair)[32][32] = reinterpret_cast<int2 (*)[32][32]>(spair_dyn);
    // programmatic dependent launch: placed while the kernel that produces the records drains; the next kernel of the
    // stream (the stage kernel) may be placed as this one retires
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y;
    const int wid = ((int)blockIdx.y * (int)gridDim.x + (int)blockIdx.x) * 4 + warp;
    (void)wid;
    long long t_stage = 0, t_walk = 0, t_mark;
    int n_it = 0;
    PP(0);
    { long long gt_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_)); PPV(8, gt_); }
    // blockIdx.x = (block of 128 rows, strip of 32 columns); a WARP owns 8 columns of the strip over all the rows: lane =
    // (group of 32 rows, column)
    const int tx = (int)blockIdx.x % tiles_x, Y0 = ((int)blockIdx.x / tiles_x) * PAINT_ROWS;
    const int X0 = tx * 32, xw = X0 + 8 * warp;  // the strip's / this warp's first column
    const int x = xw + (lane & 7), y0 = Y0 + 32 * (lane >> 3);  // this lane's column and first row
    int n = n_max;
    if (counts) n = min(n_max, counts[b]);
    const PaintRec* fr = recs + (size_t)b * rec_stride;
    const short4* frect = rects ? rects + (size_t)b * rec_stride : nullptr;
    const size_t HW = (size_t)H * W;
    // rows beyond the map count as filled
    uint32_t filled = (y0 + 32 <= H) ? 0u : (y0 >= H ? ~0u : ~((1u << (H - y0)) - 1u));
    if (x >= W) filled = ~0u;
    int np = 0;  // pairs of this lane
    // records from the largest index down, PAINT_REC_CAP at a time through shared memory
    for (int hi = n; hi > 0; hi -= PAINT_REC_CAP) {
        const int lo = max(0, hi - PAINT_REC_CAP);
        int cnt;  // records of this pass that touch the CTA's strip
        __syncthreads();
        t_mark = clock64();
        {  // thread t examines records lo + t * LU .. + LU - 1 (all loads in flight together: one memory round trip per
           // pass), keeps those that touch the strip, and writes them behind those of the threads before it
            constexpr int LU = PAINT_LU;
            short4 q[LU];
            uint32_t keep = 0u;
#pragma unroll
            for (int u = 0; u < LU; ++u) {
                const int i = lo + tid * LU + u;
                q[u] = i < hi ? (frect ? __ldg(frect + i) : __ldg(reinterpret_cast<const short4*>(fr + i)))
                              : make_short4(0, 0, 0, 0);
                if (q[u].y > q[u].x && q[u].w > q[u].z && q[u].w > X0 && q[u].z < X0 + 32 && q[u].y > Y0 &&
                    q[u].x < Y0 + PAINT_ROWS)
                    keep |= 1u << u;
            }
            const int mine = __popc(keep);
            int incl = mine;  // inclusive scan over the warp, then over the warps
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_up_sync(FULL, incl, off);
                if (lane >= off) incl += o;
            }
            if (lane == 31) swsum[warp] = incl;
            __syncthreads();
            int pos = incl - mine;
            cnt = 0;
#pragma unroll
            for (int w = 0; w < PT / 32; ++w) {
                if (w < warp) pos += swsum[w];
                cnt += swsum[w];
            }
#pragma unroll
            for (int u = 0; u < LU; ++u) {
                if (keep >> u & 1u) {
                    srect[pos] = q[u];
                    sidx[pos] = (uint16_t)(tid * LU + u);
                    ++pos;
                }
            }
        }
        __syncthreads();
        t_stage += clock64() - t_mark;
        t_mark = clock64();
        if (__all_sync(FULL, filled == ~0u)) continue;  // (still takes part in the barriers above)
        for (int base = cnt - 1; base >= 0; base -= 32) {
            // lane-parallel: this lane's record of the step (descending index over the lanes) and whether it touches
            // the warp's 8 columns
            const int i = base - lane;
            bool rel = false;
            uint32_t rows_ab = 0u, cols_ab = 0u;
            int ridx = 0;
            if (i >= 0) {
                const short4 q = srect[i];
                ridx = lo + (int)sidx[i];
                rel = q.w > xw && q.z < xw + 8;
                rows_ab = (uint32_t)(uint16_t)q.x | ((uint32_t)(uint16_t)q.y << 16);
                cols_ab = (uint32_t)(uint16_t)q.z | ((uint32_t)(uint16_t)q.w << 16);
            }
            uint32_t m = __ballot_sync(FULL, rel);
            // serial over the step's relevant records, largest index first, two per trip (their six shuffles are in
            // flight together): every lane works out the rectangle's rows inside its own group of 32 and takes those
            // that have no owner yet
            auto take = [&](uint32_t rab, uint32_t cab, int idx) {
                const int a = max((int)(rab & 0xffffu) - y0, 0), e = min((int)(rab >> 16) - y0, 32);
                uint32_t rows = 0u;
                if (e > a && x >= (int)(cab & 0xffffu) && x < (int)(cab >> 16))
                    rows = (e >= 32 ? ~0u : ((1u << e) - 1u)) & ~((1u << a) - 1u) & ~filled;
                if (rows) {
                    filled |= rows;
                    spair[warp][np][lane] = make_int2(idx, (int)rows);
                    ++np;
                }
            };
            while (m) {
                ++n_it;
                const int s0 = __ffs(m) - 1;
                m &= m - 1;
                const bool two = m != 0u;
                const int s1 = two ? __ffs(m) - 1 : s0;
                m &= m - 1;
                const uint32_t rab0 = __shfl_sync(FULL, rows_ab, s0), cab0 = __shfl_sync(FULL, cols_ab, s0);
                const int id0 = __shfl_sync(FULL, ridx, s0);
                const uint32_t rab1 = __shfl_sync(FULL, rows_ab, s1), cab1 = __shfl_sync(FULL, cols_ab, s1);
                const int id1 = __shfl_sync(FULL, ridx, s1);
                take(rab0, cab0, id0);
                if (two) take(rab1, cab1, id1);
            }
            if (__all_sync(FULL, filled == ~0u)) break;
        }
        t_walk += clock64() - t_mark;
    }
    __syncwarp();
    PP(1);
    PPV(4, t_stage);
    PPV(5, t_walk);
    PPV(6, n_it);
    // pairs -> owner per row, every lane for its own 32 pixels (each row belongs to at most one pair)
    int own[32];
#pragma unroll
    for (int r = 0; r < 32; ++r) own[r] = -1;
    for (int j = 0; j < np; ++j) {
        const int2 pr = spair[warp][j][lane];
#pragma unroll
        for (int r = 0; r < 32; ++r)
            if (((uint32_t)pr.y >> r) & 1u) own[r] = pr.x;
    }
    PP(2);
    // the lane's pixels, each written once: 8 consecutive columns (one 32-byte sector per plane) per row group and store
    float* o = out + (size_t)b * 3 * HW;
    int32_t* oo = owner_out ? owner_out + (size_t)b * HW : nullptr;
    const int rows_here = min(32, H - y0);
    if (x < W && rows_here > 0) {
        constexpr int RU = 8;  // rows per step: their 3 x RU value loads are in flight together (one L2 round trip)
#pragma unroll
        for (int r0 = 0; r0 < 32; r0 += RU) {
            float v[RU][3];
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                v[u][0] = v[u][1] = v[u][2] = 0.0f;
                if (r0 + u < rows_here && own[r0 + u] >= 0) {
                    const PaintRec* rr = fr + own[r0 + u];
                    v[u][0] = __ldg(&rr->v0);
                    v[u][1] = __ldg(&rr->v1);
                    v[u][2] = __ldg(&rr->v2);
                }
            }
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                if (r0 + u < rows_here) {
                    const size_t px = (size_t)(y0 + r0 + u) * W + x;
                    o[px] = v[u][0];
                    o[HW + px] = v[u][1];
                    o[2 * HW + px] = v[u][2];
                    if (oo) oo[px] = own[r0 + u];
                }
            }
        }
    }
    PP(3);
    { long long gt_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_)); PPV(9, gt_); }
}

constexpr size_t PAINT_DYN_SMEM = sizeof(int2) * (PT / 32) * 32 * 32;

// static + dynamic shared memory exceed 48 KB: the opt-in is process-wide state of the runtime, set once per device
static cudaError_t paint_setup() {
    static std::mutex mu;
    static std::map<int, cudaError_t> done;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> g(mu);
    auto f = done.find(dev);
    if (f != done.end()) return f->second;
    e = cudaFuncSetAttribute(paint_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PAINT_DYN_SMEM);
    done[dev] = e;
    return e;
}

cudaError_t launch_paint(const PaintRec* recs, const short4* rects, int n_per_frame, int rec_stride, const int32_t* counts,
                         int B, int H, int W, float* out, int32_t* owner, cudaStream_t st, int* launches) {
    if (cudaError_t es = paint_setup()) return es;
    const int tiles_x = (W + 31) / 32;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(tiles_x * ((H + PAINT_ROWS - 1) / PAINT_ROWS), B);
    cfg.blockDim = dim3(PT);
    cfg.dynamicSmemBytes = PAINT_DYN_SMEM;
    cfg.stream = st;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, paint_kernel, recs, rects, n_per_frame, rec_stride, counts, H, W, tiles_x, out, owner);
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// f4: frustum association of GIVEN boxes (training targets): the per-annotation loop of
// dataset/generic_dataset.py:673-687, cvtPcDepthToHeatmap per box in list order.  One warp per box, then the
// ordered paint (later boxes overwrite earlier ones, like the reference's sequential slice writes).
// -----------------------------------------------------------------------------------------------------
constexpr int BW = 8;  // warps (boxes) per CTA

__global__ void __launch_bounds__(BW * 32) box_assoc_kernel(const float* __restrict__ pc_dep, const float* __restrict__ bboxes,
                                                            const float* __restrict__ depth, const float* __restrict__ thr,
                                                            const int32_t* __restrict__ counts, int N, int H, int W,
                                                            float max_pc_dist, PaintRec* __restrict__ recs,
                                                            int32_t* __restrict__ assoc) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.x * BW + warp, b = blockIdx.y;
    if (n >= N) return;
    const size_t o = (size_t)b * N + n;
    const float* dep = pc_dep + (size_t)b * 3 * H * W;
    unsigned long long best = ~0ull;
    DetParam d;
    d.x1 = d.y1 = d.x2 = d.y2 = d.depth = d.thr = 0.0f;
    d.ra = d.rb = d.ca = d.cb = 0;
    const bool live = !counts || n < counts[b];
    if (live) {
        d.x1 = __ldg(bboxes + 4 * o);
        d.y1 = __ldg(bboxes + 4 * o + 1);
        d.x2 = __ldg(bboxes + 4 * o + 2);
        d.y2 = __ldg(bboxes + 4 * o + 3);
        d.depth = __ldg(depth + o);
        d.thr = __ldg(thr + o);
        best = warp_min_u64(roi_scan<16>(dep, H, W, d.x1, d.y1, d.x2, d.y2, d.depth, d.thr, lane, 32));
        det_set_window(d, H, W);
    }
    if (lane == 0) recs[o] = assoc_finalize(best, dep, H, W, d, max_pc_dist, assoc ? assoc + o : nullptr);
}

cudaError_t launch_box_assoc(const float* pc_dep, const float* bboxes, const float* depth, const float* thr,
                             const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, PaintRec* recs,
                             int32_t* assoc, cudaStream_t st, int* launches) {
    if (B == 0 || N == 0) return cudaSuccess;
    box_assoc_kernel<<<dim3((N + BW - 1) / BW, B), BW * 32, 0, st>>>(pc_dep, bboxes, depth, thr, counts, N, H, W,
                                                                      max_pc_dist, recs, assoc);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// f4 on the reference's own branch: the NumPy form of cvtPcDepthToHeatmap as the DataLoader calls it
// (dataset/generic_dataset.py:673-687, utils/pointcloud.py:411-481 with lib = np) under NumPy >= 2 promotion.  Boxes,
// depths and thresholds arrive as float64 (the exact container of either dtype); `flags` says which arithmetic the
// caller's dtypes select: CF_BOX_F32 -- the box was float32 (transformBbox, generic_dataset.py:507-526): centre, 0.3 * w
// and the int() of the paint window are float32 operations, else the centre is rounded to float32 and the rest is float64;
// CF_THR_F32 -- the threshold was np.float32 (what getDistanceThresh's NumPy branch returns): depth +- thr and the two
// strict compares are float32, else float64 on the exactly widened pixel.  One warp per box.
__global__ void __launch_bounds__(BW * 32) box_assoc_np_kernel(const float* __restrict__ pc_dep, const double* __restrict__ bboxes,
                                                               const double* __restrict__ depth, const double* __restrict__ thr,
                                                               const int32_t* __restrict__ counts, int N, int H, int W,
                                                               float max_pc_dist, int flags, PaintRec* __restrict__ recs,
                                                               int32_t* __restrict__ assoc) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.x * BW + warp, b = blockIdx.y;
    if (n >= N) return;
    const size_t o = (size_t)b * N + n;
    const float* dep = pc_dep + (size_t)b * 3 * H * W;
    unsigned long long best = ~0ull;
    DetParam d;
    d.x1 = d.y1 = d.x2 = d.y2 = d.depth = d.thr = 0.0f;
    d.ra = d.rb = d.ca = d.cb = 0;
    const bool live = !counts || n < counts[b];
    if (live) {
        const double x1 = __ldg(bboxes + 4 * o), y1 = __ldg(bboxes + 4 * o + 1), x2 = __ldg(bboxes + 4 * o + 2),
                     y2 = __ldg(bboxes + 4 * o + 3);
        const double dp = __ldg(depth + o), th = __ldg(thr + o);
        int ya, yb, xa, xb;
        pyslice((long long)floor(y1), (long long)ceil(y2) + 1, H, ya, yb);
        pyslice((long long)floor(x1), (long long)ceil(x2) + 1, W, xa, xb);
        double hi = __dadd_rn(dp, th), lo = __dsub_rn(dp, th);
        if (flags & CF_THR_F32) {
            hi = (double)__fadd_rn((float)dp, (float)th);
            lo = (double)__fsub_rn((float)dp, (float)th);
        }
        if (!(lo > 0.0)) lo = 0.0;  // python max(0, t)
        if (ya < yb && xa < xb) {
            const int w = xb - xa, cnt = (yb - ya) * w;
            for (int i = lane; i < cnt; i += 32) {
                const int ry = i / w, yy = ya + ry, xx = xa + (i - ry * w);
                const float v = dep[(size_t)yy * W + xx];
                if (v != 0.0f && (double)v < hi && (double)v > lo) {
                    const unsigned long long key =
                        ((unsigned long long)__float_as_uint(v) << 32) | (unsigned long long)(uint32_t)(yy * W + xx);
                    best = key < best ? key : best;
                }
            }
        }
        best = warp_min_u64(best);
        int ra, rb, ca, cb;
        if (flags & CF_BOX_F32) {
            paint_window((float)x1, (float)y1, (float)x2, (float)y2, H, W, ra, rb, ca, cb);
        } else {
            const float cx = (float)__ddiv_rn(__dadd_rn(x1, x2), 2.0), cy = (float)__ddiv_rn(__dadd_rn(y1, y2), 2.0);
            const double hw_ = __ddiv_rn(__dmul_rn(0.3, __dsub_rn(x2, x1)), 2.0);
            const double hh_ = __ddiv_rn(__dmul_rn(0.3, __dsub_rn(y2, y1)), 2.0);
            const long long wmin = (long long)__dsub_rn((double)cx, hw_), wmax = (long long)__dadd_rn((double)cx, hw_);
            const long long hmin = (long long)__dsub_rn((double)cy, hh_), hmax = (long long)__dadd_rn((double)cy, hh_);
            pyslice(hmin, hmax + 1, H, ra, rb);
            pyslice(wmin, wmax + 2, W, ca, cb);
        }
        const bool ok = ra < rb && ca < cb;
        d.ra = ok ? (int16_t)ra : 0;
        d.rb = ok ? (int16_t)rb : 0;
        d.ca = ok ? (int16_t)ca : 0;
        d.cb = ok ? (int16_t)cb : 0;
    }
    if (lane == 0) recs[o] = assoc_finalize(best, dep, H, W, d, max_pc_dist, assoc ? assoc + o : nullptr);
}

cudaError_t launch_box_assoc_np(const float* pc_dep, const double* bboxes, const double* depth, const double* thr,
                                const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, int flags,
                                PaintRec* recs, int32_t* assoc, cudaStream_t st, int* launches) {
    if (B == 0 || N == 0) return cudaSuccess;
    box_assoc_np_kernel<<<dim3((N + BW - 1) / BW, B), BW * 32, 0, st>>>(pc_dep, bboxes, depth, thr, counts, N, H, W,
                                                                         max_pc_dist, flags, recs, assoc);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// cvtPcDepthToHeatmap, one detection, in place                      reference pointcloud.py:397-481
// -----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pc2hm_single_kernel(float* __restrict__ pc_hm, const float* __restrict__ pc_dep,
                                                           int H, int W, const float* __restrict__ depth,
                                                           const float* __restrict__ bbox,
                                                           const float* __restrict__ thr, float max_pc_dist,
                                                           int32_t* __restrict__ assoc) {
    __shared__ unsigned long long wbest[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t HW = (size_t)H * W;
    const float x1 = bbox[0], y1 = bbox[1], x2 = bbox[2], y2 = bbox[3];
    unsigned long long best = warp_min_u64(roi_scan(pc_dep, H, W, x1, y1, x2, y2, depth[0], thr[0], tid, 256));
    if (lane == 0) wbest[warp] = best;
    __syncthreads();
    best = wbest[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) best = wbest[i] < best ? wbest[i] : best;
    if (best == ~0ull) {
        if (tid == 0 && assoc) assoc[0] = -1;
        return;
    }
    const uint32_t mp = (uint32_t)best;
    if (tid == 0 && assoc) assoc[0] = (int32_t)mp;
    const float v0 = __fdiv_rn(__uint_as_float((uint32_t)(best >> 32)), max_pc_dist);
    const float v1 = pc_dep[HW + mp], v2 = pc_dep[2 * HW + mp];
    int ra, rb, ca, cb;
    paint_window(x1, y1, x2, y2, H, W, ra, rb, ca, cb);
    if (ra >= rb || ca >= cb) return;
    const int w = cb - ca, n = (rb - ra) * w;
    for (int i = tid; i < n; i += 256) {
        const int ry = i / w;
        const size_t p = (size_t)(ra + ry) * W + ca + (i - ry * w);
        pc_hm[p] = v0;
        pc_hm[HW + p] = v1;
        pc_hm[2 * HW + p] = v2;
    }
}

cudaError_t launch_pc2hm_single(float* pc_hm, const float* pc_dep, int H, int W, const float* depth,
                                const float* bbox, const float* thr, float max_pc_dist, int32_t* assoc,
                                cudaStream_t st, int* launches) {
    pc2hm_single_kernel<<<1, 256, 0, st>>>(pc_hm, pc_dep, H, W, depth, bbox, thr, max_pc_dist, assoc);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// get_alpha / getDistanceThresh as stand-alone ops
// -----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) alpha_kernel(const float* __restrict__ rot, long long n,
                                                    float* __restrict__ alpha) {
    const long long i = blockIdx.x * 128ll + threadIdx.x;
    if (i >= n) return;
    float r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = rot[8 * i + j];
    alpha[i] = get_alpha(r);
}

cudaError_t launch_alpha(const float* rot, int64_t n, float* alpha, cudaStream_t st, int* launches) {
    if (n == 0) return cudaSuccess;
    alpha_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(rot, (long long)n, alpha);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

__global__ void __launch_bounds__(128) dist_thresh_kernel(const float* __restrict__ calib,
                                                          const float* __restrict__ center,
                                                          const float* __restrict__ dim,
                                                          const float* __restrict__ alpha, int B, int K,
                                                          float* __restrict__ thr) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= B * K) return;
    const int b = i / K;
    float cal[3] = {calib[12 * b + 0], calib[12 * b + 1], calib[12 * b + 2]};
    float dm[3] = {dim[3 * i], dim[3 * i + 1], dim[3 * i + 2]};
    thr[i] = dist_thresh_one(cal, center[2 * i], dm, alpha[i]);
}

cudaError_t launch_dist_thresh(const float* calib, const float* center, const float* dim, const float* alpha, int B,
                               int K, float* thr, cudaStream_t st, int* launches) {
    if (B * K == 0) return cudaSuccess;
    dist_thresh_kernel<<<(B * K + 127) / 128, 128, 0, st>>>(calib, center, dim, alpha, B, K, thr);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// f2: radar ingestion on the device -- the host-side steps of loadRadarPointCloud between the pickle and
// processPointCloud (dataset/datasets/nuscenes.py:159-205; detector.py:255-283):
//   * layout: the (rows >= 10, N) float64 array convert_nuScenes.py stores (rows 0-2 x, y, z in the camera frame,
//     rows 8, 9 the compensated velocities vx, vz) -> [X, Y, Z, vx, vz] rows (rows == 0: the input already has them);
//   * stable depth sort.  The reference sorts with np.argsort (nuscenes.py:195-197, detector.py:281), whose default
//     kind is not stable; the pillar paint is last-writer-wins, so the order is part of the result.  Here it is fixed:
//     (Z, source index) ascending == np.argsort(kind="stable"); `descending` = that order reversed, index[::-1] of
//     nuscenes.py:196-197 (PC_REVERSE false);
//   * flip augmentation (nuscenes.py:201-205): X and vx change sign (the projected u is mirrored by the pillar kernel,
//     which gets the same flags).
// The distance filter, the z offset and the in-image masks stay in pillar_rects_kernel: they drop points without
// changing the order of the rest.  One CTA per frame, rank by counting with the depths in shared memory.
// -----------------------------------------------------------------------------------------------------
constexpr int RS_T = 1024;

__global__ void __launch_bounds__(RS_T) radar_sort_kernel(const double* __restrict__ points, const int32_t* __restrict__ counts,
                                                          int rows, int Nmax, const int32_t* __restrict__ flip,
                                                          int descending, double* __restrict__ out,
                                                          int32_t* __restrict__ order) {
    extern __shared__ double sz[];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n = min(Nmax, max(0, counts[b]));
    const bool raw = rows > 0;
    const double* fp = points + (size_t)b * Nmax * (raw ? rows : 5);
    // element c in {X, Y, Z, vx, vz} of point i
    auto at = [&](int i, int c) { return raw ? fp[(size_t)(c < 3 ? c : c + 5) * Nmax + i] : fp[(size_t)i * 5 + c]; };
    double* fo = out + (size_t)b * Nmax * 5;
    const double sgn = (flip && flip[b]) ? -1.0 : 1.0;
    for (int i = tid; i < n; i += RS_T) sz[i] = at(i, 2);
    __syncthreads();
    for (int i = tid; i < Nmax; i += RS_T) {
        int rank = i;  // padding rows keep their place
        if (i < n) {
            const double zi = sz[i];
            int r0 = 0, r1 = 0;
            int j = 0;
            for (; j + 2 <= n; j += 2) {
                const double a = sz[j], c = sz[j + 1];
                r0 += (a < zi || (a == zi && j < i)) ? 1 : 0;
                r1 += (c < zi || (c == zi && j + 1 < i)) ? 1 : 0;
            }
            if (j < n) r0 += (sz[j] < zi || (sz[j] == zi && j < i)) ? 1 : 0;
            rank = r0 + r1;
            if (descending) rank = n - 1 - rank;
        }
        double* q = fo + (size_t)rank * 5;
        q[0] = i < n ? sgn * at(i, 0) : at(i, 0);
        q[1] = at(i, 1);
        q[2] = at(i, 2);
        q[3] = i < n ? sgn * at(i, 3) : at(i, 3);
        q[4] = at(i, 4);
        if (order) order[(size_t)b * Nmax + rank] = i;
    }
}

cudaError_t launch_radar_sort(const double* points, const int32_t* counts, int B, int rows, int Nmax, const int32_t* flip,
                              int descending, double* out, int32_t* order, cudaStream_t st, int* launches) {
    if (B == 0 || Nmax == 0) return cudaSuccess;
    const size_t smem = (size_t)Nmax * sizeof(double);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(radar_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    radar_sort_kernel<<<B, RS_T, smem, st>>>(points, counts, rows, Nmax, flip, descending, out, order);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// radar pillar rectangles: one thread per point, fp64 exactly like the NumPy reference
//   loadRadarPointCloud filters   dataset/datasets/nuscenes.py:172-205
//   map_pointcloud_to_image       utils/pointcloud.py:17-49 (+ devkit view_points)
//   transformPointCloud           dataset/generic_dataset.py:830-867
//   getPcPillarsSize              dataset/generic_dataset.py:869-942 (get3dBox / project3DPoints, ddd.py:8-55)
//   pillar box + np.round         dataset/generic_dataset.py:798-826
// -----------------------------------------------------------------------------------------------------
struct PillarParams {
    const double *points, *cam, *calib, *A;
    const int32_t* counts;
    const int32_t* flip;  // (B) or NULL: the frame's image is flipped (nuscenes.py:201-205); X, vx arrive negated
    int B, Nmax, img_w, img_h, H, W;
    double ox, oy, oz, max_dist, z_offset;
    PaintRec* recs;
    short4* rects;  // the rectangles once more, dense (paint_kernel's staging reads these)
};

__device__ __forceinline__ double dot3(double a0, double x, double a1, double y, double a2, double z) {
    return __dadd_rn(__dadd_rn(__dmul_rn(a0, x), __dmul_rn(a1, y)), __dmul_rn(a2, z));
}

__global__ void __launch_bounds__(128) pillar_rects_kernel(const PillarParams p) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");  // (the records are still read by the paint of the call before)
    const int i = blockIdx.x * 128 + threadIdx.x;
    const int b = blockIdx.y;
    if (i >= p.Nmax) return;
    PaintRec r;
    r.ra = r.rb = r.ca = r.cb = 0;
    r.v0 = r.v1 = r.v2 = 0.0f;
    r.pad_ = r.pad2_ = 0;
    PaintRec* dst = p.recs + (size_t)b * p.Nmax + i;
    if (i >= p.counts[b]) {
        *dst = r;
        if (p.rects) p.rects[(size_t)b * p.Nmax + i] = make_short4(0, 0, 0, 0);
        return;
    }
    const double* pt = p.points + ((size_t)b * p.Nmax + i) * 5;
    const double* Kc = p.cam + 9 * (size_t)b;
    const double* P = p.calib + 12 * (size_t)b;
    const double* Ab = p.A + 6 * (size_t)b;
    const double X = pt[0], Z = pt[2];
    double Y = pt[1];
    const int W = p.W, H = p.H;
    bool ok = !(p.max_dist > 0 && !(Z <= p.max_dist));
    Y = __dsub_rn(Y, p.z_offset);
    // a flipped frame was projected and masked BEFORE the flip (nuscenes.py:187-191, then :201-203): with the point as
    // loaded (X back to its sign), and the pixel column mirrored afterwards; the pillar below stands at the flipped X
    const bool flipped = p.flip && p.flip[b];
    const double Xp = flipped ? -X : X;
    const double uh = dot3(Kc[0], Xp, Kc[1], Y, Kc[2], Z);
    const double vh = dot3(Kc[3], Xp, Kc[4], Y, Kc[5], Z);
    const double zh = dot3(Kc[6], Xp, Kc[7], Y, Kc[8], Z);
    double u = uh / zh;
    const double v = vh / zh;
    ok = ok && (Z > 0 && u > 1 && u < p.img_w - 1 && v > 1 && v < p.img_h - 1);
    if (flipped) u = __dsub_rn((double)(p.img_w - 1), u);
    const double up = __dadd_rn(__dadd_rn(__dmul_rn(Ab[0], u), __dmul_rn(Ab[1], v)), Ab[2]);
    const double vp = __dadd_rn(__dadd_rn(__dmul_rn(Ab[3], u), __dmul_rn(Ab[4], v)), Ab[5]);
    ok = ok && (up < W && vp < H && 0 < up && 0 < vp);
    if (ok) {
        double umin = CUDART_INF, umax = -CUDART_INF, vmin = CUDART_INF, vmax = -CUDART_INF;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const double cx = __dadd_rn(X, ((c >> 1) & 1) ? -p.ox : p.ox);
            const double cy = __dadd_rn(Y, (c & 4) ? p.oy : 0.0);
            const double cz = __dadd_rn(Z, (((c + 1) >> 1) & 1) ? -p.oz : p.oz);
            const double q0 = __dadd_rn(dot3(P[0], cx, P[1], cy, P[2], cz), P[3]);
            const double q1 = __dadd_rn(dot3(P[4], cx, P[5], cy, P[6], cz), P[7]);
            const double q2 = __dadd_rn(dot3(P[8], cx, P[9], cy, P[10], cz), P[11]);
            const double uu = q0 / q2, vv = q1 / q2;
            const double ua = __dadd_rn(__dadd_rn(__dmul_rn(Ab[0], uu), __dmul_rn(Ab[1], vv)), Ab[2]);
            const double va = __dadd_rn(__dadd_rn(__dmul_rn(Ab[3], uu), __dmul_rn(Ab[4], vv)), Ab[5]);
            if (ua < umin) umin = ua;
            if (ua > umax) umax = ua;
            if (va < vmin) vmin = va;
            if (va > vmax) vmax = va;
        }
        const double pw = __dsub_rn(umax, umin), ph = __dsub_rn(vmax, vmin);
        double by1 = __dsub_rn(vp, ph);
        if (!(by1 > 0)) by1 = 0;
        double bx1 = __dsub_rn(up, pw / 2);
        if (!(bx1 > 0)) bx1 = 0;
        double bx2 = __dadd_rn(up, pw / 2);
        if (!(bx2 < W)) bx2 = W;
        long long y1 = (long long)rint(by1), y2 = (long long)rint(vp), x1 = (long long)rint(bx1),
                  x2 = (long long)rint(bx2);
        if (y2 > H) y2 = H;
        if (x2 > W) x2 = W;
        if (y1 < y2 && x1 < x2) {
            r.ra = (int16_t)y1;
            r.rb = (int16_t)y2;
            r.ca = (int16_t)x1;
            r.cb = (int16_t)x2;
        }
        r.v0 = (float)Z;
        r.v1 = (float)pt[3];
        r.v2 = (float)pt[4];
    }
    *dst = r;
    if (p.rects) p.rects[(size_t)b * p.Nmax + i] = make_short4(r.ra, r.rb, r.ca, r.cb);
}

cudaError_t launch_pillar_rects(const double* points, const int32_t* counts, const int32_t* flip, int B, int Nmax, const double* cam,
                                const double* calib, const double* A, int img_w, int img_h, int H, int W, double ph,
                                double pw, double pl, double max_dist, double z_offset, PaintRec* recs, short4* rects,
                                cudaStream_t st, int* launches) {
    PillarParams p;
    p.points = points;
    p.cam = cam;
    p.calib = calib;
    p.A = A;
    p.counts = counts;
    p.rects = rects;
    p.flip = flip;
    p.B = B;
    p.Nmax = Nmax;
    p.img_w = img_w;
    p.img_h = img_h;
    p.H = H;
    p.W = W;
    // ddd.get3dBox -> pointcloud.get3DCorners NumPy branch: fp32 corner templates scaled by fp64 dims,
    // i.e. the offsets are float32(0.5*l), float32(-h), float32(0.5*w)
    p.ox = (double)(float)(0.5 * pl);
    p.oy = (double)(float)(-1.0 * ph);
    p.oz = (double)(float)(0.5 * pw);
    p.max_dist = max_dist;
    p.z_offset = z_offset;
    p.recs = recs;
    if (Nmax == 0) return cudaSuccess;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((Nmax + 127) / 128, B);
    cfg.blockDim = dim3(128);
    cfg.stream = st;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, pillar_rects_kernel, p);
    if (e == cudaSuccess) ++*launches;
    return e;
}

}  // namespace cf

#ifdef CFD3D_PROFILE
extern "C" __attribute__((visibility("default"))) int cf_debug_read_paint_prof(long long* out, int n) {
    return (int)cudaMemcpyFromSymbol(out, cf::g_paint_prof, sizeof(long long) * 16 * n);
}
#endif
