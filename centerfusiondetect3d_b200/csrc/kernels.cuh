// kernels.cuh -- internal interface between api.cu (the C ABI) and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/cfd3d_b200.h"

namespace cf {

// ---- launch plan of the two heat-map passes for one map geometry --------------------------------------
// A frame is viewed as a (C*H) x W row matrix; a tile is R consecutive rows of it (+1 halo row each side).
struct Plan {
    int T;       // threads per CTA of the heat-map passes
    int R;       // rows per tile
    int NT;      // tiles per frame
    int Wp;      // shared-memory row stride in floats (W rounded up to 4)
    size_t smem; // dynamic shared memory of the heat-map passes
};
// returns false when the geometry is unsupported
bool make_plan(int C, int H, int W, int K, Plan* p);

constexpr int HIST_SHIFT = 20;                       // histogram bin = key >> HIST_SHIFT (sign, exponent, 3 mantissa bits)
constexpr int HIST_BINS = 1 << (32 - HIST_SHIFT);    // 4096
constexpr int CAND_CAP = 2048;                       // survivor slots per (frame, list)
constexpr int CTR_STRIDE = 4;                        // per (frame, list): [0] survivors appended, [1] bound tiles done

// per-detection frustum parameters, produced by the merge kernel, consumed by the association kernel
struct __align__(16) DetParam {
    float x1, y1, x2, y2;
    float depth, thr;
    int32_t ind;  // flat pixel of the peak
    int32_t pad_;
};

// ordered rectangle paint record: pixel rows [ra,rb) x cols [ca,cb) take (v0,v1,v2); the record with
// the LARGEST index covering a pixel wins (== the reference's sequential overwrite order)
struct __align__(16) PaintRec {
    int16_t ra, rb, ca, cb;
    float v0, v1, v2;
    int32_t pad_;
    int32_t pad2_;
};
static_assert(sizeof(PaintRec) == 32, "PaintRec");

struct Workspace {
    void* zero_base;            // [hist | ctr]: cleared by one memset at the start of every call
    size_t zero_bytes;
    uint32_t* hist;             // (B, 2, HIST_BINS) histogram of group maxima, list 0 = raw, 1 = nms'd
    uint32_t* ctr;              // (B, 2, CTR_STRIDE)
    uint32_t* tau;              // (B, 2) key threshold: every pixel of the frame's top-K has key >= tau
    unsigned long long* cand;   // (B, 2, CAND_CAP) survivors (composites), unordered
    DetParam* det;              // (B, K)
    PaintRec* recs;             // (B, K)
};
size_t workspace_layout(int B, int C, int H, int W, int K, const Plan& plan, void* base, Workspace* out);

struct FrustumHeads {
    const float *wh, *depth, *dim, *rot, *calib;
};
struct FrustumOut {
    float* thr;     // (B,K) optional
    int64_t* inds;  // (B,K) optional
};
struct TopkOut {
    float* scores;
    int64_t* inds;
    int32_t* classes;
    int64_t* ys;
    int64_t* xs;
};
struct DecodeArgs {
    cf_decode_heads_t heads;
    cf_decode_out_t out;
    int out_h, out_w, norm2d;
};

enum { LIST_RAW = 1, LIST_NMS = 2 };
enum { EPI_NONE = 0, EPI_TOPK = 1, EPI_FRUSTUM = 2, EPI_DECODE = 3 };

struct MergeJob {       // what the merge kernel does with one sorted list
    int epilogue;       // EPI_*
    TopkOut topk;       // EPI_TOPK
    FrustumHeads fh;    // EPI_FRUSTUM
    FrustumOut fo;      // EPI_FRUSTUM
    DecodeArgs dec;     // EPI_DECODE
};

// what the merge tail of the filter pass does per frame: up to two jobs (raw list, nms'd list)
struct MergeParams {
    const float* heat;
    int C, H, W, K;
    const uint32_t* ctr;
    const unsigned long long* cand;
    int list[2];  // 0 = raw, 1 = nms'd
    MergeJob job[2];
    DetParam* det;
};

// ---- launchers (each returns the cudaError_t of its launch; *launches += kernels launched) ----------
// the three steps of the peak decode ahead of the merge: clear [hist|ctr], bound pass (HBM read of the heat-map:
// group maxima -> histogram -> tau), filter pass (re-read, L2-resident at the reference batch sizes: survivors)
cudaError_t launch_peak_clear(const Workspace& ws, cudaStream_t st);
cudaError_t launch_peak_bound(const float* heat, int B, int C, int H, int W, int K, int lists, const Plan& plan,
                              const Workspace& ws, cudaStream_t st, int* launches);
cudaError_t launch_peak_filter(const float* heat, int B, int C, int H, int W, int K, int lists, const Plan& plan,
                               const Workspace& ws, const MergeJob* raw_job, const MergeJob* nms_job,
                               cudaStream_t st, int* launches);
cudaError_t launch_merge(const float* heat, int B, int C, int H, int W, int K, const Plan& plan,
                         const Workspace& ws, const MergeJob* raw_job, const MergeJob* nms_job, cudaStream_t st,
                         int* launches);
cudaError_t launch_assoc(const float* pc_dep, int B, int H, int W, int K, float max_pc_dist, const Workspace& ws,
                         int32_t* assoc, cudaStream_t st, int* launches);
cudaError_t launch_paint(const PaintRec* recs, int n_per_frame, int rec_stride, const int32_t* counts, int B, int H,
                         int W, float* out, int32_t* owner, cudaStream_t st, int* launches);
cudaError_t launch_nms(const float* heat, int B, int C, int H, int W, float* out, cudaStream_t st, int* launches);
cudaError_t launch_gather(const float* feat, int B, int ch, int64_t HW, const int64_t* inds, int M, float* out,
                          cudaStream_t st, int* launches);
cudaError_t launch_alpha(const float* rot, int64_t n, float* alpha, cudaStream_t st, int* launches);
cudaError_t launch_dist_thresh(const float* calib, const float* center, const float* dim, const float* alpha, int B,
                               int K, float* thr, cudaStream_t st, int* launches);
cudaError_t launch_pc2hm_single(float* pc_hm, const float* pc_dep, int H, int W, const float* depth,
                                const float* bbox, const float* thr, float max_pc_dist, int32_t* assoc,
                                cudaStream_t st, int* launches);
cudaError_t launch_pillar_rects(const double* points, const int32_t* counts, int B, int Nmax, const double* cam,
                                const double* calib, const double* A, int img_w, int img_h, int H, int W, double ph,
                                double pw, double pl, double max_dist, double z_offset, PaintRec* recs,
                                cudaStream_t st, int* launches);

}  // namespace cf
