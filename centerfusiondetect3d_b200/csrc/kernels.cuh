// kernels.cuh -- internal interface between api.cu (the C ABI) and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/cfd3d_b200.h"

namespace cf {

// ---- launch plan of the peak-select kernel for one map geometry ------------------------------------
struct Plan {
    int T;       // threads per CTA (512 or 1024)
    int EI;      // float4 items per thread (compile-time bound of the instantiation)
    int J;       // per-thread top-J feeding the warp bound
    int R;       // rows per band
    int S;       // bands per class plane
    int Wp;      // shared-memory row stride in floats (W rounded up to 4)
    int NB;      // candidate lists per frame and kind = C * S
    size_t smem; // dynamic shared memory of the peak-select kernel
};
// returns false when the geometry is unsupported
bool make_plan(int C, int H, int W, int K, Plan* p);

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
    unsigned long long* cand_raw;  // (B, NB, K)
    unsigned long long* cand_nms;  // (B, NB, K)
    DetParam* det;                 // (B, K)
    PaintRec* recs;                // (B, K)
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

// ---- launchers (each returns the cudaError_t of its launch; *launches += kernels launched) ----------
cudaError_t launch_peak_select(const float* heat, int B, int C, int H, int W, int K, int lists, const Plan& plan,
                               const Workspace& ws, cudaStream_t st, int* launches);
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
