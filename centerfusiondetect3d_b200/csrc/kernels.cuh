// kernels.cuh -- internal interface between api.cu (the C ABI) and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/cfd3d_b200.h"

namespace cf {

constexpr int HIST_SHIFT = 20;                       // histogram bin = key >> HIST_SHIFT (sign, exponent, 3 mantissa bits)
constexpr int HIST_BINS = 1 << (32 - HIST_SHIFT);    // 4096
constexpr int CAND_CAP = 2048;                       // survivor slots per (frame, list)
constexpr int CTR_STRIDE = 4;                        // per (frame, list): [0] survivors appended

// per-detection frustum parameters, produced by the merge kernel, consumed by the association kernel
struct __align__(16) DetParam {
    float x1, y1, x2, y2;
    float depth, thr;
    int16_t ra, rb, ca, cb;  // paint window (assoc.cuh paint_window), rb <= ra or cb <= ca: empty
};
static_assert(sizeof(DetParam) == 32, "DetParam");

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
    uint32_t* ctr;              // (B, 2, CTR_STRIDE) survivor counters, list 0 = raw, 1 = nms'd (cleared in-kernel)
    unsigned long long* cand;   // (B, 2, CAND_CAP) survivors (composites), unordered
    DetParam* det;              // (B, K)
    PaintRec* recs;             // (B, K)
};
size_t workspace_layout(int B, int K, void* base, Workspace* out);

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
// postProcess fused into a decode epilogue (row f1, post.cuh): absent when enabled == 0
struct PostFuse {
    int enabled, height, width, is_gt;
    float t[6];          // inverse output affine (2x3)
    const float* calib;  // (B,3,4) device
    cf_post_out_t out;
};
struct DecodeArgs {
    cf_decode_heads_t heads;
    cf_decode_out_t out;
    int out_h, out_w, norm2d;
    PostFuse post;
};

// head channels a merge job gathers at the K peak pixels: slot s reads base[s][(b*nch[s] + ch[s])*H*W + ind]
// (base NULL = head absent).  Frustum slots: 0-1 wh, 2 depth, 3-5 dim, 6-13 rot.  Decode slots: 0-1 reg, 2-3 wh,
// 4 depth, 5-12 rot, 13-15 dim, 16-17 amodal, 18-25 nuscenes_att, 26-28 velocity, 29 uncertainty.
constexpr int MAX_GCH = 32;
constexpr int GATH_STRIDE = 33;  // floats per detection in the shared-memory gather buffer (odd: conflict-free columns)
struct GatherTab {
    const float* base[MAX_GCH];
    unsigned char nch[MAX_GCH], ch[MAX_GCH];
    int n;
};
enum { FS_WH = 0, FS_DEPTH = 2, FS_DIM = 3, FS_ROT = 6, FS_N = 14 };
enum { DS_REG = 0, DS_WH = 2, DS_DEPTH = 4, DS_ROT = 5, DS_DIM = 13, DS_AMODAL = 16, DS_ATT = 18, DS_VEL = 26, DS_UNC = 29, DS_N = 30 };

enum { LIST_RAW = 1, LIST_NMS = 2 };
enum { EPI_NONE = 0, EPI_TOPK = 1, EPI_FRUSTUM = 2, EPI_DECODE = 3 };

struct MergeJob {       // what the merge kernel does with one sorted list
    int epilogue;       // EPI_*
    TopkOut topk;       // EPI_TOPK
    FrustumHeads fh;    // EPI_FRUSTUM
    FrustumOut fo;      // EPI_FRUSTUM
    DecodeArgs dec;     // EPI_DECODE
    GatherTab gt;       // filled by finish_job() from fh / dec.heads
};
void finish_job(MergeJob* job);

// what the merge tail of the filter pass does per frame: up to two jobs (raw list, nms'd list)
struct MergeParams {
    int prof;  // development aid (CFD3D_PROF)
    const float* heat;
    int C, H, W, K;
    const uint32_t* ctr;
    const unsigned long long* cand;
    int list[2];  // 0 = raw, 1 = nms'd
    MergeJob job[2];
    DetParam* det;
};

// ---- launchers (each returns the cudaError_t of its launch; *launches += kernels launched) ----------
// the fused stage kernel (stage.cu): bound + filter + merge jobs (+ association + paint when raw_job is a frustum
// job and pc_dep / pc_hm are given) in ONE launch, one thread-block cluster per frame
cudaError_t launch_fusion_stage(const float* heat, int B, int C, int H, int W, int K, const Workspace& ws,
                                const MergeJob* raw_job, const MergeJob* nms_job, const float* pc_dep,
                                const float* pc_vel, float max_pc_dist, float* pc_hm, long long pc_hm_stride,
                                int32_t* assoc, cudaStream_t st, int* launches);
cudaError_t launch_post_process(const cf_post_in_t& in, int B, int K, int out_h, int out_w, const float* trans_host,
                                const float* calib, int is_gt, const cf_post_out_t& out, cudaStream_t st,
                                int* launches);
cudaError_t launch_box_assoc(const float* pc_dep, const float* bboxes, const float* depth, const float* thr,
                             const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, PaintRec* recs,
                             int32_t* assoc, cudaStream_t st, int* launches);
cudaError_t launch_box_assoc_np(const float* pc_dep, const double* bboxes, const double* depth, const double* thr,
                                const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, int flags,
                                PaintRec* recs, int32_t* assoc, cudaStream_t st, int* launches);
cudaError_t launch_radar_sort(const double* points, const int32_t* counts, int B, int rows, int Nmax, const int32_t* flip,
                              int descending, double* out, int32_t* order, cudaStream_t st, int* launches);
int debug_read_prof(long long* out, int n_cta);  // -DCFD3D_PROFILE builds only
void debug_set_plan(int force_cs, int max_clusters);
void last_plan(cf_plan_info_t* out);
cudaError_t launch_paint(const PaintRec* recs, const short4* rects, int n_per_frame, int rec_stride, const int32_t* counts,
                         int B, int H, int W, float* out, int32_t* owner, cudaStream_t st, int* launches);
cudaError_t launch_nms(const float* heat, int B, int C, int H, int W, float* out, cudaStream_t st, int* launches);
cudaError_t launch_gather(const float* feat, int B, int ch, int64_t HW, const int64_t* inds, int M, float* out,
                          cudaStream_t st, int* launches);
cudaError_t launch_alpha(const float* rot, int64_t n, float* alpha, cudaStream_t st, int* launches);
cudaError_t launch_dist_thresh(const float* calib, const float* center, const float* dim, const float* alpha, int B,
                               int K, float* thr, cudaStream_t st, int* launches);
cudaError_t launch_pc2hm_single(float* pc_hm, const float* pc_dep, int H, int W, const float* depth,
                                const float* bbox, const float* thr, float max_pc_dist, int32_t* assoc,
                                cudaStream_t st, int* launches);
cudaError_t launch_pillar_rects(const double* points, const int32_t* counts, const int32_t* flip, int B, int Nmax, const double* cam,
                                const double* calib, const double* A, int img_w, int img_h, int H, int W, double ph,
                                double pw, double pl, double max_dist, double z_offset, PaintRec* recs, short4* rects,
                                cudaStream_t st, int* launches);

}  // namespace cf
