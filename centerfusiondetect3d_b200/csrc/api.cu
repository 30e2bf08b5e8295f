// api.cu -- the C ABI of libcfd3d_b200.so (include/cfd3d_b200.h): argument checks, workspace carving and
// kernel sequencing.  No allocation, no host synchronisation, no CPU compute path: every entry point
// either launches the sm_100a kernels on the caller's stream or returns an error.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "kernels.cuh"

using namespace cf;

static thread_local char g_err[512] = "";
static thread_local int g_launches = 0;

// measurement hook (cf_set_kernel_events)
static thread_local int g_ev_kernel = -1;
static thread_local cudaEvent_t g_ev_start = nullptr, g_ev_stop = nullptr;
struct EvScope {
    bool on;
    cudaStream_t st;
    EvScope(int id, cudaStream_t s) : on(id == g_ev_kernel && g_ev_start && g_ev_stop), st(s) {
        if (on) cudaEventRecord(g_ev_start, st);
    }
    ~EvScope() {
        if (on) cudaEventRecord(g_ev_stop, st);
    }
};

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
static int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return (int)e;
}
#define CF_CUDA(call, what)                         \
    do {                                            \
        cudaError_t e__ = (call);                   \
        if (e__ != cudaSuccess) return cuda_fail(e__, what); \
    } while (0)
// same, with the measurement hook around kernel `id`
#define CF_KERNEL(id, st, call, what)               \
    do {                                            \
        cudaError_t e__;                            \
        {                                           \
            EvScope ev__(id, st);                   \
            e__ = (call);                           \
        }                                           \
        if (e__ != cudaSuccess) return cuda_fail(e__, what); \
    } while (0)

static int check_dims(int B, int C, int H, int W) {
    if (B < 0 || C < 1 || H < 1 || W < 1) return fail(CF_E_SHAPE, "bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    if ((long long)C * H * W >= (1ll << 31)) return fail(CF_E_RANGE, "C*H*W does not fit 31 bits");
    if (H > 32767 || W > 32767) return fail(CF_E_RANGE, "H, W must be <= 32767 (paint windows and ROI slices are 16-bit)");
    return 0;
}
static int check_k(int K, int H, int W) {
    if (K < 1 || (long long)K > (long long)H * W)
        return fail(CF_E_K, "K=%d out of range for a %dx%d map (selected index k out of range)", K, H, W);
    if (K > CF_MAX_K) return fail(CF_E_K, "K=%d exceeds CF_MAX_K=%d", K, CF_MAX_K);
    return 0;
}
static int check_ws(void* ws, size_t ws_bytes, size_t need) {
    if (!ws) return fail(CF_E_WORKSPACE, "workspace is NULL (need %zu bytes)", need);
    if (reinterpret_cast<uintptr_t>(ws) & 255u) return fail(CF_E_WORKSPACE, "workspace must be 256-byte aligned");
    if (ws_bytes < need) return fail(CF_E_WORKSPACE, "workspace too small: %zu < %zu", ws_bytes, need);
    return 0;
}
static bool misaligned4(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 3u) != 0; }

// the peak passes + merge jobs without the frustum tail (cf_peak_topk, cf_fusion_decode)
static int run_peak_passes(const float* heat, int B, int C, int H, int W, int K, MergeJob* raw_job, MergeJob* nms_job,
                           const Workspace& w, cudaStream_t st, int* n) {
    if (raw_job) finish_job(raw_job);
    if (nms_job) finish_job(nms_job);
    CF_KERNEL(CF_KERNEL_STAGE, st,
              launch_fusion_stage(heat, B, C, H, W, K, w, raw_job, nms_job, nullptr, nullptr, 0.0f, nullptr, 0, nullptr, st, n),
              "fusion_stage");
    return 0;
}

extern "C" {

int cf_version(void) { return 100; }
const char* cf_last_error(void) { return g_err; }
int cf_last_launch_count(void) { return g_launches; }
int cf_set_kernel_events(int kernel_id, void* ev_start, void* ev_stop) {
    if (kernel_id < 0 || !ev_start || !ev_stop) {
        g_ev_kernel = -1;
        g_ev_start = g_ev_stop = nullptr;
        return 0;
    }
    g_ev_kernel = kernel_id;
    g_ev_start = (cudaEvent_t)ev_start;
    g_ev_stop = (cudaEvent_t)ev_stop;
    return 0;
}

#ifdef CFD3D_PROFILE
int cf_debug_read_prof(long long* out_host, int n_cta) { return debug_read_prof(out_host, n_cta); }
#endif

int cf_last_plan(cf_plan_info_t* out) {
    if (!out) return fail(CF_E_NULL, "cf_last_plan: out is NULL");
    last_plan(out);
    return 0;
}
int cf_debug_set_plan(int force_cluster_size, int max_clusters) {
    if (force_cluster_size < 0 || max_clusters < 0 || (force_cluster_size & (force_cluster_size - 1)) || force_cluster_size > 16)
        return fail(CF_E_SHAPE, "cf_debug_set_plan: cluster size must be 0 or a power of two <= 16");
    debug_set_plan(force_cluster_size, max_clusters);
    return 0;
}

size_t cf_workspace_bytes(int B, int C, int H, int W, int K) {
    if (B < 0 || C < 1 || H < 1 || W < 1 || K < 1 || K > CF_MAX_K) return 0;
    return workspace_layout(B, K, nullptr, nullptr) + 256;
}

static size_t pillar_rects_offset(int B, int Nmax) {
    return ((size_t)B * (size_t)Nmax * sizeof(PaintRec) + 255) / 256 * 256;
}
size_t cf_pillars_workspace_bytes(int B, int Nmax) {
    if (B < 0 || Nmax < 0) return 0;
    // paint records + the dense copy of their rectangles (8 bytes each, 256-byte aligned behind the records)
    return pillar_rects_offset(B, Nmax) + (size_t)B * (size_t)Nmax * sizeof(short4) + 256;
}

int cf_nms(const float* heat, int B, int C, int H, int W, float* out, void* stream) {
    g_launches = 0;
    if (B == 0) return 0;
    if (!heat || !out) return fail(CF_E_NULL, "cf_nms: NULL pointer");
    if (int rc = check_dims(B, C, H, W)) return rc;
    if (B == 0) return 0;
    int n = 0;
    CF_CUDA(launch_nms(heat, B, C, H, W, out, (cudaStream_t)stream, &n), "cf_nms");
    g_launches = n;
    return 0;
}

int cf_gather_nchw(const float* feat, int B, int ch, int64_t HW, const int64_t* inds, int M, float* out,
                   void* stream) {
    g_launches = 0;
    if (!feat || !inds || !out) return fail(CF_E_NULL, "cf_gather_nchw: NULL pointer");
    if (B < 0 || ch < 1 || HW < 1 || M < 0) return fail(CF_E_SHAPE, "cf_gather_nchw: bad shape");
    int n = 0;
    CF_CUDA(launch_gather(feat, B, ch, HW, inds, M, out, (cudaStream_t)stream, &n), "cf_gather_nchw");
    g_launches = n;
    return 0;
}

int cf_get_alpha(const float* rot, int64_t n_, float* alpha, void* stream) {
    g_launches = 0;
    if (!rot || !alpha) return fail(CF_E_NULL, "cf_get_alpha: NULL pointer");
    if (n_ < 0) return fail(CF_E_SHAPE, "cf_get_alpha: n < 0");
    int n = 0;
    CF_CUDA(launch_alpha(rot, n_, alpha, (cudaStream_t)stream, &n), "cf_get_alpha");
    g_launches = n;
    return 0;
}

int cf_dist_thresh(const float* calib, const float* center, const float* dim, const float* alpha, int B, int K,
                   float* thr, void* stream) {
    g_launches = 0;
    if (!calib || !center || !dim || !alpha || !thr) return fail(CF_E_NULL, "cf_dist_thresh: NULL pointer");
    if (B < 0 || K < 0) return fail(CF_E_SHAPE, "cf_dist_thresh: bad shape");
    int n = 0;
    CF_CUDA(launch_dist_thresh(calib, center, dim, alpha, B, K, thr, (cudaStream_t)stream, &n), "cf_dist_thresh");
    g_launches = n;
    return 0;
}

int cf_peak_topk(const float* heat, int B, int C, int H, int W, int K, int apply_nms, float* scores, int64_t* inds,
                 int32_t* classes, int64_t* ys, int64_t* xs, void* ws, size_t ws_bytes, void* stream) {
    g_launches = 0;
    if (B == 0 && C >= 1 && H >= 1 && W >= 1) return check_k(K, H, W);  // empty batch: data pointers may be NULL
    if (!heat) return fail(CF_E_NULL, "cf_peak_topk: heat is NULL");
    if (int rc = check_dims(B, C, H, W)) return rc;
    if (int rc = check_k(K, H, W)) return rc;
    if (misaligned4(heat)) return fail(CF_E_ALIGN, "cf_peak_topk: heat not 4-byte aligned");
    Workspace w;
    const size_t need = workspace_layout(B, K, ws, &w);
    if (int rc = check_ws(ws, ws_bytes, need)) return rc;
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    MergeJob job;
    memset(&job, 0, sizeof(job));
    job.epilogue = EPI_TOPK;
    job.topk = TopkOut{scores, inds, classes, ys, xs};
    if (int rc = run_peak_passes(heat, B, C, H, W, K, apply_nms ? nullptr : &job, apply_nms ? &job : nullptr, w, st, &n))
        return rc;
    g_launches = n;
    return 0;
}

static int check_frustum_args(const float* heat, const float* wh, const float* depth, const float* dim,
                              const float* rot, const float* pc_dep, const float* calib, float* pc_hm) {
    if (!heat || !wh || !depth || !dim || !rot || !pc_dep || !calib || !pc_hm)
        return fail(CF_E_NULL, "frustum: NULL map pointer");
    if (misaligned4(heat) || misaligned4(wh) || misaligned4(depth) || misaligned4(dim) || misaligned4(rot) ||
        misaligned4(pc_dep) || misaligned4(calib) || misaligned4(pc_hm))
        return fail(CF_E_ALIGN, "frustum: pointer not 4-byte aligned");
    return 0;
}

static int check_decode_args(const cf_decode_heads_t* h, const cf_decode_out_t* o) {
    if (!h || !o) return fail(CF_E_NULL, "decode: heads/out struct is NULL");
    if (!h->heatmap) return fail(CF_E_NULL, "decode: heatmap is NULL");
    if (!o->scores || !o->class_ids || !o->centers) return fail(CF_E_NULL, "decode: scores/class_ids/centers NULL");
    if ((h->wh && !o->bboxes) || (h->depth && !o->depth) || (h->rot && !o->rot) || (h->dim && !o->dim) ||
        (h->amodal && !o->amodal) || (h->nuscenes_att && !o->nuscenes_att) || (h->velocity && !o->velocity))
        return fail(CF_E_NULL, "decode: a head is present but its output pointer is NULL");
    if (o->bboxes && (reinterpret_cast<uintptr_t>(o->bboxes) & 15u)) return fail(CF_E_ALIGN, "decode: bboxes must be 16-byte aligned");
    return 0;
}

int cf_frustum_pc_hm(const float* heat, const float* wh, const float* depth, const float* dim, const float* rot,
                     const float* pc_dep, const float* calib, int B, int C, int H, int W, int K, float max_pc_dist,
                     float* pc_hm, int32_t* assoc, float* thr, int64_t* inds, void* ws, size_t ws_bytes,
                     void* stream) {
    return cf_fusion_stage(heat, wh, depth, dim, rot, pc_dep, calib, B, C, H, W, K, max_pc_dist, pc_hm, assoc, thr,
                           inds, nullptr, 0, 0, 0, nullptr, ws, ws_bytes, stream);
}

// postProcess fused into a decode job (row f1): argument checks + the device-side arguments
static int set_post_fuse(const cf_post_fuse_t* post, const cf_decode_heads_t* heads, PostFuse* pf) {
    memset(pf, 0, sizeof(*pf));
    if (!post) return 0;
    if (!post->trans_host) return fail(CF_E_NULL, "post: trans_host is NULL");
    if (post->height < 1 || post->width < 1) return fail(CF_E_SHAPE, "post: bad height / width");
    const cf_post_out_t& o = post->out;
    const bool has3d = heads->rot && heads->depth && heads->dim;
    if (!o.centers) return fail(CF_E_NULL, "post: centers output is NULL");
    if ((heads->wh && !o.bboxes) || (heads->rot && !o.alpha) ||
        (has3d && (!o.locations || !o.yaws || !o.bboxes3d || !post->calib)))
        return fail(CF_E_NULL, "post: a head is present but its post output (or calib) is NULL");
    pf->enabled = 1;
    pf->height = post->height;
    pf->width = post->width;
    pf->is_gt = post->is_gt;
    for (int j = 0; j < 6; ++j) pf->t[j] = post->trans_host[j];
    pf->calib = post->calib;
    pf->out = o;
    return 0;
}

int cf_fusion_decode(const cf_decode_heads_t* heads, int B, int C, int H, int W, int K, int out_h, int out_w,
                     int norm2d, const cf_decode_out_t* out, void* ws, size_t ws_bytes, void* stream) {
    return cf_fusion_decode_post(heads, B, C, H, W, K, out_h, out_w, norm2d, out, nullptr, ws, ws_bytes, stream);
}

int cf_fusion_decode_post(const cf_decode_heads_t* heads, int B, int C, int H, int W, int K, int out_h, int out_w,
                          int norm2d, const cf_decode_out_t* out, const cf_post_fuse_t* post, void* ws, size_t ws_bytes,
                          void* stream) {
    g_launches = 0;
    if (B == 0) return 0;
    if (int rc = check_decode_args(heads, out)) return rc;
    if (int rc = check_dims(B, C, H, W)) return rc;
    if (int rc = check_k(K, H, W)) return rc;
    Workspace w;
    const size_t need = workspace_layout(B, K, ws, &w);
    if (int rc = check_ws(ws, ws_bytes, need)) return rc;
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    MergeJob job;
    memset(&job, 0, sizeof(job));
    job.epilogue = EPI_DECODE;
    job.dec.heads = *heads;
    job.dec.out = *out;
    job.dec.out_h = out_h;
    job.dec.out_w = out_w;
    job.dec.norm2d = norm2d;
    if (int rc = set_post_fuse(post, heads, &job.dec.post)) return rc;
    if (int rc = run_peak_passes(heads->heatmap, B, C, H, W, K, nullptr, &job, w, st, &n)) return rc;
    g_launches = n;
    return 0;
}

int cf_fusion_stage(const float* heat, const float* wh, const float* depth, const float* dim, const float* rot,
                    const float* pc_dep, const float* calib, int B, int C, int H, int W, int K, float max_pc_dist,
                    float* pc_hm, int32_t* assoc, float* thr, int64_t* inds, const cf_decode_heads_t* dec_heads,
                    int out_h, int out_w, int norm2d, const cf_decode_out_t* dec_out, void* ws, size_t ws_bytes,
                    void* stream) {
    return cf_fusion_stage_strided(heat, wh, depth, dim, rot, pc_dep, calib, B, C, H, W, K, max_pc_dist, pc_hm,
                                   3ll * H * W, assoc, thr, inds, dec_heads, out_h, out_w, norm2d, dec_out, ws, ws_bytes,
                                   stream);
}

int cf_fusion_stage_strided(const float* heat, const float* wh, const float* depth, const float* dim, const float* rot,
                            const float* pc_dep, const float* calib, int B, int C, int H, int W, int K,
                            float max_pc_dist, float* pc_hm, int64_t pc_hm_batch_stride, int32_t* assoc, float* thr,
                            int64_t* inds, const cf_decode_heads_t* dec_heads, int out_h, int out_w, int norm2d,
                            const cf_decode_out_t* dec_out, void* ws, size_t ws_bytes, void* stream) {
    return cf_fusion_stage_split(heat, wh, depth, dim, rot, pc_dep, nullptr, calib, B, C, H, W, K, max_pc_dist, pc_hm,
                                 pc_hm_batch_stride, assoc, thr, inds, dec_heads, out_h, out_w, norm2d, dec_out, ws,
                                 ws_bytes, stream);
}

int cf_fusion_stage_split(const float* heat, const float* wh, const float* depth, const float* dim, const float* rot,
                          const float* pc_dep, const float* pc_dep_vel, const float* calib, int B, int C, int H, int W,
                          int K, float max_pc_dist, float* pc_hm, int64_t pc_hm_batch_stride, int32_t* assoc, float* thr,
                          int64_t* inds, const cf_decode_heads_t* dec_heads, int out_h, int out_w, int norm2d,
                          const cf_decode_out_t* dec_out, void* ws, size_t ws_bytes, void* stream) {
    return cf_fusion_stage_post(heat, wh, depth, dim, rot, pc_dep, pc_dep_vel, calib, B, C, H, W, K, max_pc_dist, pc_hm,
                                pc_hm_batch_stride, assoc, thr, inds, dec_heads, out_h, out_w, norm2d, dec_out, nullptr, ws,
                                ws_bytes, stream);
}

int cf_fusion_stage_post(const float* heat, const float* wh, const float* depth, const float* dim, const float* rot,
                         const float* pc_dep, const float* pc_dep_vel, const float* calib, int B, int C, int H, int W,
                         int K, float max_pc_dist, float* pc_hm, int64_t pc_hm_batch_stride, int32_t* assoc, float* thr,
                         int64_t* inds, const cf_decode_heads_t* dec_heads, int out_h, int out_w, int norm2d,
                         const cf_decode_out_t* dec_out, const cf_post_fuse_t* post, void* ws, size_t ws_bytes,
                         void* stream) {
    if (pc_hm_batch_stride < 3ll * H * W) return fail(CF_E_SHAPE, "cf_fusion_stage: pc_hm batch stride < 3*H*W");
    g_launches = 0;
    if (B == 0) return 0;
    if (int rc = check_frustum_args(heat, wh, depth, dim, rot, pc_dep, calib, pc_hm)) return rc;
    const bool with_decode = dec_heads != nullptr;
    if (with_decode) {
        if (int rc = check_decode_args(dec_heads, dec_out)) return rc;
        if (dec_heads->heatmap != heat)
            return fail(CF_E_NULL, "cf_fusion_stage: decode heatmap must be the frustum heatmap");
    } else if (post) {
        return fail(CF_E_NULL, "cf_fusion_stage_post: post needs the decode heads");
    }
    if (int rc = check_dims(B, C, H, W)) return rc;
    if (int rc = check_k(K, H, W)) return rc;
    Workspace w;
    const size_t need = workspace_layout(B, K, ws, &w);
    if (int rc = check_ws(ws, ws_bytes, need)) return rc;
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    MergeJob fj, dj;
    memset(&fj, 0, sizeof(fj));
    memset(&dj, 0, sizeof(dj));
    fj.epilogue = EPI_FRUSTUM;
    fj.fh = FrustumHeads{wh, depth, dim, rot, calib};
    fj.fo = FrustumOut{thr, inds};
    if (with_decode) {
        dj.epilogue = EPI_DECODE;
        dj.dec.heads = *dec_heads;
        dj.dec.out = *dec_out;
        dj.dec.out_h = out_h;
        dj.dec.out_w = out_w;
        dj.dec.norm2d = norm2d;
        if (int rc = set_post_fuse(post, dec_heads, &dj.dec.post)) return rc;
    }
    finish_job(&fj);
    if (with_decode) finish_job(&dj);
    CF_KERNEL(CF_KERNEL_STAGE, st,
              launch_fusion_stage(heat, B, C, H, W, K, w, &fj, with_decode ? &dj : nullptr, pc_dep, pc_dep_vel, max_pc_dist, pc_hm,
                                  (long long)pc_hm_batch_stride, assoc, st, &n),
              "fusion_stage");
    g_launches = n;
    return 0;
}

int cf_stage_depth_plane(float* pc_dep_dev, const float* pc_dep_host, int B, int H, int W, void* stream) {
    g_launches = 0;
    if (B == 0) return 0;
    if (!pc_dep_dev || !pc_dep_host) return fail(CF_E_NULL, "cf_stage_depth_plane: NULL pointer");
    if (B < 0 || H <= 0 || W <= 0) return fail(CF_E_SHAPE, "cf_stage_depth_plane: bad shape");
    const size_t plane = (size_t)H * W * sizeof(float);
    CF_CUDA(cudaMemcpy2DAsync(pc_dep_dev, 3 * plane, pc_dep_host, 3 * plane, plane, (size_t)B, cudaMemcpyHostToDevice,
                              (cudaStream_t)stream),
            "cf_stage_depth_plane");
    return 0;
}

int cf_radar_sort(const double* points, const int32_t* counts, int B, int Nmax, double* sorted, int32_t* order,
                  void* stream) {
    g_launches = 0;
    if (!counts || !sorted || (Nmax > 0 && !points)) return fail(CF_E_NULL, "cf_radar_sort: NULL pointer");
    if (B < 0 || Nmax < 0 || Nmax > 25600) return fail(CF_E_SHAPE, "cf_radar_sort: bad shape (Nmax <= 25600)");
    if (points == sorted) return fail(CF_E_NULL, "cf_radar_sort: sorted must not alias points");
    int n = 0;
    CF_CUDA(launch_radar_sort(points, counts, B, 0, Nmax, nullptr, 0, sorted, order, (cudaStream_t)stream, &n), "radar_sort");
    g_launches = n;
    return 0;
}

int cf_radar_ingest(const double* raw, const int32_t* counts, int B, int rows, int Nmax, const int32_t* flip,
                    int descending, double* sorted, int32_t* order, void* stream) {
    g_launches = 0;
    if (!counts || !sorted || (Nmax > 0 && !raw)) return fail(CF_E_NULL, "cf_radar_ingest: NULL pointer");
    if (B < 0 || Nmax < 0 || Nmax > 25600) return fail(CF_E_SHAPE, "cf_radar_ingest: bad shape (Nmax <= 25600)");
    if (rows != 0 && rows < 10) return fail(CF_E_SHAPE, "cf_radar_ingest: rows must be 0 (packed [X,Y,Z,vx,vz]) or >= 10");
    if (raw == sorted) return fail(CF_E_NULL, "cf_radar_ingest: sorted must not alias raw");
    int n = 0;
    CF_CUDA(launch_radar_sort(raw, counts, B, rows, Nmax, flip, descending ? 1 : 0, sorted, order, (cudaStream_t)stream, &n),
            "radar_ingest");
    g_launches = n;
    return 0;
}

int cf_frustum_boxes(const float* pc_dep, const float* bboxes, const float* depth, const float* thr,
                     const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, float* pc_hm, int32_t* assoc,
                     void* ws, size_t ws_bytes, void* stream) {
    g_launches = 0;
    if (!pc_dep || !pc_hm || (N > 0 && (!bboxes || !depth || !thr))) return fail(CF_E_NULL, "cf_frustum_boxes: NULL pointer");
    if (B < 0 || N < 0 || H < 1 || W < 1 || H > 32767 || W > 32767) return fail(CF_E_SHAPE, "cf_frustum_boxes: bad shape");
    const size_t need = (size_t)B * N * sizeof(PaintRec);
    if (need) {
        if (int rc = check_ws(ws, ws_bytes, need)) return rc;
    }
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    CF_KERNEL(CF_KERNEL_ASSOC, st,
              launch_box_assoc(pc_dep, bboxes, depth, thr, counts, B, N, H, W, max_pc_dist, (PaintRec*)ws, assoc, st, &n),
              "box_assoc");
    CF_KERNEL(CF_KERNEL_PAINT, st, launch_paint((const PaintRec*)ws, nullptr, N, N, counts, B, H, W, pc_hm, nullptr, st, &n), "paint");
    g_launches = n;
    return 0;
}

int cf_frustum_boxes_f64(const float* pc_dep, const double* bboxes, const double* depth, const double* thr,
                         const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, int flags, float* pc_hm,
                         int32_t* assoc, void* ws, size_t ws_bytes, void* stream) {
    g_launches = 0;
    if (!pc_dep || !pc_hm || (N > 0 && (!bboxes || !depth || !thr))) return fail(CF_E_NULL, "cf_frustum_boxes_f64: NULL pointer");
    if (B < 0 || N < 0 || H < 1 || W < 1 || H > 32767 || W > 32767) return fail(CF_E_SHAPE, "cf_frustum_boxes_f64: bad shape");
    if (flags & ~(CF_BOX_F32 | CF_THR_F32)) return fail(CF_E_SHAPE, "cf_frustum_boxes_f64: unknown flag");
    const size_t need = (size_t)B * N * sizeof(PaintRec);
    if (need) {
        if (int rc = check_ws(ws, ws_bytes, need)) return rc;
    }
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    CF_KERNEL(CF_KERNEL_ASSOC, st,
              launch_box_assoc_np(pc_dep, bboxes, depth, thr, counts, B, N, H, W, max_pc_dist, flags, (PaintRec*)ws, assoc, st, &n),
              "box_assoc_np");
    CF_KERNEL(CF_KERNEL_PAINT, st, launch_paint((const PaintRec*)ws, nullptr, N, N, counts, B, H, W, pc_hm, nullptr, st, &n), "paint");
    g_launches = n;
    return 0;
}

int cf_post_process(const cf_post_in_t* in, int B, int K, int out_h, int out_w, const float* trans_host,
                    const float* calib, int is_gt, const cf_post_out_t* out, void* stream) {
    g_launches = 0;
    if (!in || !out || !trans_host) return fail(CF_E_NULL, "cf_post_process: in/out/trans is NULL");
    if (B < 0 || K < 0 || out_h < 1 || out_w < 1) return fail(CF_E_SHAPE, "cf_post_process: bad shape");
    if (B == 0 || K == 0) return 0;
    if (!in->class_ids || !in->centers || !out->centers) return fail(CF_E_NULL, "cf_post_process: class_ids/centers NULL");
    const bool has3d = in->rot && in->depth && in->dim;
    if ((in->bboxes && !out->bboxes) || (in->rot && !out->alpha) ||
        (has3d && (!out->locations || !out->yaws || !out->bboxes3d || !calib)))
        return fail(CF_E_NULL, "cf_post_process: an input is present but its output pointer is NULL");
    int n = 0;
    CF_KERNEL(CF_KERNEL_POST, (cudaStream_t)stream,
              launch_post_process(*in, B, K, out_h, out_w, trans_host, calib, is_gt, *out, (cudaStream_t)stream, &n),
              "post_process");
    g_launches = n;
    return 0;
}

int cf_pc_dep_to_hm_single(float* pc_hm, const float* pc_dep, int H, int W, const float* depth, const float* bbox,
                           const float* thr, float max_pc_dist, int32_t* assoc, void* stream) {
    g_launches = 0;
    if (!pc_hm || !pc_dep || !depth || !bbox || !thr) return fail(CF_E_NULL, "cf_pc_dep_to_hm_single: NULL pointer");
    if (H < 1 || W < 1 || (long long)H * W >= (1ll << 31)) return fail(CF_E_SHAPE, "cf_pc_dep_to_hm_single: bad shape");
    int n = 0;
    CF_CUDA(launch_pc2hm_single(pc_hm, pc_dep, H, W, depth, bbox, thr, max_pc_dist, assoc, (cudaStream_t)stream, &n),
            "pc2hm_single");
    g_launches = n;
    return 0;
}

int cf_radar_pillars(const double* points, const int32_t* counts, int B, int Nmax, const double* cam,
                     const double* calib, const double* A, int img_w, int img_h, int H, int W,
                     const double* pillar_dims_host, double max_dist, double z_offset, float* pc_dep, int32_t* owner,
                     void* ws, size_t ws_bytes, void* stream) {
    return cf_radar_pillars_flip(points, counts, nullptr, B, Nmax, cam, calib, A, img_w, img_h, H, W, pillar_dims_host,
                                 max_dist, z_offset, pc_dep, owner, ws, ws_bytes, stream);
}

int cf_radar_pillars_flip(const double* points, const int32_t* counts, const int32_t* flip, int B, int Nmax,
                          const double* cam, const double* calib, const double* A, int img_w, int img_h, int H, int W,
                          const double* pillar_dims_host, double max_dist, double z_offset, float* pc_dep,
                          int32_t* owner, void* ws, size_t ws_bytes, void* stream) {
    g_launches = 0;
    if (!counts || !cam || !calib || !A || !pillar_dims_host || !pc_dep || (Nmax > 0 && !points))
        return fail(CF_E_NULL, "cf_radar_pillars: NULL pointer");
    if (B < 0 || Nmax < 0 || H < 1 || W < 1 || H > 32767 || W > 32767)
        return fail(CF_E_SHAPE, "cf_radar_pillars: bad shape");
    const size_t need = (size_t)B * Nmax * sizeof(PaintRec);
    if (need) {
        if (int rc = check_ws(ws, ws_bytes, need)) return rc;
    }
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    // (a workspace sized for the records alone still works: the paint then reads the rectangles out of the records)
    const size_t roff = pillar_rects_offset(B, Nmax);
    short4* rects = ws_bytes >= roff + (size_t)B * Nmax * sizeof(short4) ? reinterpret_cast<short4*>((char*)ws + roff) : nullptr;
    CF_KERNEL(CF_KERNEL_PILLAR_RECTS, st, launch_pillar_rects(points, counts, flip, B, Nmax, cam, calib, A, img_w, img_h, H, W, pillar_dims_host[0],
                                pillar_dims_host[1], pillar_dims_host[2], max_dist, z_offset, (PaintRec*)ws, rects, st, &n), "pillar_rects");
    CF_KERNEL(CF_KERNEL_PAINT, st, launch_paint((const PaintRec*)ws, rects, Nmax, Nmax, nullptr, B, H, W, pc_dep, owner, st, &n), "paint");
    g_launches = n;
    return 0;
}

}  // extern "C"

// -----------------------------------------------------------------------------------------------------
// cf_frustum_boxes_static: cf_frustum_boxes without a caller-provided workspace -- for callers that cannot allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  The paint records box_assoc_kernel hands to paint_kernel
// live in the file-scope device array below (CF_STATIC_MAX_BOXES records of 32 bytes = 1 MiB, held by every GPU that
// loads the library until the process exits).  Host code must not pass the array's NAME to a kernel (on the host it is
// a stand-in, not a device address): its device address is taken once per device with cudaGetSymbolAddress, with that
// device current, and that pointer goes to the kernels.  Nothing here allocates.
// -----------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_recs[(size_t)CF_STATIC_MAX_BOXES * sizeof(PaintRec)];

static constexpr int STATIC_MAX_DEV = 64;
static PaintRec* g_static_ptr[STATIC_MAX_DEV];  // per device: the array's address on THAT device (zero = not asked yet)
static std::mutex g_static_mu;

// device address of g_static_recs on the current device (set-up: once per device)
static int static_recs(PaintRec** out) {
    int dev = -1;
    CF_CUDA(cudaGetDevice(&dev), "cf_frustum_boxes_static: cudaGetDevice");
    if (dev < 0 || dev >= STATIC_MAX_DEV) return fail(CF_E_RANGE, "cf_frustum_boxes_static: device ordinal %d >= %d", dev, STATIC_MAX_DEV);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* ptr = nullptr;
        CF_CUDA(cudaGetSymbolAddress(&ptr, g_static_recs), "cf_frustum_boxes_static: cudaGetSymbolAddress");
        if (!ptr) return fail(CF_E_NULL, "cf_frustum_boxes_static: cudaGetSymbolAddress returned NULL");
        g_static_ptr[dev] = reinterpret_cast<PaintRec*>(ptr);
    }
    *out = g_static_ptr[dev];
    return 0;
}

extern "C" int cf_frustum_boxes_static(const float* pc_dep, const float* bboxes, const float* depth, const float* thr,
                                       const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, float* pc_hm,
                                       int32_t* assoc, void* stream) {
    g_launches = 0;
    if (!pc_dep || !pc_hm || (N > 0 && (!bboxes || !depth || !thr))) return fail(CF_E_NULL, "cf_frustum_boxes_static: NULL pointer");
    if (B < 0 || N < 0 || H < 1 || W < 1 || H > 32767 || W > 32767) return fail(CF_E_SHAPE, "cf_frustum_boxes_static: bad shape");
    if ((long long)B * N > (long long)CF_STATIC_MAX_BOXES)
        return fail(CF_E_WORKSPACE, "cf_frustum_boxes_static: B*N = %lld boxes exceed the static scratch of %d (use cf_frustum_boxes "
                                    "with a workspace)", (long long)B * N, CF_STATIC_MAX_BOXES);
    if (B == 0) return 0;
    PaintRec* recs = nullptr;
    if (int rc = static_recs(&recs)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    int n = 0;
    // (every record [0, B*N) is written by box_assoc_kernel before paint_kernel reads it: nothing has to start at zero,
    // whatever the last call left in the array)
    CF_KERNEL(CF_KERNEL_ASSOC, st,
              launch_box_assoc(pc_dep, bboxes, depth, thr, counts, B, N, H, W, max_pc_dist, recs, assoc, st, &n), "box_assoc");
    CF_KERNEL(CF_KERNEL_PAINT, st, launch_paint(recs, nullptr, N, N, counts, B, H, W, pc_hm, nullptr, st, &n), "paint");
    g_launches = n;
    return 0;
}
