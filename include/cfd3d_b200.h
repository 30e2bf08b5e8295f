/*
 * cfd3d_b200.h -- C ABI of the B200-native CenterFusion fusion stage (libcfd3d_b200.so).
 *
 * One shared library, extern "C", plain pointers and sizes, no torch types.  Every pointer is a DEVICE
 * pointer unless its name ends in `_host`.  All maps are fp32, NCHW, contiguous.  Every call is
 * stream-ordered on `stream` (a cudaStream_t passed as void*), never allocates, never synchronises the
 * host and keeps no mutable global state other than a thread-local error string and a mutex-protected cache of
 * launch plans per (device, shape): the caller owns every buffer
 * including the workspace (`cf_workspace_bytes`).  Return value: 0 = ok, >0 = the cudaError_t of a
 * failed launch, <0 = argument error (CF_E_*); `cf_last_error()` describes the last failure.
 *
 * The reference (HengWeiBin/CenterFusionDetect3D) is pure Python and has no FFI; its boundary for this
 * path is a set of Python callables.  Each entry point below names the reference callable it replaces
 * (paths relative to the reference tree).  centerfusiondetect3d_b200/ops.py is the ctypes host side
 * that restores the reference signatures on top of this ABI; INTEGRATION.md shows the binding.
 */
#ifndef CFD3D_B200_H
#define CFD3D_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CF_API __attribute__((visibility("default")))

enum {
    CF_OK = 0,
    CF_E_NULL = -1,      /* a required pointer is NULL */
    CF_E_SHAPE = -2,     /* non-positive or unsupported dimension */
    CF_E_K = -3,         /* K < 1, K > H*W (torch.topk would raise) or K > CF_MAX_K */
    CF_E_WORKSPACE = -4, /* workspace missing, misaligned or smaller than cf_workspace_bytes() */
    CF_E_ALIGN = -5,     /* a map pointer is not 4-byte aligned */
    CF_E_RANGE = -6      /* index space does not fit 32 bits (C*H*W >= 2^31) */
};

#define CF_MAX_K 1024 /* largest K the selection kernels support (reference configs use 100 and 500) */

CF_API int cf_version(void);
CF_API const char* cf_last_error(void);

/* Bytes of device workspace any entry point below needs for these dimensions (256-byte aligned base). */
CF_API size_t cf_workspace_bytes(int B, int C, int H, int W, int K);

/* ---- a1  nms(heatmap, kernel=3)                        src/lib/model/utils.py:112-128 ------------
 * out = heat * (maxpool3x3(heat) == heat), -inf padding.  out must not alias heat. */
CF_API int cf_nms(const float* heat, int B, int C, int H, int W, float* out, void* stream);

/* ---- a2  topk(heatmap, K)                               src/lib/model/utils.py:6-38 ---------------
 * Frame-wide top-K under the total order (score desc, class asc, flat index asc).  apply_nms != 0 selects
 * on nms(heat) without materialising it (what fusionDecode does, src/lib/model/decode.py:38-39).
 * Outputs (B,K), any may be NULL: scores f32, inds i64 in [0,H*W), classes i32, ys i64, xs i64. */
CF_API int cf_peak_topk(const float* heat, int B, int C, int H, int W, int K, int apply_nms, float* scores,
                        int64_t* inds, int32_t* classes, int64_t* ys, int64_t* xs, void* ws, size_t ws_bytes,
                        void* stream);

/* ---- a3  transposeAndGetFeature(feature, indices)       src/lib/model/utils.py:41-72 --------------
 * out[b,m,c] = feat[b,c,inds[b,m]]  -- straight from NCHW, no transposed copy.  inds must lie in [0,HW). */
CF_API int cf_gather_nchw(const float* feat, int B, int ch, int64_t HW, const int64_t* inds, int M, float* out,
                          void* stream);

/* ---- a4  get_alpha / getDistanceThresh                  src/lib/utils/pointcloud.py:195-328 -------
 * rot (n,8) -> alpha (n);   calib (B,3,4), center (B,K,2), dim (B,K,3), alpha (B,K) -> thr (B,K). */
CF_API int cf_get_alpha(const float* rot, int64_t n, float* alpha, void* stream);
CF_API int cf_dist_thresh(const float* calib, const float* center, const float* dim, const float* alpha, int B,
                          int K, float* thr, void* stream);

/* ---- a7  getPcFrustumHeatmap(output, pc_dep, calib, config)   src/lib/utils/pointcloud.py:331-394 -
 * heat (B,C,H,W) raw (no nms); wh (B,2,H,W); depth (B,1,H,W); dim (B,3,H,W); rot (B,8,H,W);
 * pc_dep (B,3,H,W); calib (B,3,4).  Writes pc_hm (B,3,H,W) completely (zeros where nothing is painted).
 * Optional outputs (NULL to skip): assoc (B,K) i32 = flat pixel y*W+x of the matched pc_dep pixel or -1;
 * thr (B,K) distance thresholds; inds (B,K) i64 raw top-K indices. */
CF_API int cf_frustum_pc_hm(const float* heat, const float* wh, const float* depth, const float* dim,
                            const float* rot, const float* pc_dep, const float* calib, int B, int C, int H, int W,
                            int K, float max_pc_dist, float* pc_hm, int32_t* assoc, float* thr, int64_t* inds,
                            void* ws, size_t ws_bytes, void* stream);

/* ---- a8  cvtPcDepthToHeatmap(pc_hm, pc_dep, depth, bbox, distanceThreshold, max_pc_dist)
 *                                                               src/lib/utils/pointcloud.py:397-481 -
 * One detection, IN PLACE on pc_hm (3,H,W).  depth (1), bbox (4) and thr (1) are device scalars (the
 * reference passes tensors), so the call needs no host synchronisation.  assoc (1) i32 optional. */
CF_API int cf_pc_dep_to_hm_single(float* pc_hm, const float* pc_dep, int H, int W, const float* depth,
                                  const float* bbox, const float* thr, float max_pc_dist, int32_t* assoc,
                                  void* stream);

/* ---- a9  fusionDecode(outputs, outputSize, K, norm2d)   src/lib/model/decode.py:10-174 ------------
 * Single pyramid layer.  Heads: any pointer except heatmap may be NULL (= key absent in the reference's
 * dict).  `depth` is depth2 when the model has it, else depth (decode.py:101-110); `rot` likewise
 * rotation2/rotation (:120-121) -- the host side resolves which to pass. */
typedef struct {
    const float* heatmap;      /* (B,C,H,W) */
    const float* reg;          /* (B,2,H,W) or NULL */
    const float* wh;           /* (B,2,H,W) or NULL */
    const float* depth;        /* (B,1,H,W) or NULL */
    const float* rot;          /* (B,8,H,W) or NULL */
    const float* dim;          /* (B,3,H,W) or NULL */
    const float* amodal;       /* (B,2,H,W) or NULL */
    const float* nuscenes_att; /* (B,8,H,W) or NULL */
    const float* velocity;     /* (B,3,H,W) or NULL */
    const float* uncertainty;  /* (B,1,H,W) or NULL */
} cf_decode_heads_t;

typedef struct {
    float* scores;       /* (B,K)   required */
    float* class_ids;    /* (B,K)   required, class id as fp32 like the reference */
    float* centers;      /* (B,K,2) required, normalised [x/W, y/H] */
    float* bboxes;       /* (B,K,4) iff heads.wh; 16-byte aligned (written one box per store) */
    float* depth;        /* (B,K,1) iff heads.depth */
    float* rot;          /* (B,K,8) iff heads.rot */
    float* dim;          /* (B,K,3) iff heads.dim */
    float* amodal;       /* (B,K,2) iff heads.amodal */
    float* nuscenes_att; /* (B,K,8) iff heads.nuscenes_att */
    float* velocity;     /* (B,K,3) iff heads.velocity */
    int64_t* inds;       /* (B,K)   optional: nms'd top-K flat indices */
} cf_decode_out_t;

CF_API int cf_fusion_decode(const cf_decode_heads_t* heads, int B, int C, int H, int W, int K, int out_h, int out_w,
                            int norm2d, const cf_decode_out_t* out, void* ws, size_t ws_bytes, void* stream);

/* ---- a7 + a9 in one pass over the heat-map ------------------------------------------------------------
 * What CenterFusionHead.forward (src/lib/model/networks/detectHeads.py:176-177) and the decode callers
 * (src/lib/detector.py:414-424, src/lib/model/progressBar.py:96-101) compute between them, with the
 * heat-map read once: the peak-select kernel emits the raw and the nms'd candidate lists together.
 * Frustum inputs/outputs as in cf_frustum_pc_hm (frustum heads: wh, depth, dim, rot of the FIRST stage);
 * decode heads/outputs as in cf_fusion_decode. */
CF_API int cf_fusion_stage(const float* heat, const float* wh, const float* depth, const float* dim,
                           const float* rot, const float* pc_dep, const float* calib, int B, int C, int H, int W,
                           int K, float max_pc_dist, float* pc_hm, int32_t* assoc, float* thr, int64_t* inds,
                           const cf_decode_heads_t* dec_heads, int out_h, int out_w, int norm2d,
                           const cf_decode_out_t* dec_out, void* ws, size_t ws_bytes, void* stream);

/* ---- f3  the same, painting straight into the secondary heads' input -------------------------------------------
 * ConcateCombiner.forward (src/lib/model/fusionModules.py:18-35) concatenates the 64-channel image features with the
 * 3-channel pc_hm (torch.cat: a copy of the 64 channels).  With pc_hm = &sec_feats[0][64][0][0] and
 * pc_hm_batch_stride = 67*H*W the stage writes channels 64..66 of a pre-allocated (B,67,H,W) buffer in place, so a
 * backbone that emits its features into channels 0..63 of that buffer needs no concatenation at all.  The stride is
 * in floats and must be >= 3*H*W; the three channel planes of a frame stay contiguous. */
CF_API int cf_fusion_stage_strided(const float* heat, const float* wh, const float* depth, const float* dim,
                                   const float* rot, const float* pc_dep, const float* calib, int B, int C, int H, int W,
                                   int K, float max_pc_dist, float* pc_hm, int64_t pc_hm_batch_stride, int32_t* assoc,
                                   float* thr, int64_t* inds, const cf_decode_heads_t* dec_heads, int out_h, int out_w,
                                   int norm2d, const cf_decode_out_t* dec_out, void* ws, size_t ws_bytes, void* stream);

/* ---- host-buffer callers: velocities read in place ---------------------------------------------------------------
 * cf_fusion_stage_strided with the radar map split over two memories.  The kernels read plane 0 (depth) of pc_dep in
 * full but planes 1, 2 (the velocities) only at the <= K matched pixels of a frame, and every head except the heat-map
 * only at the K peak pixels.  A caller whose inputs live in HOST memory therefore copies just the heat-map, plane 0 of
 * pc_dep and calib to the device and passes
 *   - pc_dep:      DEVICE (B,3,H,W) buffer of which only plane 0 of every frame needs to be filled,
 *   - pc_dep_vel:  the complete (B,3,H,W) tensor in page-locked, device-mapped HOST memory (cudaHostAlloc /
 *                  cudaHostRegister; under unified addressing the host pointer itself), or NULL (= pc_dep),
 *   - wh, depth, dim, rot and the decode heads: page-locked host pointers as well, if the caller wishes.
 * Pageable host memory is not readable by the device: such pointers fault in the kernel.  Results are identical to
 * cf_fusion_stage_strided (tests/test_gpu_parity.py::test_pinned_host_heads_read_in_place). */
CF_API int cf_fusion_stage_split(const float* heat, const float* wh, const float* depth, const float* dim,
                                 const float* rot, const float* pc_dep, const float* pc_dep_vel, const float* calib, int B,
                                 int C, int H, int W, int K, float max_pc_dist, float* pc_hm, int64_t pc_hm_batch_stride,
                                 int32_t* assoc, float* thr, int64_t* inds, const cf_decode_heads_t* dec_heads, int out_h,
                                 int out_w, int norm2d, const cf_decode_out_t* dec_out, void* ws, size_t ws_bytes,
                                 void* stream);

/* Plane 0 (depth) of every frame of a host (B,3,H,W) pc_dep into the same position of a device (B,3,H,W) buffer: one
 * strided asynchronous copy on `stream` (the staging step of cf_fusion_stage_split).  Page-locked host memory makes
 * it asynchronous. */
CF_API int cf_stage_depth_plane(float* pc_dep_dev, const float* pc_dep_host, int B, int H, int W, void* stream);

/* ---- f2  stable depth sort of the radar points on the device --------------------------------------------------
 * The host-side np.argsort(depth) of src/lib/dataset/datasets/nuscenes.py:195-197 / src/lib/detector.py:281 (default
 * kind: not stable, so the reference's order among equal depths is unspecified) with the order fixed to
 * (Z ascending, source index ascending) == np.argsort(kind="stable").  points (B,Nmax,5) fp64 as in cf_radar_pillars,
 * counts (B) i32; sorted (B,Nmax,5) fp64 must not alias points; rows beyond counts[b] are copied in place.
 * order (B,Nmax) i32 optional: source index of every sorted row.  Nmax <= 25600. */
CF_API int cf_radar_sort(const double* points, const int32_t* counts, int B, int Nmax, double* sorted, int32_t* order,
                         void* stream);

/* ---- f2, complete: from the stored radar array to sorted points ----------------------------------------------
 * The host-side steps of loadRadarPointCloud between the pickle and processPointCloud
 * (src/lib/dataset/datasets/nuscenes.py:159-205; src/lib/detector.py:255-283) in one launch:
 *   layout   raw (B, rows, Nmax) fp64 = per frame the (rows >= 10, N) array tools/convert_nuScenes.py:205-212,252-253
 *            stores, padded to Nmax columns: rows 0-2 x, y, z (camera frame), rows 8, 9 the compensated velocities
 *            vx, vz (what drawPcHeat writes, nuscenes.py:259-261).  rows == 0: raw is already (B,Nmax,5) [X,Y,Z,vx,vz];
 *   sort     stable by depth as in cf_radar_sort; descending != 0: that order reversed = index[::-1] of
 *            nuscenes.py:196-197 (DATASET.PC_REVERSE false); 0 = detector.py:281 / PC_REVERSE true;
 *   flip     (B) i32 or NULL: the frame's image is flipped (nuscenes.py:201-205) -> X and vx change sign; pass the
 *            same flags to cf_radar_pillars_flip, which mirrors the projected pixel column.
 * The distance filter, PC_Z_OFFSET and the in-image masks (nuscenes.py:172-191) stay in cf_radar_pillars*: they drop
 * points without changing the order of the others.  sorted (B,Nmax,5) fp64, order (B,Nmax) i32 optional. */
CF_API int cf_radar_ingest(const double* raw, const int32_t* counts, int B, int rows, int Nmax, const int32_t* flip,
                           int descending, double* sorted, int32_t* order, void* stream);

/* ---- f4  frustum association of given boxes (training targets) ---------------------------------------------------
 * The per-annotation loop of src/lib/dataset/generic_dataset.py:673-687: cvtPcDepthToHeatmap
 * (src/lib/utils/pointcloud.py:397-481) for box 0, 1, ... of every frame in list order, later boxes overwriting earlier
 * ones.  bboxes (B,N,4) [x1,y1,x2,y2] in output-map pixels, depth (B,N), thr (B,N) distance thresholds
 * (cf_dist_thresh), counts (B) i32 boxes per frame or NULL (= N).  Writes pc_hm (B,3,H,W) completely; assoc (B,N) i32
 * optional (flat matched pixel or -1).  Workspace: B*N*32 bytes, 256-byte aligned. */
CF_API int cf_frustum_boxes(const float* pc_dep, const float* bboxes, const float* depth, const float* thr,
                            const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, float* pc_hm,
                            int32_t* assoc, void* ws, size_t ws_bytes, void* stream);

/* The same loop for callers that hold NumPy arguments, as the reference's DataLoader does: the NumPy branch of
 * cvtPcDepthToHeatmap (src/lib/utils/pointcloud.py:411-481 with lib = np) under NumPy >= 2 promotion rules.  bboxes,
 * depth and thr are float64 (the exact container of a float32 or float64 value); flags state the dtypes the caller's
 * arrays had, which select the arithmetic:
 *   CF_BOX_F32  the box was float32 (transformBbox, src/lib/dataset/generic_dataset.py:507-526): centre, 0.3 * w and the
 *               int() of the paint window are float32 operations; without it the centre is rounded to float32
 *               (dtype=np.float32, :424-426) and w, 0.3 * w, centre -+ interval / 2 are float64;
 *   CF_THR_F32  the threshold was np.float32 (what getDistanceThresh's NumPy branch returns): depth +- thr and the two
 *               strict compares are float32 (the Python-float depth is weak); without it they are float64.
 * CF_BOX_F32 | CF_THR_F32 is the DataLoader's arrangement and equals cf_frustum_boxes on the float32-rounded values. */
enum { CF_BOX_F32 = 1, CF_THR_F32 = 2 };
CF_API int cf_frustum_boxes_f64(const float* pc_dep, const double* bboxes, const double* depth, const double* thr,
                                const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, int flags,
                                float* pc_hm, int32_t* assoc, void* ws, size_t ws_bytes, void* stream);

/* ---- cf_frustum_boxes for callers that cannot allocate ---------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and
 * results as cf_frustum_boxes, but no workspace argument: the paint records between its two kernels live in a
 * file-scope `static __device__` array of the library (csrc/api.cu), CF_STATIC_MAX_BOXES records of 32 bytes = 1 MiB,
 * held by every GPU that loads the library until the process exits.  B * N > CF_STATIC_MAX_BOXES is refused with
 * CF_E_WORKSPACE.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); its one piece
 * of set-up, once per device, is cudaGetSymbolAddress of the array with that device current.  Every call on a device
 * shares the array, which keeps the last call's contents -- every record a call reads it has written first, so nothing
 * needs clearing -- and therefore calls that OVERLAP on one GPU (different streams, parallel graph branches) race:
 * serialise them, or use cf_frustum_boxes with one workspace each. */
#define CF_STATIC_MAX_BOXES 32768
CF_API int cf_frustum_boxes_static(const float* pc_dep, const float* bboxes, const float* depth, const float* thr,
                                   const int32_t* counts, int B, int N, int H, int W, float max_pc_dist, float* pc_hm,
                                   int32_t* assoc, void* stream);

/* ---- f1  postProcess(y, center, scale, height, width, calibs, isGt)   src/lib/utils/postProcess.py:13-85 ----
 * Decoded detections (the cf_fusion_decode outputs) -> image-space boxes, camera-frame locations, yaw, velocity
 * re-oriented along the box and the 8 corners of the 3-D box (utils/ddd.py:8-23, :122-199).  One detection per
 * thread.  Inputs: NULL = key absent in the reference's dict.  class_ids and velocity are updated IN PLACE like the
 * reference does; everything else goes to `out`.  trans_host: the 2x3 fp32 inverse output affine
 * getAffineTransform(center, scale, 0, (width, height), inverse=True) (utils/image.py:43-83) on the HOST;
 * calib (B,3,4) on the device. */
typedef struct {
    float* class_ids;      /* (B,K)   required, in/out: += 1 */
    const float* centers;  /* (B,K,2) required: normalised [x/W, y/H] */
    const float* bboxes;   /* (B,K,4) or NULL */
    const float* depth;    /* (B,K)   or NULL */
    const float* rot;      /* (B,K,8) or NULL */
    const float* dim;      /* (B,K,3) or NULL */
    const float* amodal;   /* (B,K,2) or NULL */
    float* velocity;       /* (B,K,3) or NULL, in/out */
} cf_post_in_t;

typedef struct {
    float* centers;   /* (B,K,2)   required */
    float* bboxes;    /* (B,K,4)   iff in.bboxes */
    float* alpha;     /* (B,K)     iff in.rot */
    float* locations; /* (B,K,3)   iff in.rot && in.depth && in.dim */
    float* yaws;      /* (B,K)     same */
    float* bboxes3d;  /* (B,K,8,3) same */
} cf_post_out_t;

CF_API int cf_post_process(const cf_post_in_t* in, int B, int K, int out_h, int out_w, const float* trans_host,
                           const float* calib, int is_gt, const cf_post_out_t* out, void* stream);

/* ---- f1 fused: fusionDecode + postProcess in the decode's own launch --------------------------------------------
 * The decode callers run postProcess right after fusionDecode (src/lib/detector.py:414-424 + :341,
 * src/lib/model/progressBar.py:96-104, :190-196).  With `post` the fourth lane of every detection's lane quad in the
 * decode epilogue computes the detection's postProcess from the channels the decode has gathered anyway: no second
 * launch, no second pass over the (B,K) records.  dec_out then holds what the reference's dict holds AFTER
 * postProcess for the keys it updates in place (class_ids + 1, velocity re-oriented) and the plain decode values for
 * the rest (centers normalised, bboxes in output pixels, rot); post->out receives the replaced / new keys.  Results
 * equal cf_fusion_decode followed by cf_post_process bit for bit (same arithmetic body).  post == NULL: plain decode.
 * height / width: postProcess's arguments (the network output size). */
typedef struct {
    const float* trans_host; /* 2x3 fp32 inverse output affine, HOST (see cf_post_process) */
    const float* calib;      /* (B,3,4) device; required when rot, depth and dim heads exist */
    int height, width, is_gt;
    cf_post_out_t out;
} cf_post_fuse_t;

CF_API int cf_fusion_decode_post(const cf_decode_heads_t* heads, int B, int C, int H, int W, int K, int out_h, int out_w,
                                 int norm2d, const cf_decode_out_t* out, const cf_post_fuse_t* post, void* ws,
                                 size_t ws_bytes, void* stream);
/* cf_fusion_stage_split (declared above) with the same option */
CF_API int cf_fusion_stage_post(const float* heat, const float* wh, const float* depth, const float* dim,
                                const float* rot, const float* pc_dep, const float* pc_dep_vel, const float* calib, int B,
                                int C, int H, int W, int K, float max_pc_dist, float* pc_hm, int64_t pc_hm_batch_stride,
                                int32_t* assoc, float* thr, int64_t* inds, const cf_decode_heads_t* dec_heads, int out_h,
                                int out_w, int norm2d, const cf_decode_out_t* dec_out, const cf_post_fuse_t* post,
                                void* ws, size_t ws_bytes, void* stream);

/* ---- a5 + a6  radar pillar expansion -> pc_dep -----------------------------------------------------------
 * loadRadarPointCloud filters      src/lib/dataset/datasets/nuscenes.py:172-205
 * map_pointcloud_to_image          src/lib/utils/pointcloud.py:17-49
 * processPointCloud (pillars)      src/lib/dataset/generic_dataset.py:738-942
 * drawPcHeat                       src/lib/dataset/datasets/nuscenes.py:234-263
 * points (B,Nmax,5) fp64 [X right, Y down, Z forward, vx, vz] in the camera frame, each frame sorted
 * ascending by Z by the caller (np.argsort in the reference is unstable: order is an input contract);
 * counts (B) i32; cam (B,3,3) fp64 intrinsic; calib (B,3,4) fp64; A (B,2,3) fp64 output affine;
 * pillar_dims_host: 3 doubles (h,w,l) on the HOST.  Writes pc_dep (B,3,H,W) fp32 completely; owner (B,H,W)
 * i32 optional = index of the surviving point per pixel or -1.
 * ws: cf_pillars_workspace_bytes(B, Nmax) bytes = the paint records (32 bytes per point) + a dense 8-byte copy of their
 * rectangles that the paint stages from; a workspace of B * Nmax * 32 bytes alone is accepted too (the paint then reads
 * the rectangles out of the records: same result, slower). */
CF_API size_t cf_pillars_workspace_bytes(int B, int Nmax);
CF_API int cf_radar_pillars(const double* points, const int32_t* counts, int B, int Nmax, const double* cam,
                            const double* calib, const double* A, int img_w, int img_h, int H, int W,
                            const double* pillar_dims_host, double max_dist, double z_offset, float* pc_dep,
                            int32_t* owner, void* ws, size_t ws_bytes, void* stream);
/* ... for flipped frames (nuscenes.py:201-205): flip (B) i32 or NULL.  A flipped frame's points arrive with X and vx
 * negated (cf_radar_ingest); the kernel projects and masks the point as loaded, mirrors the pixel column
 * (img_w - 1 - u) and sizes the pillar at the flipped X, exactly the reference's order of operations. */
CF_API int cf_radar_pillars_flip(const double* points, const int32_t* counts, const int32_t* flip, int B, int Nmax,
                                 const double* cam, const double* calib, const double* A, int img_w, int img_h, int H,
                                 int W, const double* pillar_dims_host, double max_dist, double z_offset, float* pc_dep,
                                 int32_t* owner, void* ws, size_t ws_bytes, void* stream);

/* Number of kernels the last successful call on this thread launched (bench.py's gpu_launches). */
CF_API int cf_last_launch_count(void);

/* Measurement hook (bench.py's roofline leg): while set on this thread, every call records `ev_start` on the
 * call's stream right before kernel `kernel_id` and `ev_stop` right after it (both cudaEvent_t created by the
 * caller with timing enabled).  kernel_id < 0 or NULL events clear the hook.  No reference counterpart. */
enum {
    CF_KERNEL_ASSOC = 2,        /* box_assoc_kernel: per-ROI association of given boxes (cf_frustum_boxes) */
    CF_KERNEL_PAINT = 3,        /* paint_kernel: ordered pc_hm / pc_dep paint (cf_frustum_boxes, cf_radar_pillars) */
    CF_KERNEL_PILLAR_RECTS = 4, /* pillar_rects_kernel: radar point -> pillar rectangle */
    CF_KERNEL_STAGE = 6,        /* fusion_stage_kernel: one cluster per frame, the whole stage in one launch */
    CF_KERNEL_POST = 7          /* post_process_kernel */
};
CF_API int cf_set_kernel_events(int kernel_id, void* ev_start, void* ev_stop);

/* Execution plan of the last fused-stage launch on this thread (cf_peak_topk, cf_fusion_decode, cf_frustum_pc_hm,
 * cf_fusion_stage*): the host picks cluster size and residency per (device, shape) -- tests assert it so that a
 * planner change cannot silently move them off the benchmarked path. */
typedef struct {
    int cluster_size;       /* CTAs per frame (thread-block cluster) */
    int n_clusters;         /* clusters launched; a cluster takes frames b = id, id + n_clusters, ... */
    int rounds;             /* ceil(B / n_clusters) */
    int resident;           /* 1: each CTA's heat-map share stays in shared memory between the two looks; 0: ring */
    int depth_plane_staged; /* 1: pc_dep[b,0] is staged in shared memory for the ROI scans; 0: scanned from L2 */
    int chunk_rows, slots;  /* rows per bulk copy, shared-memory slots */
    int smem_bytes;         /* dynamic shared memory per CTA */
} cf_plan_info_t;
CF_API int cf_last_plan(cf_plan_info_t* out);

/* Test hook: pin the cluster size (0 = planner's choice) and cap the co-resident clusters (0 = no cap) of later
 * fused-stage launches of this process, so that tests reach every execution plan on any batch size. */
CF_API int cf_debug_set_plan(int force_cluster_size, int max_clusters);

#ifdef CFD3D_PROFILE
/* Development build only (nvcc -DCFD3D_PROFILE, `python -m centerfusiondetect3d_b200.build --prof`): the in-kernel
 * phase clocks of the last launch (48 x int64 per CTA, first n_cta CTAs) into host memory. */
CF_API int cf_debug_read_prof(long long* out_host, int n_cta);
#endif

#ifdef __cplusplus
}
#endif
#endif /* CFD3D_B200_H */
