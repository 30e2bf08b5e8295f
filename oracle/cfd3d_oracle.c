/*
 * cfd3d_oracle.c -- CPU ORACLE for the CenterFusion frustum fusion stage.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this file.  The product path
 * (centerfusiondetect3d_b200/) never links, imports or calls it and fails loudly if its CUDA
 * library is missing.
 *
 * What it is: a plain-C, scalar restatement of the algorithm the reference implements in
 * Python/PyTorch/NumPy for this path.  Each function cites the reference file:line it follows
 * (paths relative to the reference tree, src/lib/...).  No reference source is copied; the
 * reference is vectorised tensor code, this is loops.
 *
 * Parity pinning: the reference ships NO tests, golden vectors or fixtures for this path
 * (SURVEY.md section 4 / 8c), so the oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF:
 * tests/golden/make_golden.py imports the unmodified reference modules in the build container and
 * stores input/output vectors in tests/golden/ (npz files); tests/test_oracle_golden.py checks every
 * function below against them (indices / associations / painted pixels bit-exact, fp32 values to
 * 1e-6 relative where the reference's own libm (SLEEF) is not correctly rounded).
 *
 * Arithmetic conventions (documented because associations hinge on strict fp32 compares):
 *   - fp32 everywhere the reference is fp32, one rounding per written operation, no FMA
 *     contraction (compile with -ffp-contract=off);
 *   - trigonometry evaluated in fp64 libm and rounded once to fp32 (the reference's torch-CPU fp32
 *     atan2/sin/cos are 1-ulp SLEEF kernels and not shape-stable, see SURVEY.md "Hard parts");
 *   - radar pillar geometry in fp64 exactly like the NumPy reference, rint() = half-to-even;
 *   - top-K total order: score descending, then class ascending, then flat index ascending
 *     (torch.topk leaves tie order unspecified; parity inputs are tie-free in the top-K region).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define CFO_API __attribute__((visibility("default")))

static int g_threads = 1;

CFO_API void cfo_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
CFO_API int cfo_get_threads(void) { return g_threads; }
CFO_API int cfo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------------------------
 * a1  nms(heatmap, kernel=3)                               model/utils.py:112-128
 * hmax = max_pool2d(3x3, stride 1, pad 1, -inf padding); keep = (hmax == heat); out = heat*keep.
 * ------------------------------------------------------------------------------------------ */
static inline float nms_pixel(const float* p, int H, int W, int y, int x) {
    float v = p[(size_t)y * W + x];
    float m = -INFINITY;
    for (int dy = -1; dy <= 1; ++dy) {
        int yy = y + dy;
        if (yy < 0 || yy >= H) continue;
        for (int dx = -1; dx <= 1; ++dx) {
            int xx = x + dx;
            if (xx < 0 || xx >= W) continue;
            float n = p[(size_t)yy * W + xx];
            if (n > m) m = n;
        }
    }
    float keep = (m == v) ? 1.0f : 0.0f;
    return v * keep;
}

CFO_API void cfo_nms(const float* heat, int B, int C, int H, int W, float* out) {
    size_t HW = (size_t)H * W;
#pragma omp parallel for num_threads(g_threads) schedule(static)
    for (long pc = 0; pc < (long)B * C; ++pc) {
        const float* p = heat + pc * HW;
        float* o = out + pc * HW;
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) o[(size_t)y * W + x] = nms_pixel(p, H, W, y, x);
    }
}

/* ------------------------------------------------------------------------------------------
 * a2  topk(heatmap, K)                                      model/utils.py:6-38
 * Reference: per-(b,c) torch.topk over H*W, then torch.topk over the C*K concatenation; class =
 * position // K; inds/ys/xs gathered.  Under a total order this equals the frame-wide top-K of
 * all C*H*W values, sorted descending.  Order: (score desc, class asc, index asc).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    float s;
    int32_t c;
    int32_t i;
} cand_t;

/* returns 1 when a ranks strictly before b (a is "better") */
static inline int cand_before(const cand_t* a, const cand_t* b) {
    if (a->s != b->s) return a->s > b->s;
    if (a->c != b->c) return a->c < b->c;
    return a->i < b->i;
}

static int cand_cmp_qsort(const void* pa, const void* pb) {
    const cand_t* a = (const cand_t*)pa;
    const cand_t* b = (const cand_t*)pb;
    if (cand_before(a, b)) return -1;
    if (cand_before(b, a)) return 1;
    return 0;
}

/* heap with the WORST kept candidate at the root */
static void heap_sift_down(cand_t* h, int n, int i) {
    for (;;) {
        int l = 2 * i + 1, r = l + 1, w = i;
        if (l < n && cand_before(&h[w], &h[l])) w = l;
        if (r < n && cand_before(&h[w], &h[r])) w = r;
        if (w == i) return;
        cand_t t = h[i];
        h[i] = h[w];
        h[w] = t;
        i = w;
    }
}

static void topk_frame(const float* heat, int C, int H, int W, int K, int apply_nms, cand_t* heap) {
    int n = 0;
    size_t HW = (size_t)H * W;
    for (int c = 0; c < C; ++c) {
        const float* p = heat + (size_t)c * HW;
        for (int y = 0; y < H; ++y) {
            for (int x = 0; x < W; ++x) {
                float v = apply_nms ? nms_pixel(p, H, W, y, x) : p[(size_t)y * W + x];
                cand_t cd = {v, c, (int32_t)(y * W + x)};
                if (n < K) {
                    heap[n++] = cd;
                    if (n == K)
                        for (int i = K / 2 - 1; i >= 0; --i) heap_sift_down(heap, K, i);
                } else if (cand_before(&cd, &heap[0])) {
                    heap[0] = cd;
                    heap_sift_down(heap, K, 0);
                }
            }
        }
    }
    qsort(heap, (size_t)n, sizeof(cand_t), cand_cmp_qsort);
}

CFO_API int cfo_topk(const float* heat, int B, int C, int H, int W, int K, int apply_nms, float* scores,
                     int64_t* inds, int32_t* classes, int64_t* ys, int64_t* xs) {
    if ((long)K > (long)H * W) return -1; /* torch.topk would raise */
    size_t CHW = (size_t)C * H * W;
#pragma omp parallel for num_threads(g_threads) schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        cand_t* heap = (cand_t*)malloc(sizeof(cand_t) * (size_t)K);
        topk_frame(heat + b * CHW, C, H, W, K, apply_nms, heap);
        for (int k = 0; k < K; ++k) {
            size_t o = (size_t)b * K + k;
            scores[o] = heap[k].s;
            inds[o] = heap[k].i;
            classes[o] = heap[k].c;
            ys[o] = heap[k].i / W;
            xs[o] = heap[k].i % W;
        }
        free(heap);
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * a3  transposeAndGetFeature(feature, indices)              model/utils.py:41-72
 * out[b,m,c] = feature[b,c,ind[b,m]]
 * ------------------------------------------------------------------------------------------ */
CFO_API void cfo_gather_nchw(const float* feat, int B, int ch, long HW, const int64_t* inds, int M, float* out) {
    for (int b = 0; b < B; ++b)
        for (int m = 0; m < M; ++m) {
            int64_t ind = inds[(size_t)b * M + m];
            for (int c = 0; c < ch; ++c) out[((size_t)b * M + m) * ch + c] = feat[((size_t)b * ch + c) * HW + ind];
        }
}

/* ------------------------------------------------------------------------------------------
 * a4  get_alpha                 utils/pointcloud.py:195-211
 *     cvtAlphaToYaw             utils/pointcloud.py:214-236
 *     get3DCorners              utils/pointcloud.py:239-296   (only the z row matters)
 *     getDistanceThresh         utils/pointcloud.py:299-328
 * ------------------------------------------------------------------------------------------ */
static const float PI_F = 3.14159265358979323846f; /* float(torch.pi) */

static inline float f32_atan2(float y, float x) { return (float)atan2((double)y, (double)x); }

CFO_API float cfo_get_alpha(const float* rot /*8*/) {
    /* pointcloud.py:207-210: blend a1*idx + a2*(~idx) with idx = rot[1] > rot[5] */
    float idx = rot[1] > rot[5] ? 1.0f : 0.0f;
    float nidx = 1.0f - idx;
    float a1 = f32_atan2(rot[2], rot[3]) + (-0.5f * PI_F);
    float a2 = f32_atan2(rot[6], rot[7]) + (0.5f * PI_F);
    float t1 = a1 * idx;
    float t2 = a2 * nidx;
    return t1 + t2;
}

CFO_API float cfo_dist_thresh_one(const float* calib /*3x4*/, float cx, const float* dim /*h,w,l*/, float alpha) {
    /* pointcloud.py:230-235 -- note cx is in OUTPUT-map pixels, calib in image pixels (kept) */
    float yaw = alpha + f32_atan2(cx - calib[2], calib[0]);
    if (yaw > PI_F) yaw -= 2.0f * PI_F;
    if (yaw < -PI_F) yaw += 2.0f * PI_F;
    /* pointcloud.py:268-295: z row of R*[x;y;z], R[2,:] = [-s, 0, c];  x = +-l/2, z = +-w/2 */
    float c = (float)cos((double)yaw), s = (float)sin((double)yaw);
    float hx = 0.5f * dim[2], hz = 0.5f * dim[1];
    float ns = -s;
    float zmax = -INFINITY, zmin = INFINITY;
    for (int i = 0; i < 4; ++i) {
        float xc = (i & 1) ? -hx : hx;
        float zc = (i & 2) ? -hz : hz;
        float a = ns * xc;
        float b = c * zc;
        float z = a + b;
        if (z > zmax) zmax = z;
        if (z < zmin) zmin = z;
    }
    /* pointcloud.py:325-327: max - min/2  (operator precedence as written in the reference) */
    float half_min = zmin / 2.0f;
    return zmax - half_min;
}

CFO_API void cfo_get_alpha_batch(const float* rot, long n, float* alpha) {
    for (long i = 0; i < n; ++i) alpha[i] = cfo_get_alpha(rot + 8 * i);
}

CFO_API void cfo_dist_thresh(const float* calib /*B,3,4*/, const float* center /*B,K,2*/, const float* dim /*B,K,3*/,
                             const float* alpha /*B,K*/, int B, int K, float* thr) {
    for (int b = 0; b < B; ++b)
        for (int k = 0; k < K; ++k) {
            size_t o = (size_t)b * K + k;
            thr[o] = cfo_dist_thresh_one(calib + 12 * b, center[2 * o], dim + 3 * o, alpha[o]);
        }
}

/* ------------------------------------------------------------------------------------------
 * a8  cvtPcDepthToHeatmap       utils/pointcloud.py:397-481
 * Python slice semantics reproduced by pyslice().
 * ------------------------------------------------------------------------------------------ */
static inline void pyslice(long lo, long hi, long n, long* a, long* b) {
    if (lo < 0) {
        lo += n;
        if (lo < 0) lo = 0;
    } else if (lo > n)
        lo = n;
    if (hi < 0) {
        hi += n;
        if (hi < 0) hi = 0;
    } else if (hi > n)
        hi = n;
    *a = lo;
    *b = hi; /* empty when lo >= hi */
}

/* returns flat index y*W+x of the matched pc_dep pixel, or -1 */
CFO_API int32_t cfo_pc_dep_to_hm(float* pc_hm /*3,H,W in place*/, const float* pc_dep /*3,H,W*/, int H, int W,
                                 float depth, const float* bbox /*x1,y1,x2,y2*/, float thr, float max_pc_dist) {
    size_t HW = (size_t)H * W;
    float x1 = bbox[0], y1 = bbox[1], x2 = bbox[2], y2 = bbox[3];
    /* :424-426 */
    float sx = x1 + x2, sy = y1 + y2;
    float cx = sx / 2.0f, cy = sy / 2.0f;
    /* :427-437 */
    long bx0 = (long)floorf(x1), by0 = (long)floorf(y1), bx1 = (long)ceilf(x2), by1 = (long)ceilf(y2);
    long ya, yb, xa, xb;
    pyslice(by0, by1 + 1, H, &ya, &yb);
    pyslice(bx0, bx1 + 1, W, &xa, &xb);
    /* :444-461 */
    float hi = depth + thr;
    float lo = depth - thr;
    if (!(lo > 0.0f)) lo = 0.0f; /* python max(0, t) */
    int found = 0;
    float best = 0.f;
    long byy = 0, bxx = 0;
    for (long yy = ya; yy < yb; ++yy)
        for (long xx = xa; xx < xb; ++xx) {
            float v = pc_dep[(size_t)yy * W + xx];
            if (v != 0.0f && v < hi && v > lo && (!found || v < best)) {
                found = 1;
                best = v;
                byy = yy;
                bxx = xx;
            }
        }
    if (!found) return -1;
    size_t mp = (size_t)byy * W + bxx;
    float dist = best / max_pc_dist; /* :462-466 */
    float vx = pc_dep[HW + mp], vz = pc_dep[2 * HW + mp];
    /* :468-476 */
    float w = x2 - x1;
    float wi = 0.3f * w;
    float hw_ = wi / 2.0f;
    long wmin = (long)(cx - hw_), wmax = (long)(cx + hw_); /* int(): trunc toward zero */
    float h = y2 - y1;
    float hi_ = 0.3f * h;
    float hh_ = hi_ / 2.0f;
    long hmin = (long)(cy - hh_), hmax = (long)(cy + hh_);
    /* :478-481  rows h_min:h_max+1, cols w_min:w_max+1+1 */
    long ra, rb, ca, cb;
    pyslice(hmin, hmax + 1, H, &ra, &rb);
    pyslice(wmin, wmax + 2, W, &ca, &cb);
    for (long yy = ra; yy < rb; ++yy)
        for (long xx = ca; xx < cb; ++xx) {
            size_t p = (size_t)yy * W + xx;
            pc_hm[p] = dist;
            pc_hm[HW + p] = vx;
            pc_hm[2 * HW + p] = vz;
        }
    return (int32_t)mp;
}

/* f4: the NumPy branch of cvtPcDepthToHeatmap as the DataLoader calls it (dataset/generic_dataset.py:673-687;
 * utils/pointcloud.py:411-481 with lib = np), under NumPy >= 2 promotion rules (NEP 50; the container has 2.3):
 *   - pc_dep / pc_hm are float32 maps; depth is a Python float (weak).  getDistanceThresh's NumPy branch returns
 *     np.float32 (its rotation matrix and corner arrays are float32, pointcloud.py:268-295), so in the DataLoader
 *     depth +- thr is float32 arithmetic and the compares are float32 (thr_f32 != 0); a float64 threshold
 *     (thr_f32 == 0) makes them float64 on the exactly widened pixel;
 *   - bbox is float32 in the DataLoader (transformBbox, generic_dataset.py:507-526, builds it from float32 scalars):
 *     centre, 0.3 * w and the int() of the paint window are float32 arithmetic (Python floats are weak), exactly the
 *     torch branch (bbox_f32 != 0);  a float64 bbox (bbox_f32 == 0) keeps the centre in float32 (dtype=np.float32)
 *     but computes w, 0.3 * w and centre - interval / 2 in float64;
 *   - dist /= max_pc_dist stays a float32 division.
 * Returns the matched flat pixel or -1. */
CFO_API int32_t cfo_pc_dep_to_hm_np(float* pc_hm, const float* pc_dep, int H, int W, double depth,
                                    const double* bbox /*x1,y1,x2,y2*/, double thr, float max_pc_dist, int bbox_f32, int thr_f32) {
    size_t HW = (size_t)H * W;
    double x1 = bbox[0], y1 = bbox[1], x2 = bbox[2], y2 = bbox[3];
    float cx, cy;
    if (bbox_f32) {
        cx = ((float)x1 + (float)x2) / 2.0f;
        cy = ((float)y1 + (float)y2) / 2.0f;
    } else {
        cx = (float)((x1 + x2) / 2.0);
        cy = (float)((y1 + y2) / 2.0);
    }
    long bx0 = (long)floor(x1), by0 = (long)floor(y1), bx1 = (long)ceil(x2), by1 = (long)ceil(y2);
    long ya, yb, xa, xb;
    pyslice(by0, by1 + 1, H, &ya, &yb);
    pyslice(bx0, bx1 + 1, W, &xa, &xb);
    double hi = depth + thr;
    double lo = depth - thr;
    if (thr_f32) { /* Python float (weak) +- np.float32 -> float32 arithmetic, float32 compares */
        hi = (double)((float)depth + (float)thr);
        lo = (double)((float)depth - (float)thr);
    }
    if (!(lo > 0.0)) lo = 0.0; /* python max(0, t) */
    int found = 0;
    float best = 0.f;
    long byy = 0, bxx = 0;
    for (long yy = ya; yy < yb; ++yy)
        for (long xx = xa; xx < xb; ++xx) {
            float v = pc_dep[(size_t)yy * W + xx];
            if (v != 0.0f && (double)v < hi && (double)v > lo && (!found || v < best)) {
                found = 1;
                best = v;
                byy = yy;
                bxx = xx;
            }
        }
    if (!found) return -1;
    size_t mp = (size_t)byy * W + bxx;
    float dist = best / max_pc_dist;
    float vx = pc_dep[HW + mp], vz = pc_dep[2 * HW + mp];
    long wmin, wmax, hmin, hmax;
    if (bbox_f32) {
        float w = (float)x2 - (float)x1, h = (float)y2 - (float)y1;
        float hw_ = (0.3f * w) / 2.0f, hh_ = (0.3f * h) / 2.0f;
        wmin = (long)(cx - hw_), wmax = (long)(cx + hw_);
        hmin = (long)(cy - hh_), hmax = (long)(cy + hh_);
    } else {
        double hw_ = (0.3 * (x2 - x1)) / 2.0, hh_ = (0.3 * (y2 - y1)) / 2.0;
        wmin = (long)((double)cx - hw_), wmax = (long)((double)cx + hw_);
        hmin = (long)((double)cy - hh_), hmax = (long)((double)cy + hh_);
    }
    long ra, rb, ca, cb;
    pyslice(hmin, hmax + 1, H, &ra, &rb);
    pyslice(wmin, wmax + 2, W, &ca, &cb);
    for (long yy = ra; yy < rb; ++yy)
        for (long xx = ca; xx < cb; ++xx) {
            size_t p = (size_t)yy * W + xx;
            pc_hm[p] = dist;
            pc_hm[HW + p] = vx;
            pc_hm[2 * HW + p] = vz;
        }
    return (int32_t)mp;
}

/* ------------------------------------------------------------------------------------------
 * a7  getPcFrustumHeatmap       utils/pointcloud.py:331-394
 * topk on the RAW heat-map (no nms, :347), gathers (:353-371), boxes (:359-367), alpha, centre,
 * distance threshold, then the ordered B x K association loop (:383-392).
 * Optional extra outputs (may be NULL): assoc (B,K) int32 matched flat pixel or -1; thr (B,K).
 * ------------------------------------------------------------------------------------------ */
CFO_API int cfo_frustum_pc_hm(const float* heat, const float* wh, const float* depth, const float* dim,
                              const float* rot, const float* pc_dep, const float* calib, int B, int C, int H,
                              int W, int K, float max_pc_dist, float* pc_hm, int32_t* assoc, float* thr_out,
                              int64_t* inds_out) {
    size_t HW = (size_t)H * W;
    if ((long)K > (long)HW) return -1;
    memset(pc_hm, 0, sizeof(float) * 3 * HW * (size_t)B);
    int rc = 0;
#pragma omp parallel for num_threads(g_threads) schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        cand_t* heap = (cand_t*)malloc(sizeof(cand_t) * (size_t)K);
        topk_frame(heat + (size_t)b * C * HW, C, H, W, K, 0, heap);
        for (int k = 0; k < K; ++k) {
            size_t ind = (size_t)heap[k].i;
            float x = (float)(heap[k].i % W) + 0.5f; /* :348-349 */
            float y = (float)(heap[k].i / W) + 0.5f;
            float w = wh[((size_t)b * 2 + 0) * HW + ind];
            float h = wh[((size_t)b * 2 + 1) * HW + ind];
            if (w < 0.0f) w = 0.0f; /* :358 */
            if (h < 0.0f) h = 0.0f;
            float hw2 = w / 2.0f, hh2 = h / 2.0f;
            float bbox[4] = {x - hw2, y - hh2, x + hw2, y + hh2}; /* :359-367 */
            float r[8], dm[3];
            for (int j = 0; j < 8; ++j) r[j] = rot[((size_t)b * 8 + j) * HW + ind];
            for (int j = 0; j < 3; ++j) dm[j] = dim[((size_t)b * 3 + j) * HW + ind];
            float d = depth[(size_t)b * HW + ind];
            float alpha = cfo_get_alpha(r);
            float sx = bbox[0] + bbox[2];
            float cx = sx / 2.0f; /* :373-379 */
            float t = cfo_dist_thresh_one(calib + 12 * (size_t)b, cx, dm, alpha);
            int32_t a = cfo_pc_dep_to_hm(pc_hm + 3 * HW * (size_t)b, pc_dep + 3 * HW * (size_t)b, H, W, d, bbox, t,
                                         max_pc_dist);
            size_t o = (size_t)b * K + k;
            if (assoc) assoc[o] = a;
            if (thr_out) thr_out[o] = t;
            if (inds_out) inds_out[o] = heap[k].i;
        }
        free(heap);
    }
    return rc;
}

/* ------------------------------------------------------------------------------------------
 * a9  fusionDecode(outputs, outputSize, K, norm2d)           model/decode.py:10-174
 * Single pyramid layer (the reference's DLASeg never emits more, dla.py:579).  Any head pointer
 * may be NULL (= key absent).  `depth` is depth2 when present else depth (:101-110); `rot` is
 * rotation2 when present else rotation (:120-121) -- the caller resolves which pointer to pass.
 * ------------------------------------------------------------------------------------------ */
CFO_API int cfo_fusion_decode(const float* heat, const float* reg, const float* wh, const float* depth,
                              const float* rot, const float* dim, const float* amodal, const float* natt,
                              const float* vel, const float* unc, int B, int C, int H, int W, int K, int out_h,
                              int out_w, int norm2d, float* scores, float* class_ids, float* centers,
                              float* bboxes, float* o_depth, float* o_rot, float* o_dim, float* o_amodal,
                              float* o_natt, float* o_vel, int64_t* inds_out) {
    size_t HW = (size_t)H * W;
    if ((long)K > (long)HW) return -1;
#pragma omp parallel for num_threads(g_threads) schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        cand_t* heap = (cand_t*)malloc(sizeof(cand_t) * (size_t)K);
        topk_frame(heat + (size_t)b * C * HW, C, H, W, K, 1, heap); /* :38-39 nms then topk */
        for (int k = 0; k < K; ++k) {
            size_t o = (size_t)b * K + k;
            size_t ind = (size_t)heap[k].i;
            float sc = heap[k].s;
            if (unc) { /* :80-85 */
                float u = unc[(size_t)b * HW + ind];
                sc = sc * expf(-expf(u));
            }
            scores[o] = sc;
            class_ids[o] = (float)heap[k].c;
            float xn = (float)(heap[k].i % W) / (float)W; /* :40-41 */
            float yn = (float)(heap[k].i / W) / (float)H;
            centers[2 * o + 0] = xn;
            centers[2 * o + 1] = yn;
            float xs = xn * (float)out_w, ys = yn * (float)out_h; /* :132-133 */
            if (reg) {
                xs = xs + reg[((size_t)b * 2 + 0) * HW + ind];
                ys = ys + reg[((size_t)b * 2 + 1) * HW + ind];
            } else {
                xs = xs + 0.5f;
                ys = ys + 0.5f;
            }
            if (wh) { /* :148-163 */
                float w = wh[((size_t)b * 2 + 0) * HW + ind];
                float h = wh[((size_t)b * 2 + 1) * HW + ind];
                if (w < 0.0f) w = 0.0f;
                if (h < 0.0f) h = 0.0f;
                if (norm2d) {
                    w = w * (float)out_w;
                    h = h * (float)out_h;
                }
                float hw2 = w / 2.0f, hh2 = h / 2.0f;
                bboxes[4 * o + 0] = xs - hw2;
                bboxes[4 * o + 1] = ys - hh2;
                bboxes[4 * o + 2] = xs + hw2;
                bboxes[4 * o + 3] = ys + hh2;
            }
            if (depth) o_depth[o] = depth[(size_t)b * HW + ind];
            if (rot)
                for (int j = 0; j < 8; ++j) o_rot[8 * o + j] = rot[((size_t)b * 8 + j) * HW + ind];
            if (dim)
                for (int j = 0; j < 3; ++j) o_dim[3 * o + j] = dim[((size_t)b * 3 + j) * HW + ind];
            if (amodal) { /* :165 denormalise */
                float ax = amodal[((size_t)b * 2 + 0) * HW + ind];
                float ay = amodal[((size_t)b * 2 + 1) * HW + ind];
                if (norm2d) {
                    ax = ax * (float)out_w;
                    ay = ay * (float)out_h;
                }
                o_amodal[2 * o + 0] = ax;
                o_amodal[2 * o + 1] = ay;
            }
            if (natt)
                for (int j = 0; j < 8; ++j) o_natt[8 * o + j] = natt[((size_t)b * 8 + j) * HW + ind];
            if (vel)
                for (int j = 0; j < 3; ++j) o_vel[3 * o + j] = vel[((size_t)b * 3 + j) * HW + ind];
            if (inds_out) inds_out[o] = heap[k].i;
        }
        free(heap);
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * a5  loadRadarPointCloud filters      dataset/datasets/nuscenes.py:172-205
 *     map_pointcloud_to_image          utils/pointcloud.py:17-49 (+ devkit view_points)
 * a6  transformPointCloud              dataset/generic_dataset.py:830-867
 *     getPcPillarsSize                 dataset/generic_dataset.py:869-942
 *     get3dBox / project3DPoints       utils/ddd.py:8-55  (NumPy branch; corner offsets are fp32)
 *     processPointCloud pillar loop    dataset/generic_dataset.py:798-826
 *     drawPcHeat                       dataset/datasets/nuscenes.py:259-261
 *
 * points: (B, Nmax, 5) fp64 [X right, Y down, Z forward, vx, vz] in the camera frame, ALREADY
 * sorted ascending by Z per frame by the caller (np.argsort is unstable; order is an input
 * contract).  counts (B).  cam (B,3,3) intrinsic used by view_points; calib (B,3,4) used for the
 * pillar corners; A (B,2,3) output affine.  pc_dep (B,3,H,W) fp32 out.  owner (B,H,W) int32
 * out (optional): index of the surviving point per pixel, or -1.
 * ------------------------------------------------------------------------------------------ */
CFO_API void cfo_radar_pillars(const double* points, const int32_t* counts, const int32_t* flip /*B or NULL*/, int B,
                               int Nmax, const double* cam, const double* calib, const double* A, int img_w, int img_h,
                               int H, int W, const double* pillar_dims /*h,w,l*/, double max_dist, double z_offset,
                               float* pc_dep, int32_t* owner) {
    size_t HW = (size_t)H * W;
    memset(pc_dep, 0, sizeof(float) * 3 * HW * (size_t)B);
    if (owner)
        for (size_t i = 0; i < HW * (size_t)B; ++i) owner[i] = -1;
    /* ddd.get3dBox -> pointcloud.get3DCorners NumPy branch: fp32 arrays scaled in place by fp64
       dims, i.e. offsets are float32(0.5*l), -h, float32(0.5*w) */
    const double ox = (double)(float)(0.5 * pillar_dims[2]);
    const double oy = (double)(float)(-1.0 * pillar_dims[0]);
    const double oz = (double)(float)(0.5 * pillar_dims[1]);
#pragma omp parallel for num_threads(g_threads) schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        const double* Kc = cam + 9 * (size_t)b;
        const double* P = calib + 12 * (size_t)b;
        const double* Ab = A + 6 * (size_t)b;
        float* dep = pc_dep + 3 * HW * (size_t)b;
        int n = counts[b];
        for (int i = 0; i < n; ++i) {
            const double* pt = points + ((size_t)b * Nmax + i) * 5;
            double X = pt[0], Y = pt[1], Z = pt[2];
            if (max_dist > 0 && !(Z <= max_dist)) continue; /* nuscenes.py:177-180 */
            Y -= z_offset;                                  /* :183-184 */
            /* a flipped frame (nuscenes.py:201-205) arrives with X and vx negated; the reference projects and masks
               the point as loaded (:187-191), THEN mirrors the pixel column (:202) and the 3-D X (:203) */
            const int flipped = flip && flip[b];
            const double Xp = flipped ? -X : X;
            /* view_points, normalize=True */
            double uh = Kc[0] * Xp + Kc[1] * Y + Kc[2] * Z;
            double vh = Kc[3] * Xp + Kc[4] * Y + Kc[5] * Z;
            double zh = Kc[6] * Xp + Kc[7] * Y + Kc[8] * Z;
            double u = uh / zh, v = vh / zh;
            /* pointcloud.py:40-45 */
            if (!(Z > 0 && u > 1 && u < img_w - 1 && v > 1 && v < img_h - 1)) continue;
            if (flipped) u = (double)(img_w - 1) - u; /* img.shape[1] - 1 - pc_2d[0] */
            /* transformPointCloud :853-862 */
            double up = Ab[0] * u + Ab[1] * v + Ab[2];
            double vp = Ab[3] * u + Ab[4] * v + Ab[5];
            if (!(up < W && vp < H && 0 < up && 0 < vp)) continue;
            /* getPcPillarsSize */
            double umin = INFINITY, umax = -INFINITY, vmin = INFINITY, vmax = -INFINITY;
            for (int c = 0; c < 8; ++c) {
                /* pointcloud.py:280-292 sign pattern */
                double cx = X + (((c >> 1) & 1) ? -ox : ox);
                double cy = Y + ((c & 4) ? oy : 0.0);
                double cz = Z + ((((c + 1) >> 1) & 1) ? -oz : oz);
                double q0 = P[0] * cx + P[1] * cy + P[2] * cz + P[3];
                double q1 = P[4] * cx + P[5] * cy + P[6] * cz + P[7];
                double q2 = P[8] * cx + P[9] * cy + P[10] * cz + P[11];
                double uu = q0 / q2, vv = q1 / q2;
                double ua = Ab[0] * uu + Ab[1] * vv + Ab[2];
                double va = Ab[3] * uu + Ab[4] * vv + Ab[5];
                if (ua < umin) umin = ua;
                if (ua > umax) umax = ua;
                if (va < vmin) vmin = va;
                if (va > vmax) vmax = va;
            }
            double pw = umax - umin, ph = vmax - vmin;
            /* generic_dataset.py:804-809, :822 */
            double by1 = vp - ph;
            if (!(by1 > 0)) by1 = 0; /* python max(a, 0) -> a if a > 0 else 0 ... */
            double bx1 = up - pw / 2;
            if (!(bx1 > 0)) bx1 = 0;
            double bx2 = up + pw / 2;
            if (!(bx2 < W)) bx2 = W;
            long y1 = (long)rint(by1), y2 = (long)rint(vp), x1 = (long)rint(bx1), x2 = (long)rint(bx2);
            if (y2 > H) y2 = H; /* numpy slice clamps */
            if (x2 > W) x2 = W;
            float fd = (float)Z, fvx = (float)pt[3], fvz = (float)pt[4];
            for (long yy = y1; yy < y2; ++yy)
                for (long xx = x1; xx < x2; ++xx) {
                    size_t p = (size_t)yy * W + xx;
                    dep[p] = fd;
                    dep[HW + p] = fvx;
                    dep[2 * HW + p] = fvz;
                    if (owner) owner[HW * (size_t)b + p] = i;
                }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * f1  postProcess               utils/postProcess.py:13-85
 *     affineTransform           utils/image.py:86-109   (2x3 fp32 matrix, points (x, y, 1))
 *     cvtImgToCamCoord          utils/ddd.py:176-199 with project2DTo3D :146-172, alpha2rot_y :122-142
 *     get3dBox / get3DCorners   utils/ddd.py:8-23, utils/pointcloud.py:239-296
 * One detection per iteration; any input pointer except class_ids / centers may be NULL (= key absent in
 * the reference's dict).  class_ids and velocity are updated in place like the reference does.
 * Arithmetic: fp32, one rounding per written operation, sums left to right (the reference's torch.mm /
 * einsum leave the order to the BLAS: parity with the reference is a tolerance, SURVEY.md "Hard parts").
 * ------------------------------------------------------------------------------------------ */
static inline void affine_pt(const float* t, float x, float y, float* ox, float* oy) {
    float a = t[0] * x, b = t[1] * y;
    float s = a + b;
    *ox = s + t[2];
    a = t[3] * x;
    b = t[4] * y;
    s = a + b;
    *oy = s + t[5];
}

CFO_API void cfo_post_process(float* class_ids, const float* centers_in, const float* bboxes_in, const float* depth,
                              const float* rot, const float* dim, const float* amodal, float* velocity, int B, int K,
                              int out_h, int out_w, const float* trans /*2x3*/, const float* calib /*B,3,4*/, int is_gt,
                              float* centers, float* bboxes, float* alpha_out, float* locations, float* yaws,
                              float* bboxes3d) {
    const int has3d = rot && depth && dim;
    for (long i = 0; i < (long)B * K; ++i) {
        const float* cal = calib + 12 * (i / K);
        class_ids[i] += 1.0f;                                            /* :34 */
        float cx = centers_in[2 * i] * (float)out_w, cy = centers_in[2 * i + 1] * (float)out_h; /* :35-37 */
        float bb[4] = {0, 0, 0, 0};
        if (bboxes_in) {                                                 /* :39-42 */
            affine_pt(trans, bboxes_in[4 * i], bboxes_in[4 * i + 1], &bb[0], &bb[1]);
            affine_pt(trans, bboxes_in[4 * i + 2], bboxes_in[4 * i + 3], &bb[2], &bb[3]);
            for (int j = 0; j < 4; ++j) bboxes[4 * i + j] = bb[j];
        }
        float alpha = 0.0f;
        if (rot) {                                                       /* :47-48 */
            alpha = cfo_get_alpha(rot + 8 * i);
            alpha_out[i] = alpha;
        }
        if (has3d) {
            if (!is_gt && amodal) {                                      /* :54-59 */
                float ax = cx + amodal[2 * i], ay = cy + amodal[2 * i + 1];
                affine_pt(trans, ax, ay, &cx, &cy);
            } else if (!is_gt && bboxes_in) {                            /* :62-64 */
                float sx = bb[0] + bb[2], sy = bb[1] + bb[3];
                cx = sx / 2.0f;
                cy = sy / 2.0f;
            }
            /* project2DTo3D, ddd.py:160-166 */
            const float d = depth[i];
            float z = d - cal[11];
            float x = cx * d;
            x = x - cal[3];
            float t = cal[2] * z;
            x = x - t;
            x = x / cal[0];
            float y = cy * d;
            y = y - cal[7];
            t = cal[6] * z;
            y = y - t;
            y = y / cal[5];
            float hh = dim[3 * i] / 2.0f;                                /* ddd.py:197 */
            y = y + hh;
            float yaw = alpha + f32_atan2(cx - cal[2], cal[0]);          /* ddd.py:137-141 */
            if (yaw > PI_F) yaw -= 2.0f * PI_F;
            if (yaw < -PI_F) yaw += 2.0f * PI_F;
            locations[3 * i] = x;
            locations[3 * i + 1] = y;
            locations[3 * i + 2] = z;
            yaws[i] = yaw;
            const float c = (float)cos((double)yaw), s = (float)sin((double)yaw);
            if (!is_gt && velocity) {                                    /* :71-76 */
                float v0 = velocity[3 * i], v2 = velocity[3 * i + 2];
                float q0 = v0 * v0, q2 = v2 * v2;
                float V = sqrtf(q0 + q2);
                velocity[3 * i] = c * V;
                velocity[3 * i + 2] = (-s) * V;
            }
            /* get3DCorners + location, zeroed when any dimension <= 0  (:78-81) */
            const float h = dim[3 * i], w = dim[3 * i + 1], l = dim[3 * i + 2];
            const int dead = h <= 0.0f || w <= 0.0f || l <= 0.0f;
            for (int k = 0; k < 8; ++k) {
                float xc = ((k & 2) ? -0.5f : 0.5f) * l;
                float yc = (k & 4) ? h * -1.0f : 0.0f;
                float zc = (((k + 1) & 2) ? -0.5f : 0.5f) * w;
                float a = c * xc, b2 = s * zc;
                float X = a + b2;
                a = (-s) * xc;
                b2 = c * zc;
                float Z = a + b2;
                float* o = bboxes3d + (24 * i + 3 * k);
                o[0] = dead ? 0.0f : X + x;
                o[1] = dead ? 0.0f : yc + y;
                o[2] = dead ? 0.0f : Z + z;
            }
        }
        centers[2 * i] = cx;
        centers[2 * i + 1] = cy;
    }
}
