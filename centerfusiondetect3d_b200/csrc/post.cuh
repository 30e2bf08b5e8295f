// post.cuh -- postProcess of ONE decoded detection by one thread: the arithmetic shared by post_process_kernel (post.cu)
// and the decode epilogue of the fused kernels (merge.cuh), which runs it on the fourth lane of a detection's lane quad
// so that the detections leave the launch final (row f1: one launch fewer, no second pass over the (B,K) records).
//
// Replaces (reference, src/lib/): utils/postProcess.py:13-85; utils/image.py:86-109 affineTransform;
// utils/ddd.py:122-199 alpha2rot_y / project2DTo3D / cvtImgToCamCoord; utils/ddd.py:8-23 get3dBox with
// utils/pointcloud.py:239-296 get3DCorners.
//
// Arithmetic contract (shared with oracle/cfd3d_oracle.c cfo_post_process): one rounding per written operation, sums
// left to right, trigonometry in fp64 rounded once.  The reference leaves the summation order of its 2x3 and 3x3
// products to the BLAS, so parity with the reference itself is the 1e-5 tolerance north_star states.
#pragma once
#include "geometry.cuh"
#include "kernels.cuh"

namespace cf {

// inputs of one detection (what the reference's dict holds for it after fusionDecode); `has` = PD_* bits of the keys
// that exist
enum { PD_BBOX = 1, PD_ROT = 2, PD_DEPTH = 4, PD_DIM = 8, PD_AMODAL = 16, PD_VEL = 32 };
struct PostDet {
    float cx, cy;  // centre, normalised [x/W, y/H]
    float bb[4], rot[8], depth, dim[3], amodal[2];
    float vel[3];  // in/out (:71-76)
    unsigned has;
};

__device__ __forceinline__ void affine_pt(const float* t, float x, float y, float& ox, float& oy) {
    ox = __fadd_rn(__fadd_rn(__fmul_rn(t[0], x), __fmul_rn(t[1], y)), t[2]);
    oy = __fadd_rn(__fadd_rn(__fmul_rn(t[3], x), __fmul_rn(t[4], y)), t[5]);
}

// Writes every `o` output of detection i; d.vel is updated in place (the caller stores it where the reference's
// in-place update lands).  cal = the frame's 3x4 calibration (read only when rot, depth and dim exist).  Inlined into
// callers that are out of line themselves (merge.cuh epilogue_post) or kernels of their own (post.cu).
__device__ __forceinline__ void post_detection(const PostFuse& pf, const float* cal, long long i, PostDet& d) {
    const cf_post_out_t& o = pf.out;
    const float* t = pf.t;
    const bool has3d = (d.has & (PD_ROT | PD_DEPTH | PD_DIM)) == (PD_ROT | PD_DEPTH | PD_DIM);
    float cx = __fmul_rn(d.cx, (float)pf.width), cy = __fmul_rn(d.cy, (float)pf.height);  // :36
    float bb[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (d.has & PD_BBOX) {                                                                // :39-42
        affine_pt(t, d.bb[0], d.bb[1], bb[0], bb[1]);
        affine_pt(t, d.bb[2], d.bb[3], bb[2], bb[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j) o.bboxes[4 * i + j] = bb[j];
    }
    float alpha = 0.0f;
    if (d.has & PD_ROT) {                                                                 // :47-48
        alpha = get_alpha(d.rot);
        o.alpha[i] = alpha;
    }
    if (has3d) {
        if (!pf.is_gt && (d.has & PD_AMODAL)) {                                           // :54-59
            const float ax = __fadd_rn(cx, d.amodal[0]), ay = __fadd_rn(cy, d.amodal[1]);
            affine_pt(t, ax, ay, cx, cy);
        } else if (!pf.is_gt && (d.has & PD_BBOX)) {                                      // :62-64
            cx = __fdiv_rn(__fadd_rn(bb[0], bb[2]), 2.0f);
            cy = __fdiv_rn(__fadd_rn(bb[1], bb[3]), 2.0f);
        }
        // project2DTo3D, ddd.py:160-166
        const float dp = d.depth;
        const float z = __fsub_rn(dp, cal[11]);
        float x = __fsub_rn(__fmul_rn(cx, dp), cal[3]);
        x = __fdiv_rn(__fsub_rn(x, __fmul_rn(cal[2], z)), cal[0]);
        float y = __fsub_rn(__fmul_rn(cy, dp), cal[7]);
        y = __fdiv_rn(__fsub_rn(y, __fmul_rn(cal[6], z)), cal[5]);
        const float h = d.dim[0], w = d.dim[1], l = d.dim[2];
        y = __fadd_rn(y, __fdiv_rn(h, 2.0f));                                             // ddd.py:197
        const float yaw = yaw_wrap(alpha, f32_atan2(__fsub_rn(cx, cal[2]), cal[0]));      // ddd.py:137-141
        o.locations[3 * i] = x;
        o.locations[3 * i + 1] = y;
        o.locations[3 * i + 2] = z;
        o.yaws[i] = yaw;
        const float c = (float)cos((double)yaw), s = (float)sin((double)yaw);
        const float ns = -s;
        if (!pf.is_gt && (d.has & PD_VEL)) {                                              // :71-76
            const float v0 = d.vel[0], v2 = d.vel[2];
            const float V = __fsqrt_rn(__fadd_rn(__fmul_rn(v0, v0), __fmul_rn(v2, v2)));
            d.vel[0] = __fmul_rn(c, V);
            d.vel[2] = __fmul_rn(ns, V);
        }
        // get3DCorners + location, zeroed when any dimension <= 0  (:78-81)
        const bool dead = h <= 0.0f || w <= 0.0f || l <= 0.0f;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float xc = __fmul_rn((k & 2) ? -0.5f : 0.5f, l);
            const float yc = (k & 4) ? __fmul_rn(h, -1.0f) : 0.0f;
            const float zc = __fmul_rn(((k + 1) & 2) ? -0.5f : 0.5f, w);
            const float X = __fadd_rn(__fmul_rn(c, xc), __fmul_rn(s, zc));
            const float Z = __fadd_rn(__fmul_rn(ns, xc), __fmul_rn(c, zc));
            float* q = o.bboxes3d + (24 * i + 3 * k);
            q[0] = dead ? 0.0f : __fadd_rn(X, x);
            q[1] = dead ? 0.0f : __fadd_rn(yc, y);
            q[2] = dead ? 0.0f : __fadd_rn(Z, z);
        }
    }
    o.centers[2 * i] = cx;
    o.centers[2 * i + 1] = cy;
}

}  // namespace cf
