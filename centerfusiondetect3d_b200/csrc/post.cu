// post.cu -- postProcess: decoded detections -> image-space boxes, camera-frame locations, yaw, re-oriented
// velocity and 8-corner 3-D boxes.  One thread per detection; pure elementwise fp32 on (B,K) records.
//
// Replaces (reference, src/lib/): utils/postProcess.py:13-85; utils/image.py:86-109 affineTransform;
// utils/ddd.py:122-199 alpha2rot_y / project2DTo3D / cvtImgToCamCoord; utils/ddd.py:8-23 get3dBox with
// utils/pointcloud.py:239-296 get3DCorners.
//
// Arithmetic contract (shared with oracle/cfd3d_oracle.c cfo_post_process): one rounding per written operation, sums
// left to right, trigonometry in fp64 rounded once.  The reference leaves the summation order of its 2x3 and 3x3
// products to the BLAS, so parity with the reference itself is the 1e-5 tolerance north_star states.
#include <math_constants.h>

#include "geometry.cuh"
#include "kernels.cuh"

namespace cf {

struct PostParams {
    cf_post_in_t in;
    cf_post_out_t out;
    int B, K, out_h, out_w, is_gt;
    float t[6];
    const float* calib;
};

__device__ __forceinline__ void affine_pt(const float* t, float x, float y, float& ox, float& oy) {
    ox = __fadd_rn(__fadd_rn(__fmul_rn(t[0], x), __fmul_rn(t[1], y)), t[2]);
    oy = __fadd_rn(__fadd_rn(__fmul_rn(t[3], x), __fmul_rn(t[4], y)), t[5]);
}

__global__ void __launch_bounds__(128) post_process_kernel(const PostParams p) {
    const long long i = blockIdx.x * 128ll + threadIdx.x;
    if (i >= (long long)p.B * p.K) return;
    const float* cal = p.calib + 12 * (i / p.K);
    const cf_post_in_t& in = p.in;
    const cf_post_out_t& o = p.out;
    const bool has3d = in.rot && in.depth && in.dim;
    in.class_ids[i] = __fadd_rn(in.class_ids[i], 1.0f);                                   // :34
    float cx = __fmul_rn(in.centers[2 * i], (float)p.out_w), cy = __fmul_rn(in.centers[2 * i + 1], (float)p.out_h);
    float bb[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (in.bboxes) {                                                                      // :39-42
        affine_pt(p.t, in.bboxes[4 * i], in.bboxes[4 * i + 1], bb[0], bb[1]);
        affine_pt(p.t, in.bboxes[4 * i + 2], in.bboxes[4 * i + 3], bb[2], bb[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j) o.bboxes[4 * i + j] = bb[j];
    }
    float alpha = 0.0f;
    if (in.rot) {                                                                         // :47-48
        float r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = in.rot[8 * i + j];
        alpha = get_alpha(r);
        o.alpha[i] = alpha;
    }
    if (has3d) {
        if (!p.is_gt && in.amodal) {                                                      // :54-59
            const float ax = __fadd_rn(cx, in.amodal[2 * i]), ay = __fadd_rn(cy, in.amodal[2 * i + 1]);
            affine_pt(p.t, ax, ay, cx, cy);
        } else if (!p.is_gt && in.bboxes) {                                               // :62-64
            cx = __fdiv_rn(__fadd_rn(bb[0], bb[2]), 2.0f);
            cy = __fdiv_rn(__fadd_rn(bb[1], bb[3]), 2.0f);
        }
        // project2DTo3D, ddd.py:160-166
        const float d = in.depth[i];
        const float z = __fsub_rn(d, cal[11]);
        float x = __fsub_rn(__fmul_rn(cx, d), cal[3]);
        x = __fdiv_rn(__fsub_rn(x, __fmul_rn(cal[2], z)), cal[0]);
        float y = __fsub_rn(__fmul_rn(cy, d), cal[7]);
        y = __fdiv_rn(__fsub_rn(y, __fmul_rn(cal[6], z)), cal[5]);
        const float h = in.dim[3 * i], w = in.dim[3 * i + 1], l = in.dim[3 * i + 2];
        y = __fadd_rn(y, __fdiv_rn(h, 2.0f));                                             // ddd.py:197
        const float yaw = yaw_wrap(alpha, f32_atan2(__fsub_rn(cx, cal[2]), cal[0]));      // ddd.py:137-141
        o.locations[3 * i] = x;
        o.locations[3 * i + 1] = y;
        o.locations[3 * i + 2] = z;
        o.yaws[i] = yaw;
        const float c = (float)cos((double)yaw), s = (float)sin((double)yaw);
        const float ns = -s;
        if (!p.is_gt && in.velocity) {                                                    // :71-76
            const float v0 = in.velocity[3 * i], v2 = in.velocity[3 * i + 2];
            const float V = __fsqrt_rn(__fadd_rn(__fmul_rn(v0, v0), __fmul_rn(v2, v2)));
            in.velocity[3 * i] = __fmul_rn(c, V);
            in.velocity[3 * i + 2] = __fmul_rn(ns, V);
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

cudaError_t launch_post_process(const cf_post_in_t& in, int B, int K, int out_h, int out_w, const float* trans_host,
                                const float* calib, int is_gt, const cf_post_out_t& out, cudaStream_t st,
                                int* launches) {
    PostParams p;
    p.in = in;
    p.out = out;
    p.B = B;
    p.K = K;
    p.out_h = out_h;
    p.out_w = out_w;
    p.is_gt = is_gt;
    for (int j = 0; j < 6; ++j) p.t[j] = trans_host[j];
    p.calib = calib;
    const long long n = (long long)B * K;
    if (n == 0) return cudaSuccess;
    post_process_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

}  // namespace cf
