// post.cu -- postProcess as a launch of its own: decoded detections -> image-space boxes, camera-frame locations, yaw,
// re-oriented velocity and 8-corner 3-D boxes.  One thread per detection; the arithmetic lives in post.cuh, shared
// with the decode epilogue of the fused kernels (cf_fusion_decode_post / cf_fusion_stage_post run it there).
//
// Replaces (reference, src/lib/): utils/postProcess.py:13-85 (see post.cuh for the callees).
#include <math_constants.h>

#include "post.cuh"

namespace cf {

struct PostParams {
    cf_post_in_t in;
    PostFuse pf;
    int B, K;
};

__global__ void __launch_bounds__(128) post_process_kernel(const PostParams p) {
    const long long i = blockIdx.x * 128ll + threadIdx.x;
    if (i >= (long long)p.B * p.K) return;
    const cf_post_in_t& in = p.in;
    in.class_ids[i] = __fadd_rn(in.class_ids[i], 1.0f);  // :34
    PostDet d;
    d.has = (in.bboxes ? PD_BBOX : 0) | (in.rot ? PD_ROT : 0) | (in.depth ? PD_DEPTH : 0) | (in.dim ? PD_DIM : 0) |
            (in.amodal ? PD_AMODAL : 0) | (in.velocity ? PD_VEL : 0);
    d.cx = in.centers[2 * i];
    d.cy = in.centers[2 * i + 1];
#pragma unroll
    for (int j = 0; j < 4; ++j) d.bb[j] = in.bboxes ? in.bboxes[4 * i + j] : 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) d.rot[j] = in.rot ? in.rot[8 * i + j] : 0.0f;
    d.depth = in.depth ? in.depth[i] : 0.0f;
#pragma unroll
    for (int j = 0; j < 3; ++j) d.dim[j] = in.dim ? in.dim[3 * i + j] : 0.0f;
#pragma unroll
    for (int j = 0; j < 2; ++j) d.amodal[j] = in.amodal ? in.amodal[2 * i + j] : 0.0f;
#pragma unroll
    for (int j = 0; j < 3; ++j) d.vel[j] = in.velocity ? in.velocity[3 * i + j] : 0.0f;
    post_detection(p.pf, p.pf.calib + 12 * (i / p.K), i, d);
    if (in.velocity) {  // in place, like the reference (:71-76); unchanged values when the branch was not taken
        in.velocity[3 * i] = d.vel[0];
        in.velocity[3 * i + 2] = d.vel[2];
    }
}

cudaError_t launch_post_process(const cf_post_in_t& in, int B, int K, int out_h, int out_w, const float* trans_host,
                                const float* calib, int is_gt, const cf_post_out_t& out, cudaStream_t st,
                                int* launches) {
    PostParams p;
    p.in = in;
    p.pf.enabled = 1;
    p.pf.out = out;
    p.B = B;
    p.K = K;
    p.pf.height = out_h;
    p.pf.width = out_w;
    p.pf.is_gt = is_gt;
    for (int j = 0; j < 6; ++j) p.pf.t[j] = trans_host[j];
    p.pf.calib = calib;
    const long long n = (long long)B * K;
    if (n == 0) return cudaSuccess;
    post_process_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

}  // namespace cf
