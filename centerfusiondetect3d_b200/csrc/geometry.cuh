// geometry.cuh -- per-detection frustum geometry in the fixed fp32 operation order the parity contract
// documents (oracle/cfd3d_oracle.c "Arithmetic conventions"): one rounding per written fp32 operation
// (explicit _rn intrinsics: never contracted into FMA), trigonometry evaluated in fp64 and rounded once.
#pragma once
#include "common.cuh"

namespace cf {

__device__ constexpr float PI_F = 3.14159265358979323846f;  // float(torch.pi)

__device__ __forceinline__ float f32_atan2(float y, float x) { return (float)atan2((double)y, (double)x); }

// get_alpha                                   reference src/lib/utils/pointcloud.py:195-211
// blend a1*idx + a2*(~idx) with idx = rot[1] > rot[5], exactly as written there
__device__ __forceinline__ float get_alpha(const float* rot) {
    const float idx = rot[1] > rot[5] ? 1.0f : 0.0f;
    const float nidx = __fsub_rn(1.0f, idx);
    const float a1 = __fadd_rn(f32_atan2(rot[2], rot[3]), -0.5f * PI_F);
    const float a2 = __fadd_rn(f32_atan2(rot[6], rot[7]), 0.5f * PI_F);
    return __fadd_rn(__fmul_rn(a1, idx), __fmul_rn(a2, nidx));
}

// cvtAlphaToYaw + get3DCorners (z row) + getDistanceThresh   reference pointcloud.py:214-328
// cx is in OUTPUT-map pixels while calib is in image pixels (kept); thr = max(z) - min(z)/2 (kept).
__device__ __forceinline__ float dist_thresh_one(const float* calib, float cx, const float* dim, float alpha) {
    float yaw = __fadd_rn(alpha, f32_atan2(__fsub_rn(cx, calib[2]), calib[0]));
    if (yaw > PI_F) yaw = __fsub_rn(yaw, 2.0f * PI_F);
    if (yaw < -PI_F) yaw = __fadd_rn(yaw, 2.0f * PI_F);
    const float c = (float)cos((double)yaw), s = (float)sin((double)yaw);
    const float hx = __fmul_rn(0.5f, dim[2]), hz = __fmul_rn(0.5f, dim[1]);
    const float ns = -s;
    float zmax = -INFINITY, zmin = INFINITY;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float xc = (i & 1) ? -hx : hx;
        const float zc = (i & 2) ? -hz : hz;
        const float z = __fadd_rn(__fmul_rn(ns, xc), __fmul_rn(c, zc));
        if (z > zmax) zmax = z;
        if (z < zmin) zmin = z;
    }
    return __fsub_rn(zmax, __fdiv_rn(zmin, 2.0f));
}

}  // namespace cf
