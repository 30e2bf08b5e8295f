// geometry.cuh -- per-detection frustum geometry in the fixed fp32 operation order the parity contract
// documents (oracle/cfd3d_oracle.c "Arithmetic conventions"): one rounding per written fp32 operation
// (explicit _rn intrinsics: never contracted into FMA), trigonometry evaluated in fp64 and rounded once.
#pragma once
#include "common.cuh"

namespace cf {

__device__ constexpr float PI_F = 3.14159265358979323846f;  // float(torch.pi)

__device__ __forceinline__ float f32_atan2(float y, float x) { return (float)atan2((double)y, (double)x); }

// get_alpha                                   reference src/lib/utils/pointcloud.py:195-211
// blend a1*idx + a2*(~idx) with idx = rot[1] > rot[5], exactly as written there.  at1 = atan2(rot[2],rot[3]),
// at2 = atan2(rot[6],rot[7]) (the fused kernel evaluates the two on different lanes).
__device__ __forceinline__ float alpha_blend(float rot1, float rot5, float at1, float at2) {
    const float idx = rot1 > rot5 ? 1.0f : 0.0f;
    const float nidx = __fsub_rn(1.0f, idx);
    const float a1 = __fadd_rn(at1, -0.5f * PI_F);
    const float a2 = __fadd_rn(at2, 0.5f * PI_F);
    return __fadd_rn(__fmul_rn(a1, idx), __fmul_rn(a2, nidx));
}
__device__ __forceinline__ float get_alpha(const float* rot) {
    return alpha_blend(rot[1], rot[5], f32_atan2(rot[2], rot[3]), f32_atan2(rot[6], rot[7]));
}

// cvtAlphaToYaw + get3DCorners (z row) + getDistanceThresh   reference pointcloud.py:214-328
// cx is in OUTPUT-map pixels while calib is in image pixels (kept); thr = max(z) - min(z)/2 (kept).
// ray = atan2(cx - calib[0,2], calib[0,0]);  c, s = cos / sin of the wrapped yaw.
__device__ __forceinline__ float yaw_wrap(float alpha, float ray) {
    float yaw = __fadd_rn(alpha, ray);
    if (yaw > PI_F) yaw = __fsub_rn(yaw, 2.0f * PI_F);
    if (yaw < -PI_F) yaw = __fadd_rn(yaw, 2.0f * PI_F);
    return yaw;
}
__device__ __forceinline__ float thr_from_cs(float c, float s, float dim_w, float dim_l) {
    const float hx = __fmul_rn(0.5f, dim_l), hz = __fmul_rn(0.5f, dim_w);
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
__device__ __forceinline__ float dist_thresh_one(const float* calib, float cx, const float* dim, float alpha) {
    const float yaw = yaw_wrap(alpha, f32_atan2(__fsub_rn(cx, calib[2]), calib[0]));
    return thr_from_cs((float)cos((double)yaw), (float)sin((double)yaw), dim[1], dim[2]);
}

}  // namespace cf
