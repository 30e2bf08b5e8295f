// peak_ops.cu -- workspace layout, the gather tables of the merge jobs, and the stand-alone nms / gather kernels
// (callers that use the primitives individually; the fused stage kernel in stage.cu never materialises either).
//
// Replaces (reference, src/lib/): model/utils.py:112-128 nms, :41-72 getFeature / transposeAndGetFeature.
#include <math_constants.h>

#include <cstring>

#include "kernels.cuh"

namespace cf {

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

size_t workspace_layout(int B, int K, void* base, Workspace* out) {
    size_t off = 0;
    char* b = (char*)base;
    const size_t ctr = align_up((size_t)B * 2 * CTR_STRIDE * sizeof(uint32_t), 256);
    const size_t cand = align_up((size_t)B * 2 * CAND_CAP * sizeof(unsigned long long), 256);
    const size_t det = align_up((size_t)B * K * sizeof(DetParam), 256);
    const size_t rec = align_up((size_t)B * K * sizeof(PaintRec), 256);
    if (out) out->ctr = (uint32_t*)(b + off);
    off += ctr;
    if (out) out->cand = (unsigned long long*)(b + off);
    off += cand;
    if (out) out->det = (DetParam*)(b + off);
    off += det;
    if (out) out->recs = (PaintRec*)(b + off);
    off += rec;
    return off;
}

static void tab_set(GatherTab& gt, int slot0, const float* base, int nch) {
    for (int c = 0; c < nch; ++c) {
        gt.base[slot0 + c] = base;
        gt.nch[slot0 + c] = (unsigned char)nch;
        gt.ch[slot0 + c] = (unsigned char)c;
    }
}

// the head channels a job gathers at its K peaks (slot layout: kernels.cuh)
void finish_job(MergeJob* job) {
    GatherTab& gt = job->gt;
    memset(&gt, 0, sizeof(gt));
    if (job->epilogue == EPI_FRUSTUM) {
        tab_set(gt, FS_WH, job->fh.wh, 2);
        tab_set(gt, FS_DEPTH, job->fh.depth, 1);
        tab_set(gt, FS_DIM, job->fh.dim, 3);
        tab_set(gt, FS_ROT, job->fh.rot, 8);
        // channels the frustum geometry never reads are not gathered (their slots read 0): the height of the box
        // (getDistanceThresh takes the corners' x / z extent only, pointcloud.py:214-236) and the two bin scores
        // get_alpha does not compare (rot[:, 0], rot[:, 4]; pointcloud.py:195-211)
        gt.base[FS_DIM + 0] = nullptr;
        gt.base[FS_ROT + 0] = nullptr;
        gt.base[FS_ROT + 4] = nullptr;
        gt.n = FS_N;
    } else if (job->epilogue == EPI_DECODE) {
        const cf_decode_heads_t& h = job->dec.heads;
        tab_set(gt, DS_REG, h.reg, 2);
        tab_set(gt, DS_WH, h.wh, 2);
        tab_set(gt, DS_DEPTH, h.depth, 1);
        tab_set(gt, DS_ROT, h.rot, 8);
        tab_set(gt, DS_DIM, h.dim, 3);
        tab_set(gt, DS_AMODAL, h.amodal, 2);
        tab_set(gt, DS_ATT, h.nuscenes_att, 8);
        tab_set(gt, DS_VEL, h.velocity, 3);
        tab_set(gt, DS_UNC, h.uncertainty, 1);
        gt.n = DS_N;
    }
}

// -----------------------------------------------------------------------------------------------------
// stand-alone nms (materialised heat-map) and NCHW gather
// -----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nms_kernel(const float* __restrict__ heat, int H, int W, float* __restrict__ out,
                                                  long long total) {
    const long long HW = (long long)H * W;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256ll) {
        const long long pc = i / HW;
        const int ind = (int)(i - pc * HW);
        const int y = ind / W, x = ind - y * W;
        const float* pl = heat + pc * HW;
        const float v = __ldg(pl + ind);
        float m = -CUDART_INF_F;
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy) {
            const int yy = y + dy;
            if (yy < 0 || yy >= H) continue;
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                const int xx = x + dx;
                if (xx < 0 || xx >= W) continue;
                m = fmaxf(m, __ldg(pl + (size_t)yy * W + xx));
            }
        }
        out[i] = __fmul_rn(v, (m == v) ? 1.0f : 0.0f);
    }
}

cudaError_t launch_nms(const float* heat, int B, int C, int H, int W, float* out, cudaStream_t st, int* launches) {
    const long long total = (long long)B * C * H * W;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    nms_kernel<<<(int)blocks, 256, 0, st>>>(heat, H, W, out, total);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

__global__ void __launch_bounds__(256) gather_kernel(const float* __restrict__ feat, int ch, long long HW,
                                                     const int64_t* __restrict__ inds, int M, float* __restrict__ out,
                                                     long long total) {
    for (long long t = blockIdx.x * 256ll + threadIdx.x; t < total; t += (long long)gridDim.x * 256ll) {
        const int c = (int)(t % ch);
        const long long bm = t / ch;
        const long long b = bm / M;
        // (torch.gather raises a device-side assert on an index outside [0, HW); here the element becomes NaN and
        // nothing outside the tensor is read)
        const long long ix = inds[bm];
        out[t] = (ix >= 0 && ix < HW) ? __ldg(feat + ((size_t)b * ch + c) * HW + ix) : __int_as_float(0x7fc00000);
    }
}

cudaError_t launch_gather(const float* feat, int B, int ch, int64_t HW, const int64_t* inds, int M, float* out,
                          cudaStream_t st, int* launches) {
    const long long total = (long long)B * M * ch;
    if (total == 0) return cudaSuccess;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    gather_kernel<<<(int)blocks, 256, 0, st>>>(feat, ch, (long long)HW, inds, M, out, total);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
}

}  // namespace cf


