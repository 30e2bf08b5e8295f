// paint.cuh -- ordered rectangle paint of one band of rows by one CTA: every pixel takes the covering record with
// the LARGEST index (== the reference's sequential overwrite order), resolved on a shared-memory owner map, after
// which every output pixel of the band is written to HBM exactly once, coalesced, zeros included.
//
// Replaces (reference, src/lib/): the write half of utils/pointcloud.py:478-481 under the B x K loop of :383-392,
// and dataset/datasets/nuscenes.py:234-263 drawPcHeat under the pillar loop of dataset/generic_dataset.py:798-826.
#pragma once
#include "common.cuh"
#include "kernels.cuh"

namespace cf {

// rows [y0,y1) of one frame: fr = the frame's records (SHARED = true: a shared-memory copy; false: global memory),
// out = the frame's (3,H,W) map, owner = the frame's (H,W) owner map or NULL,
// sown = shared-memory owner map holding max_rows x W ints.  Whole CTA of T threads; ends with a barrier.
// PREINIT: the caller has already set the owner map of the FIRST piece of the band to -1 (and passed a CTA barrier).
template <int T, bool SHARED, bool PREINIT = false>
__device__ __forceinline__ void paint_band(const PaintRec* __restrict__ fr, int n, int H, int W, int y0, int y1,
                                           float* __restrict__ out, int32_t* __restrict__ owner, int* sown,
                                           int max_rows) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t HW = (size_t)H * W;
    const bool vec = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0) &&
                     (!owner || (reinterpret_cast<uintptr_t>(owner) & 15u) == 0);
    for (int ys = y0; ys < y1; ys += max_rows) {
        const int ye = min(y1, ys + max_rows);
        const int npx = (ye - ys) * W;
        if (!(PREINIT && ys == y0)) {
            for (int i = tid; i < npx; i += T) sown[i] = -1;
            __syncthreads();
        }
        if (SHARED) {
            // records in shared memory: record k on warp k % (T/32), so that all warps paint (a broadcast read each)
            for (int k = warp; k < n; k += T / 32) {
                const short4 q = *reinterpret_cast<const short4*>(fr + k);
                if (!(q.y > q.x && q.w > q.z && q.x < ye && q.y > ys)) continue;
                const int qa = max((int)q.x, ys), qb = min((int)q.y, ye);
                const int xa = q.z, w = q.w - q.z;
                const int cnt = (qb - qa) * w;
                for (int t = lane; t < cnt; t += 32) {
                    const int ry = t / w;
                    atomicMax(&sown[(qa - ys + ry) * W + xa + (t - ry * w)], k);
                }
            }
        } else {
            // records in global memory: 32 coalesced loads per warp, then the hits one after the other
            for (int base = warp * 32; base < n; base += T) {
                const int k = base + lane;
                int ra = 0, rb = 0, ca = 0, cb = 0;
                if (k < n) {
                    const short4 q = __ldg(reinterpret_cast<const short4*>(fr + k));
                    ra = q.x;
                    rb = q.y;
                    ca = q.z;
                    cb = q.w;
                }
                const bool hit = rb > ra && cb > ca && ra < ye && rb > ys;
                uint32_t m = __ballot_sync(FULL, hit);
                while (m) {
                    const int bit = __ffs(m) - 1;
                    m &= m - 1;
                    const int qa = max(__shfl_sync(FULL, ra, bit), ys), qb = min(__shfl_sync(FULL, rb, bit), ye);
                    const int xa = __shfl_sync(FULL, ca, bit), xb = __shfl_sync(FULL, cb, bit);
                    const int w = xb - xa;
                    const int cnt = (qb - qa) * w;
                    for (int t = lane; t < cnt; t += 32) {
                        const int ry = t / w;
                        atomicMax(&sown[(qa - ys + ry) * W + xa + (t - ry * w)], base + bit);
                    }
                }
            }
        }
        __syncthreads();

        float* o0 = out + (size_t)ys * W;
        int32_t* oo = owner ? owner + (size_t)ys * W : nullptr;
        if (vec) {
            for (int i = tid * 4; i < npx; i += T * 4) {
                const int4 ow = *reinterpret_cast<const int4*>(sown + i);
                const int os[4] = {ow.x, ow.y, ow.z, ow.w};
                float v[3][4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    v[0][j] = v[1][j] = v[2][j] = 0.0f;
                    if (os[j] >= 0) {
                        const PaintRec* r = fr + os[j];
                        v[0][j] = (SHARED ? r->v0 : __ldg(&r->v0));
                        v[1][j] = (SHARED ? r->v1 : __ldg(&r->v1));
                        v[2][j] = (SHARED ? r->v2 : __ldg(&r->v2));
                    }
                }
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    *reinterpret_cast<float4*>(o0 + c * HW + i) = make_float4(v[c][0], v[c][1], v[c][2], v[c][3]);
                if (oo) *reinterpret_cast<int4*>(oo + i) = ow;
            }
        } else {
            for (int i = tid; i < npx; i += T) {
                const int ow = sown[i];
                float v0 = 0.0f, v1 = 0.0f, v2 = 0.0f;
                if (ow >= 0) {
                    const PaintRec* r = fr + ow;
                    v0 = (SHARED ? r->v0 : __ldg(&r->v0));
                    v1 = (SHARED ? r->v1 : __ldg(&r->v1));
                    v2 = (SHARED ? r->v2 : __ldg(&r->v2));
                }
                o0[i] = v0;
                o0[HW + i] = v1;
                o0[2 * HW + i] = v2;
                if (oo) oo[i] = ow;
            }
        }
        __syncthreads();
    }
}

}  // namespace cf
