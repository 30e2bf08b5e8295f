// peak_select.cu -- heat-map peak decode: K1 peak_select (one pass over the heat-map, raw and 3x3-nms'd
// candidate lists together), K2 merge (frame-wide sorted top-K + frustum / decode epilogues), and the
// stand-alone nms / gather kernels.
//
// Replaces (reference, src/lib/): model/utils.py:112-128 nms, :6-38 topk, :41-72 getFeature /
// transposeAndGetFeature; the gathers + box/alpha/threshold block of utils/pointcloud.py:347-380; and
// model/decode.py:10-174 fusionDecode.
//
// K1 layout: grid = B*C*S CTAs, one per (frame, class, row band).  The band (+1 halo row each side) is
// contiguous in NCHW memory, so ONE 1-D TMA bulk copy (cp.async.bulk, mbarrier complete_tx) stages it in
// shared memory.  Each thread owns EI float4 items: vertical 3-max from three LDS.128, horizontal 3-max
// via two warp shuffles, keep-bit per pixel.  Exact top-K' of the band for both lists by block_select
// (select.cuh); K' composites per list go to the workspace.  HBM traffic: the heat-map once.
#include <math_constants.h>

#include <cstdio>
#include <cstdlib>

#include "geometry.cuh"
#include "kernels.cuh"
#include "select.cuh"

namespace cf {

// -----------------------------------------------------------------------------------------------------
// plan + workspace
// -----------------------------------------------------------------------------------------------------
static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static size_t peak_smem_bytes(int R, int Wp) {
    return align_up((size_t)(R + 2) * Wp * sizeof(float), 16) + sizeof(SelScratch) + 16;
}

bool make_plan(int C, int H, int W, int K, Plan* p) {
    const int Wp = (W + 3) & ~3;
    const int W4 = Wp / 4;
    int T, EI, J;
    if (K <= 256) {
        T = 512;
        EI = 4;
        J = 1;
    } else {
        T = 1024;
        EI = 8;
        J = K <= 512 ? 1 : 2;
    }
    const char* e = getenv("CFD3D_PLAN");  // tuning override: "T,EI" with T in {512,1024}, EI in {4,8}
    if (e) {
        int t = 0, ei = 0;
        if (sscanf(e, "%d,%d", &t, &ei) == 2 && (t == 512 || t == 1024) && (ei == 4 || ei == 8)) {
            T = t;
            EI = ei;
            const int nw = T / 32;
            J = 1;
            while ((K + nw - 1) / nw > 16 * J && J < 4) J *= 2;
        }
    }
    const int max_items = T * EI;
    if (W4 > max_items || W4 > 65535 || H > 32767) return false;
    int R = max_items / W4;
    if (R > H) R = H;
    const size_t smem_cap = 227 * 1024;
    while (R > 1 && peak_smem_bytes(R, Wp) > smem_cap) --R;
    if (peak_smem_bytes(R, Wp) > smem_cap) return false;
    const int S = (H + R - 1) / R;
    R = (H + S - 1) / S;
    p->T = T;
    p->EI = EI;
    p->J = J;
    p->R = R;
    p->S = (H + R - 1) / R;
    p->Wp = Wp;
    p->NB = C * p->S;
    p->smem = peak_smem_bytes(R, Wp);
    return true;
}

size_t workspace_layout(int B, int C, int H, int W, int K, const Plan& plan, void* base, Workspace* out) {
    (void)C;
    (void)H;
    (void)W;
    size_t off = 0;
    const size_t cand = align_up((size_t)B * plan.NB * K * sizeof(unsigned long long), 256);
    const size_t det = align_up((size_t)B * K * sizeof(DetParam), 256);
    const size_t rec = align_up((size_t)B * K * sizeof(PaintRec), 256);
    char* b = (char*)base;
    if (out) out->cand_raw = (unsigned long long*)(b + off);
    off += cand;
    if (out) out->cand_nms = (unsigned long long*)(b + off);
    off += cand;
    if (out) out->det = (DetParam*)(b + off);
    off += det;
    if (out) out->recs = (PaintRec*)(b + off);
    off += rec;
    return off;
}

// -----------------------------------------------------------------------------------------------------
// K1  peak_select
// -----------------------------------------------------------------------------------------------------
struct PeakParams {
    const float* heat;
    int C, H, W, Wp, K, S, R, lists, use_tma;
    unsigned long long* cand_raw;
    unsigned long long* cand_nms;
};

template <int T, int EI>
struct TileSrc {
    const float4* t4;
    int W4, W;
    uint32_t gbase;  // class*H*W + r0*W
    uint32_t keep;   // bit 4k+j: pixel j of item k is a 3x3 local maximum
    uint32_t rc[EI]; // (row << 16) | float4 column, 0xffffffff = no item
    bool nms;
    template <class F>
    __device__ __forceinline__ void for_each(F&& f) const {
#pragma unroll
        for (int k = 0; k < EI; ++k) {
            const bool v = rc[k] != 0xffffffffu;
            const int r = v ? (int)(rc[k] >> 16) : 0;
            const int c4 = v ? (int)(rc[k] & 0xffffu) : 0;
            const float4 ce = t4[(r + 1) * W4 + c4];
            const float vals[4] = {ce.x, ce.y, ce.z, ce.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int x = 4 * c4 + j;
                const bool ok = v && x < W;
                uint32_t key = f2key(vals[j]);
                if (nms && !((keep >> (4 * k + j)) & 1u)) key = KEY_ZERO;
                f(ok ? make_comp(key, gbase + (uint32_t)(r * W + x)) : 0ull);
            }
        }
    }
};

template <int T, int EI, int J>
__global__ void __launch_bounds__(T) peak_select_kernel(const PeakParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31;
    const int band = blockIdx.x % p.S;
    const int pc = blockIdx.x / p.S;  // b*C + c
    const int c = pc % p.C;
    const int b = pc / p.C;
    const int H = p.H, W = p.W, Wp = p.Wp, W4 = Wp >> 2;
    const int r0 = band * p.R;
    const int Rb = min(p.R, H - r0);

    float* tile = reinterpret_cast<float*>(smem_raw);  // (Rb+2) rows x Wp; row 0 / Rb+1 = halo
    const size_t tile_bytes = (((size_t)(p.R + 2) * Wp * sizeof(float)) + 15) & ~(size_t)15;
    SelScratch& sc = *reinterpret_cast<SelScratch*>(smem_raw + tile_bytes);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + tile_bytes + sizeof(SelScratch));

    const float* plane = p.heat + (size_t)pc * H * W;
    const int g0 = max(r0 - 1, 0), g1 = min(r0 + Rb + 1, H);
    if (p.use_tma) {
        if (tid == 0) {
            mbar_init(bar, 1);
            mbar_fence_init();
        }
        __syncthreads();
        if (tid == 0) {
            const uint32_t bytes = (uint32_t)(g1 - g0) * (uint32_t)W * 4u;
            mbar_expect_tx(bar, bytes);
            tma_bulk_g2s(tile + (size_t)(g0 - (r0 - 1)) * Wp, plane + (size_t)g0 * W, bytes, bar);
        }
        if (r0 == 0)
            for (int x = tid; x < Wp; x += T) tile[x] = -CUDART_INF_F;
        if (r0 + Rb == H)
            for (int x = tid; x < Wp; x += T) tile[(size_t)(Rb + 1) * Wp + x] = -CUDART_INF_F;
        mbar_wait(bar, 0);
        __syncthreads();
    } else {
        const int n = (Rb + 2) * Wp;
        for (int i = tid; i < n; i += T) {
            const int rr = i / Wp, x = i - rr * Wp;
            const int gr = r0 - 1 + rr;
            tile[i] = (gr >= 0 && gr < H && x < W) ? __ldg(plane + (size_t)gr * W + x) : -CUDART_INF_F;
        }
        __syncthreads();
    }

    // ---- 3x3 local-maximum bits ------------------------------------------------------------------------
    TileSrc<T, EI> src;
    src.t4 = reinterpret_cast<const float4*>(tile);
    src.W4 = W4;
    src.W = W;
    src.gbase = (uint32_t)c * (uint32_t)(H * W) + (uint32_t)(r0 * W);
    src.keep = 0u;
    const int n_items = Rb * W4;
    const bool need_nms = (p.lists & LIST_NMS) != 0;
#pragma unroll
    for (int k = 0; k < EI; ++k) {
        const int i = tid + k * T;
        const bool valid = i < n_items;
        const int ii = valid ? i : 0;
        const int r = ii / W4, c4 = ii - r * W4;
        src.rc[k] = valid ? (((uint32_t)r << 16) | (uint32_t)c4) : 0xffffffffu;
        if (need_nms) {
            const float4 up = src.t4[r * W4 + c4];
            const float4 ce = src.t4[(r + 1) * W4 + c4];
            const float4 dn = src.t4[(r + 2) * W4 + c4];
            float4 vm;
            vm.x = fmax3(up.x, ce.x, dn.x);
            vm.y = fmax3(up.y, ce.y, dn.y);
            vm.z = fmax3(up.z, ce.z, dn.z);
            vm.w = fmax3(up.w, ce.w, dn.w);
            float lft = __shfl_up_sync(FULL, vm.w, 1);
            float rgt = __shfl_down_sync(FULL, vm.x, 1);
            if (lane == 0 && c4 > 0) {
                const int x = 4 * c4 - 1;
                lft = fmax3(tile[r * Wp + x], tile[(r + 1) * Wp + x], tile[(r + 2) * Wp + x]);
            }
            if (lane == 31 && c4 < W4 - 1) {
                const int x = 4 * c4 + 4;
                rgt = fmax3(tile[r * Wp + x], tile[(r + 1) * Wp + x], tile[(r + 2) * Wp + x]);
            }
            if (c4 == 0) lft = -CUDART_INF_F;
            if (c4 == W4 - 1) rgt = -CUDART_INF_F;
            const float m0 = fmax3(lft, vm.x, vm.y), m1 = fmax3(vm.x, vm.y, vm.z);
            const float m2 = fmax3(vm.y, vm.z, vm.w), m3 = fmax3(vm.z, vm.w, rgt);
            const uint32_t kb = (m0 == ce.x ? 1u : 0u) | (m1 == ce.y ? 2u : 0u) | (m2 == ce.z ? 4u : 0u) |
                                (m3 == ce.w ? 8u : 0u);
            if (valid) src.keep |= kb << (4 * k);
        }
    }

    // ---- exact band-local top-K' for each requested list -------------------------------------------------
    const int Kp = min(p.K, Rb * W);
    const size_t list_off = ((size_t)b * (p.C * p.S) + (size_t)(c * p.S + band)) * p.K;
    if (p.lists & LIST_RAW) {
        unsigned long long* out = p.cand_raw + list_off;
        src.nms = false;
        block_select<T, J>(src, Kp, sc, [&](uint32_t slot, unsigned long long cc) { out[slot] = cc; });
        for (int i = Kp + tid; i < p.K; i += T) out[i] = 0ull;
    }
    if (p.lists & LIST_NMS) {
        unsigned long long* out = p.cand_nms + list_off;
        src.nms = true;
        block_select<T, J>(src, Kp, sc, [&](uint32_t slot, unsigned long long cc) { out[slot] = cc; });
        for (int i = Kp + tid; i < p.K; i += T) out[i] = 0ull;
    }
}

template <int T, int EI, int J>
static cudaError_t launch_peak_inst(const PeakParams& pp, int grid, size_t smem, cudaStream_t st) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(peak_select_kernel<T, EI, J>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    peak_select_kernel<T, EI, J><<<grid, T, smem, st>>>(pp);
    return cudaGetLastError();
}

cudaError_t launch_peak_select(const float* heat, int B, int C, int H, int W, int K, int lists, const Plan& plan,
                               const Workspace& ws, cudaStream_t st, int* launches) {
    PeakParams pp;
    pp.heat = heat;
    pp.C = C;
    pp.H = H;
    pp.W = W;
    pp.Wp = plan.Wp;
    pp.K = K;
    pp.S = plan.S;
    pp.R = plan.R;
    pp.lists = lists;
    pp.use_tma = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0);
    pp.cand_raw = ws.cand_raw;
    pp.cand_nms = ws.cand_nms;
    const int grid = B * C * plan.S;
    cudaError_t e;
#define CF_PEAK_CASE(T_, EI_, J_) \
    if (plan.T == T_ && plan.EI == EI_ && plan.J == J_) e = launch_peak_inst<T_, EI_, J_>(pp, grid, plan.smem, st); else
    CF_PEAK_CASE(512, 4, 1)
    CF_PEAK_CASE(512, 4, 2)
    CF_PEAK_CASE(512, 4, 4)
    CF_PEAK_CASE(512, 8, 1)
    CF_PEAK_CASE(512, 8, 2)
    CF_PEAK_CASE(512, 8, 4)
    CF_PEAK_CASE(1024, 4, 1)
    CF_PEAK_CASE(1024, 4, 2)
    CF_PEAK_CASE(1024, 8, 1)
    CF_PEAK_CASE(1024, 8, 2)
    e = cudaErrorInvalidConfiguration;
#undef CF_PEAK_CASE
    if (e == cudaSuccess) ++*launches;
    return e;
}

// -----------------------------------------------------------------------------------------------------
// K2  merge: frame-wide sorted top-K of the NB*K candidates + epilogue
// -----------------------------------------------------------------------------------------------------
constexpr int MT = 1024;

struct GlobSrc {
    const unsigned long long* cand;
    int n_in, n_round;
    template <class F>
    __device__ __forceinline__ void for_each(F&& f) const {
        for (int i = threadIdx.x; i < n_round; i += MT) f(i < n_in ? __ldg(cand + i) : 0ull);
    }
};

struct MergeParams {
    const float* heat;
    int C, H, W, K, NB;
    const unsigned long long* cand[2];
    MergeJob job[2];
    DetParam* det;
};

__device__ __forceinline__ void epilogue_frustum(const MergeParams& p, const MergeJob& jb, int b, int k, int ind,
                                                 int x, int y) {
    const size_t HW = (size_t)p.H * p.W;
    const FrustumHeads& fh = jb.fh;
    // reference src/lib/utils/pointcloud.py:348-367
    const float xf = __fadd_rn((float)x, 0.5f), yf = __fadd_rn((float)y, 0.5f);
    float w = __ldg(fh.wh + ((size_t)b * 2 + 0) * HW + ind);
    float h = __ldg(fh.wh + ((size_t)b * 2 + 1) * HW + ind);
    if (w < 0.0f) w = 0.0f;
    if (h < 0.0f) h = 0.0f;
    const float hw2 = __fdiv_rn(w, 2.0f), hh2 = __fdiv_rn(h, 2.0f);
    DetParam d;
    d.x1 = __fsub_rn(xf, hw2);
    d.y1 = __fsub_rn(yf, hh2);
    d.x2 = __fadd_rn(xf, hw2);
    d.y2 = __fadd_rn(yf, hh2);
    float r[8], dm[3];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = __ldg(fh.rot + ((size_t)b * 8 + j) * HW + ind);
#pragma unroll
    for (int j = 0; j < 3; ++j) dm[j] = __ldg(fh.dim + ((size_t)b * 3 + j) * HW + ind);
    d.depth = __ldg(fh.depth + (size_t)b * HW + ind);
    const float alpha = get_alpha(r);                                   // :372
    const float cx = __fdiv_rn(__fadd_rn(d.x1, d.x2), 2.0f);            // :373-379
    d.thr = dist_thresh_one(fh.calib + 12 * (size_t)b, cx, dm, alpha);  // :380
    d.ind = ind;
    d.pad_ = 0;
    const size_t o = (size_t)b * p.K + k;
    p.det[o] = d;
    if (jb.fo.thr) jb.fo.thr[o] = d.thr;
    if (jb.fo.inds) jb.fo.inds[o] = ind;
}

__device__ __forceinline__ void epilogue_decode(const MergeParams& p, const MergeJob& jb, int b, int k, float score,
                                                int cls, int ind, int x, int y) {
    // reference src/lib/model/decode.py:38-41, :80-85, :132-163
    const size_t HW = (size_t)p.H * p.W;
    const cf_decode_heads_t& hd = jb.dec.heads;
    const cf_decode_out_t& o = jb.dec.out;
    const size_t ok = (size_t)b * p.K + k;
    if (hd.uncertainty) {
        const float u = __ldg(hd.uncertainty + (size_t)b * HW + ind);
        score = __fmul_rn(score, expf(-expf(u)));
    }
    o.scores[ok] = score;
    o.class_ids[ok] = (float)cls;
    const float xn = __fdiv_rn((float)x, (float)p.W), yn = __fdiv_rn((float)y, (float)p.H);
    o.centers[2 * ok + 0] = xn;
    o.centers[2 * ok + 1] = yn;
    float xs = __fmul_rn(xn, (float)jb.dec.out_w), ys = __fmul_rn(yn, (float)jb.dec.out_h);
    if (hd.reg) {
        xs = __fadd_rn(xs, __ldg(hd.reg + ((size_t)b * 2 + 0) * HW + ind));
        ys = __fadd_rn(ys, __ldg(hd.reg + ((size_t)b * 2 + 1) * HW + ind));
    } else {
        xs = __fadd_rn(xs, 0.5f);
        ys = __fadd_rn(ys, 0.5f);
    }
    if (hd.wh) {
        float w = __ldg(hd.wh + ((size_t)b * 2 + 0) * HW + ind);
        float h = __ldg(hd.wh + ((size_t)b * 2 + 1) * HW + ind);
        if (w < 0.0f) w = 0.0f;
        if (h < 0.0f) h = 0.0f;
        if (jb.dec.norm2d) {
            w = __fmul_rn(w, (float)jb.dec.out_w);
            h = __fmul_rn(h, (float)jb.dec.out_h);
        }
        const float hw2 = __fdiv_rn(w, 2.0f), hh2 = __fdiv_rn(h, 2.0f);
        float4 bb;
        bb.x = __fsub_rn(xs, hw2);
        bb.y = __fsub_rn(ys, hh2);
        bb.z = __fadd_rn(xs, hw2);
        bb.w = __fadd_rn(ys, hh2);
        o.bboxes[4 * ok + 0] = bb.x;
        o.bboxes[4 * ok + 1] = bb.y;
        o.bboxes[4 * ok + 2] = bb.z;
        o.bboxes[4 * ok + 3] = bb.w;
    }
    if (hd.amodal) {  // :165 de-normalise
        float ax = __ldg(hd.amodal + ((size_t)b * 2 + 0) * HW + ind);
        float ay = __ldg(hd.amodal + ((size_t)b * 2 + 1) * HW + ind);
        if (jb.dec.norm2d) {
            ax = __fmul_rn(ax, (float)jb.dec.out_w);
            ay = __fmul_rn(ay, (float)jb.dec.out_h);
        }
        o.amodal[2 * ok + 0] = ax;
        o.amodal[2 * ok + 1] = ay;
    }
    if (hd.depth) o.depth[ok] = __ldg(hd.depth + (size_t)b * HW + ind);
    if (o.inds) o.inds[ok] = ind;
}

// plain channel copies of the decode heads, spread over the whole block: out[b,k,ch] = head[b,ch,ind_k]
__device__ __forceinline__ void copy_head(const float* head, float* out, int nch, int b, int K, size_t HW,
                                          const int* sind) {
    if (!head) return;
    for (int t = threadIdx.x; t < K * nch; t += MT) {
        const int k = t / nch, ch = t - k * nch;
        out[((size_t)b * K + k) * nch + ch] = __ldg(head + ((size_t)b * nch + ch) * HW + sind[k]);
    }
}

__global__ void __launch_bounds__(MT) merge_kernel(const MergeParams p) {
    __shared__ SelScratch sc;
    __shared__ unsigned long long win[CF_MAX_K];
    __shared__ unsigned long long sorted[CF_MAX_K];
    __shared__ int sind[CF_MAX_K];
    const int tid = threadIdx.x;
    const int b = blockIdx.x, li = blockIdx.y;
    const MergeJob& jb = p.job[li];
    const int K = p.K;
    const int HWi = p.H * p.W;

    GlobSrc src;
    src.n_in = p.NB * K;
    src.n_round = (src.n_in + MT - 1) / MT * MT;
    src.cand = p.cand[li] + (size_t)b * src.n_in;
    // NW = 32 warps: rounds = ceil(K/32) <= 32 always (K <= 1024), J = 1 suffices for the bound
    block_select<MT, 1>(src, K, sc, [&](uint32_t slot, unsigned long long c) { win[slot] = c; });

    // rank sort (composites are distinct: ranks are a permutation)
    if (tid < K) {
        const unsigned long long me = win[tid];
        int rank = 0;
        for (int j = 0; j < K; ++j) rank += (win[j] > me) ? 1 : 0;
        sorted[rank] = me;
    }
    __syncthreads();

    if (tid < K) {
        const unsigned long long c = sorted[tid];
        const uint32_t g = comp_gidx(c), key = comp_key(c);
        const int cls = (int)(g / (uint32_t)HWi);
        const int ind = (int)(g - (uint32_t)cls * (uint32_t)HWi);
        const int y = ind / p.W, x = ind - y * p.W;
        sind[tid] = ind;
        float score = key2f(key);
        if (key == KEY_ZERO) score = __fmul_rn(__ldg(p.heat + ((size_t)b * p.C + cls) * HWi + ind), 0.0f);
        const size_t o = (size_t)b * K + tid;
        if (jb.epilogue == EPI_TOPK) {
            if (jb.topk.scores) jb.topk.scores[o] = score;
            if (jb.topk.inds) jb.topk.inds[o] = ind;
            if (jb.topk.classes) jb.topk.classes[o] = cls;
            if (jb.topk.ys) jb.topk.ys[o] = y;
            if (jb.topk.xs) jb.topk.xs[o] = x;
        } else if (jb.epilogue == EPI_FRUSTUM) {
            epilogue_frustum(p, jb, b, tid, ind, x, y);
        } else if (jb.epilogue == EPI_DECODE) {
            epilogue_decode(p, jb, b, tid, score, cls, ind, x, y);
        }
    }
    if (jb.epilogue == EPI_DECODE) {
        __syncthreads();
        const size_t HW = (size_t)HWi;
        copy_head(jb.dec.heads.rot, jb.dec.out.rot, 8, b, K, HW, sind);
        copy_head(jb.dec.heads.dim, jb.dec.out.dim, 3, b, K, HW, sind);
        copy_head(jb.dec.heads.nuscenes_att, jb.dec.out.nuscenes_att, 8, b, K, HW, sind);
        copy_head(jb.dec.heads.velocity, jb.dec.out.velocity, 3, b, K, HW, sind);
    }
}

cudaError_t launch_merge(const float* heat, int B, int C, int H, int W, int K, const Plan& plan,
                         const Workspace& ws, const MergeJob* raw_job, const MergeJob* nms_job, cudaStream_t st,
                         int* launches) {
    MergeParams mp;
    mp.heat = heat;
    mp.C = C;
    mp.H = H;
    mp.W = W;
    mp.K = K;
    mp.NB = plan.NB;
    mp.det = ws.det;
    int n = 0;
    if (raw_job) {
        mp.cand[n] = ws.cand_raw;
        mp.job[n] = *raw_job;
        ++n;
    }
    if (nms_job) {
        mp.cand[n] = ws.cand_nms;
        mp.job[n] = *nms_job;
        ++n;
    }
    if (n == 0) return cudaSuccess;
    if (n == 1) {
        mp.cand[1] = mp.cand[0];
        mp.job[1] = mp.job[0];
    }
    merge_kernel<<<dim3(B, n), MT, 0, st>>>(mp);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) ++*launches;
    return e;
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
        out[t] = __ldg(feat + ((size_t)b * ch + c) * HW + inds[bm]);
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
