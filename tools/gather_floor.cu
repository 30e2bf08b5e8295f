// gather_floor.cu -- stand-alone floor of the tail's scattered head gathers: n_cta CTAs x 512 threads, each CTA reads
// `per_cta` random 4-byte values (one 32-byte sector each) out of a `mb`-megabyte region that is cold in L2, once per
// launch.  Reports the in-kernel time from the first issue to the last arrival (clock64, max over CTAs) and the kernel time.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/gather_floor tools/gather_floor.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(512) gather(const float* __restrict__ src, const unsigned* __restrict__ idx, int per_cta,
                                              float* out, long long* clk, int mode) {
    __shared__ float buf[4096];
    const int tid = threadIdx.x;
    const unsigned* my = idx + (size_t)blockIdx.x * per_cta;
    __syncthreads();
    const long long t0 = clock64();
    float acc = 0.f;
    if (mode == 0) {  // plain loads, all of a thread's loads in flight together
        float v[4] = {0, 0, 0, 0};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = tid + u * 512;
            if (t < per_cta) v[u] = __ldg(src + my[t]);
        }
        acc = v[0] + v[1] + v[2] + v[3];
    } else {  // cp.async 4-byte copies
        for (int t = tid; t < per_cta; t += 512)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(buf + t)), "l"(src + my[t]) : "memory");
        asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
        acc = buf[tid];
    }
    __syncthreads();
    const long long t1 = clock64();
    if (acc == 12345.678f) out[0] = acc;
    if (tid == 0) clk[blockIdx.x] = t1 - t0;
}

__global__ void flush(float* p, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = p[i] * 1.0001f + 1.f;
}

int main() {
    const size_t region = 512ull << 20;  // heads of several input sets
    float* src;
    cudaMalloc(&src, region);
    cudaMemset(src, 0, region);
    float* fl;
    const size_t fl_n = (256ull << 20) / 4;
    cudaMalloc(&fl, fl_n * 4);
    float* out;
    cudaMalloc(&out, 4);
    long long* clk;
    cudaMalloc(&clk, 148 * 8);
    const int n_cta = 128;
    for (int mode = 0; mode < 2; ++mode)
        for (int per_cta : {128, 350, 700, 1100, 2048}) {
            std::vector<unsigned> h((size_t)n_cta * per_cta);
            unsigned* idx;
            cudaMalloc(&idx, h.size() * 4);
            float best_k = 1e9, sum_k = 0, best_c = 1e9, sum_c = 0;
            const int reps = 10;
            for (int r = 0; r < reps; ++r) {
                for (auto& v : h) v = (unsigned)((((unsigned long long)rand() << 16) ^ rand()) % (region / 4));
                cudaMemcpy(idx, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
                flush<<<296, 512>>>(fl, fl_n);
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0);
                cudaEventCreate(&e1);
                cudaEventRecord(e0);
                gather<<<n_cta, 512>>>(src, idx, per_cta, out, clk, mode);
                cudaEventRecord(e1);
                cudaDeviceSynchronize();
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                long long hc[148];
                cudaMemcpy(hc, clk, n_cta * 8, cudaMemcpyDeviceToHost);
                long long mx = 0;
                for (int i = 0; i < n_cta; ++i) mx = hc[i] > mx ? hc[i] : mx;
                const float us_c = mx / 1965.0f;
                best_k = ms * 1e3f < best_k ? ms * 1e3f : best_k;
                sum_k += ms * 1e3f;
                best_c = us_c < best_c ? us_c : best_c;
                sum_c += us_c;
            }
            printf("%s  %5d sectors/CTA x %d CTAs (%6d total): in-kernel issue->last arrival min %.2f mean %.2f us | kernel min %.2f mean %.2f us\n",
                   mode ? "cp.async" : "ld.nc   ", per_cta, n_cta, per_cta * n_cta, best_c, sum_c / reps, best_k, sum_k / reps);
            cudaFree(idx);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
