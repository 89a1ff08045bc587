// Probe: cycles per depthwise board (dw_board / dw_board2 of hm_trunk.cuh) with 8 warps per SM
// (2 per sub-partition, as in trunk_kernel) and nothing else running on the SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o probe_dw probe_dw.cu
#include <cstdio>
#include "../../hivemind_b200/csrc/hm_trunk.cuh"
using namespace hm;
template <int KS, int V>
__global__ void __launch_bounds__(256, 1) k(const float* in, const float* w, uint32_t* out, int iters, long long* cyc) {
    float y[64], ww[KS * KS];
    for (int i = 0; i < 64; ++i) y[i] = in[i * 256 + threadIdx.x];
    for (int i = 0; i < KS * KS; ++i) ww[i] = w[i * 256 + threadIdx.x];
    uint32_t acc = 0;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        uint32_t p[32];
        if (V == 0) dw_board<KS>(y, ww, w[threadIdx.x], p); else dw_board2<KS>(y, ww, w[threadIdx.x], p);
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= p[i];
        y[it & 63] += __uint_as_float(acc & 0x3f800000u) * 1e-9f;   // keep a loop-carried dependence
    }
    long long t1 = clock64();
    out[blockIdx.x * 256 + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    float *in, *w; uint32_t* o; long long* c;
    cudaMalloc(&in, 64 * 256 * 4); cudaMalloc(&w, 32 * 256 * 4); cudaMalloc(&o, 148 * 256 * 4); cudaMalloc(&c, 148 * 8);
    cudaMemset(in, 0, 64 * 256 * 4); cudaMemset(w, 0, 32 * 256 * 4);
    const int iters = 200;
    auto run = [&](const char* name, auto kern) {
        for (int rep = 0; rep < 2; ++rep) { kern<<<148, 256>>>(in, w, o, iters, c); cudaDeviceSynchronize(); }
        long long h[148]; cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0; for (auto v : h) if (v > mx) mx = v;
        printf("%s: %.0f cycles per board (8 warps/SM)\n", name, double(mx) / iters);
    };
    run("3x3 scalar FFMA ", k<3, 0>); run("3x3 packed FFMA2", k<3, 1>);
    run("5x5 scalar FFMA ", k<5, 0>); run("5x5 packed FFMA2", k<5, 1>);
    return 0;
}
