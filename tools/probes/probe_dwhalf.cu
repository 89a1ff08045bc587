// Probe: cycles of the four dw_row<KS> calls of a chunk (hm_trunk.cuh) alone, 16 warps per CTA on every SM: is the in-kernel depthwise
// (5x5: ~4300 cycles per chunk, 3x3: ~1560) slow by itself or by interference?  FMA-pipe floors: 2448 / 1056.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o probe_dwhalf probe_dwhalf.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../hivemind_b200/csrc/hm_trunk.cuh"
using namespace hm;

template <int KS>
__global__ void __launch_bounds__(640, 1) k(const float* in, float* out, long long* cyc, int iters) {
    constexpr int NR = 4 + KS / 2;
    float y[NR * 8], w[KS * KS];
    for (int i = 0; i < NR * 8; ++i) y[i] = in[(threadIdx.x * 61 + i) & 1023];
    for (int i = 0; i < KS * KS; ++i) w[i] = in[(threadIdx.x * 7 + i * 3) & 1023];
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        uint32_t packed[4][4];
#pragma unroll
        for (int lo = 0; lo < 4; ++lo) dw_row<KS>(lo, y, w, in[it & 1023], packed[lo]);
#pragma unroll
        for (int j = 0; j < 16; ++j) acc ^= packed[j >> 2][j & 3];
        y[it & 7] += __uint_as_float(acc & 0x3f800000u) * 1e-9f;   // keep the loop honest
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = (t1 - t0) / iters;
    out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float(acc);
}

int main() {
    float *in, *out; long long* cyc;
    cudaMalloc(&in, 4096); cudaMalloc(&out, 148 * 640 * 4); cudaMalloc(&cyc, 8);
    float h[1024]; for (int i = 0; i < 1024; ++i) h[i] = (i % 13) * 0.1f - 0.5f;
    cudaMemcpy(in, h, 4096, cudaMemcpyHostToDevice);
    for (int threads : {128, 256, 512}) {
        long long c3, c5;
        k<3><<<148, threads>>>(in, out, cyc, 200); cudaDeviceSynchronize(); cudaMemcpy(&c3, cyc, 8, cudaMemcpyDeviceToHost);
        k<5><<<148, threads>>>(in, out, cyc, 200); cudaDeviceSynchronize(); cudaMemcpy(&c5, cyc, 8, cudaMemcpyDeviceToHost);
        printf("threads %3d (%d warps/SMSP): dw_half<3> %lld cycles, dw_half<5> %lld cycles per call (%s)\n", threads, threads / 128, c3, c5, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
