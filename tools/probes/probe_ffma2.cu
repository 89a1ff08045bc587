// Probe: FP32 FMA throughput per SM with scalar FFMA vs packed FFMA2 (fma.rn.f32x2), as a function of
// warps per SM.  Tells whether the in-register depthwise is bound by the FMA pipe or by issue slots.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o probe_ffma2 probe_ffma2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long pk(float a, float b) {
    unsigned long long r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
template <int MODE>
__global__ void k(float* out, float w, int iters, long long* cyc) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    unsigned long long p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) p[i] = pk(a[2 * i], a[2 * i + 1]);
    const unsigned long long ww = pk(w, w);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], w, 0.5f);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %0;" : "+l"(p[i]) : "l"(ww));
        }
    }
    long long t1 = clock64();
    float s = 0;
    if (MODE == 0) { for (int i = 0; i < 16; ++i) s += a[i]; }
    else { for (int i = 0; i < 8; ++i) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(p[i])); s += x + y; } }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    float* o; long long* c;
    cudaMalloc(&o, 148 * 1024 * 4); cudaMalloc(&c, 148 * 8);
    const int iters = 4096;
    for (int mode = 0; mode < 2; ++mode)
        for (int warps : {4, 8, 12, 16, 32}) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) k<0><<<148, warps * 32>>>(o, 0.999f, iters, c); else k<1><<<148, warps * 32>>>(o, 0.999f, iters, c);
                cudaDeviceSynchronize();
            }
            long long h[148]; cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
            double mx = 0; for (auto v : h) if (v > mx) mx = v;
            double fma = double(iters) * 16 * warps * 32;
            printf("%s warps/SM %2d: %.1f FMA/clk/SM (%.0f cycles)\n", mode ? "FFMA2" : "FFMA ", warps, fma / mx, mx);
        }
    return 0;
}
