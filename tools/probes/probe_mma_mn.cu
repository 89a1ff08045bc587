// Probe: tcgen05.mma kind::f16 with A K-major (SW128) and B MN-major (SW128), M=128 N=128 K=64.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o probe_mma_mn probe_mma_mn.cu
#include <cuda.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../hivemind_b200/csrc/hm_ptx.cuh"
using namespace hm;

// A[m][k] (128x64) and B[k][n] (64x128) given in plain row-major global memory
__global__ void probe(const __half* A, const __half* B, float* D, uint32_t lbo, uint32_t sbo, uint32_t kstep) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32768);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // A: K-major SW128: row m at m*128 B, 16-byte chunk j (8 k's) at (j ^ (m&7))*16
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) {
        int m = i / 64, k = i % 64;
        reinterpret_cast<__half*>(smem + m * 128 + (((k / 8) ^ (m & 7)) * 16))[k % 8] = A[i];
    }
    // B: MN-major SW128: [n-atom (2)][k (64)][64 n]: row k at atom*8192 + k*128, chunk j at (j ^ (k&7))*16
    for (int i = threadIdx.x; i < 64 * 128; i += blockDim.x) {
        int k = i / 128, n = i % 128;
        int atom = n / 64, nn = n % 64;
        reinterpret_cast<__half*>(smem + 16384 + atom * 8192 + k * 128 + (((nn / 8) ^ (k & 7)) * 16))[nn % 8] = B[i];
    }
    fence_proxy_async_smem();
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc(slot, 128); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = umma_idesc_16_bmn(128, 128, 0);
        const uint32_t a_addr = smem_u32(smem);
        for (int k = 0; k < 4; ++k) {
            uint64_t a_desc = umma_desc_sw128(a_addr) + 2 * k;
            uint64_t b_desc = umma_desc_mn_sw128(a_addr + 16384 + k * kstep, lbo, sbo);
            umma_f16(tmem, a_desc, b_desc, idesc, k > 0);
        }
        umma_commit(bar);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(tmem + ((uint32_t)(warp * 32) << 16) + c0, acc);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * 128 + c0 + j] = __uint_as_float(acc[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 128); }
}

int main(int argc, char** argv) {
    uint32_t lbo = argc > 1 ? atoi(argv[1]) : 8192, sbo = argc > 2 ? atoi(argv[2]) : 1024, kstep = argc > 3 ? atoi(argv[3]) : 2048;
    std::vector<__half> A(128 * 64), B(64 * 128);
    std::vector<float> Af(128 * 64), Bf(64 * 128);
    srand(1);
    for (size_t i = 0; i < A.size(); ++i) { Af[i] = (rand() % 17 - 8) / 8.0f; A[i] = __float2half(Af[i]); }
    for (size_t i = 0; i < B.size(); ++i) { Bf[i] = (rand() % 13 - 6) / 4.0f; B[i] = __float2half(Bf[i]); }
    __half *dA, *dB; float* dD;
    cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, 128 * 128 * 4);
    cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 + 64);
    probe<<<1, 128, 32768 + 64>>>(dA, dB, dD, lbo, sbo, kstep);
    cudaError_t e = cudaDeviceSynchronize();
    printf("lbo %u sbo %u kstep %u: %s\n", lbo, sbo, kstep, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<float> D(128 * 128);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0; double maxd = 0;
    for (int m = 0; m < 128; ++m) for (int n = 0; n < 128; ++n) {
        float want = 0; for (int k = 0; k < 64; ++k) want += Af[m * 64 + k] * Bf[k * 128 + n];
        double d = fabs(want - D[m * 128 + n]); if (d > maxd) maxd = d;
        if (d > 1e-3 && bad++ < 5) printf("  mismatch m%d n%d got %f want %f\n", m, n, D[m * 128 + n], want);
    }
    printf("  mismatches %d max|d| %g\n", bad, maxd);
    return 0;
}
