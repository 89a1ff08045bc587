// Probe: steady-state rate of a TMA weight ring (16 KB boxes [128 rows x 64 K] from an L2-resident
// matrix) as a function of ring depth and grid size, with a trivial consumer.  Tells how many
// bytes must be in flight per SM to stream weights at the rate the tensor core consumes them.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o probe_ring probe_ring.cu -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../hivemind_b200/csrc/hm_ptx.cuh"
using namespace hm;

__global__ void ring(const __grid_constant__ CUtensorMap tm, const uint8_t* flat, int mode, int slots, int n_stages, int rows_total, int consume_cycles,
                     long long* cycles_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + slots * 16384);
    uint64_t* empty = full + 16;
    if (threadIdx.x == 0) {
        for (int i = 0; i < slots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        fence_barrier_init();
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5;
    long long t0 = clock64();
    if (warp == 0 && (threadIdx.x & 31) == 0) {
        int rb = (blockIdx.x * 7) % (rows_total / 128);
        for (int s = 0; s < n_stages; ++s) {
            const int sl = s % slots;
            mbar_wait(&empty[sl], ((s / slots) & 1) ^ 1);
            mbar_arrive_expect_tx(&full[sl], 16384);
            if (mode == 1) bulk_load_1d(smem + sl * 16384, flat + ((size_t)rb * 6 + (s % 6)) * 16384, 16384, &full[sl]);
            else tma_load_2d(smem + sl * 16384, &tm, &full[sl], (s % 6) * 64, rb * 128);
            if (s % 6 == 5) rb = (rb + 1) % (rows_total / 128);
        }
    } else if (warp == 1 && (threadIdx.x & 31) == 0) {
        for (int s = 0; s < n_stages; ++s) {
            const int sl = s % slots;
            mbar_wait(&full[sl], (s / slots) & 1);
            if (mode == 2) mbar_arrive(&empty[sl]);
            long long t = clock64();
            while (clock64() - t < consume_cycles) {}
            if (mode != 2) mbar_arrive(&empty[sl]);
        }
        cycles_out[blockIdx.x] = clock64() - t0;
    }
}

int main() {
    const int rows = 81 * 128, K = 384;  // W1cat-sized matrix (7.9 MB, L2-resident)
    __half* d;
    cudaMalloc(&d, (size_t)rows * K * 2);
    cudaMemset(d, 0, (size_t)rows * K * 2);
    CUtensorMap tm;
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t es[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, dims, strides, box, es,
        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r) { printf("encode %d\n", (int)r); return 1; }
    long long* dc;
    cudaMalloc(&dc, 1024 * 8);
    cudaFuncSetAttribute(ring, cudaFuncAttributeMaxDynamicSharedMemorySize, 13 * 16384 + 512);
    const int n_stages = 1200;
    printf("grid slots consume  cycles/stage  GB/s/SM(@1.9GHz)\n");
    printf("mode 0 = 2-D tensor box, 1 = 1-D bulk copy of a pre-tiled 16 KB block, 2 = tensor box with early slot release\n");
    for (int mode : {0, 1, 2}) for (int grid : {148}) for (int consume : {0, 128, 256, 512}) for (int slots : {2, 4, 8}) {
        for (int rep = 0; rep < 2; ++rep) {
            ring<<<grid, 64, slots * 16384 + 512>>>(tm, reinterpret_cast<const uint8_t*>(d), mode, slots, n_stages, rows, consume, dc);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("err %s\n", cudaGetErrorString(e)); return 1; }
        }
        std::vector<long long> c(grid);
        cudaMemcpy(c.data(), dc, grid * 8, cudaMemcpyDeviceToHost);
        double mx = 0; for (auto v : c) if (v > mx) mx = v;
        double per = mx / n_stages;
        printf("m%d %4d %5d %7d  %10.1f  %8.1f\n", mode, grid, slots, consume, per, 16384.0 / per * 1.9);
    }
    return 0;
}
