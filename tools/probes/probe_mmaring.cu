// Probe: weight ring -> tcgen05.mma pipeline rate without any epilogue work.  A tiles [128 x 64] K-major
// stream through a ring (TMA), B is a fixed MN-major block in shared memory, D accumulates in TMEM.
// Reports cycles per 16 KB stage (4 MMAs of M128 N128 K16 = 256 tensor cycles) for several ring depths
// and MMAs-per-barrier groupings.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o probe_mmaring probe_mmaring.cu -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../hivemind_b200/csrc/hm_ptx.cuh"
using namespace hm;

// group = stages per full/empty barrier (1 -> 16 KB per barrier, 2 -> 32 KB ...)
__global__ void __launch_bounds__(128, 1) mmaring(const __grid_constant__ CUtensorMap tm, int slots, int group, int n_stages,
                                                  int rows_total, int do_mma, long long* cycles_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* bop = smem;                       // 16 KB B block
    uint8_t* ringp = smem + 16384;
    uint64_t* full = reinterpret_cast<uint64_t*>(ringp + slots * 16384);
    uint64_t* empty = full + 16;
    uint32_t* slot = reinterpret_cast<uint32_t*>(empty + 16);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nb = slots / group;  // barriers
    if (threadIdx.x == 0) {
        for (int i = 0; i < nb; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc(slot, 128); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) reinterpret_cast<uint32_t*>(bop)[i] = 0;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    const int n_groups = n_stages / group;
    long long t0 = clock64();
    if (warp == 0 && lane == 0) {
        int rb = (blockIdx.x * 7) % (rows_total / 128);
        int s = 0;
        for (int g = 0; g < n_groups; ++g) {
            const int b = g % nb;
            mbar_wait(&empty[b], ((g / nb) & 1) ^ 1);
            mbar_arrive_expect_tx(&full[b], 16384 * group);
            for (int j = 0; j < group; ++j, ++s) {
                tma_load_2d(ringp + (b * group + j) * 16384, &tm, &full[b], (s % 6) * 64, rb * 128);
                if (s % 6 == 5) rb = (rb + 1) % (rows_total / 128);
            }
        }
    } else if (warp == 1 && lane == 0) {
        const uint32_t idesc = umma_idesc_16_bmn(128, 128, 0);
        for (int g = 0; g < n_groups; ++g) {
            const int b = g % nb;
            mbar_wait(&full[b], (g / nb) & 1);
            tc_fence_after();
            if (do_mma) {
                for (int j = 0; j < group; ++j) {
                    const uint32_t a_addr = smem_u32(ringp + (b * group + j) * 16384);
                    const uint64_t a_desc = umma_desc_sw128(a_addr);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(tmem, a_desc + 2 * k, umma_desc_mn_sw128(smem_u32(bop) + k * 2048, 8192, 1024), idesc, 1u);
                }
                umma_commit(&empty[b]);
            } else {
                mbar_arrive(&empty[b]);
            }
        }
        if (do_mma) { umma_commit(&full[0]); }
        cycles_out[blockIdx.x] = clock64() - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 128); }
}


// "fast" variant: one lap over the ring is fully unrolled (slot addresses are compile-time offsets),
// B descriptors are hoisted, only the low descriptor word is recomputed with one 32-bit add.
template <int SLOTS, int GROUP>
__global__ void __launch_bounds__(128, 1) mmaring_fast(const __grid_constant__ CUtensorMap tm, int n_laps, int rows_total,
                                                       long long* cycles_out, int stagger, int mode,
                                                       const __grid_constant__ CUtensorMap tm2, const uint8_t* flat,
                                                       const __grid_constant__ CUtensorMap tm3) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* bop = smem;
    uint8_t* ringp = smem + 16384;
    constexpr int NB = SLOTS / GROUP;
    uint64_t* full = reinterpret_cast<uint64_t*>(ringp + SLOTS * 16384);
    uint64_t* empty = full + 16;
    uint32_t* slot = reinterpret_cast<uint32_t*>(empty + 16);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NB; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc(slot, 128); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) reinterpret_cast<uint32_t*>(bop)[i] = 0;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    long long t0 = clock64();
    if (warp == 0 && lane == 0) {
        int rb = stagger ? (blockIdx.x * 7) % (rows_total / 128) : 0;
        uint32_t par = 1;
        for (int lap = 0; lap < n_laps; ++lap) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                mbar_wait(&empty[b], par);
                mbar_arrive_expect_tx(&full[b], 16384 * GROUP);
                if (mode == 2) {
                    bulk_load_1d(ringp + b * GROUP * 16384, flat + ((size_t)rb * 6 + b * GROUP) * 16384, 16384 * GROUP, &full[b]);
                } else {
#pragma unroll
                    for (int j = 0; j < GROUP; ++j) {
                        const int s = b * GROUP + j;
                        if (mode == 3) tma_load_2d(ringp + s * 16384, &tm3, &full[b], 0, (rb * 6 + s) * 128);
                        else if (mode == 1) tma_load_2d(ringp + s * 16384, &tm2, &full[b], (s & 1) * 64, (rb * 3 + (s >> 1)) * 128);
                        else tma_load_2d(ringp + s * 16384, &tm, &full[b], (s % 6) * 64, rb * 128);
                    }
                }
            }
            par ^= 1;
            rb = rb + 1 == rows_total / 128 ? 0 : rb + 1;
        }
    } else if (warp == 1 && lane == 0) {
        const uint32_t idesc = umma_idesc_16_bmn(128, 128, 0);
        const uint64_t a0 = umma_desc_sw128(smem_u32(ringp));
        uint64_t bd[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) bd[k] = umma_desc_mn_sw128(smem_u32(bop) + k * 2048, 8192, 1024);
        uint32_t par = 0;
        for (int lap = 0; lap < n_laps; ++lap) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                mbar_wait(&full[b], par);
                tc_fence_after();
#pragma unroll
                for (int j = 0; j < GROUP; ++j) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(tmem, a0 + ((b * GROUP + j) * 1024 + 2 * k), bd[k], idesc, 1u);
                }
                umma_commit(&empty[b]);
            }
            par ^= 1;
        }
        umma_commit(&full[0]);
        cycles_out[blockIdx.x] = clock64() - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 128); }
}
static CUtensorMap g_tm2, g_tm3;
static const uint8_t* g_flat;
template <int SLOTS, int GROUP>
static void run_fast(const CUtensorMap& tm, int rows, long long* dc) {
    cudaFuncSetAttribute(mmaring_fast<SLOTS, GROUP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 13 * 16384 + 1024);
    const int n_laps = 2400 / SLOTS;
    for (int mode : {0, 1, 2, 3}) for (int stagger : {0}) for (int grid : {148}) {
        for (int rep = 0; rep < 2; ++rep) {
            mmaring_fast<SLOTS, GROUP><<<grid, 128, (SLOTS + 1) * 16384 + 1024>>>(tm, n_laps, rows, dc, stagger, mode, g_tm2, g_flat, g_tm3);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("err %s\n", cudaGetErrorString(e)); exit(1); }
        }
        std::vector<long long> c(grid);
        cudaMemcpy(c.data(), dc, grid * 8, cudaMemcpyDeviceToHost);
        double mx = 0; for (auto v : c) if (v > mx) mx = v;
        printf("fast mode=%d %4d %5d %5d  %10.1f\n", mode, grid, SLOTS, GROUP, mx / (n_laps * SLOTS));
    }
}

int main() {
    const int rows = 81 * 128, K = 384;
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
    {   // W2-like view of the same memory: [rows*3][128] fp16, 256-byte pitch
        cuuint64_t d2[2] = {128, (cuuint64_t)rows * 3};
        cuuint64_t s2[1] = {256};
        r = cuTensorMapEncodeTiled(&g_tm2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, d2, s2, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode2 %d\n", (int)r); return 1; }
        g_flat = reinterpret_cast<const uint8_t*>(d);
        cuuint64_t d3[2] = {64, (cuuint64_t)rows * 6};
        cuuint64_t s3[1] = {128};
        r = cuTensorMapEncodeTiled(&g_tm3, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, d3, s3, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode3 %d\n", (int)r); return 1; }
    }
    long long* dc;
    cudaMalloc(&dc, 1024 * 8);
    cudaFuncSetAttribute(mmaring, cudaFuncAttributeMaxDynamicSharedMemorySize, 13 * 16384 + 1024);
    const int n_stages = 2400;
    printf("mma grid slots group  cycles/stage (tensor floor 256)\n");
    for (int do_mma : {1}) for (int grid : {148}) for (int group : {2}) for (int slots : {4}) {
        if (slots % group || slots / group < 2) continue;
        for (int rep = 0; rep < 2; ++rep) {
            mmaring<<<grid, 128, (slots + 1) * 16384 + 1024>>>(tm, slots, group, n_stages, rows, do_mma, dc);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("err %s\n", cudaGetErrorString(e)); return 1; }
        }
        std::vector<long long> c(grid);
        cudaMemcpy(c.data(), dc, grid * 8, cudaMemcpyDeviceToHost);
        double mx = 0; for (auto v : c) if (v > mx) mx = v;
        printf("%3d %4d %5d %5d  %10.1f\n", do_mma, grid, slots, group, mx / n_stages);
    }
    run_fast<6, 2>(tm, rows, dc);
    run_fast<6, 3>(tm, rows, dc);
    run_fast<8, 2>(tm, rows, dc);
    return 0;
}
