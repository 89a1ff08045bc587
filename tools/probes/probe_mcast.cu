// Probe: does TMA multicast across a thread-block cluster lift the ~60 B/clk per-SM L2->SM ingest limit
// of the weight ring (probe_mmaring.cu)?  CSZ CTAs of a cluster consume the SAME 16 KB A tiles; each CTA
// loads 1/CSZ of every tile (a slice of rows) and multicasts it to all CTAs of the cluster.  A slot may
// be refilled only after every CTA's MMAs on it retired: the empty barriers count CSZ multicast commits.
// Reports cycles per 16 KB stage (4 MMAs M128 N128 K16 = 256 tensor cycles).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o probe_mcast probe_mcast.cu -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../hivemind_b200/csrc/hm_ptx.cuh"
using namespace hm;

__device__ __forceinline__ void tma_load_2d_mc(void* dst, const void* tm, uint64_t* bar, int c0, int c1, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
        : "memory");
}
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)), "h"(mask)
                 : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

__global__ void fill_random(__half* p, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t h = (uint32_t)i * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        p[i] = __float2half(((h & 0xffff) / 65536.0f - 0.5f) * 0.25f);
    }
}
static int g_long = 0;   // >0: long runs timed in milliseconds (power-throttled wall clock), random operands
template <int CSZ, int SLOTS, int GROUP, bool NOTMA, int PROD>
__global__ void __launch_bounds__(128, 1) mcring(const __grid_constant__ CUtensorMap tm, int n_laps, int rows_total,
                                                 long long* cycles_out, int layout, const uint8_t* flat) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* bop = smem;
    uint8_t* ringp = smem + 16384;
    constexpr int NB = SLOTS / GROUP;
    constexpr int ROWS = 128 / CSZ;  // rows of each tile this CTA loads
    uint64_t* full = reinterpret_cast<uint64_t*>(ringp + SLOTS * 16384);
    uint64_t* empty = full + 16;
    uint32_t* slot = reinterpret_cast<uint32_t*>(empty + 16);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = CSZ > 1 ? cluster_ctarank() : 0;
    const uint16_t mask = static_cast<uint16_t>((1u << CSZ) - 1);
    if (threadIdx.x == 0) {
        for (int i = 0; i < NB; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], CSZ); }
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc(slot, 128); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) reinterpret_cast<uint32_t*>(bop)[i] = (layout & 16) ? (0x2c002c00u ^ ((i * 2654435761u) & 0x83ff83ffu)) : 0;
    layout &= 15;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (CSZ > 1) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = *slot;
    long long t0 = clock64();
    if (warp == 0 || (PROD == 2 && warp == 2)) {
        const bool leader = elect_one();
        const int tiles = rows_total / 128;
        const int me = warp >> 1;
        uint32_t par = 1;
        for (int lap = 0; lap < n_laps; ++lap) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                if (PROD == 2 && (b & 1) != me) continue;
                mbar_wait(&empty[b], par);
                if (leader) {
                    if (NOTMA) {
                        mbar_arrive(&full[b]);
                    } else {
                        mbar_arrive_expect_tx(&full[b], 16384 * GROUP);
#pragma unroll
                        for (int j = 0; j < GROUP; ++j) {
                            const int s = b * GROUP + j;
                            const int tile = (lap * SLOTS + s) % tiles;
                            uint8_t* dst = ringp + s * 16384 + rank * ROWS * 128;
                            if (CSZ > 1) tma_load_2d_mc(dst, &tm, &full[b], 0, tile * 128 + rank * ROWS, mask);
                            else if (layout == 3) bulk_load_1d(dst, flat + (size_t)tile * 16384, 16384, &full[b]);
                            else if (layout == 0) tma_load_2d(dst, &tm, &full[b], 0, tile * 128);
                            else if (layout == 1) tma_load_2d(dst, &tm, &full[b], (tile % 6) * 64, (tile / 6) * 128);
                            else tma_load_2d(dst, &tm, &full[b], (tile % 2) * 64, (tile / 2) * 128);
                        }
                    }
                }
                __syncwarp();
            }
            par ^= 1;
        }
    } else if (warp == 1) {
        const bool leader = elect_one();
        const uint32_t idesc = umma_idesc_16_bmn(128, 128, 0);
        const uint64_t a0d = umma_desc_sw128(smem_u32(ringp));
        const uint32_t a0 = static_cast<uint32_t>(a0d), a_hi = static_cast<uint32_t>(a0d >> 32);
        const uint64_t bd0 = umma_desc_mn_sw128(smem_u32(bop), 8192, 1024);
        const uint32_t b0 = static_cast<uint32_t>(bd0), b_hi = static_cast<uint32_t>(bd0 >> 32);
        uint32_t par = 0;
        for (int lap = 0; lap < n_laps; ++lap) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                mbar_wait(&full[b], par);
                tc_fence_after();
                if (leader) {
#pragma unroll
                    for (int j = 0; j < GROUP; ++j) {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16_lh(tmem, a0 + ((b * GROUP + j) * 1024 + 2 * k), a_hi, b0 + k * 128, b_hi, idesc, 1u);
                    }
                    if (CSZ > 1) umma_commit_mc(&empty[b], mask);
                    else umma_commit(&empty[b]);
                }
                __syncwarp();
            }
            par ^= 1;
        }
        if (leader) umma_commit(&full[0]);
        mbar_wait(&full[0], par);
        if (leader) cycles_out[blockIdx.x] = clock64() - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (CSZ > 1) cluster_sync_all();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 128); }
}


// MMA thread alone (no producer, operands fixed in smem): what do commit / wait / fence cost per group?
// MODE 0: MMAs only; 1: + commit per group (nobody waits); 2: + wait on the barrier committed NB groups ago
// (ring reuse handshake without a producer); 3: as 2 + tcgen05.fence::after_thread_sync after the wait
template <int GROUP, int NB, int MODE>
__global__ void __launch_bounds__(128, 1) mmaonly(int n_groups, long long* cycles_out, int rec_from) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* bop = smem;
    uint8_t* ringp = smem + 16384;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ringp + NB * GROUP * 16384);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 32);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 32; ++i) mbar_init(&bars[i], 1);
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc(slot, 128); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 4096 * (1 + NB * GROUP); i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (warp == 1 && lane == 0) {
        const uint32_t idesc = umma_idesc_16_bmn(128, 128, 0);
        const uint64_t a0 = umma_desc_sw128(smem_u32(ringp));
        uint64_t bd[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) bd[k] = umma_desc_mn_sw128(smem_u32(bop) + k * 2048, 8192, 1024);
        long long t0 = clock64();
        uint32_t par = 1;
        for (int g = 0; g < n_groups; g += NB) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                const bool rec = g + b >= rec_from && g + b < rec_from + 12;
                long long ta = rec ? clock64() : 0;
                if (MODE >= 2) { mbar_wait(&bars[b], par); if (MODE >= 3) tc_fence_after(); }
                long long tb = rec ? clock64() : 0;
#pragma unroll
                for (int j = 0; j < GROUP; ++j)
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_f16(tmem, a0 + ((b * GROUP + j) * 1024 + 2 * k), bd[k], idesc, 1u);
                long long tc = rec ? clock64() : 0;
                if (MODE >= 1) umma_commit(&bars[b]);
                if (rec) {
                    long long* o = cycles_out + 8 + (g + b - rec_from) * 4;
                    o[0] = ta - t0; o[1] = tb - t0; o[2] = tc - t0; o[3] = clock64() - t0;
                }
            }
            par ^= 1;
        }
        umma_commit(&bars[31]);
        mbar_wait(&bars[31], 0);
        cycles_out[0] = clock64() - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 128); }
}
template <int GROUP, int NB, int MODE>
static void run_only(long long* dc) {
    auto kern = mmaonly<GROUP, NB, MODE>;
    const int smem = (NB * GROUP + 1) * 16384 + 1024;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int n_groups = (2400 / (GROUP * NB)) * NB;
    for (int rep = 0; rep < 2; ++rep) {
        kern<<<1, 128, smem>>>(n_groups, dc, 600 / GROUP);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("err %s\n", cudaGetErrorString(e)); exit(1); }
    }
    long long c;
    cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
    printf("mma-only group=%d nb=%d mode=%d  cycles/stage %8.1f  (per group %8.1f)\n", GROUP, NB, MODE, (double)c / (n_groups * GROUP),
           (double)c / n_groups);
    if (MODE == 2) {
        long long t[48];
        cudaMemcpy(t, dc + 8, sizeof(t), cudaMemcpyDeviceToHost);
        for (int i = 0; i < 12; ++i)
            printf("   group %2d: wait-begin %7lld  wait %4lld  issue %4lld  commit %4lld\n", i, t[i * 4], t[i * 4 + 1] - t[i * 4],
                   t[i * 4 + 2] - t[i * 4 + 1], t[i * 4 + 3] - t[i * 4 + 2]);
    }
}
static const uint8_t* g_flat;
template <int CSZ, int SLOTS, int GROUP, bool NOTMA = false, int PROD = 1>
static void run(const CUtensorMap& tm, int rows, long long* dc, int layout = 0) {
    auto kern = mcring<CSZ, SLOTS, GROUP, NOTMA, PROD>;
    const int smem = (SLOTS + 1) * 16384 + 1024;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (CSZ > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    const int n_laps = (2400 / SLOTS) * (g_long ? g_long : 1);
    for (int grid : {CSZ == 4 ? 132 : CSZ == 8 ? 120 : 148}) {
        if (grid % CSZ) continue;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(128);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = CSZ; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int max_clusters = 0;
        cudaOccupancyMaxActiveClusters(&max_clusters, kern, &cfg);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        float ms = 0;
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tm, n_laps, rows, dc, layout | (g_long ? 16 : 0), g_flat);
            cudaEventRecord(e1);
            if (e == cudaSuccess) e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("err %s\n", cudaGetErrorString(e)); exit(1); }
            cudaEventElapsedTime(&ms, e0, e1);
        }
        if (g_long) printf("   wall %.3f ms -> %.1f TF/s over %d CTAs, effective clock %.0f MHz\n", ms,
                           grid * (double)n_laps * SLOTS * 4 * 2.0 * 128 * 128 * 16 / (ms * 1e-3) / 1e12, grid,
                           0.0);
        std::vector<long long> c(grid);
        cudaMemcpy(c.data(), dc, grid * 8, cudaMemcpyDeviceToHost);
        double mx = 0, mn = 1e30; for (auto v : c) { if (v > mx) mx = v; if (v < mn) mn = v; }
        printf("prod %d layout %d %s csz=%d slots=%d group=%d grid=%4d max_active_clusters=%3d  cycles/stage max %8.1f min %8.1f\n", PROD, layout, NOTMA ? "no-tma" : "tma   ", CSZ, SLOTS, GROUP,
               grid, max_clusters, mx / (n_laps * SLOTS), mn / (n_laps * SLOTS));
    }
}

int main() {
    const int rows = 486 * 128;  // 486 tiles of 16 KB = 7.6 MB (L2-resident)
    __half* d;
    cudaMalloc(&d, (size_t)rows * 64 * 2);
    cudaMemset(d, 0, (size_t)rows * 64 * 2);
    long long* dc;
    cudaMalloc(&dc, 1024 * 8);
    cuuint64_t dims[2] = {64, (cuuint64_t)rows};
    cuuint64_t strides[1] = {128};
    cuuint32_t es[2] = {1, 1};
    auto make = [&](int box_rows) {
        CUtensorMap tm;
        cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
        CUresult r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode %d\n", (int)r); exit(1); }
        return tm;
    };
    CUtensorMap t128 = make(128), t64 = make(64), t32 = make(32), t16 = make(16);
    printf("tensor floor: 256 cycles per 16 KB stage; unicast ingest floor measured ~271\n");
    CUtensorMap tw1, tw2;
    {
        cuuint32_t box[2] = {64, 128};
        cuuint64_t d1[2] = {384, (cuuint64_t)rows / 6};
        cuuint64_t s1[1] = {768};
        CUresult r = cuTensorMapEncodeTiled(&tw1, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, d1, s1, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode w1 %d\n", (int)r); return 1; }
        cuuint64_t d2[2] = {128, (cuuint64_t)rows / 2};
        cuuint64_t s2[1] = {256};
        r = cuTensorMapEncodeTiled(&tw2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, d2, s2, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode w2 %d\n", (int)r); return 1; }
    }
    g_flat = reinterpret_cast<const uint8_t*>(d);
    fill_random<<<1024, 256>>>(d, (size_t)rows * 64);
    cudaDeviceSynchronize();
    g_long = 60;
    run<1, 6, 2, true>(t128, rows, dc, 0);
    run<1, 6, 2>(t128, rows, dc, 0);
    run<1, 6, 2>(tw1, rows, dc, 1);
    run<2, 6, 2>(t64, rows, dc, 0);
    run<4, 6, 2>(t32, rows, dc, 0);
    run<8, 6, 2>(t16, rows, dc, 0);
    run<1, 6, 2, true>(t128, rows, dc, 0);
    run<1, 6, 2>(t128, rows, dc, 0);
    return 0;
}
