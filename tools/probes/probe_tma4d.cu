// Probe: what does a 4-D TMA box (x=8, y=8, c=64, pos=2) with SWIZZLE_128B write to shared memory?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o probe_tma4d probe_tma4d.cu -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../hivemind_b200/csrc/hm_ptx.cuh"
using namespace hm;

__global__ void probe(const __grid_constant__ CUtensorMap tm, uint16_t* out, int dx, int dy) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384);
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
        mbar_arrive_expect_tx(bar, 16384);
        if (dx == 99) tma_load_3d(smem, &tm, bar, 8 * dy, 0, 0);
        else tma_load_4d(smem, &tm, bar, dx, dy, 0, 0);
    }
    __syncthreads();
    mbar_wait(bar, 0);
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) out[i] = reinterpret_cast<uint16_t*>(smem)[i];
}

int main(int argc, char** argv) {
    const int mode = argc > 1 ? atoi(argv[1]) : 0;
    const int n = 2, C = 74;
    std::vector<__half> h(n * C * 64);
    // value encodes (pos, c, sq): pos*4096 + c*64 + sq  (exact in fp16 up to 2048 only) -> use bits directly
    std::vector<uint16_t> hb(n * C * 64);
    for (int p = 0; p < n; ++p) for (int c = 0; c < C; ++c) for (int s = 0; s < 64; ++s) hb[(p * C + c) * 64 + s] = 1 + p * 8192 + c * 64 + s;
    uint16_t *d, *dout;
    cudaMalloc(&d, hb.size() * 2); cudaMemcpy(d, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice);
    cudaMalloc(&dout, 16384);
    CUtensorMap tm;
    cuuint64_t dims[4] = {8, 8, (cuuint64_t)C, (cuuint64_t)n};
    cuuint64_t strides[3] = {16, 128, (cuuint64_t)C * 128};
    cuuint32_t box[4] = {8, 8, 64, 2};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, d, dims, strides, box, es,
        CU_TENSOR_MAP_INTERLEAVE_NONE, mode == 1 ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (mode == 2) {
        cuuint64_t d3[3] = {64, (cuuint64_t)C, (cuuint64_t)n};
        cuuint64_t s3[2] = {128, (cuuint64_t)C * 128};
        cuuint32_t b3[3] = {64, 64, 2};
        r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, d, d3, s3, b3, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    printf("mode %d encode: %d\n", mode, (int)r);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 + 64);
    for (int t = 0; t < 2; ++t) {
        int dx = t ? 1 : 0, dy = t ? -1 : 0;
        if (mode == 2) dx = 99;
        probe<<<1, 128, 16384 + 64>>>(tm, dout, dx, dy);
        cudaError_t e = cudaDeviceSynchronize();
        printf("shift (%d,%d): %s\n", dx, dy, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        std::vector<uint16_t> o(8192);
        cudaMemcpy(o.data(), dout, 16384, cudaMemcpyDeviceToHost);
        // expected: element (pos p, ch c, sq s) at byte p*8192 + c*128 + ((s/8) ^ (c&7))*16 + (s%8)*2
        int bad = 0;
        for (int p = 0; p < 2; ++p) for (int c = 0; c < 64; ++c) for (int s = 0; s < 64; ++s) {
            int y = s / 8 + dy, x = s % 8 + (dx == 99 ? 0 : dx);
            uint16_t want = (x < 0 || x > 7 || y < 0 || y > 7) ? 0 : 1 + p * 8192 + c * 64 + y * 8 + x;
            uint16_t got = o[(p * 8192 + c * 128 + (((s / 8) ^ (c & 7)) * 16) + (s % 8) * 2) / 2];
            if (got != want && bad++ < 5) printf("  mismatch p%d c%d s%d got %u want %u\n", p, c, s, got, want);
        }
        printf("  mismatches: %d\n", bad);
        if (bad) { printf("  first 64 elems:"); for (int i = 0; i < 64; ++i) printf(" %u", o[i]); printf("\n  elems 64..127:"); for (int i = 64; i < 128; ++i) printf(" %u", o[i]); printf("\n"); }
    }
    return 0;
}
