// tcgen05.mma issue cost vs completion latency for the shapes of the rollout kernel (one CTA of 512 threads per SM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I eam_rl4co_b200/csrc -o profiles/ubench_mma.bin profiles/ubench_mma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"

using namespace r4::tc;

// MODE 0: one thread, NMMA TS-form MMAs (M128 N16 K16) into ONE accumulator
// MODE 1: one thread, NMMA TS-form MMAs alternating between TWO accumulators
// MODE 2: one thread, NMMA SS-form MMAs (M128 N112 K16) into one accumulator
// MODE 3: four threads (warps 0..3), each NMMA TS-form MMAs into its own accumulator
// MODE 4: one thread, NMMA TS-form MMAs (M128 N112 K16) into one accumulator (logits shape)
template <int MODE, bool ELECT = false>
__global__ void __launch_bounds__(512, 1) k(long long* out, int nmma) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 512) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // 1.0h
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(&slot, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = slot;
    {   // some finite A operand in TMEM columns [0,128)
        uint32_t w8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) w8[i] = 0x3c003c00u;
        const uint32_t lb = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(warp >> 2) * 32u;
        for (int c = 0; c < 32; c += 8) tmem_st8(lb + c, w8);
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t sb = smem_u32(smem);
    long long t_issue = 0, t_done = 0;
    bool me;
    if (ELECT) {
        me = false;
        if ((MODE == 3) ? (warp < 4) : (warp == 0)) me = elect_one();   // warp-uniform branch, then elect.sync
    } else {
        me = (MODE == 3) ? (warp < 4 && lane == 0) : (threadIdx.x == 0);
    }
    if (me) {
        const uint32_t idN16 = idesc_f16(128, 16), idN112 = idesc_f16(128, 112);
        const uint64_t b16 = smem_desc(sb, 128, 256);            // 16 x 16 B tile
        const uint64_t b112 = smem_desc(sb, 128, 2048);          // 112 x 16 B tile
        const uint64_t a128 = smem_desc(sb + 32768, 128, 256);   // 128 x 16 A tile
        const uint32_t d0 = tmem + 256u + 32u * (MODE == 3 ? warp : 0), d1 = tmem + 256u + 16u;
        const long long t0 = clock64();
        for (int i = 0; i < nmma; ++i) {
            if (MODE == 0 || MODE == 3) mma_ts(d0, tmem + (i & 7) * 8u, b16, idN16, i > 0);
            if (MODE == 1) mma_ts((i & 1) ? d1 : d0, tmem + (i & 7) * 8u, b16, idN16, i > 1);
            if (MODE == 2) mma_ss(tmem + 256u, a128, b112, idN112, i > 0);
            if (MODE == 4) mma_ts(tmem + 256u, tmem + (i & 7) * 8u, b112, idN112, i > 0);
        }
        tc_commit(&bar[MODE == 3 ? warp : 0]);
        const long long t1 = clock64();
        mbar_wait(&bar[MODE == 3 ? warp : 0], 0);
        const long long t2 = clock64();
        t_issue = t1 - t0;
        t_done = t2 - t0;
        out[blockIdx.x * 8 + (MODE == 3 ? warp : 0) * 2] = t_issue;
        out[blockIdx.x * 8 + (MODE == 3 ? warp : 0) * 2 + 1] = t_done;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int MODE, bool ELECT = false>
static void run(const char* name, int nmma) {
    long long* out;
    cudaMalloc(&out, 148 * 8 * sizeof(long long));
    cudaMemset(out, 0, 148 * 8 * sizeof(long long));
    cudaFuncSetAttribute(k<MODE, ELECT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int rep = 0; rep < 2; ++rep) k<MODE, ELECT><<<148, 512, 64 * 1024>>>(out, nmma);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    long long h[8];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-52s n=%3d  issue %6lld cyc (%5.1f / MMA)   issue+complete %6lld cyc (%5.1f / MMA)\n", name, nmma, h[0],
           (double)h[0] / nmma, h[1], (double)h[1] / nmma);
    cudaFree(out);
}

int main() {
    for (int n : {3, 21, 63}) {
        run<0>("1 thread, TS M128 N16 K16, one accumulator", n);
        run<1>("1 thread, TS M128 N16 K16, two accumulators", n);
        run<3>("4 threads, TS M128 N16 K16, own accumulators", n);
        run<0, true>("elect.sync: 1 thread, TS N16, one accumulator", n);
        run<3, true>("elect.sync: 4 threads, TS N16, own accumulators", n);
    }
    for (int n : {3, 12, 24}) {
        run<2>("1 thread, SS M128 N112 K16, one accumulator", n);
        run<4>("1 thread, TS M128 N112 K16, one accumulator", n);
        run<4, true>("elect.sync: TS M128 N112 K16, one accumulator", n);
        run<2, true>("elect.sync: SS M128 N112 K16, one accumulator", n);
    }
    return 0;
}
