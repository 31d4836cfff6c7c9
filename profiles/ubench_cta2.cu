// Probe for the next round's two-half-instances-per-SM design: where does tcgen05.mma.cta_group::2 with M = 128
// (64 rows per CTA of the pair) put its accumulator in each CTA's TMEM, and how are the B rows split over the pair?
// A[r][0] = r, A[r][1] = 1, B[n][0] = 1, B[n][1] = n / 256  =>  D[r][n] = r + n / 256 (exact in fp16 x fp16 -> fp32),
// so every TMEM cell that the MMA wrote names its (row, column).  Each CTA holds ITS 64 rows of A and ITS N / 2 rows of
// B at the same shared-memory offsets; the leader issues one MMA; both CTAs dump lanes 0..127 x N columns.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I eam_rl4co_b200/csrc -o ubench_cta2.bin profiles/ubench_cta2.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"

using namespace r4::tc;

constexpr int kN = 32;            // MMA N (both CTAs together)
constexpr uint32_t kLBO = 128, kSBO = 256;   // K = 16 halves = two 16-byte chunks per 8-row group

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void put(unsigned char* tile, int r, int k, float v) {
    *reinterpret_cast<__half*>(tile + (r >> 3) * kSBO + (k >> 3) * kLBO + (r & 7) * 16 + (k & 7) * 2) = __float2half(v);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) probe(float* out, int swap_b, int ts_form) {
    __shared__ __align__(128) unsigned char a_tile[8 * kSBO];          // 64 rows x 16 halves
    __shared__ __align__(128) unsigned char b_tile[(kN / 16) * kSBO];  // N/2 rows x 16 halves
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_ptr;
    const uint32_t rank = cluster_rank();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (int)sizeof(a_tile); i += blockDim.x) a_tile[i] = 0;
    for (int i = threadIdx.x; i < (int)sizeof(b_tile); i += blockDim.x) b_tile[i] = 0;
    __syncthreads();
    if (threadIdx.x < 64) {
        const int r = threadIdx.x;
        put(a_tile, r, 0, (float)(64 * rank + r));      // global row index
        put(a_tile, r, 1, 1.0f);
    }
    if (threadIdx.x < kN / 2) {
        const int nl = threadIdx.x;
        const int n = (kN / 2) * (swap_b ? 1 - rank : rank) + nl;      // global B row (= D column) this CTA provides
        put(b_tile, nl, 0, 1.0f);
        put(b_tile, nl, 1, (float)n / 256.0f);
    }
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(64u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = tmem_ptr;
    // poison the columns so untouched cells are recognisable
    {
        const uint32_t z[8] = {0xffc00000u, 0xffc00000u, 0xffc00000u, 0xffc00000u, 0xffc00000u, 0xffc00000u, 0xffc00000u, 0xffc00000u};
        for (int c = 0; c < 64; c += 8) tmem_st8(tmem + ((uint32_t)(32 * warp) << 16) + c, z);
        tc_wait_st();
    }
    if (ts_form) {
        // A operand in TMEM columns [32, 40): every lane offers a row; lanes 0..63 say "row 64 rank + lane", lanes 64..127
        // say "500 + ..." so that D tells which lanes the pair-MMA reads its A rows from
        const int l = threadIdx.x;
        const float v = (l < 64) ? (float)(64 * rank + l) : (float)(500 + 64 * rank + (l - 64));
        const __half2 w0 = __floats2half2_rn(v, 1.0f);
        uint32_t a8[8] = {*reinterpret_cast<const uint32_t*>(&w0), 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        tmem_st8(tmem + ((uint32_t)(32 * warp) << 16) + 32u, a8);
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    if (rank == 0 && threadIdx.x == 0 && ts_form) {
        const uint32_t id = idesc_f16(128, kN);
        const uint64_t bd = smem_desc(smem_u32(b_tile), kLBO, kSBO);
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem),
            "r"(tmem + 32u), "l"(bd), "r"(id), "r"(0u)
            : "memory");
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                         smem_u32(&bar)),
                     "h"((uint16_t)3)
                     : "memory");
    }
    if (rank == 0 && threadIdx.x == 0 && !ts_form) {
        const uint32_t id = idesc_f16(128, kN);
        const uint64_t ad = smem_desc(smem_u32(a_tile), kLBO, kSBO);
        const uint64_t bd = smem_desc(smem_u32(b_tile), kLBO, kSBO);
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
            "l"(ad), "l"(bd), "r"(id), "r"(0u)
            : "memory");
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                         smem_u32(&bar)),
                     "h"((uint16_t)3)
                     : "memory");
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < 64; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16) + c0, v);
        tc_wait_ld();
        for (int i = 0; i < 16; ++i) out[(rank * 128 + 32 * warp + lane) * 64 + c0 + i] = __uint_as_float(v[i]);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u) : "memory");
}

int main() {
    float* d;
    cudaMalloc(&d, 2 * 128 * 64 * sizeof(float));
    for (int mode = 0; mode < 3; ++mode) {
        const int swap_b = mode == 1, ts_form = mode == 2;
        cudaMemset(d, 0, 2 * 128 * 64 * sizeof(float));
        probe<<<2, 128>>>(d, swap_b, ts_form);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            printf("swap_b=%d: %s\n", swap_b, cudaGetErrorString(e));
            return 1;
        }
        std::vector<float> h(2 * 128 * 64);
        cudaMemcpy(h.data(), d, h.size() * sizeof(float), cudaMemcpyDeviceToHost);
        printf("== %s; B rows [0,N/2) provided by CTA %d ==\n", ts_form ? "A from TMEM columns [32,40)" : "A from shared memory", swap_b);
        for (int cta = 0; cta < 2; ++cta) {
            printf("CTA %d: lane -> (row of D, first columns as D column index; '.' = untouched)\n", cta);
            for (int lane = 0; lane < 128; lane += 1) {
                if (!(lane % 32 < 3 || lane % 32 == 15 || lane % 32 == 16 || lane % 32 == 31)) continue;
                printf("  lane %3d:", lane);
                for (int c = 0; c < 28; c += (c < 2 ? 1 : 13)) {
                    const float v = h[(cta * 128 + lane) * 64 + c];
                    if (v != v) printf("  c%-2d .      ", c);
                    else printf("  c%-2d r%3d n%3d", c, (int)v, (int)((v - (int)v) * 256.f + 0.5f));
                }
                printf("\n");
            }
        }
    }
    return 0;
}
