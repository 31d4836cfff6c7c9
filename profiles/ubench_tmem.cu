// Micro-benchmarks behind DESIGN.md §4.2: what one SM sustains for the resources the rollout kernel leans on,
// in the kernel's own configuration (one CTA of 512 threads = 16 warps per SM, all 512 TMEM columns allocated).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I eam_rl4co_b200/csrc -o ubench_tmem profiles/ubench_tmem.cu
//   ./ubench_tmem
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"

using namespace r4::tc;

constexpr int kIters = 256;

template <int MODE>
__global__ void __launch_bounds__(512, 1) bench_kernel(long long* out, float* sink, int nwarps) {
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = tmem_slot + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(warp >> 2) * 112u;
    float acc = 0.f;
    uint32_t v[16], w8[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w8[i] = threadIdx.x * 8 + i;
    {   // initialise this thread's columns
        for (int c = 0; c < 112; c += 8) tmem_st8(base + c, w8);
        tc_wait_st();
    }
    __syncthreads();
    const long long t0 = clock64();
    if (warp < nwarps) {
        if (MODE == 0) {            // tcgen05.ld 32x32b.x16, 7 loads in flight then one wait
            for (int it = 0; it < kIters; ++it) {
#pragma unroll
                for (int c = 0; c < 7; ++c) {
                    tmem_ld16(base + 16u * c, v);
                    tc_wait_ld();
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc += __uint_as_float(v[i]);
                }
            }
        } else if (MODE == 1) {     // x16 loads, double buffered like the kernel
            uint32_t u[16];
            for (int it = 0; it < kIters; ++it) {
                tmem_ld16(base, v);
#pragma unroll
                for (int c = 0; c < 6; c += 2) {
                    tc_wait_ld();
                    tmem_ld16(base + 16u * (c + 1), u);
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc += __uint_as_float(v[i]);
                    tc_wait_ld();
                    tmem_ld16(base + 16u * (c + 2), v);
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc += __uint_as_float(u[i]);
                }
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 16; ++i) acc += __uint_as_float(v[i]);
            }
        } else if (MODE == 2) {     // tcgen05.st x8, 14 stores then one wait
            for (int it = 0; it < kIters; ++it) {
#pragma unroll
                for (int c = 0; c < 14; ++c) tmem_st8(base + 8u * c, w8);
                tc_wait_st();
            }
        } else if (MODE == 3) {     // MUFU.EX2: 16 independent chains
            float x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = -0.001f * (lane + i);
            for (int it = 0; it < kIters; ++it) {
#pragma unroll
                for (int r = 0; r < 7; ++r) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) x[i] = ex2_approx(x[i]);
                }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) acc += x[i];
        } else if (MODE == 4) {     // split2 (F2FP + 2 FHADD + F2FP per pair): 8 pairs x 7
            float x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = 0.37f * (lane + i);
            uint32_t hs = 0, ls = 0;
            for (int it = 0; it < kIters; ++it) {
#pragma unroll
                for (int r = 0; r < 7; ++r) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        uint32_t h, l;
                        split2(x[2 * i], x[2 * i + 1], h, l);
                        hs ^= h;
                        ls ^= l;
                        x[2 * i] += 1.0f;
                        x[2 * i + 1] += 1.0f;
                    }
                }
            }
            acc += __uint_as_float(hs ^ ls);
        } else if (MODE == 5) {     // MUFU.EX2 + 7 FFMA per element (independent): do the pipes overlap?
            float x[16], y[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) { x[i] = -0.001f * (lane + i); y[i] = 0.5f * i; }
            for (int it = 0; it < kIters; ++it) {
#pragma unroll
                for (int r = 0; r < 7; ++r) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        x[i] = ex2_approx(x[i]);
#pragma unroll
                        for (int k = 0; k < 7; ++k) y[i] = fmaf(y[i], 0.999f, 0.001f);
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) acc += x[i] + y[i];
        } else if (MODE == 6) {     // loads (double buffered) + 16 MUFU + stores per chunk: the softmax skeleton
            uint32_t u[16];
            for (int it = 0; it < kIters; ++it) {
                tmem_ld16(base, v);
#pragma unroll
                for (int c = 0; c < 7; ++c) {
                    tc_wait_ld();
                    if (c + 1 < 7) tmem_ld16(base + 16u * (c + 1), u);
                    uint32_t h8[8], l8[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float p0 = ex2_approx(__uint_as_float(v[2 * i]) * 1e-30f);
                        const float p1 = ex2_approx(__uint_as_float(v[2 * i + 1]) * 1e-30f);
                        acc += p0 + p1;
                        split2(p0, p1, h8[i], l8[i]);
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = u[i];
                    tmem_st8(base + 16u * c, h8);
                    tmem_st8(base + 16u * c + 8u, l8);
                }
                tc_wait_st();
            }
        }
    }
    const long long t1 = clock64();
    if (lane == 0) out[blockIdx.x * 16 + warp] = t1 - t0;
    sink[blockIdx.x * 512 + threadIdx.x] = acc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_slot, 512);
}

template <int MODE>
static void run(const char* name, double units_per_warp_iter, const char* unit) {
    long long* out;
    float* sink;
    cudaMalloc(&out, 148 * 16 * sizeof(long long));
    cudaMalloc(&sink, 148 * 512 * sizeof(float));
    for (int nw : {1, 4, 8, 16}) {
        cudaMemset(out, 0, 148 * 16 * sizeof(long long));
        bench_kernel<MODE><<<148, 512>>>(out, sink, nw);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
        long long h[16];
        cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int w = 0; w < nw; ++w) mx = h[w] > mx ? h[w] : mx;
        printf("%-44s warps=%2d  %9lld cycles  %8.2f %s/clk/SM\n", name, nw, mx,
               units_per_warp_iter * kIters * nw / (double)mx, unit);
    }
    cudaFree(out);
    cudaFree(sink);
}

int main() {
    run<0>("tcgen05.ld 32x32b.x16, wait after each", 7 * 2048.0, "B");
    run<1>("tcgen05.ld 32x32b.x16, double buffered", 7 * 2048.0, "B");
    run<2>("tcgen05.st 32x32b.x8, 14 then wait", 14 * 1024.0, "B");
    run<3>("MUFU.EX2 (16 chains)", 7 * 16 * 32.0, "lane-op");
    run<4>("split2 (4 instr / pair)", 7 * 8 * 4.0, "warp-instr");
    run<5>("EX2 + 7 FFMA per element", 7 * 16 * 8.0, "warp-instr");
    run<6>("ld x16 + 16 EX2 + split + st (softmax skeleton)", 7 * 16.0, "elem-warp");
    return 0;
}
