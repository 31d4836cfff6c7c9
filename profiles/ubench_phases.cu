// The CUDA-core phases of the tensor-path rollout kernel timed in isolation (softmax_inplace, select_row) in the
// kernel's configuration: 1 CTA of 512 threads per SM, NPAD = 112, n = 100.  Scores / logits are synthetic.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I eam_rl4co_b200/csrc -I include \
//        -o profiles/ubench_phases.bin profiles/ubench_phases.cu && ./profiles/ubench_phases.bin
#include <cstdio>
#include "../eam_rl4co_b200/csrc/rollout_tc.cu"

using namespace r4::tc;

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// MODE 0: refill only; 1: refill + softmax_inplace; 2: refill + select_row (greedy)
template <int MODE, bool SPECB = false>
__global__ void __launch_bounds__(kThreads, 1) phase_kernel(long long* out, float* sink, int reps, int feasible_pct, int big_logits,
                                                            rl4co_rollout_args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    Smem* s = reinterpret_cast<Smem*>(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    Ctx x;
    x.smem = smem;
    x.npad = 112;
    if (warp == 0) tmem_alloc(&s->tmem_ptr, kTmemCols);
    const int n = 100;
    for (int w = 0; w < 4; ++w) {
        uint32_t bits = 0;
        for (int i = 0; i < 32; ++i) {
            const int node = 32 * w + i;
            if (node < n && (int)(hash32(row * 131u + node + blockIdx.x * 7919u) % 100u) < feasible_pct) bits |= 1u << i;
        }
        if (j == 0) s->maskw[w][row] = bits;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    x.tmem = s->tmem_ptr;
    const uint32_t lane_base = x.tmem + ((uint32_t)(32 * q) << 16);
    float acc = 0.f;
    int err = 0;
    const uint32_t small_bits = big_logits ? 0x3f800000u : 0x3e000000u;   // logits in [2,4) or [0.25,0.5)
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
        // refill: scores of this thread's (row, head j) / logits quarter
        const uint32_t dst = (MODE == 2) ? lane_base + kColLG + 28u * (uint32_t)j
                                         : lane_base + kColSP + (uint32_t)(j * x.npad);
        const int ncol = (MODE == 2) ? 32 : 112;   // 28 logit columns per thread (8-column stores; overlaps rewritten)
        uint32_t lcg = (uint32_t)(row * 977 + r * 31 + j * 13) * 2654435761u;
        for (int c = 0; c < ncol; c += 8) {
            uint32_t w8[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {   // cheap synthetic scores in [1, 3)
                lcg = lcg * 1664525u + 1013904223u;
                w8[i] = (MODE == 2 ? small_bits : 0x3f800000u) | (lcg >> 9);
            }
            tmem_st8(dst + c, w8);
        }
        tc_wait_st();
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        if (MODE == 1) {
            float m;
            acc += softmax_inplace(x, lane_base + kColSP + (uint32_t)(j * x.npad), row, nullptr, n, SPECB, -8.0f, m);
            tc_fence_before();
            __syncthreads();
        } else if (MODE == 2) {
            const RowSel sel = select_row(x, a, n, q, j, row, row < 100, row, r, 100, 0u, -1, err);
            acc += sel.logp + sel.action;
            __syncthreads();
        }
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    sink[blockIdx.x * kThreads + threadIdx.x] = acc + err;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(x.tmem, kTmemCols);
}

template <int MODE, bool SPECB = false>
static double run(const char* name, int pct, double base, int big = 0) {
    long long* out;
    float* sink;
    cudaMalloc(&out, 148 * sizeof(long long));
    cudaMalloc(&sink, 148 * kThreads * sizeof(float));
    rl4co_rollout_args a = {};
    a.mode = RL4CO_MODE_GREEDY;
    a.tanh_clipping = 10.f;
    a.temperature = 1.f;
    const int reps = 200;
    const size_t smem = kOffTiles;
    cudaFuncSetAttribute(phase_kernel<MODE, SPECB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    phase_kernel<MODE, SPECB><<<148, kThreads, smem>>>(out, sink, reps, pct, big, a);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return 0; }
    long long h;
    cudaMemcpy(&h, out, sizeof(h), cudaMemcpyDeviceToHost);
    const double per = (double)h / reps;
    printf("%-28s feasible=%3d%%  %8.0f cycles/call  (net of refill %8.0f)\n", name, pct, per, per - base);
    cudaFree(out);
    cudaFree(sink);
    return per;
}

int main() {
    for (int pct : {100, 50, 15}) {
        const double b1 = run<0>("refill (112 cols)", pct, 0);
        run<1>("softmax exact (online)", pct, b1);
        run<1, true>("softmax speculative", pct, b1);
        run<2>("select_row greedy |v|<1", pct, b1 * 0.5);
        run<2>("select_row greedy |v|>1", pct, b1 * 0.5, 1);
    }
    return 0;
}
