// Instance normalisation of the encoder (nn/ops.py:48-49: nn.InstanceNorm1d over the node axis of [B, N, E], affine, biased
// variance, eps inside the square root) with its backward, for the TRAINING path: the module-by-module PyTorch statement is ~9
// small kernels forward and ~20 backward per call (12 calls per POMO encode: 74 reductions + ~150 elementwise launches, ~1.5 ms
// of the c5 step); here it is one kernel each way.  (Inference runs the fused encoder kernels of encoder_tc.cu.)
// CTA = instance b, thread = (node group g of 8, channel e): the <= 16 values of a thread stay in registers, the statistics of
// a channel are combined over the 8 groups through shared memory.  fp32 throughout, two-pass variance (mean first).
#include "common.cuh"

namespace r4 {
namespace nrm {

constexpr int kGroups = 8;
constexpr int kPer = 16;                                  // nodes per thread: N <= 128
constexpr int kThreads = kGroups * kE;                    // 1024

// sum of v over the 8 node groups of channel e (all threads call)
__device__ __forceinline__ float group_sum(float v, float* s_red, int g, int e) {
    __syncthreads();                                      // s_red may still be read from the previous call
    s_red[g * kE + e] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < kGroups; ++k) t += s_red[k * kE + e];
    return t;
}

__global__ void __launch_bounds__(kThreads)
instnorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                    float* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int N, float eps) {
    __shared__ float s_red[kGroups * kE];
    const int e = threadIdx.x & (kE - 1), g = threadIdx.x >> 7;
    const int64_t base = (int64_t)blockIdx.x * N * kE + e;
    float v[kPer];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int nd = g + kGroups * k;
        v[k] = nd < N ? x[base + (int64_t)nd * kE] : 0.f;
        s += v[k];
    }
    const float mean = group_sum(s, s_red, g, e) / (float)N;
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int nd = g + kGroups * k;
        const float c = nd < N ? v[k] - mean : 0.f;
        v[k] = c;
        sq = fmaf(c, c, sq);
    }
    const float rstd = rsqrtf(group_sum(sq, s_red, g, e) / (float)N + eps);
    const float we = w[e] * rstd, be = bias[e];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int nd = g + kGroups * k;
        if (nd < N) y[base + (int64_t)nd * kE] = fmaf(v[k], we, be);
    }
    if (g == 0) {
        mean_out[(int64_t)blockIdx.x * kE + e] = mean;
        rstd_out[(int64_t)blockIdx.x * kE + e] = rstd;
    }
}

// dx = w rstd (dy - mean(dy) - xhat mean(dy xhat)); dw[e] += sum dy xhat, db[e] += sum dy (over nodes here, over b by atomics)
__global__ void __launch_bounds__(kThreads)
instnorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ mean_in,
                    const float* __restrict__ rstd_in, const float* __restrict__ dy, float* __restrict__ dx,
                    float* __restrict__ dw, float* __restrict__ db, int N) {
    __shared__ float s_red[kGroups * kE];
    const int e = threadIdx.x & (kE - 1), g = threadIdx.x >> 7;
    const int64_t base = (int64_t)blockIdx.x * N * kE + e;
    const float mean = mean_in[(int64_t)blockIdx.x * kE + e], rstd = rstd_in[(int64_t)blockIdx.x * kE + e];
    float xh[kPer], gy[kPer];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int nd = g + kGroups * k;
        const bool ok = nd < N;
        xh[k] = ok ? (x[base + (int64_t)nd * kE] - mean) * rstd : 0.f;
        gy[k] = ok ? dy[base + (int64_t)nd * kE] : 0.f;
        s1 += gy[k];
        s2 = fmaf(gy[k], xh[k], s2);
    }
    s1 = group_sum(s1, s_red, g, e);
    s2 = group_sum(s2, s_red, g, e);
    const float m1 = s1 / (float)N, m2 = s2 / (float)N, wr = w[e] * rstd;
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int nd = g + kGroups * k;
        if (nd < N) dx[base + (int64_t)nd * kE] = wr * (gy[k] - m1 - xh[k] * m2);
    }
    if (g == 0) {
        atomicAdd(dw + e, s2);
        atomicAdd(db + e, s1);
    }
}

}  // namespace nrm
}  // namespace r4

using namespace r4;

extern "C" int32_t rl4co_instnorm_max_nodes(void) { return nrm::kGroups * nrm::kPer; }

extern "C" int rl4co_instnorm_fwd(const float* x, const float* w, const float* bias, float* y, float* mean, float* rstd,
                                  int64_t B, int32_t N, float eps, void* stream) {
    R4_CHECK_ARG(x && w && bias && y && mean && rstd, "rl4co_instnorm_fwd: null pointer");
    R4_CHECK_ARG(B >= 0 && B <= 2147483647LL && N >= 1 && N <= rl4co_instnorm_max_nodes(), "rl4co_instnorm_fwd: B=%lld N=%d (N <= %d)",
                 (long long)B, N, rl4co_instnorm_max_nodes());
    if (B == 0) return RL4CO_OK;
    nrm::instnorm_fwd_kernel<<<(unsigned)B, nrm::kThreads, 0, (cudaStream_t)stream>>>(x, w, bias, y, mean, rstd, N, eps);
    R4_CHECK_LAUNCH("rl4co_instnorm_fwd");
    return RL4CO_OK;
}

extern "C" int rl4co_instnorm_bwd(const float* x, const float* w, const float* mean, const float* rstd, const float* dy,
                                  float* dx, float* dw, float* db, int64_t B, int32_t N, void* stream) {
    R4_CHECK_ARG(x && w && mean && rstd && dy && dx && dw && db, "rl4co_instnorm_bwd: null pointer");
    R4_CHECK_ARG(B >= 0 && B <= 2147483647LL && N >= 1 && N <= rl4co_instnorm_max_nodes(), "rl4co_instnorm_bwd: B=%lld N=%d (N <= %d)",
                 (long long)B, N, rl4co_instnorm_max_nodes());
    if (B == 0) return RL4CO_OK;
    nrm::instnorm_bwd_kernel<<<(unsigned)B, nrm::kThreads, 0, (cudaStream_t)stream>>>(x, w, mean, rstd, dy, dx, dw, db, N);
    R4_CHECK_LAUNCH("rl4co_instnorm_bwd");
    return RL4CO_OK;
}
