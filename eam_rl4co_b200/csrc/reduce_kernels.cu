// Best-of-starts / best-of-augmentations reduction that follows the rollout at evaluation time
// (pomo/model.py:115-145).  Rows are start-major over the augmented batch:
// row(b, a, s) = s * (A * B) + a * B + b  (ops.py:13-56 batchify order).  One CTA per instance reads its A * S rewards
// once, takes the first maximum over starts per augmentation and then over augmentations (torch.max tie rule), and
// copies only the winning tours - the [B, A, S, T] regrouping of the actions is never materialised.
#include "common.cuh"

namespace r4 {

constexpr int kMaxAug = 32;

__global__ void __launch_bounds__(128) best_of_starts_kernel(const float* __restrict__ reward,
                                                             const int64_t* __restrict__ actions, int64_t B, int A,
                                                             int S, int T, float* __restrict__ max_reward,
                                                             int64_t* __restrict__ best_start,
                                                             int64_t* __restrict__ best_ms_actions,
                                                             float* __restrict__ max_aug_reward,
                                                             int64_t* __restrict__ best_aug,
                                                             int64_t* __restrict__ best_aug_actions) {
    __shared__ float s_val[kMaxAug];
    __shared__ int s_idx[kMaxAug];
    __shared__ int s_best_a;
    const int64_t b = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t AB = (int64_t)A * B;
    for (int a = warp; a < A; a += 4) {
        float v = -INFINITY;
        int idx = 0x7fffffff;
        for (int s = lane; s < S; s += 32) {
            const float r = reward[(int64_t)s * AB + (int64_t)a * B + b];
            if (r > v || idx == 0x7fffffff) {   // ascending s + strict > : first maximum; the first element always enters
                v = r;
                idx = s;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(kFull, v, o);
            const int oi = __shfl_xor_sync(kFull, idx, o);
            if (oi != 0x7fffffff && (idx == 0x7fffffff || ov > v || (ov == v && oi < idx))) {
                v = ov;
                idx = oi;
            }
        }
        if (lane == 0) {
            s_val[a] = v;
            s_idx[a] = idx;
            max_reward[b * A + a] = v;
            best_start[b * A + a] = idx;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int ba = 0;
        for (int a = 1; a < A; ++a)
            if (s_val[a] > s_val[ba]) ba = a;
        s_best_a = ba;
        if (max_aug_reward != nullptr) max_aug_reward[b] = s_val[ba];
        if (best_aug != nullptr) best_aug[b] = ba;
    }
    __syncthreads();
    if (actions == nullptr) return;
    for (int a = 0; a < A; ++a) {
        const int64_t row = (int64_t)s_idx[a] * AB + (int64_t)a * B + b;
        for (int t = threadIdx.x; t < T; t += blockDim.x) {
            const int64_t v = actions[row * T + t];
            best_ms_actions[(b * A + a) * T + t] = v;
            if (a == s_best_a && best_aug_actions != nullptr) best_aug_actions[b * T + t] = v;
        }
    }
}

}  // namespace r4

extern "C" int rl4co_best_of_starts(const float* reward, const int64_t* actions, int64_t B, int32_t n_aug,
                                    int32_t n_start, int32_t T, float* max_reward, int64_t* best_start,
                                    int64_t* best_multistart_actions, float* max_aug_reward, int64_t* best_aug,
                                    int64_t* best_aug_actions, void* stream) {
    R4_CHECK_ARG(reward && max_reward && best_start && B >= 0, "best_of_starts: null pointer");
    R4_CHECK_ARG(n_aug >= 1 && n_aug <= r4::kMaxAug && n_start >= 1 && T >= 0,
                 "best_of_starts: 1 <= n_aug <= 32, n_start >= 1 (got %d, %d)", n_aug, n_start);
    R4_CHECK_ARG(actions == nullptr || best_multistart_actions != nullptr, "best_of_starts: actions need an output");
    if (B == 0) return RL4CO_OK;
    r4::best_of_starts_kernel<<<(unsigned)B, 128, 0, (cudaStream_t)stream>>>(
        reward, actions, B, n_aug, n_start, T, max_reward, best_start, best_multistart_actions, max_aug_reward, best_aug,
        best_aug_actions);
    R4_CHECK_LAUNCH("best_of_starts");
    return RL4CO_OK;
}
