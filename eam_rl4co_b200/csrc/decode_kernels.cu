// Step-granular decoder kernels: rl4co_select, rl4co_decode_step, rl4co_noise_gumbel.
#include "decode_core.cuh"

namespace r4 {

constexpr int kStepWarps = 8;

// ---- process_logits + strategy step on given logits (decoding.py:140-190, 346-465) ------
__global__ void select_kernel(const float* __restrict__ logits, const uint8_t* __restrict__ mask,
                              float tanh_clip, float temperature, int mode,
                              const float* __restrict__ gumbel, uint32_t seed, int64_t inst_offset,
                              int64_t B, int t, const int64_t* __restrict__ action_in,
                              float* __restrict__ logp_all, int64_t* __restrict__ action_out,
                              float* __restrict__ logp_sel, int32_t* err, int64_t rows, int n) {
    const int lane = threadIdx.x & 31;
    const int64_t row = blockIdx.x * (int64_t)kStepWarps + (threadIdx.x >> 5);
    if (row >= rows) return;
    uint32_t maskw[kNB];
    load_mask_bits(mask + row * n, n, lane, maskw);
    float lg[kNB], gum[kNB], lp[kNB];
    const uint32_t key = noise_key(seed, inst_offset + row % B, row / B, t);
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        const int node = blk * 32 + lane;
        lg[blk] = (node < n) ? logits[row * n + node] : -INFINITY;
        gum[blk] = 0.f;
        if (mode == RL4CO_MODE_SAMPLING && node < n)
            gum[blk] = gumbel ? gumbel[row * n + node] : noise_gumbel(key, node);
    }
    int given = 0;
    if (mode == RL4CO_MODE_EVALUATE) given = (int)action_in[row];
    bool nan_seen;
    Selection sel = warp_select(lg, maskw, tanh_clip, temperature, mode, gum, given, lane, lp, nan_seen);
    if (logp_all) {
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) {
            const int node = blk * 32 + lane;
            if (node < n) logp_all[row * n + node] = lp[blk];
        }
    }
    if (lane == 0) {
        action_out[row] = sel.action;
        logp_sel[row] = sel.logp;
        if (mode != RL4CO_MODE_EVALUATE && !bit_set(maskw, sel.action))
            raise_err(err, RL4CO_ERR_INFEASIBLE, row);
    }
}

__global__ void noise_kernel(uint32_t seed, int64_t inst_offset, int64_t B, int t,
                             float* __restrict__ out, int64_t rows, int n) {
    int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (idx >= rows * n) return;
    int64_t row = idx / n;
    int node = (int)(idx % n);
    out[idx] = noise_gumbel(noise_key(seed, inst_offset + row % B, row / B, t), node);
}

// ---- AttentionModelDecoder.forward for all rows (am/decoder.py:161-198) -----------------
// Returns RAW logits for every node (the reference masks later, in process_logits).
__global__ void decode_step_kernel(int env, rl4co_cache c, const int64_t* __restrict__ first_node,
                                   const int64_t* __restrict__ current_node,
                                   const int64_t* __restrict__ i_step,
                                   const float* __restrict__ used, const float* __restrict__ cap,
                                   const uint8_t* __restrict__ mask, float* __restrict__ logits_out,
                                   int32_t* err, int64_t rows) {
    const int lane = threadIdx.x & 31;
    const int64_t row = blockIdx.x * (int64_t)kStepWarps + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int n = c.n_nodes;
    const int64_t b = row % c.B;
    uint32_t maskw[kNB], allw[kNB];
    load_mask_bits(mask + row * n, n, lane, maskw);
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        int rem = n - blk * 32;
        allw[blk] = rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u));
    }
    float4 q4;
    if (env == RL4CO_ENV_TSP) {
        // context.py:124 tests only i of the FIRST batch element; through the public API all
        // rows share the same i, so i_step[0] is read like the reference does.
        q4 = row_query(env, c, b, (int)first_node[row], (int)current_node[row], i_step[0] < 1, 0.f, lane);
    } else {
        q4 = row_query(env, c, b, 0, (int)current_node[row], false, cap[row] - used[row], lane);
    }
    InstPtrs ip{c.gk + b * (int64_t)n * kE, c.gv + b * (int64_t)n * kE, c.lk + b * (int64_t)n * kE};
    const float4 hd = warp_glimpse(ip, q4, maskw, lane);
    float lg[kNB];
    warp_logits(ip, hd, allw, lane, lg);
    bool nan_local = false;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        const int node = blk * 32 + lane;
        if (node < n) {
            logits_out[row * n + node] = lg[blk];
            nan_local |= (lg[blk] != lg[blk]);
        }
    }
    if (__any_sync(kFull, nan_local) && lane == 0) raise_err(err, RL4CO_ERR_NAN_LOGITS, row);  // attention.py:303
}

}  // namespace r4

using namespace r4;

namespace r4 {
int check_cache(const rl4co_cache* c, int env, const char* who) {
    R4_CHECK_ARG(c != nullptr, "%s: null cache", who);
    R4_CHECK_ARG(c->embed_dim == kE && c->num_heads == kH, "%s: only embed_dim=128, num_heads=8 (got %d, %d)",
                 who, c->embed_dim, c->num_heads);
    R4_CHECK_ARG(c->n_nodes >= 2 && c->n_nodes <= kMaxNodes, "%s: n_nodes=%d outside [2,%d]", who,
                 c->n_nodes, kMaxNodes);
    R4_CHECK_ARG(c->B > 0 && c->gk && c->gv && c->lk && c->pb, "%s: cache pointers missing", who);
    if (env == RL4CO_ENV_TSP) R4_CHECK_ARG(c->pa && c->q0, "%s: TSP cache needs pa and q0", who);
    if (env == RL4CO_ENV_CVRP) R4_CHECK_ARG(c->wcap, "%s: CVRP cache needs wcap", who);
    return RL4CO_OK;
}
}  // namespace r4

extern "C" int rl4co_select(const float* logits, const uint8_t* mask, float tanh_clipping,
                            float temperature, int32_t mode, const float* gumbel, uint32_t seed,
                            int64_t inst_offset, int64_t B, int32_t t, const int64_t* action_in,
                            float* logp_all, int64_t* action_out, float* logp_sel, int32_t* err,
                            int64_t rows, int32_t n_nodes, void* stream) {
    R4_CHECK_ARG(rows >= 0 && n_nodes >= 1 && n_nodes <= kMaxNodes, "select: n_nodes=%d outside [1,%d]",
                 n_nodes, kMaxNodes);
    R4_CHECK_ARG(mode >= 0 && mode <= 2, "select: bad mode %d", mode);
    if (rows == 0) return RL4CO_OK;
    R4_CHECK_ARG(logits && mask && action_out && logp_sel, "select: null pointer");
    R4_CHECK_ARG(mode != RL4CO_MODE_EVALUATE || action_in, "select: evaluate needs action_in");
    R4_CHECK_ARG(B > 0 && temperature > 0.f, "select: bad B / temperature");
    unsigned grid = (unsigned)((rows + kStepWarps - 1) / kStepWarps);
    select_kernel<<<grid, kStepWarps * 32, 0, (cudaStream_t)stream>>>(
        logits, mask, tanh_clipping, temperature, mode, gumbel, seed, inst_offset, B, t, action_in,
        logp_all, action_out, logp_sel, err, rows, n_nodes);
    R4_CHECK_LAUNCH("select");
    return RL4CO_OK;
}

extern "C" int rl4co_noise_gumbel(uint32_t seed, int64_t inst_offset, int64_t B, int32_t t, float* out,
                                  int64_t rows, int32_t n_nodes, void* stream) {
    R4_CHECK_ARG(rows >= 0 && n_nodes >= 1 && B > 0 && out, "noise_gumbel: bad args");
    if (rows == 0) return RL4CO_OK;
    int64_t total = rows * n_nodes;
    noise_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, inst_offset, B, t,
                                                                                   out, rows, n_nodes);
    R4_CHECK_LAUNCH("noise_gumbel");
    return RL4CO_OK;
}

extern "C" int rl4co_decode_step(int32_t env, const rl4co_cache* cache, const int64_t* first_node,
                                 const int64_t* current_node, const int64_t* i_step,
                                 const float* used_capacity, const float* vehicle_capacity,
                                 const uint8_t* mask, float* logits_out, int32_t* err, int64_t rows,
                                 void* stream) {
    R4_CHECK_ARG(env == RL4CO_ENV_TSP || env == RL4CO_ENV_CVRP, "decode_step: bad env %d", env);
    int rc = check_cache(cache, env, "decode_step");
    if (rc) return rc;
    if (rows == 0) return RL4CO_OK;
    R4_CHECK_ARG(rows > 0 && current_node && mask && logits_out, "decode_step: null pointer");
    if (env == RL4CO_ENV_TSP) R4_CHECK_ARG(first_node && i_step, "decode_step: TSP needs first_node, i");
    if (env == RL4CO_ENV_CVRP)
        R4_CHECK_ARG(used_capacity && vehicle_capacity, "decode_step: CVRP needs used/vehicle capacity");
    unsigned grid = (unsigned)((rows + kStepWarps - 1) / kStepWarps);
    decode_step_kernel<<<grid, kStepWarps * 32, 0, (cudaStream_t)stream>>>(
        env, *cache, first_node, current_node, i_step, used_capacity, vehicle_capacity, mask, logits_out,
        err, rows);
    R4_CHECK_LAUNCH("decode_step");
    return RL4CO_OK;
}
