// Environment transition / mask / reward kernels (step-granular API).
// One warp per row; lanes own nodes lane, lane+32, ...; masks are built with
// __ballot_sync / __popc and materialised as the reference's bool tensors only here.
#include "common.cuh"

namespace r4 {

constexpr int kWarpsPerBlock = 8;

// ---- any(i == 0) over the batch: tsp/env.py:65  `td["i"].all() == 0` ------------------
__global__ void any_zero_kernel(const int64_t* __restrict__ i_in, int64_t rows, int32_t* flag) {
    int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    bool z = (r < rows) && (i_in[r] == 0);
    if (__any_sync(kFull, z) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

// ---- TSPEnv._step: tsp/env.py:62-88 --------------------------------------------------
__global__ void tsp_step_kernel(const int64_t* __restrict__ action, const uint8_t* __restrict__ mask_in,
                                const int64_t* __restrict__ i_in, const int64_t* __restrict__ first_in,
                                uint8_t* __restrict__ mask_out, int64_t* __restrict__ cur_out,
                                int64_t* __restrict__ first_out, int64_t* __restrict__ i_out,
                                uint8_t* __restrict__ done_out, uint8_t* __restrict__ reward_out,
                                const int32_t* __restrict__ any_zero, int64_t rows, int32_t n) {
    const int lane = threadIdx.x & 31;
    const int64_t row = blockIdx.x * (int64_t)kWarpsPerBlock + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int64_t a = action[row];
    const uint8_t* mi = mask_in + row * n;
    uint8_t* mo = mask_out + row * n;
    int remaining = 0;
    for (int base = 0; base < n; base += 32) {
        int node = base + lane;
        bool avail = false;
        if (node < n) {
            avail = (mi[node] != 0) && (node != a);  // scatter(-1, action, 0)  :68-70
            mo[node] = avail ? 1 : 0;
        }
        remaining += __popc(__ballot_sync(kFull, avail));
    }
    if (lane == 0) {
        cur_out[row] = a;
        first_out[row] = (*any_zero) ? a : first_in[row];  // :65
        i_out[row] = i_in[row] + 1;
        done_out[row] = (remaining == 0);                  // :73
        reward_out[row] = 0;                               // :76 zeros_like(done) -> bool
    }
}

// ---- CVRP mask for one row, lane-parallel: cvrp/env.py:132-144 ------------------------
// visited/mask pointers are the row's [n] arrays; returns nothing, writes mask_out.
__device__ __forceinline__ void cvrp_mask_row(const float* __restrict__ demand_row, float used, float cap,
                                              const uint8_t* visited_row, int64_t cur,
                                              uint8_t* __restrict__ mask_row, int n, int lane) {
    const float lim = cap + 1e-5f;  // fp32 add, as `vehicle_capacity + 1e-5` on an fp32 tensor (:135)
    int n_free = 0;
    for (int base = 0; base < n; base += 32) {
        int node = base + lane;
        bool masked = true;
        if (node >= 1 && node < n) {
            bool exceeds = (demand_row[node - 1] + used) > lim;  // :135
            masked = (visited_row[node] != 0) || exceeds;        // :138
            mask_row[node] = masked ? 0 : 1;                     // ~mask_loc   :144
        }
        n_free += __popc(__ballot_sync(kFull, !masked));
    }
    if (lane == 0) {
        bool mask_depot = (cur == 0) && (n_free > 0);            // :141-143
        mask_row[0] = mask_depot ? 0 : 1;
    }
}

// ---- CVRPEnv._step + get_action_mask: cvrp/env.py:68-100 -----------------------------
__global__ void cvrp_step_kernel(const int64_t* __restrict__ action, const float* __restrict__ demand,
                                 int64_t demand_rows, const float* __restrict__ used_in,
                                 const float* __restrict__ cap, const uint8_t* __restrict__ visited_in,
                                 float* __restrict__ used_out, uint8_t* visited_out,
                                 int64_t* __restrict__ cur_out, uint8_t* __restrict__ done_out,
                                 uint8_t* __restrict__ reward_out, uint8_t* __restrict__ mask_out,
                                 int64_t rows, int32_t n) {
    const int lane = threadIdx.x & 31;
    const int64_t row = blockIdx.x * (int64_t)kWarpsPerBlock + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int64_t a = action[row];
    const float* drow = demand + (row % demand_rows) * (int64_t)(n - 1);
    int64_t di = a - 1;                          // clamp(current_node - 1, 0, n_loc - 1)  :73-75
    di = di < 0 ? 0 : (di > n - 2 ? n - 2 : di);
    const float sel = drow[di];
    const float used = (used_in[row] + sel) * ((a != 0) ? 1.0f : 0.0f);  // :78-80 (same two fp32 ops)
    const uint8_t* vi = visited_in + row * n;
    uint8_t* vo = visited_out + row * n;
    int count = 0;
    for (int base = 0; base < n; base += 32) {
        int node = base + lane;
        bool v = false;
        if (node < n) {
            v = (vi[node] != 0) || (node == a);  // scatter(-1, current_node, 1)  :84
            vo[node] = (node == a) ? 1 : vi[node];
        }
        count += __popc(__ballot_sync(kFull, v));
    }
    __syncwarp();
    if (lane == 0) {
        used_out[row] = used;
        cur_out[row] = a;
        // visited.sum(-1) == visited.size(-1): uint8 sum; with 0/1 entries equals the popcount (:87)
        done_out[row] = (count == n);
        reward_out[row] = 0;
    }
    cvrp_mask_row(drow, used, cap[row], vo, a, mask_out + row * n, n, lane);
}

__global__ void cvrp_mask_kernel(const float* __restrict__ demand, int64_t demand_rows,
                                 const float* __restrict__ used, const float* __restrict__ cap,
                                 const uint8_t* __restrict__ visited, const int64_t* __restrict__ cur,
                                 uint8_t* __restrict__ mask_out, int64_t rows, int32_t n) {
    const int lane = threadIdx.x & 31;
    const int64_t row = blockIdx.x * (int64_t)kWarpsPerBlock + (threadIdx.x >> 5);
    if (row >= rows) return;
    cvrp_mask_row(demand + (row % demand_rows) * (int64_t)(n - 1), used[row], cap[row],
                  visited + row * n, cur[row], mask_out + row * n, n, lane);
}

// ---- teacher-forced state replay (the training caller) --------------------------------
// Replays actions [rows, T] (start-major rows, row = s*B + b) through _step / get_action_mask
// (tsp/env.py:62-88, cvrp/env.py:68-100,132-144) and writes what the decoder sees BEFORE step t = 1..T-1, directly
// in instance-major order [B, S, T-1, ...]: current node, feasibility mask, remaining capacity (CVRP).  One warp
// per row, lanes own nodes (visited flags in a register bit per 32-node block), one launch instead of T env steps.
__global__ void replay_states_kernel(int env, const int64_t* __restrict__ actions, int64_t rows, int32_t T,
                                     int64_t B, int32_t n, const float* __restrict__ demand,
                                     const float* __restrict__ cap, int64_t* __restrict__ cur_out,
                                     uint8_t* __restrict__ mask_out, float* __restrict__ rem_out) {
    const int lane = threadIdx.x & 31;
    const int64_t row = blockIdx.x * (int64_t)kWarpsPerBlock + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int64_t b = row % B, sidx = row / B, S = rows / B;
    const int64_t orow = (b * S + sidx) * (int64_t)(T - 1);     // first [b, s, 0] element
    const int64_t* arow = actions + row * (int64_t)T;
    const float* drow = (env == RL4CO_ENV_CVRP) ? demand + b * (int64_t)(n - 1) : nullptr;
    const float capv = (env == RL4CO_ENV_CVRP) ? cap[b] : 1.f;
    const float lim = capv + 1e-5f;                             // cvrp/env.py:135, fp32 add
    uint32_t vis = 0u;                                          // bit k: node 32k + lane visited
    float used = 0.f;
    float dm[4] = {0.f, 0.f, 0.f, 0.f};                         // demand of this lane's nodes
    if (env == RL4CO_ENV_CVRP) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int node = 32 * k + lane;
            if (node >= 1 && node < n) dm[k] = drow[node - 1];
        }
    }
    for (int t = 1; t < T; ++t) {
        const int64_t a64 = arow[t - 1];
        const int a = (int)a64;
        if ((a & 31) == lane && a >= 0 && a < n) vis |= 1u << (a >> 5);     // scatter(-1, action, .)
        if (env == RL4CO_ENV_CVRP) {
            int di = a - 1;                                     // clamp(current_node - 1, 0, n_loc - 1)  :73-75
            di = di < 0 ? 0 : (di > n - 2 ? n - 2 : di);
            used = (used + drow[di]) * ((a != 0) ? 1.0f : 0.0f);            // :78-80 (same two fp32 ops)
        }
        uint8_t* mrow = mask_out + (orow + (t - 1)) * (int64_t)n;
        int n_free = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int node = 32 * k + lane;
            bool feas = false;
            if (node < n) {
                if (env == RL4CO_ENV_TSP) {
                    feas = !((vis >> k) & 1u);                  // tsp/env.py:68-70
                    mrow[node] = feas ? 1 : 0;
                } else if (node >= 1) {
                    const bool exceeds = (dm[k] + used) > lim;  // cvrp/env.py:135
                    feas = !(((vis >> k) & 1u) || exceeds);     // :138
                    mrow[node] = feas ? 1 : 0;
                }
            }
            if (32 * k < n) n_free += __popc(__ballot_sync(kFull, feas && node >= 1));
        }
        if (lane == 0) {
            if (env == RL4CO_ENV_CVRP) {
                mrow[0] = ((a == 0) && (n_free > 0)) ? 0 : 1;   // :141-144
                rem_out[orow + (t - 1)] = capv - used;
            }
            cur_out[orow + (t - 1)] = a64;
        }
    }
}

// ---- _get_reward + check_solution_validity --------------------------------------------
// tsp/env.py:152-168, cvrp/env.py:146-185, ops.py:82-95.  One warp per row; lanes own
// tour positions; edge lengths summed with a warp tree.
__global__ void tour_reward_kernel(int env, const float* __restrict__ locs, int64_t loc_rows,
                                   const int64_t* __restrict__ actions, int64_t astride, int T,
                                   const float* __restrict__ demand, int64_t demand_rows,
                                   const float* __restrict__ cap, float* __restrict__ reward,
                                   int check, int32_t* err, int64_t rows, int n) {
    const int lane = threadIdx.x & 31;
    const int64_t row = blockIdx.x * (int64_t)kWarpsPerBlock + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float2* L = reinterpret_cast<const float2*>(locs) + (row % loc_rows) * (int64_t)n;
    const int64_t* A = actions + row * astride;
    // ordered tour: TSP a_0..a_{T-1}; CVRP depot, a_0..a_{T-1}   (cvrp/env.py:148-154)
    const int len = (env == RL4CO_ENV_CVRP) ? T + 1 : T;
    float acc = 0.f;
    bool bad = false;
    for (int p = lane; p < len; p += 32) {
        int q = (p + 1 == len) ? 0 : p + 1;  // roll(-1)   ops.py:94
        int64_t a0, a1;
        if (env == RL4CO_ENV_CVRP) {
            a0 = (p == 0) ? 0 : A[p - 1];
            a1 = (q == 0) ? 0 : A[q - 1];
        } else {
            a0 = A[p];
            a1 = A[q];
        }
        if (a0 < 0 || a0 >= n || a1 < 0 || a1 >= n) {
            bad = true;
            continue;
        }
        float2 x = L[a0], y = L[a1];
        float dx = y.x - x.x, dy = y.y - x.y;
        acc += sqrtf(dx * dx + dy * dy);     // norm(p=2)   ops.py:84
    }
    acc = warp_sum(acc);
    if (lane == 0) reward[row] = -acc;
    if (!check) return;
    // validity: every customer exactly once (TSP: every node exactly once, T == n)
    int first_node = (env == RL4CO_ENV_CVRP) ? 1 : 0;
    if (env == RL4CO_ENV_TSP && T != n) bad = true;
    for (int base = 0; base < n; base += 32) {
        int node = base + lane;
        int cnt = 0;
        if (node < n)
            for (int t = 0; t < T; ++t) cnt += (A[t] == node);
        if (node >= first_node && node < n && cnt != 1) bad = true;
    }
    if (__any_sync(kFull, bad)) {
        if (lane == 0) raise_err(err, RL4CO_ERR_INVALID_TOUR, row);
        return;
    }
    if (env == RL4CO_ENV_CVRP && lane == 0) {
        // capacity replay, cvrp/env.py:172-185
        const float* drow = demand + (row % demand_rows) * (int64_t)(n - 1);
        const float c = cap[row % demand_rows];
        float used = 0.f;
        for (int t = 0; t < T; ++t) {
            int64_t a = A[t];
            used += (a == 0) ? -c : drow[a - 1];
            if (used < 0.f) used = 0.f;
            if (!(used <= c + 1e-5f)) {
                raise_err(err, RL4CO_ERR_CAPACITY, row);
                break;
            }
        }
    }
}

}  // namespace r4

using namespace r4;

extern "C" int rl4co_tsp_step(const int64_t* action, const uint8_t* mask_in, const int64_t* i_in,
                              const int64_t* first_in, uint8_t* mask_out, int64_t* cur_out,
                              int64_t* first_out, int64_t* i_out, uint8_t* done_out,
                              uint8_t* reward_out, int32_t* scratch_flag, int64_t rows,
                              int32_t n_nodes, void* stream) {
    R4_CHECK_ARG(rows >= 0 && n_nodes > 0, "tsp_step: bad sizes rows=%lld n=%d", (long long)rows, n_nodes);
    if (rows == 0) return RL4CO_OK;
    R4_CHECK_ARG(action && mask_in && i_in && first_in && mask_out && cur_out && first_out && i_out &&
                     done_out && reward_out && scratch_flag, "tsp_step: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    cudaMemsetAsync(scratch_flag, 0, sizeof(int32_t), st);
    any_zero_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(i_in, rows, scratch_flag);
    unsigned grid = (unsigned)((rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
    tsp_step_kernel<<<grid, kWarpsPerBlock * 32, 0, st>>>(action, mask_in, i_in, first_in, mask_out,
                                                          cur_out, first_out, i_out, done_out,
                                                          reward_out, scratch_flag, rows, n_nodes);
    R4_CHECK_LAUNCH("tsp_step");
    return RL4CO_OK;
}

extern "C" int rl4co_cvrp_step_mask(const int64_t* action, const float* demand, int64_t demand_rows,
                                    const float* used_in, const float* vehicle_capacity,
                                    const uint8_t* visited_in, float* used_out, uint8_t* visited_out,
                                    int64_t* cur_out, uint8_t* done_out, uint8_t* reward_out,
                                    uint8_t* mask_out, int64_t rows, int32_t n_nodes, void* stream) {
    R4_CHECK_ARG(rows >= 0 && n_nodes > 1 && demand_rows > 0, "cvrp_step: bad sizes");
    if (rows == 0) return RL4CO_OK;
    R4_CHECK_ARG(action && demand && used_in && vehicle_capacity && visited_in && used_out &&
                     visited_out && cur_out && done_out && reward_out && mask_out, "cvrp_step: null pointer");
    unsigned grid = (unsigned)((rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
    cvrp_step_kernel<<<grid, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        action, demand, demand_rows, used_in, vehicle_capacity, visited_in, used_out, visited_out,
        cur_out, done_out, reward_out, mask_out, rows, n_nodes);
    R4_CHECK_LAUNCH("cvrp_step");
    return RL4CO_OK;
}

extern "C" int rl4co_cvrp_action_mask(const float* demand, int64_t demand_rows, const float* used,
                                      const float* vehicle_capacity, const uint8_t* visited,
                                      const int64_t* current_node, uint8_t* mask_out, int64_t rows,
                                      int32_t n_nodes, void* stream) {
    R4_CHECK_ARG(rows >= 0 && n_nodes > 1 && demand_rows > 0, "cvrp_action_mask: bad sizes");
    if (rows == 0) return RL4CO_OK;
    R4_CHECK_ARG(demand && used && vehicle_capacity && visited && current_node && mask_out,
                 "cvrp_action_mask: null pointer");
    unsigned grid = (unsigned)((rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
    cvrp_mask_kernel<<<grid, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        demand, demand_rows, used, vehicle_capacity, visited, current_node, mask_out, rows, n_nodes);
    R4_CHECK_LAUNCH("cvrp_action_mask");
    return RL4CO_OK;
}

extern "C" int rl4co_tour_reward(int32_t env, const float* locs, int64_t loc_rows,
                                 const int64_t* actions, int64_t actions_row_stride, int32_t T,
                                 const float* demand, int64_t demand_rows,
                                 const float* vehicle_capacity, float* reward, int32_t check,
                                 int32_t* err, int64_t rows, int32_t n_nodes, void* stream) {
    R4_CHECK_ARG(env == RL4CO_ENV_TSP || env == RL4CO_ENV_CVRP, "tour_reward: bad env %d", env);
    R4_CHECK_ARG(rows >= 0 && n_nodes > 0 && T > 0 && loc_rows > 0, "tour_reward: bad sizes");
    if (rows == 0) return RL4CO_OK;
    R4_CHECK_ARG(locs && actions && reward, "tour_reward: null pointer");
    if (check && env == RL4CO_ENV_CVRP)
        R4_CHECK_ARG(demand && vehicle_capacity && demand_rows > 0 && err,
                     "tour_reward: check on CVRP needs demand, capacity and err");
    if (check) R4_CHECK_ARG(err, "tour_reward: check needs err");
    unsigned grid = (unsigned)((rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
    tour_reward_kernel<<<grid, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        env, locs, loc_rows, actions, actions_row_stride, T, demand, demand_rows, vehicle_capacity,
        reward, check, err, rows, n_nodes);
    R4_CHECK_LAUNCH("tour_reward");
    return RL4CO_OK;
}

extern "C" int rl4co_replay_states(int32_t env, const int64_t* actions, int64_t rows, int32_t T, int64_t B,
                                   int32_t n_nodes, const float* demand, const float* vehicle_capacity,
                                   int64_t* cur_out, uint8_t* mask_out, float* rem_out, void* stream) {
    R4_CHECK_ARG(env == RL4CO_ENV_TSP || env == RL4CO_ENV_CVRP, "replay_states: bad env %d", env);
    R4_CHECK_ARG(rows >= 0 && B > 0 && rows % B == 0 && T >= 1 && n_nodes > 1 && n_nodes <= 128,
                 "replay_states: bad sizes (n_nodes <= 128)");
    if (rows == 0 || T == 1) return RL4CO_OK;
    R4_CHECK_ARG(actions && cur_out && mask_out, "replay_states: null pointer");
    if (env == RL4CO_ENV_CVRP) R4_CHECK_ARG(demand && vehicle_capacity && rem_out, "replay_states: CVRP needs demand, capacity, rem_out");
    unsigned grid = (unsigned)((rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
    replay_states_kernel<<<grid, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        env, actions, rows, T, B, n_nodes, demand, vehicle_capacity, cur_out, mask_out, rem_out);
    R4_CHECK_LAUNCH("replay_states");
    return RL4CO_OK;
}
