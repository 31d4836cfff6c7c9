// Persistent fused rollout, fp32 CUDA-core path (v1):
// ConstructivePolicy.forward's main loop (constructive/base.py:221-263) for every row, in one
// launch and without host synchronisation.  One warp runs one row's whole rollout.
//   kSmem = true  (multistart, S rows share an instance): one CTA per instance; the instance's
//                 glimpse key / value / logit key tiles (3 x n x 512 B) are brought into shared
//                 memory once with TMA bulk copies (cp.async.bulk + mbarrier) and reused by all
//                 S rows for all T steps.
//   kSmem = false (single start): K/V/L stream from global memory/L2 as coalesced 512-byte rows.
#include "decode_core.cuh"

namespace r4 {

constexpr int kRolloutWarps = 20;  // 640 threads, <= 102 registers/thread
constexpr int kGlobalWarps = 8;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

struct RowOut {
    float ll;
    float len;
    int steps;
};

// One row's complete rollout by one warp.
template <int ENV>
__device__ __forceinline__ void run_row(const rl4co_cache& c, const rl4co_rollout_args& a, const InstPtrs& ip,
                                        int64_t b, int s, int lane) {
    const int n = c.n_nodes;
    const int64_t row = (int64_t)s * c.B + b;  // start-major (ops.py:13-16)
    const int T = a.max_steps;
    const bool multistart = a.num_starts > 1;
    const int ncust = (ENV == RL4CO_ENV_TSP) ? n : n - 1;
    const int smod = (a.start_mod > 0 && a.start_mod <= ncust) ? a.start_mod : ncust;   // ops.py:147-161
    const float2* L = reinterpret_cast<const float2*>(a.locs) + b * (int64_t)n;

    // ---- reset (tsp/env.py:90-115, cvrp/env.py:102-130) ----
    uint32_t maskw[kNB], visw[kNB];
    float dm[kNB];
    float used = 0.f, cap = 1.f;
    int cur = 0, first = 0;
    bool started = false;  // TSP: i > 0
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        int rem = n - blk * 32;
        maskw[blk] = rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u));
        visw[blk] = 0u;
        dm[blk] = 0.f;
    }
    if (ENV == RL4CO_ENV_CVRP) {
        cap = a.vehicle_capacity[b];
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) {
            const int node = blk * 32 + lane;
            if (node >= 1 && node < n) dm[blk] = a.demand[b * (int64_t)(n - 1) + node - 1];
        }
        cvrp_mask_bits(dm, visw, used, cap, cur, n, lane, maskw);
    }

    float ll = 0.f, len = 0.f;
    int t = 0;
    int prev = -1, first_action = 0;
    bool done = false;
    int err_bits = 0;

    while (!done) {
        if (t >= T) {
            err_bits |= RL4CO_ERR_MAX_STEPS;
            break;
        }
        int act;
        float lpa = 0.f;
        if (multistart && t == 0) {
            // forced first action (decoding.py:305-327; select_start_nodes ops.py:147-161)
            if (a.mode == RL4CO_MODE_EVALUATE)
                act = (int)a.actions_in[row * a.actions_in_stride];
            else
                act = (ENV == RL4CO_ENV_TSP) ? (s % smod) : (s % smod) + 1;
        } else {
            float4 q4 = row_query(ENV, c, b, first, cur, (ENV == RL4CO_ENV_TSP) && !started, cap - used, lane);
            const float4 hd = warp_glimpse(ip, q4, maskw, lane);
            float lg[kNB], gum[kNB], lp[kNB];
            warp_logits(ip, hd, maskw, lane, lg);
            int given = 0;
#pragma unroll
            for (int blk = 0; blk < kNB; ++blk) gum[blk] = 0.f;
            if (a.mode == RL4CO_MODE_SAMPLING) {
                const uint32_t key = noise_key(a.seed, a.inst_offset + b, s, t);
#pragma unroll
                for (int blk = 0; blk < kNB; ++blk) {
                    const int node = blk * 32 + lane;
                    if (node < n)
                        gum[blk] = a.gumbel ? a.gumbel[(row * T + t) * (int64_t)n + node] : noise_gumbel(key, node);
                }
            } else if (a.mode == RL4CO_MODE_EVALUATE) {
                given = (t < a.actions_in_T) ? (int)a.actions_in[row * a.actions_in_stride + t] : 0;
                if (given < 0 || given >= n) {
                    err_bits |= RL4CO_ERR_INFEASIBLE;
                    given = 0;
                }
            }
            bool nan_seen;
            Selection sel = warp_select(lg, maskw, a.tanh_clipping, a.temperature, a.mode, gum, given, lane, lp,
                                        nan_seen);
            if (nan_seen) err_bits |= RL4CO_ERR_NAN_LOGITS;
            act = sel.action;
            lpa = sel.logp;
            if (a.mode != RL4CO_MODE_EVALUATE && !bit_set(maskw, act)) err_bits |= RL4CO_ERR_INFEASIBLE;
            if (!(lpa > -1000.f)) err_bits |= RL4CO_ERR_NEG_INF_LOGP;  // decoding.py:56-58
            if (a.logp_all) {
#pragma unroll
                for (int blk = 0; blk < kNB; ++blk) {
                    const int node = blk * 32 + lane;
                    if (node < n) a.logp_all[(row * T + t) * (int64_t)n + node] = lp[blk];
                }
            }
        }
        if (lane == 0) {
            a.actions[row * T + t] = act;
            a.logp[row * T + t] = lpa;
        }
        ll += lpa;
        // tour length so far (ops.py:82-95): edge prev -> act; CVRP starts at the depot
        {
            const int from = (prev < 0) ? ((ENV == RL4CO_ENV_CVRP) ? 0 : act) : prev;
            const float2 x = L[from], y = L[act];
            const float dx = y.x - x.x, dy = y.y - x.y;
            len += sqrtf(dx * dx + dy * dy);
            if (prev < 0) first_action = act;
            prev = act;
        }
        // ---- env step ----
        if (ENV == RL4CO_ENV_TSP) {  // tsp/env.py:62-88
            if (!started) first = act;
            started = true;
            cur = act;
            bit_clear(maskw, act);
            bool any = false;
#pragma unroll
            for (int blk = 0; blk < kNB; ++blk) any |= (maskw[blk] != 0u);
            done = !any;
        } else {  // cvrp/env.py:68-100
            int di = act - 1;
            di = di < 0 ? 0 : (di > n - 2 ? n - 2 : di);
            float sel_d = 0.f;
#pragma unroll
            for (int blk = 0; blk < kNB; ++blk)
                if (((di + 1) >> 5) == blk) sel_d = dm[blk];
            sel_d = __shfl_sync(kFull, sel_d, (di + 1) & 31);
            used = (used + sel_d) * ((act != 0) ? 1.0f : 0.0f);
            bit_put(visw, act);
            cur = act;
            int cnt = 0;
#pragma unroll
            for (int blk = 0; blk < kNB; ++blk) cnt += __popc(visw[blk]);
            done = (cnt == n);
            cvrp_mask_bits(dm, visw, used, cap, cur, n, lane, maskw);
        }
        ++t;
    }
    // closing edge (roll(-1) wraps to the first element: TSP first action, CVRP depot)
    if (prev >= 0) {
        const int to = (ENV == RL4CO_ENV_CVRP) ? 0 : first_action;
        const float2 x = L[prev], y = L[to];
        const float dx = y.x - x.x, dy = y.y - x.y;
        len += sqrtf(dx * dx + dy * dy);
    }
    if (lane == 0) {
        a.reward[row] = -len;
        a.log_likelihood[row] = ll;
        a.lengths[row] = t;
        atomicMax(a.max_len, t);
        if (err_bits) raise_err(a.err, err_bits, row);
    }
}

template <int ENV>
__global__ void __launch_bounds__(kRolloutWarps * 32, 1)
    rollout_smem_kernel(rl4co_cache c, rl4co_rollout_args a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t bar;
    const int n = c.n_nodes;
    const uint32_t tile_bytes = (uint32_t)n * kE * sizeof(float);
    float* sk = reinterpret_cast<float*>(smem_raw);
    float* sv = sk + n * kE;
    float* sl = sv + n * kE;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t parity = 0;
    for (int64_t b = blockIdx.x; b < c.B; b += gridDim.x) {
        if (threadIdx.x == 0) {
            // order prior generic-proxy reads of the tiles before the async-proxy overwrite
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&bar, 3u * tile_bytes);
            const int64_t o = b * (int64_t)n * kE;
            bulk_g2s(sk, c.gk + o, tile_bytes, &bar);
            bulk_g2s(sv, c.gv + o, tile_bytes, &bar);
            bulk_g2s(sl, c.lk + o, tile_bytes, &bar);
        }
        mbar_wait(&bar, parity);
        parity ^= 1u;
        const InstPtrs ip{sk, sv, sl};
        for (int s = warp; s < a.num_starts; s += kRolloutWarps) run_row<ENV>(c, a, ip, b, s, lane);
        __syncthreads();  // all rows of this instance finished before the tiles are overwritten
    }
}

#ifndef R4_GLOBAL_MINBLOCKS
#define R4_GLOBAL_MINBLOCKS 4   // 32 warps per SM (64 registers): c2 fits one wave, c3 two equal ones; measured 1.39 / 10.1 ms against 1.50 / 11.6 ms at 3 CTAs per SM
#endif
template <int ENV>
__global__ void __launch_bounds__(kGlobalWarps * 32, R4_GLOBAL_MINBLOCKS)
    rollout_global_kernel(rl4co_cache c, rl4co_rollout_args a) {
    const int lane = threadIdx.x & 31;
    const int S = a.num_starts > 1 ? a.num_starts : 1;
    const int64_t rows = c.B * S;
    const int64_t nw = (int64_t)gridDim.x * kGlobalWarps;
    for (int64_t w = blockIdx.x * (int64_t)kGlobalWarps + (threadIdx.x >> 5); w < rows; w += nw) {
        // instance-major work order so that the S rows of an instance run close together (L2 reuse)
        const int64_t b = w / S;
        const int s = (int)(w % S);
        const int64_t o = b * (int64_t)c.n_nodes * kE;
        const InstPtrs ip{c.gk + o, c.gv + o, c.lk + o};
        run_row<ENV>(c, a, ip, b, s, lane);
    }
}

}  // namespace r4

using namespace r4;

namespace r4 {
int check_cache(const rl4co_cache* c, int env, const char* who);
bool tc_supported(const rl4co_cache* c, int S);
int launch_rollout_tc(const rl4co_cache* cache, const rl4co_rollout_args& a, cudaStream_t st);
bool coop_supported(const rl4co_cache* c);
int launch_rollout_coop(const rl4co_cache* cache, const rl4co_rollout_args& a, cudaStream_t st);
}

extern "C" int rl4co_rollout(const rl4co_cache* cache, const rl4co_rollout_args* args, void* stream) {
    R4_CHECK_ARG(args != nullptr, "rollout: null args");
    const int env = args->env;
    R4_CHECK_ARG(env == RL4CO_ENV_TSP || env == RL4CO_ENV_CVRP, "rollout: bad env %d", env);
    int rc = check_cache(cache, env, "rollout");
    if (rc) return rc;
    R4_CHECK_ARG(args->mode >= 0 && args->mode <= 2, "rollout: bad mode %d", args->mode);
    R4_CHECK_ARG(args->max_steps > 0 && args->temperature > 0.f, "rollout: bad max_steps / temperature");
    R4_CHECK_ARG(args->locs && args->actions && args->logp && args->reward && args->log_likelihood &&
                     args->lengths && args->max_len && args->err, "rollout: null output/input pointer");
    if (env == RL4CO_ENV_CVRP) R4_CHECK_ARG(args->demand && args->vehicle_capacity, "rollout: CVRP needs demand/capacity");
    if (args->mode == RL4CO_MODE_EVALUATE)
        R4_CHECK_ARG(args->actions_in && args->actions_in_T > 0 && args->actions_in_stride >= args->actions_in_T,
                     "rollout: evaluate needs actions_in");
    const int S = args->num_starts > 1 ? args->num_starts : 1;
    const int max_starts = (env == RL4CO_ENV_TSP) ? cache->n_nodes : cache->n_nodes - 1;
    R4_CHECK_ARG(args->mode == RL4CO_MODE_EVALUATE || S <= max_starts || S == 1,
                 "rollout: num_starts=%d > %d start nodes", S, max_starts);
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    rl4co_rollout_args a = *args;
    a.num_starts = S;
    R4_CHECK_ARG(a.impl >= RL4CO_IMPL_AUTO && a.impl <= RL4CO_IMPL_COOP, "rollout: bad impl %d", a.impl);
    if (a.impl == RL4CO_IMPL_TC)
        R4_CHECK_ARG(tc_supported(cache, S), "rollout: tensor path needs tc_img, 8 <= S=%d <= 128, n_nodes=%d <= 112",
                     S, cache->n_nodes);
    if (a.impl != RL4CO_IMPL_FP32 && a.impl != RL4CO_IMPL_STREAM && tc_supported(cache, S))
        return launch_rollout_tc(cache, a, st);
    // Single-start rows.  A batch that gives the streaming kernel at least ~8 warps per SM is served best by it (one
    // warp per row, 64 rows per SM in flight hide every latency; measured c2 1.9 ms / c3 13.5 ms against 3.5 / 29 ms for
    // the cooperative kernel).  Small batches (c1: 512 rows on 148 SMs) are latency-bound there: the cooperative kernel
    // (CTA per instance, tables in shared memory, the warps of the CTA share one row's step) halves their time.
    if (S == 1 && a.impl != RL4CO_IMPL_STREAM && coop_supported(cache) &&
        (a.impl == RL4CO_IMPL_COOP || cache->B < (int64_t)sms * 8))
        return launch_rollout_coop(cache, a, st);
    const size_t smem = 3ull * cache->n_nodes * kE * sizeof(float);
    const bool use_smem = S >= 8;
    if (use_smem) {
        auto kern = (env == RL4CO_ENV_TSP) ? rollout_smem_kernel<RL4CO_ENV_TSP> : rollout_smem_kernel<RL4CO_ENV_CVRP>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("rollout: cudaFuncSetAttribute(%zu B smem): %s", smem, cudaGetErrorString(e));
            return RL4CO_E_LAUNCH;
        }
        unsigned grid = (unsigned)(cache->B < sms ? cache->B : sms);
        kern<<<grid, kRolloutWarps * 32, smem, st>>>(*cache, a);
    } else {
        auto kern = (env == RL4CO_ENV_TSP) ? rollout_global_kernel<RL4CO_ENV_TSP> : rollout_global_kernel<RL4CO_ENV_CVRP>;
        const int64_t rows = cache->B * S;
        int64_t blocks = (rows + kGlobalWarps - 1) / kGlobalWarps;
        // A row is one long unit of work (its whole rollout) and the kernel is bound by HBM latency x bytes in flight:
        // a last, nearly empty wave of CTAs runs at a fraction of the bandwidth.  The grid is sized to the resident
        // capacity and the rows are split into EQUAL waves (each warp walks `waves` rows).
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kGlobalWarps * 32, 0) != cudaSuccess || occ < 1)
            occ = 2;
        const int64_t cap = (int64_t)sms * occ;
        if (blocks > cap) {
            const int64_t waves = (blocks + cap - 1) / cap;
            blocks = (blocks + waves - 1) / waves;
        }
        kern<<<(unsigned)blocks, kGlobalWarps * 32, 0, st>>>(*cache, a);
    }
    R4_CHECK_LAUNCH("rollout");
    return RL4CO_OK;
}
