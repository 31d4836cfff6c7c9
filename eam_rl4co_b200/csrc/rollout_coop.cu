// Persistent fused rollout for SINGLE-START rows (AttentionModel greedy / sampling / evaluate, S = 1):
// ConstructivePolicy.forward's loop (constructive/base.py:221-263) with one CTA per instance.
//
// The reference re-reads the instance's cached glimpse key / value / logit key (3 x N' x 512 B) from memory at
// every decoding step (am/decoder.py:161-198, attention.py:282-328).  Here the three tables are brought into
// shared memory ONCE per instance by TMA bulk copies and stay there for all T steps, and the W = ceil(N'/8)
// warps of the CTA share the work of ONE row's step:
//   warp w owns nodes [8w, 8w+8): partial masked glimpse (online softmax over its feasible nodes, four nodes per
//   iteration: independent dot products, one rescale), partials combined through shared memory;
//   every warp then holds the 8 heads, computes the logits of its own nodes and publishes them;
//   selection (tanh clip, mask, log-softmax, argmax / Gumbel-max / given action) and the env step are computed
//   redundantly by every warp from the same shared logits, so no broadcast of the action is needed.
// Two CTA barriers per step.  The row state (mask words, capacity, current node) is warp-uniform and replicated.
#include "decode_core.cuh"

namespace r4 {
namespace coop {

constexpr int kNPW = 8;                       // nodes per warp
constexpr int kMaxWarps = kMaxNodes / kNPW;   // 16

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0, spins = 0;
    while (!ok) {
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (!ok && ++spins > (1u << 24)) __trap();   // a copy that never lands must not hang the box
    }
}
// TMA 1-D bulk copy global -> shared, completion counted on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// partial masked glimpse over the feasible nodes of `sub` (bit i <=> node node0 + i): attention.py:308-322 restricted
// to this warp's nodes, as (running max m, sum l, un-normalised heads acc); decode_core.cuh::glimpse_accum
__device__ __forceinline__ void glimpse_part(const float* __restrict__ sk, const float* __restrict__ sv, float4 qs,
                                             uint32_t sub, int node0, int lane, float& m, float& l, float4& acc) {
    m = -INFINITY;
    l = 0.f;
    acc = make_float4(0.f, 0.f, 0.f, 0.f);
    glimpse_accum(sk, sv, qs, sub, node0, lane, m, l, acc);
}

// logits of this warp's nodes in `sub`: heads . lk[n] / sqrt(E) (attention.py:296-301, project_out folded),
// four independent warp reductions per iteration; lane 0 publishes them
__device__ __forceinline__ void logits_part(const float* __restrict__ sl, float4 hd, uint32_t sub, int node0, int lane,
                                            float* __restrict__ out) {
    const int off = 4 * lane;
    while (sub) {
        int j[4];
        bool ok[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            ok[u] = sub != 0u;
            j[u] = ok[u] ? (__ffs(sub) - 1) : 0;
            sub &= sub - 1u;
        }
        float d[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) d[u] = dot4(hd, ld4(sl + (node0 + j[u]) * kE + off));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int u = 0; u < 4; ++u) d[u] += __shfl_xor_sync(kFull, d[u], o);
        }
        if (lane == 0) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) out[node0 + j[u]] = d[u] / 11.313708498984761f;   // / math.sqrt(128)
        }
    }
}

struct Tables {
    const float *gk, *gv, *lk;   // shared memory
    const float *pa, *pb;        // shared or global memory, instance-relative
    float* part;                 // [W][6][32]
    float* logits;               // [kMaxNodes]
};

// One instance's complete rollout by the W warps of the CTA.
template <int ENV>
__device__ __forceinline__ void run_instance(const rl4co_cache& c, const rl4co_rollout_args& a, const Tables& tb,
                                             int64_t b, int warp, int lane, int W) {
    const int n = c.n_nodes;
    const int64_t row = b;                        // S = 1: row = instance
    const int T = a.max_steps;
    const float2* L = reinterpret_cast<const float2*>(a.locs) + b * (int64_t)n;
    const int node0 = kNPW * warp;

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
    const int off = 4 * lane;

    while (!done) {
        if (t >= T) {
            err_bits |= RL4CO_ERR_MAX_STEPS;
            break;
        }
        // ---- query: context embedding + graph context (context.py:118-140 / :70-74,155-157) ----
        float4 q4;
        if (ENV == RL4CO_ENV_TSP) {
            if (!started) {
                q4 = ld4(c.q0 + b * kE + off);
            } else {
                const float4 x = ld4(tb.pa + first * kE + off), y = ld4(tb.pb + cur * kE + off);
                q4 = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
            }
        } else {
            const float4 y = ld4(tb.pb + cur * kE + off), wc = ld4(c.wcap + off);
            const float rem = cap - used;
            q4 = make_float4(fmaf(rem, wc.x, y.x), fmaf(rem, wc.y, y.y), fmaf(rem, wc.z, y.z), fmaf(rem, wc.w, y.w));
        }
        const float4 qs = make_float4(q4.x * 0.25f, q4.y * 0.25f, q4.z * 0.25f, q4.w * 0.25f);
        // this warp's 8 nodes of the (warp-uniform) feasibility words
        uint32_t word = 0u;
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk)
            if ((node0 >> 5) == blk) word = maskw[blk];
        const uint32_t sub = (word >> (node0 & 31)) & ((1u << kNPW) - 1u);
        {
            float m, l;
            float4 acc;
            glimpse_part(tb.gk, tb.gv, qs, sub, node0, lane, m, l, acc);
            float* p = tb.part + warp * (6 * 32) + lane;
            p[0] = m;
            p[32] = l;
            p[64] = acc.x;
            p[96] = acc.y;
            p[128] = acc.z;
            p[160] = acc.w;
        }
        __syncthreads();
        // combine the W partial softmaxes (every warp, redundantly -> every warp holds the heads)
        float4 hd;
        {
            float M = -INFINITY;
            for (int w = 0; w < W; ++w) M = fmaxf(M, tb.part[w * (6 * 32) + lane]);
            float lsum = 0.f;
            float4 A = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int w = 0; w < W; ++w) {
                const float* p = tb.part + w * (6 * 32) + lane;
                const float cw = expf(p[0] - M);  // a warp without feasible nodes: exp(-inf) = 0
                lsum = fmaf(p[32], cw, lsum);
                A.x = fmaf(p[64], cw, A.x);
                A.y = fmaf(p[96], cw, A.y);
                A.z = fmaf(p[128], cw, A.z);
                A.w = fmaf(p[160], cw, A.w);
            }
            const float inv = 1.0f / lsum;        // nothing feasible: NaN, flagged below like the reference's assert
            hd = make_float4(A.x * inv, A.y * inv, A.z * inv, A.w * inv);
        }
        logits_part(tb.lk, hd, sub, node0, lane, tb.logits);
        __syncthreads();

        float lg[kNB], gum[kNB], lp[kNB];
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) {
            const int node = blk * 32 + lane;
            lg[blk] = (node < n && ((maskw[blk] >> lane) & 1u)) ? tb.logits[node] : -INFINITY;
            gum[blk] = 0.f;
        }
        int given = 0;
        if (a.mode == RL4CO_MODE_SAMPLING) {
            const uint32_t key = noise_key(a.seed, a.inst_offset + b, 0, t);
#pragma unroll
            for (int blk = 0; blk < kNB; ++blk) {
                const int node = blk * 32 + lane;
                if (node < n) gum[blk] = a.gumbel ? a.gumbel[(row * T + t) * (int64_t)n + node] : noise_gumbel(key, node);
            }
        } else if (a.mode == RL4CO_MODE_EVALUATE) {
            given = (t < a.actions_in_T) ? (int)a.actions_in[row * a.actions_in_stride + t] : 0;
            if (given < 0 || given >= n) {
                err_bits |= RL4CO_ERR_INFEASIBLE;
                given = 0;
            }
        }
        bool nan_seen;
        const Selection sel = warp_select(lg, maskw, a.tanh_clipping, a.temperature, a.mode, gum, given, lane, lp,
                                          nan_seen);
        if (nan_seen) err_bits |= RL4CO_ERR_NAN_LOGITS;
        const int act = sel.action;
        const float lpa = sel.logp;
        if (a.mode != RL4CO_MODE_EVALUATE && !bit_set(maskw, act)) err_bits |= RL4CO_ERR_INFEASIBLE;
        if (!(lpa > -1000.f)) err_bits |= RL4CO_ERR_NEG_INF_LOGP;  // decoding.py:56-58
        if (warp == 0) {
            if (a.logp_all) {
#pragma unroll
                for (int blk = 0; blk < kNB; ++blk) {
                    const int node = blk * 32 + lane;
                    if (node < n) a.logp_all[(row * T + t) * (int64_t)n + node] = lp[blk];
                }
            }
            if (lane == 0) {
                a.actions[row * T + t] = act;
                a.logp[row * T + t] = lpa;
            }
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
    if (warp == 0 && lane == 0) {
        a.reward[row] = -len;
        a.log_likelihood[row] = ll;
        a.lengths[row] = t;
        atomicMax(a.max_len, t);
        if (err_bits) raise_err(a.err, err_bits, row);
    }
}

// dynamic shared memory: [gk][gv][lk]([pa][pb] | [pb]) [part W x 6 x 32][logits 128][mbarrier]
template <int ENV>
__global__ void __launch_bounds__(kMaxWarps * 32)
    rollout_coop_kernel(rl4co_cache c, rl4co_rollout_args a, int ctx_in_smem) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int n = c.n_nodes;
    const int W = blockDim.x >> 5;
    const uint32_t tile_bytes = (uint32_t)n * kE * sizeof(float);
    float* sk = reinterpret_cast<float*>(smem_raw);
    float* sv = sk + n * kE;
    float* sl = sv + n * kE;
    float* sp = sl + n * kE;
    const int nctx = ctx_in_smem ? ((ENV == RL4CO_ENV_TSP) ? 2 : 1) : 0;
    float* part = sp + (size_t)nctx * n * kE;
    float* logits = part + W * 6 * 32;
    uint64_t* bar = reinterpret_cast<uint64_t*>(logits + kMaxNodes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t parity = 0;
    for (int64_t b = blockIdx.x; b < c.B; b += gridDim.x) {
        const int64_t o = b * (int64_t)n * kE;
        if (threadIdx.x == 0) {
            // order prior generic-proxy reads of the tiles before the async-proxy overwrite
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(bar, (3u + (uint32_t)nctx) * tile_bytes);
            bulk_g2s(sk, c.gk + o, tile_bytes, bar);
            bulk_g2s(sv, c.gv + o, tile_bytes, bar);
            bulk_g2s(sl, c.lk + o, tile_bytes, bar);
            if (nctx == 2) {
                bulk_g2s(sp, c.pa + o, tile_bytes, bar);
                bulk_g2s(sp + n * kE, c.pb + o, tile_bytes, bar);
            } else if (nctx == 1) {
                bulk_g2s(sp, c.pb + o, tile_bytes, bar);
            }
        }
        mbar_wait(bar, parity);
        parity ^= 1u;
        Tables tb;
        tb.gk = sk;
        tb.gv = sv;
        tb.lk = sl;
        tb.pa = (ENV == RL4CO_ENV_TSP) ? (nctx ? sp : c.pa + o) : nullptr;
        tb.pb = nctx ? (ENV == RL4CO_ENV_TSP ? sp + n * kE : sp) : c.pb + o;
        tb.part = part;
        tb.logits = logits;
        run_instance<ENV>(c, a, tb, b, warp, lane, W);
        __syncthreads();  // the instance is finished in every warp before the tiles are overwritten
    }
}

}  // namespace coop

// S = 1 rows: CTA per instance with the tables in shared memory.  Returns false when the shape does not fit.
bool coop_supported(const rl4co_cache* c) { return c->n_nodes <= kMaxNodes && c->n_nodes >= 2; }

int launch_rollout_coop(const rl4co_cache* cache, const rl4co_rollout_args& a, cudaStream_t st) {
    const int n = cache->n_nodes;
    const int W = (n + coop::kNPW - 1) / coop::kNPW;
    const size_t tile = (size_t)n * kE * sizeof(float);
    const size_t scratch = (size_t)W * 6 * 32 * sizeof(float) + kMaxNodes * sizeof(float) + 16;
    const int nctx_full = (a.env == RL4CO_ENV_TSP) ? 2 : 1;
    int dev = 0, sms = 148, max_smem = 227 * 1024;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    const size_t per_sm = 228 * 1024;
    auto ctas_per_sm = [&](size_t smem) {
        if (smem > (size_t)max_smem) return 0;
        int k = (int)(per_sm / (smem + 1024));
        const int by_threads = 2048 / (32 * W);
        return k < by_threads ? k : by_threads;
    };
    const size_t smem_lo = 3 * tile + scratch, smem_hi = (3 + nctx_full) * tile + scratch;
    const int k_lo = ctas_per_sm(smem_lo), k_hi = ctas_per_sm(smem_hi);
    if (k_lo == 0) {
        set_error("rollout: n_nodes=%d does not fit the shared-memory rollout", n);
        return RL4CO_E_ARG;
    }
    // the context tables (pa / pb) ride along in shared memory when that does not cost resident instances
    // (more instances in flight beat the L2 latency of two row gathers per step)
    const int64_t need = cache->B;
    const bool ctx = k_hi > 0 && (k_hi == k_lo || (int64_t)k_hi * sms >= need);
    const size_t smem = ctx ? smem_hi : smem_lo;
    const int k = ctx ? k_hi : k_lo;
    int64_t grid = (int64_t)k * sms;
    if (grid > cache->B) grid = cache->B;
    cudaError_t e;
    if (a.env == RL4CO_ENV_TSP) {
        e = cudaFuncSetAttribute(coop::rollout_coop_kernel<RL4CO_ENV_TSP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess)
            coop::rollout_coop_kernel<RL4CO_ENV_TSP><<<(unsigned)grid, W * 32, smem, st>>>(*cache, a, ctx ? 1 : 0);
    } else {
        e = cudaFuncSetAttribute(coop::rollout_coop_kernel<RL4CO_ENV_CVRP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess)
            coop::rollout_coop_kernel<RL4CO_ENV_CVRP><<<(unsigned)grid, W * 32, smem, st>>>(*cache, a, ctx ? 1 : 0);
    }
    if (e != cudaSuccess) {
        set_error("rollout_coop: cudaFuncSetAttribute(%zu B smem): %s", smem, cudaGetErrorString(e));
        return RL4CO_E_LAUNCH;
    }
    R4_CHECK_LAUNCH("rollout_coop");
    return RL4CO_OK;
}

}  // namespace r4
