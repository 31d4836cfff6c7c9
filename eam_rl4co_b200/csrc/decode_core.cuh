// Warp-level building blocks of the decoder step (fp32 CUDA-core path).
//
// One warp owns one row (instance x start).  Two lane mappings are used:
//   * "lane owns channels": lane L holds embedding channels 4L..4L+3, i.e. head L/4
//     (attention.py:324: head h = channels [16h, 16h+16)).  K/V/logit-key rows are read as
//     one coalesced 512-byte line per node (128-bit per lane), from global or shared memory.
//   * "lane owns nodes": lane L holds nodes L, L+32, L+64, L+96 for the masked log-softmax,
//     the argmax / Gumbel-max selection and the mask construction (ballot).
// Feasibility masks are bit-packed: word `blk` bit `j` <=> node 32*blk+j feasible, replicated
// in every lane so that all mask-driven control flow is warp-uniform.
#pragma once
#include "common.cuh"

namespace r4 {

struct InstPtrs {
    const float* gk;  // [n, E] glimpse key      (global or shared)
    const float* gv;  // [n, E] glimpse value
    const float* lk;  // [n, E] logit key folded with project_out
};

__device__ __forceinline__ float dot4(const float4& a, const float4& b) {
    return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
}

// Masked 8-head glimpse (attention.py:308-322: SDPA with attn_mask, scale 1/sqrt(16)) as an online softmax over the
// feasible nodes only.  Four nodes per iteration: their eight 512-byte rows are requested before any of them is used
// (one warp then keeps 4 KB in flight instead of 1 KB - the streaming kernels are bound by latency x bytes in flight,
// measured 61-65 % of the HBM peak with one node per iteration), the four dot products and their two shuffle levels are
// independent, and the running maximum is rescaled once per group.
// glimpse_accum continues the state (m, l, acc) over the nodes of `sub` (bit i <=> node node0 + i); qs = q / sqrt(d).
__device__ __forceinline__ void glimpse_accum(const float* __restrict__ gk, const float* __restrict__ gv, float4 qs,
                                              uint32_t sub, int node0, int lane, float& m, float& l, float4& acc) {
    const int off = 4 * lane;
    while (sub) {
        int j[4];
        bool ok[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            ok[u] = sub != 0u;
            j[u] = ok[u] ? (__ffs(sub) - 1) : 0;
            sub &= sub - 1u;                      // 0 stays 0
        }
#pragma unroll
        for (int u = 1; u < 4; ++u) j[u] = ok[u] ? j[u] : j[0];   // unused slots re-read slot 0's rows (cache hits)
        float4 k4[4], v4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            k4[u] = ld4(gk + (node0 + j[u]) * kE + off);
            v4[u] = ld4(gv + (node0 + j[u]) * kE + off);
        }
        float s[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) s[u] = dot4(qs, k4[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u) s[u] += __shfl_xor_sync(kFull, s[u], 1);
#pragma unroll
        for (int u = 0; u < 4; ++u) s[u] += __shfl_xor_sync(kFull, s[u], 2);   // the 4 lanes of a head hold its score
#pragma unroll
        for (int u = 1; u < 4; ++u) s[u] = ok[u] ? s[u] : -INFINITY;
        const float mn = fmaxf(fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3])), m);
        const float c = expf(m - mn);             // first group: exp(-inf) = 0
        float p[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) p[u] = expf(s[u] - mn);   // exp(-inf) = 0 for the unused slots
        l = fmaf(l, c, (p[0] + p[1]) + (p[2] + p[3]));
        acc.x = fmaf(acc.x, c, fmaf(p[0], v4[0].x, p[1] * v4[1].x) + fmaf(p[2], v4[2].x, p[3] * v4[3].x));
        acc.y = fmaf(acc.y, c, fmaf(p[0], v4[0].y, p[1] * v4[1].y) + fmaf(p[2], v4[2].y, p[3] * v4[3].y));
        acc.z = fmaf(acc.z, c, fmaf(p[0], v4[0].z, p[1] * v4[1].z) + fmaf(p[2], v4[2].z, p[3] * v4[3].z));
        acc.w = fmaf(acc.w, c, fmaf(p[0], v4[0].w, p[1] * v4[1].w) + fmaf(p[2], v4[2].w, p[3] * v4[3].w));
        m = mn;
    }
}

// Returns this lane's 4 channels of `heads`.
__device__ __forceinline__ float4 warp_glimpse(const InstPtrs& ip, float4 q4, const uint32_t (&maskw)[kNB],
                                               int lane) {
    const float4 qs = make_float4(q4.x * 0.25f, q4.y * 0.25f, q4.z * 0.25f, q4.w * 0.25f);
    float m = -INFINITY, l = 0.f;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) glimpse_accum(ip.gk, ip.gv, qs, maskw[blk], blk * 32, lane, m, l, acc);
    const float inv = 1.0f / l;  // l == 0 (nothing feasible) -> inf * 0 = NaN, flagged by caller
    return make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
}

// logits[n] = heads . lk[n] / sqrt(E)   (attention.py:296-301 with project_out folded, :328).
// Output in "lane owns nodes" form: lg[blk] is the logit of node 32*blk+lane.
// Nodes outside `which` keep -inf.  Four nodes per iteration (four independent warp reductions).
__device__ __forceinline__ void warp_logits(const InstPtrs& ip, float4 hd, const uint32_t (&which)[kNB],
                                            int lane, float (&lg)[kNB]) {
    const int off = 4 * lane;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        lg[blk] = -INFINITY;
        uint32_t w = which[blk];
        while (w) {
            int j[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                ok[u] = w != 0u;
                j[u] = ok[u] ? (__ffs(w) - 1) : 0;
                w &= w - 1u;
            }
#pragma unroll
            for (int u = 1; u < 4; ++u) j[u] = ok[u] ? j[u] : j[0];
            float d[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) d[u] = dot4(hd, ld4(ip.lk + (blk * 32 + j[u]) * kE + off));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int u = 0; u < 4; ++u) d[u] += __shfl_xor_sync(kFull, d[u], o);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u] && lane == j[u]) lg[blk] = d[u] / 11.313708498984761f;  // / math.sqrt(128)
        }
    }
}

struct Selection {
    int action;
    float logp;  // log-prob of the selected action
};

// process_logits (decoding.py:140-190) + Greedy/Sampling/Evaluate._step (:391-465) for one row.
// lg[blk]: raw logits ("lane owns nodes").  lp[blk] receives the log-softmax.
// gum[blk]: Gumbel noise for SAMPLING (ignored otherwise).  given: action for EVALUATE.
__device__ __forceinline__ Selection warp_select(const float (&lg)[kNB], const uint32_t (&maskw)[kNB],
                                                 float tanh_clip, float temperature, int mode,
                                                 const float (&gum)[kNB], int given, int lane,
                                                 float (&lp)[kNB], bool& nan_seen) {
    float x[kNB];
    float mx = -INFINITY;
    bool nan_local = false;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        const bool feas = (maskw[blk] >> lane) & 1u;
        float v = lg[blk];
        if (feas && v != v) nan_local = true;
        if (tanh_clip > 0.f) v = tanhf(v) * tanh_clip;   // :171-172
        v = feas ? v : -INFINITY;                         // :177
        v = v / temperature;                              // :179
        x[blk] = v;
        mx = fmaxf(mx, v);
    }
    nan_seen = __any_sync(kFull, nan_local);
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) sum += expf(x[blk] - mx);  // exp(-inf) = 0
    sum = warp_sum(sum);
    const float logsum = logf(sum);
    float best = -INFINITY;
    int best_i = 0x7fffffff;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        lp[blk] = (x[blk] - mx) - logsum;                 // log_softmax  :190
        float cand = lp[blk];
        if (mode == RL4CO_MODE_SAMPLING) cand = lp[blk] + gum[blk];
        const int node = blk * 32 + lane;
        // first max index wins (torch.argmax): ascending blk with strict >, NaN never wins
        if (cand > best) {
            best = cand;
            best_i = node;
        }
    }
    Selection sel;
    if (mode == RL4CO_MODE_EVALUATE) {
        sel.action = given;
    } else {
        warp_argmax(best, best_i);
        sel.action = (best_i == 0x7fffffff) ? 0 : best_i;
    }
    // gather_by_index(logprobs, selected)  :384-385
    float mine = -INFINITY;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk)
        if ((sel.action >> 5) == blk) mine = lp[blk];
    sel.logp = __shfl_sync(kFull, mine, sel.action & 31);
    return sel;
}

__device__ __forceinline__ bool bit_set(const uint32_t (&w)[kNB], int node) {
    bool r = false;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk)
        if ((node >> 5) == blk) r = (w[blk] >> (node & 31)) & 1u;
    return r;
}
__device__ __forceinline__ void bit_clear(uint32_t (&w)[kNB], int node) {
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk)
        if ((node >> 5) == blk) w[blk] &= ~(1u << (node & 31));
}
__device__ __forceinline__ void bit_put(uint32_t (&w)[kNB], int node) {
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk)
        if ((node >> 5) == blk) w[blk] |= (1u << (node & 31));
}

// bool [n] row -> bit-packed words (ballot), replicated in all lanes
__device__ __forceinline__ void load_mask_bits(const uint8_t* __restrict__ row, int n, int lane,
                                               uint32_t (&w)[kNB]) {
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        const int node = blk * 32 + lane;
        const bool f = (node < n) && (row[node] != 0);
        w[blk] = __ballot_sync(kFull, f);
    }
}

// CVRP feasibility (cvrp/env.py:132-144) from bit-packed visited + per-lane demand registers.
// dm[blk] = demand of node 32*blk+lane (node >= 1), visw = visited bits.
__device__ __forceinline__ void cvrp_mask_bits(const float (&dm)[kNB], const uint32_t (&visw)[kNB],
                                               float used, float cap, int cur, int n, int lane,
                                               uint32_t (&maskw)[kNB]) {
    const float lim = cap + 1e-5f;
    bool any_free = false;
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        const int node = blk * 32 + lane;
        const bool ok = (node >= 1) && (node < n) && !((visw[blk] >> lane) & 1u) && !((dm[blk] + used) > lim);
        maskw[blk] = __ballot_sync(kFull, ok);
        any_free |= (maskw[blk] != 0u);
    }
    const bool mask_depot = (cur == 0) && any_free;
    if (!mask_depot) maskw[0] |= 1u;
}

// ---- query for one row: context embedding + graph context -------------------------------
// TSP  (context.py:118-140): i == 0 -> placeholder (q0), else pa[first] + pb[cur]
// CVRP (context.py:70-74,155-157): pb[cur] + (capacity - used) * wcap
__device__ __forceinline__ float4 row_query(int env, const rl4co_cache& c, int64_t b, int first, int cur,
                                            bool placeholder, float remaining, int lane) {
    const int off = 4 * lane;
    const int64_t base = b * (int64_t)c.n_nodes;
    if (env == RL4CO_ENV_TSP) {
        if (placeholder) return ld4(c.q0 + b * kE + off);
        const float4 a = ld4(c.pa + (base + first) * kE + off);
        const float4 p = ld4(c.pb + (base + cur) * kE + off);
        return make_float4(a.x + p.x, a.y + p.y, a.z + p.z, a.w + p.w);
    } else {
        const float4 p = ld4(c.pb + (base + cur) * kE + off);
        const float4 wc = ld4(c.wcap + off);
        return make_float4(fmaf(remaining, wc.x, p.x), fmaf(remaining, wc.y, p.y),
                           fmaf(remaining, wc.z, p.z), fmaf(remaining, wc.w, p.w));
    }
}

}  // namespace r4
