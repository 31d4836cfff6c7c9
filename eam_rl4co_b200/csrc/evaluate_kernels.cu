// Teacher-forced evaluate (decoding.py:452-465, constructive/base.py:204-229) of ALL decoding steps at once, forward and
// backward: log p(action_t | prefix) for every (instance, start, step) row and its gradient w.r.t. the glimpse query and the
// per-instance glimpse key / glimpse value / logit key tables.  SURVEY.md §8(b): rl4co_eval_ll_fwd / rl4co_eval_ll_bwd.
//
// With the actions known, the rows (instance b, start s, step t) are independent: row r of instance b has its query q[b,r,:]
// (context.py:118-157 gathers, built by the caller with autograd), its feasibility mask and its taken action.  All rows of an
// instance share the three [n,128] tables, which a CTA keeps in shared memory; one warp owns one row at a time:
//   glimpse  (attention.py:308-322): 8-head masked softmax attention, lane L holds channels 4L..4L+3 (head L/4),
//   pointer  (attention.py:296-301,328; decoding.py:140-190): logits = heads . lkf[n] / sqrt(E), tanh clip, mask,
//            temperature, log_softmax, gather at the taken action.
// Only FEASIBLE nodes are visited (masked nodes have probability exactly 0 in both softmaxes and contribute no gradient).
// The backward recomputes the probabilities from the saved log-sum-exps (flash-attention style) in two passes:
//   row pass  (a warp per row)            : dheads, dq and du = d loss / d raw logit - everything that reduces over NODES;
//   node pass (a warp per 7 nodes, all rows): dgk = dS^T Q, dgv = P^T dO, dlkf = dU^T O - the reductions over ROWS, with the
//             accumulators in registers (fp32 shared-memory atomics are CAS loops on sm_100: a first version that
//             accumulated the table gradients in shared memory spent 25 of 31 ms of the c5 evaluate in ATOMS.CAST.SPIN).
//
// fp32 CUDA-core arithmetic; exp of the glimpse softmax on MUFU ex2 (as the tensor-path rollout), tanhf / expf / logf in the
// pointer.
#include "decode_core.cuh"

namespace r4 {
namespace ev {

constexpr int kThreads = 512;
constexpr float kInvSqrtE = 0.08838834764831845f;   // 1 / sqrt(128)

__device__ __forceinline__ void load_mask_words(const uint8_t* __restrict__ m, int n, int lane, uint32_t (&w)[kNB]) {
#pragma unroll
    for (int blk = 0; blk < kNB; ++blk) {
        const int i = 32 * blk + lane;
        w[blk] = __ballot_sync(kFull, i < n && m[i] != 0);
    }
}

__device__ __forceinline__ void copy_table(float* dst, const float* __restrict__ src, int n) {
    for (int i = threadIdx.x; i < n * (kE / 4); i += blockDim.x)
        reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
}
__device__ __forceinline__ float sel4(const float (&v)[kNB], int k) {
    return k == 0 ? v[0] : (k == 1 ? v[1] : (k == 2 ? v[2] : v[3]));
}

__device__ __forceinline__ uint32_t sel4u(const uint32_t (&v)[kNB], int k) {
    return k == 0 ? v[0] : (k == 1 ? v[1] : (k == 2 ? v[2] : v[3]));
}

// rows [r0, r1) of the instance for CTA (blockIdx.x = instance, blockIdx.y = row split)
__device__ __forceinline__ void row_range(int64_t R, int64_t& r0, int64_t& r1) {
    const int64_t per = (R + gridDim.y - 1) / gridDim.y;
    r0 = (int64_t)blockIdx.y * per;
    r1 = r0 + per < R ? r0 + per : R;
}

// exp on the MUFU ex2 approximation (relative error ~2^-22, the arithmetic of the tensor-path rollout kernels)
__device__ __forceinline__ float fexp(float x) { return __expf(x); }

// glimpse_accum (decode_core.cuh) with fexp: online masked softmax over the feasible nodes of `sub`, four per iteration
__device__ __forceinline__ void glimpse_accum_f(const float* gk, const float* gv, float4 qs, uint32_t sub, int node0,
                                                int lane, float& m, float& l, float4& acc) {
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
#pragma unroll
        for (int u = 1; u < 4; ++u) j[u] = ok[u] ? j[u] : j[0];
        float4 k4[4], v4[4];
        float s[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            k4[u] = ld4(gk + (node0 + j[u]) * kE + off);
            v4[u] = ld4(gv + (node0 + j[u]) * kE + off);
            s[u] = dot4(qs, k4[u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) s[u] += __shfl_xor_sync(kFull, s[u], 1);
#pragma unroll
        for (int u = 0; u < 4; ++u) s[u] += __shfl_xor_sync(kFull, s[u], 2);
#pragma unroll
        for (int u = 1; u < 4; ++u) s[u] = ok[u] ? s[u] : -INFINITY;
        const float mn = fmaxf(fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3])), m);
        const float c = fexp(m - mn);
        float p[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) p[u] = fexp(s[u] - mn);
        l = fmaf(l, c, (p[0] + p[1]) + (p[2] + p[3]));
        acc.x = fmaf(acc.x, c, fmaf(p[0], v4[0].x, p[1] * v4[1].x) + fmaf(p[2], v4[2].x, p[3] * v4[3].x));
        acc.y = fmaf(acc.y, c, fmaf(p[0], v4[0].y, p[1] * v4[1].y) + fmaf(p[2], v4[2].y, p[3] * v4[3].y));
        acc.z = fmaf(acc.z, c, fmaf(p[0], v4[0].z, p[1] * v4[1].z) + fmaf(p[2], v4[2].z, p[3] * v4[3].z));
        acc.w = fmaf(acc.w, c, fmaf(p[0], v4[0].w, p[1] * v4[1].w) + fmaf(p[2], v4[2].w, p[3] * v4[3].w));
        m = mn;
    }
}

// warp_logits (decode_core.cuh) with half the shuffles: the four partial dot products of a group are reduce-scattered
// (xor 16 keeps two of them per half warp, xor 8 one per quarter) before the last three butterfly levels, and the logit
// is scaled by a multiplication.  lg[blk] = heads . lkf[32 blk + lane] / sqrt(E) on the nodes of `which`, -inf elsewhere.
__device__ __forceinline__ void warp_logits_f(const float* lk, float4 hd, const uint32_t (&which)[kNB], int lane,
                                              float (&lg)[kNB]) {
    const int off = 4 * lane;
    const bool h16 = lane & 16, h8 = lane & 8;
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
            for (int u = 0; u < 4; ++u) d[u] = dot4(hd, ld4(lk + (blk * 32 + j[u]) * kE + off));
            float a0 = h16 ? d[2] : d[0], a1 = h16 ? d[3] : d[1];
            a0 += __shfl_xor_sync(kFull, h16 ? d[0] : d[2], 16);
            a1 += __shfl_xor_sync(kFull, h16 ? d[1] : d[3], 16);
            float t = h8 ? a1 : a0;
            t += __shfl_xor_sync(kFull, h8 ? a0 : a1, 8);
            t += __shfl_xor_sync(kFull, t, 4);
            t += __shfl_xor_sync(kFull, t, 2);
            t += __shfl_xor_sync(kFull, t, 1);        // lanes [8 u, 8 u + 8) hold the dot product of node u
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float v = __shfl_sync(kFull, t, 8 * u);
                if (ok[u] && lane == j[u]) lg[blk] = v * kInvSqrtE;
            }
        }
    }
}

// ---- forward: logp[b,r]; saved for the backward: heads [B,R,E], lse_h [B,R,H], lse_z [B,R], maskw [B,R,4] -----------------
__global__ void __launch_bounds__(kThreads, 1)
    eval_ll_fwd_kernel(const float* __restrict__ q, const float* __restrict__ gk, const float* __restrict__ gv,
                       const float* __restrict__ lkf, const uint8_t* __restrict__ mask, const int64_t* __restrict__ taken,
                       float* __restrict__ logp, float* __restrict__ heads, float* __restrict__ lse_h,
                       float* __restrict__ lse_z, uint32_t* __restrict__ maskw, int64_t R, int n, float clip,
                       float temperature) {
    extern __shared__ __align__(16) float sm[];
    float* s_gk = sm;
    float* s_gv = sm + n * kE;
    float* s_lk = sm + 2 * n * kE;
    const int64_t b = blockIdx.x;
    copy_table(s_gk, gk + b * n * kE, n);
    copy_table(s_gv, gv + b * n * kE, n);
    copy_table(s_lk, lkf + b * n * kE, n);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    int64_t r0, r1;
    row_range(R, r0, r1);
    for (int64_t r = r0 + warp; r < r1; r += nw) {
        const int64_t row = b * R + r;
        const float4 q4 = ld4(q + row * kE + 4 * lane);
        uint32_t w[kNB];
        load_mask_words(mask + row * n, n, lane, w);
        if (lane < kNB) maskw[row * kNB + lane] = sel4u(w, lane);
        const float4 qs = make_float4(q4.x * 0.25f, q4.y * 0.25f, q4.z * 0.25f, q4.w * 0.25f);
        float m = -INFINITY, l = 0.f;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) glimpse_accum_f(s_gk, s_gv, qs, w[blk], blk * 32, lane, m, l, acc);
        const bool any = l > 0.f;
        const float inv = any ? 1.0f / l : 0.f;
        const float4 o4 = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
        *reinterpret_cast<float4*>(heads + row * kE + 4 * lane) = o4;
        if ((lane & 3) == 0) lse_h[row * kH + (lane >> 2)] = any ? m + logf(l) : 0.f;
        float lg[kNB];
        warp_logits_f(s_lk, o4, w, lane, lg);           // raw logits / sqrt(E), lane owns nodes; -inf off the mask
        float z[kNB], mx = -INFINITY;
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) {
            const bool f = (w[blk] >> lane) & 1u;
            float v = lg[blk];
            if (clip > 0.f) v = tanhf(v) * clip;        // decoding.py:171-172
            z[blk] = f ? v / temperature : -INFINITY;   // :177, :180
            mx = fmaxf(mx, z[blk]);
        }
        mx = warp_max(mx);
        float sum = 0.f;
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) sum += (z[blk] == -INFINITY) ? 0.f : expf(z[blk] - mx);
        sum = warp_sum(sum);
        const float lz = any ? mx + logf(sum) : 0.f;
        const int a = (int)taken[row];
        const float za = __shfl_sync(kFull, sel4(z, (a >> 5) & 3), a & 31);
        if (lane == 0) {
            logp[row] = any ? za - lz : 0.f;            // :190 log_softmax, :462 gather
            lse_z[row] = lz;
        }
    }
}

// du_t [B,R,128]: d loss / d (raw logit of node n), stored so that the nodes one warp of the node pass owns
// (n = w, w + 16, ..., w + 96) are the eight contiguous floats [8 w, 8 w + 8)
__device__ __forceinline__ int du_index(int node) { return (node & 15) * 8 + (node >> 4); }

// ---- backward, row pass (one warp per row, no atomics): dheads, dq, du_t ---------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
    eval_row_bwd_kernel(const float* __restrict__ q, const float* __restrict__ gk, const float* __restrict__ gv,
                        const float* __restrict__ lkf, const uint32_t* __restrict__ maskw, const int64_t* __restrict__ taken,
                        const float* __restrict__ heads, const float* __restrict__ lse_h, const float* __restrict__ lse_z,
                        const float* __restrict__ dlogp, float* __restrict__ dheads, float* __restrict__ dq,
                        float* __restrict__ du_t, float* __restrict__ dd_out, int64_t R, int n, float clip,
                        float temperature) {
    extern __shared__ __align__(16) float sm[];
    float* s_gk = sm;
    float* s_gv = sm + n * kE;
    float* s_lk = sm + 2 * n * kE;
    const int64_t b = blockIdx.x;
    copy_table(s_gk, gk + b * n * kE, n);
    copy_table(s_gv, gv + b * n * kE, n);
    copy_table(s_lk, lkf + b * n * kE, n);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const int off = 4 * lane;
    int64_t r0, r1;
    row_range(R, r0, r1);
    for (int64_t r = r0 + warp; r < r1; r += nw) {
        const int64_t row = b * R + r;
        const float4 o4 = ld4(heads + row * kE + off);
        const float g = dlogp[row], lz = lse_z[row];
        const int a = (int)taken[row];
        uint32_t w[kNB];
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) w[blk] = maskw[row * kNB + blk];
        // pointer: du[n] = g (1[n = a] - softmax(z)[n]) dz/du / sqrt(E), lane owns nodes
        float lg[kNB], du[kNB];
        warp_logits_f(s_lk, o4, w, lane, lg);
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) {
            const bool f = (w[blk] >> lane) & 1u;
            const float uv = f ? lg[blk] : 0.f;
            float zv, dzu;
            if (clip > 0.f) {
                const float t = tanhf(uv);
                zv = t * clip / temperature;
                dzu = clip * (1.0f - t * t) / temperature;
            } else {
                zv = uv / temperature;
                dzu = 1.0f / temperature;
            }
            const float pz = expf(zv - lz);
            du[blk] = f ? g * ((32 * blk + lane == a ? 1.0f : 0.0f) - pz) * dzu * kInvSqrtE : 0.f;
            du_t[row * kE + du_index(32 * blk + lane)] = du[blk];
        }
        // dheads = sum_n du[n] lkf[n]
        float4 d4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) {
            uint32_t wb = w[blk];
            while (wb) {
                const int j = __ffs(wb) - 1;
                wb &= wb - 1u;
                const float dn = __shfl_sync(kFull, du[blk], j);
                const float4 l4 = ld4(s_lk + (blk * 32 + j) * kE + off);
                d4.x = fmaf(dn, l4.x, d4.x);
                d4.y = fmaf(dn, l4.y, d4.y);
                d4.z = fmaf(dn, l4.z, d4.z);
                d4.w = fmaf(dn, l4.w, d4.w);
            }
        }
        *reinterpret_cast<float4*>(dheads + row * kE + off) = d4;
        // glimpse: dq = 0.25 sum_n dS[n] k[n], dS = P (dP - D), P recomputed from the saved log-sum-exp
        const float4 q4 = ld4(q + row * kE + off);
        const float4 qs = make_float4(q4.x * 0.25f, q4.y * 0.25f, q4.z * 0.25f, q4.w * 0.25f);
        const float lh = lse_h[row * kH + (lane >> 2)];
        float dd = dot4(d4, o4);
        dd += __shfl_xor_sync(kFull, dd, 1);
        dd += __shfl_xor_sync(kFull, dd, 2);
        if ((lane & 3) == 0) dd_out[row * kH + (lane >> 2)] = dd;
        float4 dqs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
        for (int blk = 0; blk < kNB; ++blk) {
            uint32_t wb = w[blk];
            while (wb) {
                int j[4];
                bool ok[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    ok[u] = wb != 0u;
                    j[u] = ok[u] ? (__ffs(wb) - 1) : 0;
                    wb &= wb - 1u;
                }
#pragma unroll
                for (int u = 1; u < 4; ++u) j[u] = ok[u] ? j[u] : j[0];
                float4 k4[4];
                float s[4], dp[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    k4[u] = ld4(s_gk + (blk * 32 + j[u]) * kE + off);
                    const float4 v4 = ld4(s_gv + (blk * 32 + j[u]) * kE + off);
                    s[u] = dot4(qs, k4[u]);
                    dp[u] = dot4(d4, v4);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    s[u] += __shfl_xor_sync(kFull, s[u], 1);
                    dp[u] += __shfl_xor_sync(kFull, dp[u], 1);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    s[u] += __shfl_xor_sync(kFull, s[u], 2);
                    dp[u] += __shfl_xor_sync(kFull, dp[u], 2);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float ds = ok[u] ? fexp(s[u] - lh) * (dp[u] - dd) : 0.f;
                    dqs.x = fmaf(ds, k4[u].x, dqs.x);
                    dqs.y = fmaf(ds, k4[u].y, dqs.y);
                    dqs.z = fmaf(ds, k4[u].z, dqs.z);
                    dqs.w = fmaf(ds, k4[u].w, dqs.w);
                }
            }
        }
        *reinterpret_cast<float4*>(dq + row * kE + off) =
            make_float4(dqs.x * 0.25f, dqs.y * 0.25f, dqs.z * 0.25f, dqs.w * 0.25f);
    }
}

// ---- backward, node pass: the GEMM-shaped reductions over rows with the accumulators in REGISTERS -------------------------------
// Warp w of the CTA owns nodes w, w + 16, ..., w + 96 and visits every row of the CTA's row range:
//   blockIdx.z = 0 (pointer): dlkf[n] += du[row, n] heads[row]                          (dL = dU^T O)
//   blockIdx.z = 1 (glimpse): P, dS of (row, its nodes) recomputed; dgv[n] += P dheads[row], dgk[n] += dS 0.25 q[row]
// Lane L holds channels 4L..4L+3 of every accumulator; a CTA's sums reach global memory with one red per element.
constexpr int kOwn = 7;   // 16 warps x 7 nodes = 112 = rl4co_eval_ll_max_nodes()

struct NodeRow {   // what the glimpse part of the node pass reads per row
    float4 q4, d4;
    float lh, dd;
    uint4 mw;
};
__device__ __forceinline__ NodeRow load_node_row(const float* __restrict__ q, const float* __restrict__ dheads,
                                                 const float* __restrict__ lse_h, const float* __restrict__ dd,
                                                 const uint32_t* __restrict__ maskw, int64_t row, int lane) {
    NodeRow x;
    x.q4 = ld4(q + row * kE + 4 * lane);
    x.d4 = ld4(dheads + row * kE + 4 * lane);
    x.lh = lse_h[row * kH + (lane >> 2)];
    x.dd = dd[row * kH + (lane >> 2)];
    x.mw = *reinterpret_cast<const uint4*>(maskw + row * kNB);
    return x;
}

__global__ void __launch_bounds__(kThreads, 1)
    eval_node_bwd_kernel(const float* __restrict__ q, const float* __restrict__ gk, const float* __restrict__ gv,
                         const uint32_t* __restrict__ maskw, const float* __restrict__ heads, const float* __restrict__ lse_h,
                         const float* __restrict__ dheads, const float* __restrict__ du_t, const float* __restrict__ dd,
                         float* __restrict__ dgk, float* __restrict__ dgv, float* __restrict__ dlkf, int64_t R, int n) {
    extern __shared__ __align__(16) float sm[];
    const int64_t b = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int off = 4 * lane;
    int64_t r0, r1;
    row_range(R, r0, r1);
    if (r0 >= r1) return;
    if (blockIdx.z == 0) {
        float4 acc[kOwn];
#pragma unroll
        for (int i = 0; i < kOwn; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        // the next row's loads are in flight while a row is accumulated (every warp of the CTA streams the same rows)
        float4 o4 = ld4(heads + (b * R + r0) * kE + off);
        float4 da = ld4(du_t + (b * R + r0) * kE + 8 * warp);
        float4 db = ld4(du_t + (b * R + r0) * kE + 8 * warp + 4);
        for (int64_t r = r0; r < r1; ++r) {
            const int64_t nxt = b * R + (r + 1 < r1 ? r + 1 : r);
            const float4 o4n = ld4(heads + nxt * kE + off);
            const float4 dan = ld4(du_t + nxt * kE + 8 * warp);
            const float4 dbn = ld4(du_t + nxt * kE + 8 * warp + 4);
            const float dn[8] = {da.x, da.y, da.z, da.w, db.x, db.y, db.z, db.w};
#pragma unroll
            for (int i = 0; i < kOwn; ++i) {
                acc[i].x = fmaf(dn[i], o4.x, acc[i].x);
                acc[i].y = fmaf(dn[i], o4.y, acc[i].y);
                acc[i].z = fmaf(dn[i], o4.z, acc[i].z);
                acc[i].w = fmaf(dn[i], o4.w, acc[i].w);
            }
            o4 = o4n;
            da = dan;
            db = dbn;
        }
#pragma unroll
        for (int i = 0; i < kOwn; ++i) {
            const int node = warp + 16 * i;
            if (node < n) {
                float* dst = dlkf + (b * n + node) * kE + off;
                atomicAdd(dst + 0, acc[i].x);
                atomicAdd(dst + 1, acc[i].y);
                atomicAdd(dst + 2, acc[i].z);
                atomicAdd(dst + 3, acc[i].w);
            }
        }
        return;
    }
    float* s_gk = sm;
    float* s_gv = sm + n * kE;
    copy_table(s_gk, gk + b * n * kE, n);
    copy_table(s_gv, gv + b * n * kE, n);
    __syncthreads();
    float4 ak[kOwn], av[kOwn];
#pragma unroll
    for (int i = 0; i < kOwn; ++i) {
        ak[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        av[i] = ak[i];
    }
    NodeRow x = load_node_row(q, dheads, lse_h, dd, maskw, b * R + r0, lane);
    for (int64_t r = r0; r < r1; ++r) {
        const NodeRow xn = load_node_row(q, dheads, lse_h, dd, maskw, b * R + (r + 1 < r1 ? r + 1 : r), lane);
        // feasibility of the warp's nodes: node w + 16 i = bit (w + 16 (i & 1)) of word i / 2
        const uint32_t m0 = (x.mw.x >> warp) & 0x10001u, m1 = (x.mw.y >> warp) & 0x10001u,
                       m2 = (x.mw.z >> warp) & 0x10001u, m3 = (x.mw.w >> warp) & 0x10001u;
        const uint32_t own = (m0 & 1u) | ((m0 >> 15) & 2u) | ((m1 & 1u) << 2) | ((m1 >> 13) & 8u) | ((m2 & 1u) << 4) |
                             ((m2 >> 11) & 32u) | ((m3 & 1u) << 6);
        const float4 qs = make_float4(x.q4.x * 0.25f, x.q4.y * 0.25f, x.q4.z * 0.25f, x.q4.w * 0.25f);
#pragma unroll
        for (int i = 0; i < kOwn; ++i) {
            if (!((own >> i) & 1u)) continue;            // warp-uniform
            const int node = warp + 16 * i;
            const float4 k4 = ld4(s_gk + node * kE + off);
            const float4 v4 = ld4(s_gv + node * kE + off);
            float s = dot4(qs, k4), dp = dot4(x.d4, v4);
            s += __shfl_xor_sync(kFull, s, 1);
            dp += __shfl_xor_sync(kFull, dp, 1);
            s += __shfl_xor_sync(kFull, s, 2);
            dp += __shfl_xor_sync(kFull, dp, 2);
            const float p = fexp(s - x.lh);
            const float ds = p * (dp - x.dd);
            ak[i].x = fmaf(ds, qs.x, ak[i].x);
            ak[i].y = fmaf(ds, qs.y, ak[i].y);
            ak[i].z = fmaf(ds, qs.z, ak[i].z);
            ak[i].w = fmaf(ds, qs.w, ak[i].w);
            av[i].x = fmaf(p, x.d4.x, av[i].x);
            av[i].y = fmaf(p, x.d4.y, av[i].y);
            av[i].z = fmaf(p, x.d4.z, av[i].z);
            av[i].w = fmaf(p, x.d4.w, av[i].w);
        }
        x = xn;
    }
#pragma unroll
    for (int i = 0; i < kOwn; ++i) {
        const int node = warp + 16 * i;
        if (node < n) {
            float* dk = dgk + (b * n + node) * kE + off;
            float* dv = dgv + (b * n + node) * kE + off;
            atomicAdd(dk + 0, ak[i].x);
            atomicAdd(dk + 1, ak[i].y);
            atomicAdd(dk + 2, ak[i].z);
            atomicAdd(dk + 3, ak[i].w);
            atomicAdd(dv + 0, av[i].x);
            atomicAdd(dv + 1, av[i].y);
            atomicAdd(dv + 2, av[i].z);
            atomicAdd(dv + 3, av[i].w);
        }
    }
}

// Row split per instance: `units` equal CTAs per split step over sms * ctas_per_sm slots, the split in [1, 32] with the
// fewest waves per unit of work (64 instances x 2 = 128 CTAs would leave 20 SMs idle; x 9 = 576 CTAs fill 3.9 of 4 waves).
static int grid_split(int64_t units, int ctas_per_sm) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t slots = (int64_t)sms * ctas_per_sm;
    int best = 1;
    double best_cost = 1e30;
    for (int s = 1; s <= 32; ++s) {
        const double cost = (double)((units * s + slots - 1) / slots) / s;
        if (cost < best_cost * 0.97) {   // prefer the smaller split unless the larger one wins by > 3 %
            best_cost = cost;
            best = s;
        }
    }
    return best;
}

template <typename K>
static int set_smem(K kern, size_t bytes, const char* who) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) {
        set_error("%s: cudaFuncSetAttribute(%zu B smem): %s", who, bytes, cudaGetErrorString(e));
        return RL4CO_E_LAUNCH;
    }
    return RL4CO_OK;
}

}  // namespace ev
}  // namespace r4

using namespace r4;

extern "C" int32_t rl4co_eval_ll_max_nodes(void) { return 16 * ev::kOwn; }   // 112: node pass = 16 warps x 7 nodes

static int eval_check(const char* who, int64_t B, int64_t R, int32_t n_nodes, float temperature) {
    R4_CHECK_ARG(B >= 0 && R >= 0 && B <= 65535, "%s: bad sizes B=%lld R=%lld", who, (long long)B, (long long)R);
    R4_CHECK_ARG(n_nodes >= 2 && n_nodes <= rl4co_eval_ll_max_nodes(), "%s: n_nodes=%d outside [2, %d]", who, n_nodes,
                 rl4co_eval_ll_max_nodes());
    R4_CHECK_ARG(temperature > 0.f, "%s: temperature must be > 0", who);
    return RL4CO_OK;
}

extern "C" int rl4co_eval_ll_fwd(const float* q, const float* gk, const float* gv, const float* lkf, const uint8_t* mask,
                                 const int64_t* taken, float* logp, float* heads, float* lse_h, float* lse_z,
                                 uint32_t* maskw, int64_t B, int64_t R, int32_t n_nodes, float tanh_clipping,
                                 float temperature, void* stream) {
    R4_CHECK_ARG(q && gk && gv && lkf && mask && taken && logp && heads && lse_h && lse_z && maskw,
                 "eval_ll_fwd: null pointer");
    int rc;
    if ((rc = eval_check("eval_ll_fwd", B, R, n_nodes, temperature))) return rc;
    if (B == 0 || R == 0) return RL4CO_OK;
    const size_t smem = 3ull * n_nodes * kE * sizeof(float);
    if ((rc = ev::set_smem(ev::eval_ll_fwd_kernel, smem, "eval_ll_fwd"))) return rc;
    dim3 grid((unsigned)B, (unsigned)ev::grid_split(B, 1));
    ev::eval_ll_fwd_kernel<<<grid, ev::kThreads, smem, (cudaStream_t)stream>>>(q, gk, gv, lkf, mask, taken, logp, heads, lse_h,
                                                                              lse_z, maskw, R, n_nodes, tanh_clipping,
                                                                              temperature);
    R4_CHECK_LAUNCH("eval_ll_fwd");
    return RL4CO_OK;
}

extern "C" int rl4co_eval_ll_bwd(const float* q, const float* gk, const float* gv, const float* lkf, const uint32_t* maskw,
                                 const int64_t* taken, const float* heads, const float* lse_h, const float* lse_z,
                                 const float* dlogp, float* dheads, float* du_t, float* dd, float* dq, float* dgk,
                                 float* dgv, float* dlkf, int64_t B, int64_t R, int32_t n_nodes, float tanh_clipping, float temperature,
                                 void* stream) {
    R4_CHECK_ARG(q && gk && gv && lkf && maskw && taken && heads && lse_h && lse_z && dlogp && dheads && du_t && dd && dq &&
                     dgk && dgv && dlkf,
                 "eval_ll_bwd: null pointer");
    int rc;
    if ((rc = eval_check("eval_ll_bwd", B, R, n_nodes, temperature))) return rc;
    if (B == 0 || R == 0) return RL4CO_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t table = (size_t)n_nodes * kE * sizeof(float);
    if ((rc = ev::set_smem(ev::eval_row_bwd_kernel, 3 * table, "eval_ll_bwd (row pass)"))) return rc;
    if ((rc = ev::set_smem(ev::eval_node_bwd_kernel, 2 * table, "eval_ll_bwd (node pass)"))) return rc;
    dim3 g1((unsigned)B, (unsigned)ev::grid_split(B, 1));
    ev::eval_row_bwd_kernel<<<g1, ev::kThreads, 3 * table, st>>>(q, gk, gv, lkf, maskw, taken, heads, lse_h, lse_z, dlogp,
                                                                dheads, dq, du_t, dd, R, n_nodes, tanh_clipping, temperature);
    R4_CHECK_LAUNCH("eval_ll_bwd (row pass)");
    // pointer part (z = 0) and glimpse part (z = 1) of the node pass share the grid: twice grid_split CTAs per instance
    dim3 g2((unsigned)B, (unsigned)ev::grid_split(2 * B, 1), 2u);
    ev::eval_node_bwd_kernel<<<g2, ev::kThreads, 2 * table, st>>>(q, gk, gv, maskw, heads, lse_h, dheads, du_t, dd, dgk, dgv,
                                                                 dlkf, R, n_nodes);
    R4_CHECK_LAUNCH("eval_ll_bwd (node pass)");
    return RL4CO_OK;
}
