// The glimpse query of every teacher-forced row and its gradient (the step context of context.py:118-157 with the graph
// context of am/decoder.py:223-224, for all B x S x (T-1) rows of a multistart batch at once):
//   q[b,s,t,:] = pb[b, cur[b,s,t], :]  (+ pa[b, first[b,s], :]   TSP: project_context applied to [first, current])
//                                      (+ rem[b,s,t] * wcap[:]   CVRP: the remaining-capacity column of project_context)
//                                      (+ ctx[b, :]              project_fixed_context of the graph mean)
// where pa / pb = h W_a^T / h W_b^T are the per-instance [n,128] products (project_context is linear, so the projection is
// applied to the n node embeddings once and the rows gather the result).  The forward is one streaming kernel (a warp per
// row, a lane per 16 bytes); PyTorch builds the same tensor in five passes over [rows,128].
//
// The backward is a reduction over ROWS with data-dependent destinations: dpb[b,m] = sum of dq over the rows with cur = m.
// Shared-memory fp32 atomics are CAS loops on sm_100 and global ones collide heavily (100 starts revisit the same nodes), so a
// CTA counting-sorts the row indices of its row range by node (integer shared-memory atomics, 16-bit row ids), cuts the
// sorted list into items of at most 64 rows of one node, and its warps sum the rows of an item in registers (a lane owns
// four channels, a row is one coalesced 512-byte load, four rows in flight) and add the 128 sums to the table with one
// red.global each.  The same kernel with key = first (one key per start) gives dpa; sum of rem * dq (dwcap) is accumulated
// in registers on the way.  dctx[b] = sum over m of dpb[b,m] is left to the caller (every row has exactly one current node).
#include "common.cuh"

namespace r4 {
namespace evq {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kPassRows = 8192;                                 // rows sorted at a time (16-bit ids, 16 KB)
constexpr int kItemRows = 64;
constexpr int kMaxItems = kPassRows / kItemRows + kMaxNodes;    // 256

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void add4(float4& a, const float4& b) {
    a.x += b.x;
    a.y += b.y;
    a.z += b.z;
    a.w += b.w;
}
__device__ __forceinline__ void fma4(float4& a, float s, const float4& b) {
    a.x = fmaf(s, b.x, a.x);
    a.y = fmaf(s, b.y, a.y);
    a.z = fmaf(s, b.z, a.z);
    a.w = fmaf(s, b.w, a.w);
}
__device__ __forceinline__ int clamp_node(int64_t k, int n) { return k < 0 ? 0 : (k >= n ? n - 1 : (int)k); }

__global__ void __launch_bounds__(256)
eval_query_fwd_kernel(const float* __restrict__ pb, const float* __restrict__ pa, const float* __restrict__ wcap,
                      const float* __restrict__ ctx, const int64_t* __restrict__ cur, const int64_t* __restrict__ first,
                      const float* __restrict__ rem, float* __restrict__ q, int64_t rows, int64_t R, int64_t Tm, int n) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float4 wc = wcap ? ld4(wcap + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < rows; row += nw) {
        const int64_t b = row / R;
        float4 v = ld4(pb + (b * n + clamp_node(cur[row], n)) * kE + 4 * lane);
        if (pa) add4(v, ld4(pa + (b * n + clamp_node(first[row / Tm], n)) * kE + 4 * lane));
        if (wcap) fma4(v, rem[row], wc);
        if (ctx) add4(v, ld4(ctx + b * kE + 4 * lane));
        __stcs(reinterpret_cast<float4*>(q + row * kE + 4 * lane), v);
    }
}

// grid (B, row split).  key[(b R + r) / key_div] is the destination node of row r of instance b.
__global__ void __launch_bounds__(kThreads)
eval_query_bwd_kernel(const float* __restrict__ dq, const int64_t* __restrict__ key, int64_t key_div,
                      const float* __restrict__ rem, float* __restrict__ dtab, float* __restrict__ dwcap, int64_t R, int n) {
    __shared__ int s_hist[kMaxNodes], s_off[kMaxNodes], s_ioff[kMaxNodes + 1];
    __shared__ uint32_t s_items[kMaxItems];
    __shared__ uint16_t s_list[kPassRows];
    __shared__ __align__(16) float s_red[kWarps][kE];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t b = blockIdx.x;
    const int64_t per = (R + gridDim.y - 1) / gridDim.y;
    const int64_t r0 = (int64_t)blockIdx.y * per, r1 = r0 + per < R ? r0 + per : R;
    float4 wsum = make_float4(0.f, 0.f, 0.f, 0.f);

    for (int64_t base = r0; base < r1; base += kPassRows) {
        const int cnt = (int)(r1 - base < kPassRows ? r1 - base : kPassRows);
        const int64_t g0 = b * R + base;
        for (int i = tid; i < kMaxNodes; i += kThreads) s_hist[i] = 0;
        __syncthreads();
        for (int i = tid; i < cnt; i += kThreads) atomicAdd(&s_hist[clamp_node(key[(g0 + i) / key_div], n)], 1);
        __syncthreads();
        if (warp == 0) {   // exclusive scans of the node counts and of the per-node item counts (4 nodes per lane)
            int c[4], rows_l = 0, items_l = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                c[j] = s_hist[4 * lane + j];
                rows_l += c[j];
                items_l += (c[j] + kItemRows - 1) / kItemRows;
            }
            int rows_i = rows_l, items_i = items_l;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int a = __shfl_up_sync(kFull, rows_i, d), e = __shfl_up_sync(kFull, items_i, d);
                if (lane >= d) {
                    rows_i += a;
                    items_i += e;
                }
            }
            int ro = rows_i - rows_l, io = items_i - items_l;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                s_off[4 * lane + j] = ro;
                s_ioff[4 * lane + j] = io;
                ro += c[j];
                io += (c[j] + kItemRows - 1) / kItemRows;
            }
            if (lane == 31) s_ioff[kMaxNodes] = io;
        }
        __syncthreads();
        for (int m = tid; m < kMaxNodes; m += kThreads) {   // items: node | first list slot << 8 | rows << 24
            const int c = s_hist[m], o = s_off[m], io = s_ioff[m];
            for (int j = 0; j * kItemRows < c; ++j) {
                const int len = c - j * kItemRows < kItemRows ? c - j * kItemRows : kItemRows;
                s_items[io + j] = (uint32_t)m | ((uint32_t)(o + j * kItemRows) << 8) | ((uint32_t)len << 24);
            }
        }
        __syncthreads();   // s_off becomes the fill cursor below
        for (int i = tid; i < cnt; i += kThreads)
            s_list[atomicAdd(&s_off[clamp_node(key[(g0 + i) / key_div], n)], 1)] = (uint16_t)i;
        __syncthreads();
        const int nitems = s_ioff[kMaxNodes];
        for (int it = warp; it < nitems; it += kWarps) {
            const uint32_t e = s_items[it];
            const int m = e & 255, st = (e >> 8) & 0xffff, len = e >> 24;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
            for (int j = 0; j < len; ++j) {
                const int64_t g = g0 + s_list[st + j];
                const float4 v = ld4(dq + g * kE + 4 * lane);
                add4(acc, v);
                if (rem) fma4(wsum, __ldg(rem + g), v);
            }
            float* d = dtab + (b * n + m) * kE + 4 * lane;
            atomicAdd(d, acc.x);
            atomicAdd(d + 1, acc.y);
            atomicAdd(d + 2, acc.z);
            atomicAdd(d + 3, acc.w);
        }
        __syncthreads();
    }
    if (dwcap) {
        *reinterpret_cast<float4*>(&s_red[warp][4 * lane]) = wsum;
        __syncthreads();
        if (tid < kE) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) s += s_red[w][tid];
            atomicAdd(dwcap + tid, s);
        }
    }
}

static int split_for(int64_t B, int ctas_per_sm) {   // the rule of evaluate_kernels.cu: fewest waves per unit of work
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t slots = (int64_t)sms * ctas_per_sm;
    int best = 1;
    double best_cost = 1e30;
    for (int s = 1; s <= 32; ++s) {
        const double cost = (double)((B * s + slots - 1) / slots) / s;
        if (cost < best_cost * 0.97) {
            best_cost = cost;
            best = s;
        }
    }
    return best;
}

}  // namespace evq
}  // namespace r4

using namespace r4;

static int check_query_sizes(const char* who, int64_t B, int64_t S, int64_t Tm, int32_t n_nodes) {
    R4_CHECK_ARG(B >= 0 && S >= 0 && Tm >= 0 && B <= 65535, "%s: bad sizes B=%lld S=%lld Tm=%lld", who, (long long)B,
                 (long long)S, (long long)Tm);
    R4_CHECK_ARG(n_nodes >= 1 && n_nodes <= kMaxNodes, "%s: n_nodes=%d outside [1, %d]", who, n_nodes, kMaxNodes);
    return RL4CO_OK;
}

extern "C" int rl4co_eval_query_fwd(const float* pb, const float* pa, const float* wcap, const float* ctx, const int64_t* cur,
                                    const int64_t* first, const float* rem, float* q, int64_t B, int64_t S, int64_t Tm,
                                    int32_t n_nodes, void* stream) {
    int rc = check_query_sizes("rl4co_eval_query_fwd", B, S, Tm, n_nodes);
    if (rc) return rc;
    R4_CHECK_ARG(pb && cur && q, "rl4co_eval_query_fwd: null pb / cur / q");
    R4_CHECK_ARG((pa == nullptr) == (first == nullptr), "rl4co_eval_query_fwd: pa and first come together");
    R4_CHECK_ARG((wcap == nullptr) == (rem == nullptr), "rl4co_eval_query_fwd: wcap and rem come together");
    const int64_t rows = B * S * Tm;
    if (rows == 0) return RL4CO_OK;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t want = (rows + 7) / 8;
    const unsigned grid = (unsigned)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
    evq::eval_query_fwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(pb, pa, wcap, ctx, cur, first, rem, q, rows, S * Tm, Tm,
                                                                      n_nodes);
    R4_CHECK_LAUNCH("rl4co_eval_query_fwd");
    return RL4CO_OK;
}

extern "C" int rl4co_eval_query_bwd(const float* dq, const int64_t* cur, const int64_t* first, const float* rem, float* dpb,
                                    float* dpa, float* dwcap, int64_t B, int64_t S, int64_t Tm, int32_t n_nodes, void* stream) {
    int rc = check_query_sizes("rl4co_eval_query_bwd", B, S, Tm, n_nodes);
    if (rc) return rc;
    R4_CHECK_ARG(dq && cur && dpb, "rl4co_eval_query_bwd: null dq / cur / dpb");
    R4_CHECK_ARG((dpa == nullptr) == (first == nullptr), "rl4co_eval_query_bwd: dpa and first come together");
    R4_CHECK_ARG((dwcap == nullptr) == (rem == nullptr), "rl4co_eval_query_bwd: dwcap and rem come together");
    const int64_t R = S * Tm;
    if (B == 0 || R == 0) return RL4CO_OK;
    dim3 grid((unsigned)B, (unsigned)evq::split_for(B, 2));
    evq::eval_query_bwd_kernel<<<grid, evq::kThreads, 0, (cudaStream_t)stream>>>(dq, cur, 1, rem, dpb, dwcap, R, n_nodes);
    R4_CHECK_LAUNCH("rl4co_eval_query_bwd (current node)");
    if (dpa) {
        evq::eval_query_bwd_kernel<<<grid, evq::kThreads, 0, (cudaStream_t)stream>>>(dq, first, Tm, nullptr, dpa, nullptr, R,
                                                                                    n_nodes);
        R4_CHECK_LAUNCH("rl4co_eval_query_bwd (first node)");
    }
    return RL4CO_OK;
}
