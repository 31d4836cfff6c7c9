// Graph-attention encoder of the AttentionModel on tcgen05 (the caller of the rollout path).
//
// One encoder layer (nn/graph/attnnet.py:16-62, nn/attention.py:66-136, nn/ops.py:32-56, nn/mlp.py):
//   qkv = h Wqkv^T + b                                  enc_linear_kernel<EPI_PLAIN>      [n,128] x [128,384]
//   a   = softmax(q k^T / sqrt(16)) v   per head        enc_attention_kernel              8 heads of [n,n]
//   h   = IN(h + a Wo^T + bo)                           enc_linear_kernel<EPI_RES_NORM>   [n,128] x [128,128]
//   f   = relu(h W1^T + b1)                             enc_linear_kernel<EPI_RELU>       [n,128] x [128,512]
//   h   = IN(h + f W2^T + b2)                           enc_linear_kernel<EPI_RES_NORM>   [n,512] x [512,128]
// IN = InstanceNorm1d over the node axis with affine parameters: its statistics are per (instance, channel), so a CTA
// that owns all nodes of an instance (the M = 128 rows of the MMA tile, n <= 128) normalises in its epilogue.
//
// Precision follows the rollout kernel: fp32 operands are split into fp16 hi + lo and every product is three
// kind::f16 MMAs accumulated in fp32 in TMEM (~2^-22 relative), so the embeddings match the fp32 torch encoder
// to ~1e-6 and the rollout's decisions are unchanged up to true near-ties.
//
// Weights are packed once per forward into the shared-memory image of their MMA tiles (enc_pack_weight_kernel) and
// streamed with TMA bulk copies, double buffered; activations are converted by the CUDA cores on the way in.
#include "common.cuh"
#include "tc_ptx.cuh"

namespace r4 {
namespace enc {

using namespace r4::tc;

constexpr int kThreads = 256;
constexpr int kRows = 128;                  // MMA M: nodes of one instance, padded
constexpr uint32_t kLBO = 128;              // K-major, no swizzle: 8 rows x 16 B core matrices, K-chunks contiguous
constexpr uint32_t kSBO = 16 * kLBO;        // 16 chunks = 128 channels per 8-row group
constexpr uint32_t kHalf = (kRows / 8) * kSBO;   // 32 KB: the hi (or lo) half of a 128 x 128 tile
constexpr uint32_t kTile = 2 * kHalf;       // 64 KB
constexpr int kStage = 132;                 // fp32 staging row stride (floats): 16-byte row writes are conflict-free

enum { EPI_PLAIN = 0, EPI_RELU = 1, EPI_RES_NORM = 2 };

struct Smem {
    uint64_t tma_bar[2];
    uint64_t mma_bar[2];
    uint32_t tmem_ptr;
    uint32_t pad_;
    float mean[kRows];
    float rstd[kRows];
};
constexpr uint32_t kOffX = 0, kOffW = kTile, kOffS = 3 * kTile;
constexpr size_t kSmemBytes = kOffS + sizeof(Smem);

// ---- weights: W [n_out, K] fp32 (nn.Linear layout) -> tiles (kb, nb) of 128 x 128, hi then lo, in consumption order
__global__ void __launch_bounds__(256) enc_pack_weight_kernel(const float* __restrict__ W, int n_out, int K,
                                                              unsigned char* __restrict__ out) {
    const int nN = n_out / 128;
    const int t = blockIdx.x;                 // tile = kb * nN + nb
    const int kb = t / nN, nb = t % nN;
    unsigned char* o = out + (size_t)t * kTile;
    for (int it = threadIdx.x; it < 128 * 16; it += blockDim.x) {
        const int r = it >> 4, c = it & 15;
        const float* src = W + (size_t)(nb * 128 + r) * K + kb * 128 + 8 * c;
        const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
        uint32_t h[4], l[4];
        split2(a.x, a.y, h[0], l[0]);
        split2(a.z, a.w, h[1], l[1]);
        split2(b.x, b.y, h[2], l[2]);
        split2(b.z, b.w, h[3], l[3]);
        const uint32_t off = (uint32_t)(r >> 3) * kSBO + (uint32_t)c * kLBO + (uint32_t)(r & 7) * 16u;
        *reinterpret_cast<uint4*>(o + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(o + kHalf + off) = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

// ---- activations: rows [0,n) of X[b] (fp32, row stride K), channels [128 kb, 128 kb + 128) -> A tile hi/lo ------
// A warp pass covers 8 rows x 4 chunks: lanes write 512 contiguous bytes (conflict-free), read whole 32-byte sectors.
__device__ __forceinline__ void convert_x(const float* __restrict__ Xb, int n, int K, int kb, unsigned char* xs) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        float4 a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int v = (half * 4 + i) * 8 + warp;          // 64 warp passes over 8 warps
            const int r = 8 * (v >> 2) + (lane & 7), c = 4 * (v & 3) + (lane >> 3);
            a[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            b[i] = a[i];
            if (r < n) {
                const float* src = Xb + (size_t)r * K + kb * 128 + 8 * c;
                a[i] = *reinterpret_cast<const float4*>(src);
                b[i] = *reinterpret_cast<const float4*>(src + 4);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int v = (half * 4 + i) * 8 + warp;
            const int rg = v >> 2, c = 4 * (v & 3) + (lane >> 3);
            uint32_t h[4], l[4];
            split2(a[i].x, a[i].y, h[0], l[0]);
            split2(a[i].z, a[i].w, h[1], l[1]);
            split2(b[i].x, b[i].y, h[2], l[2]);
            split2(b[i].z, b[i].w, h[3], l[3]);
            const uint32_t off = (uint32_t)rg * kSBO + (uint32_t)c * kLBO + (uint32_t)(lane & 7) * 16u;
            *reinterpret_cast<uint4*>(xs + off) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(xs + kHalf + off) = make_uint4(l[0], l[1], l[2], l[3]);
        }
    }
}

// 128 x 128 x 128 block product into TMEM columns [dcol, dcol + 128): 8 k-steps x (hi*hi + hi*lo + lo*hi)
__device__ __forceinline__ void issue_tile(uint32_t tmem, uint32_t dcol, uint32_t xs, uint32_t ws, bool first) {
    const uint32_t id = idesc_f16(kRows, 128);
    uint64_t a_hi = smem_desc(xs, kLBO, kSBO), a_lo = smem_desc(xs + kHalf, kLBO, kSBO);
    uint64_t b_hi = smem_desc(ws, kLBO, kSBO), b_lo = smem_desc(ws + kHalf, kLBO, kSBO);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
        mma_ss(tmem + dcol, a_hi, b_hi, id, (first && ks == 0) ? 0u : 1u);
        mma_ss(tmem + dcol, a_hi, b_lo, id, 1u);
        mma_ss(tmem + dcol, a_lo, b_hi, id, 1u);
        a_hi += (2u * kLBO) >> 4;
        a_lo += (2u * kLBO) >> 4;
        b_hi += (2u * kLBO) >> 4;
        b_lo += (2u * kLBO) >> 4;
    }
}

// ---- Y[b] = epilogue(X[b] W^T + bias): one CTA per instance at a time, persistent over the batch ----------------
// Tiles are walked K-block outer, N-block inner (the order enc_pack_weight_kernel stores them); the accumulators of
// all N blocks stay in TMEM (n_out <= 512 columns), so X is read once.
template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
    enc_linear_kernel(const float* __restrict__ X, const unsigned char* __restrict__ Wp, const float* __restrict__ bias,
                      const float* __restrict__ R, const float* __restrict__ gamma, const float* __restrict__ beta,
                      float eps, float* __restrict__ Y, int64_t B, int n, int K, int n_out) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* xs = smem + kOffX;
    unsigned char* wsm = smem + kOffW;
    Smem* s = reinterpret_cast<Smem*>(smem + kOffS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s->tma_bar[i], 1);
            mbar_init(&s->mma_bar[i], 1);
        }
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(&s->tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s->tmem_ptr;
    const uint32_t xs_a = smem_u32(xs), ws_a = smem_u32(wsm);
    const int nK = K / 128, nN = n_out / 128, T = nK * nN;
    uint32_t use_tma[2] = {0u, 0u}, use_mma[2] = {0u, 0u};   // phases consumed per buffer (same in every thread)
    const int q = warp & 3, hlf = warp >> 2, row = 32 * q + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);

    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        const float* Xb = X + b * (int64_t)n * K;
        if (warp == 0) {
            if (elect_one()) {
                mbar_expect_tx(&s->tma_bar[0], kTile);
                bulk_g2s(wsm, Wp, kTile, &s->tma_bar[0]);
            }
            __syncwarp();
        }
        for (int t = 0; t < T; ++t) {
            const int kb = t / nN, nb = t - kb * nN, buf = t & 1;
            if (nb == 0) {
                if (kb > 0) {   // the MMAs of the previous K block still read the A tile
                    // (barrier first: a parity wait is only meaningful for a thread at most one phase behind, and
                    // the warps that issue nothing run ahead of the issuing warp)
                    __syncthreads();
                    mbar_wait(&s->mma_bar[buf ^ 1], (use_mma[buf ^ 1] - 1u) & 1u);
                    tc_fence_after();
                }
                convert_x(Xb, n, K, kb, xs);
                fence_proxy_async();
                __syncthreads();
            }
            if (warp == 0) {
                if (elect_one()) {
                    if (t + 1 < T) {   // next W tile into the other buffer once its previous reader (tile t-1) is done
                        if (t >= 1) mbar_wait(&s->mma_bar[buf ^ 1], (use_mma[buf ^ 1] - 1u) & 1u);
                        mbar_expect_tx(&s->tma_bar[buf ^ 1], kTile);
                        bulk_g2s(wsm + (size_t)(buf ^ 1) * kTile, Wp + (size_t)(t + 1) * kTile, kTile,
                                 &s->tma_bar[buf ^ 1]);
                    }
                    mbar_wait(&s->tma_bar[buf], use_tma[buf] & 1u);
                    tc_fence_after();
                    issue_tile(tmem, (uint32_t)nb * 128u, xs_a, ws_a + (uint32_t)buf * kTile, kb == 0);
                    tc_commit(&s->mma_bar[buf]);
                }
                __syncwarp();
            }
            use_tma[buf]++;
            use_mma[buf]++;
        }
        // every MMA of the instance is complete when the last commit arrives (barrier first, as above)
        __syncthreads();
        mbar_wait(&s->mma_bar[(T - 1) & 1], (use_mma[(T - 1) & 1] - 1u) & 1u);
        tc_fence_after();

        if (EPI != EPI_RES_NORM) {
            float* Yb = Y + b * (int64_t)n * n_out;
            for (int nb = 0; nb < nN; ++nb) {
#pragma unroll
                for (int cc = 0; cc < 64; cc += 16) {
                    const int col = nb * 128 + hlf * 64 + cc;
                    uint32_t v[16];
                    tmem_ld16(lane_base + (uint32_t)col, v);
                    tc_wait_ld();
                    if (row < n) {
#pragma unroll
                        for (int i = 0; i < 16; i += 4) {
                            const float4 bb = *reinterpret_cast<const float4*>(bias + col + i);
                            float4 o = make_float4(__uint_as_float(v[i]) + bb.x, __uint_as_float(v[i + 1]) + bb.y,
                                                   __uint_as_float(v[i + 2]) + bb.z, __uint_as_float(v[i + 3]) + bb.w);
                            if (EPI == EPI_RELU)
                                o = make_float4(fmaxf(o.x, 0.f), fmaxf(o.y, 0.f), fmaxf(o.z, 0.f), fmaxf(o.w, 0.f));
                            *reinterpret_cast<float4*>(Yb + (size_t)row * n_out + col + i) = o;
                        }
                    }
                }
            }
        } else {
            // residual + InstanceNorm over the node axis (nn/ops.py:48-49; biased variance, eps inside the sqrt)
            float* st = reinterpret_cast<float*>(wsm);   // [128][kStage] fp32: the W buffers are idle now
            const float* Rb = R + b * (int64_t)n * 128;
#pragma unroll
            for (int cc = 0; cc < 64; cc += 16) {
                const int col = hlf * 64 + cc;
                uint32_t v[16];
                tmem_ld16(lane_base + (uint32_t)col, v);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (row < n) {
                        const float4 bb = *reinterpret_cast<const float4*>(bias + col + i);
                        const float4 rr = *reinterpret_cast<const float4*>(Rb + (size_t)row * 128 + col + i);
                        o = make_float4(rr.x + (__uint_as_float(v[i]) + bb.x), rr.y + (__uint_as_float(v[i + 1]) + bb.y),
                                        rr.z + (__uint_as_float(v[i + 2]) + bb.z), rr.w + (__uint_as_float(v[i + 3]) + bb.w));
                    }
                    *reinterpret_cast<float4*>(st + row * kStage + col + i) = o;
                }
            }
            __syncthreads();
            if (threadIdx.x < 128) {
                const int c = threadIdx.x;
                float sum = 0.f;
                for (int r = 0; r < n; ++r) sum += st[r * kStage + c];
                const float mean = sum / (float)n;
                float sq = 0.f;
                for (int r = 0; r < n; ++r) {
                    const float d = st[r * kStage + c] - mean;
                    sq = fmaf(d, d, sq);
                }
                s->mean[c] = mean;
                s->rstd[c] = rsqrtf(sq / (float)n + eps);
            }
            __syncthreads();
            float* Yb = Y + b * (int64_t)n * 128;
            for (int idx = threadIdx.x; idx < n * 32; idx += kThreads) {
                const int r = idx >> 5, c4 = (idx & 31) * 4;
                const float4 xv = *reinterpret_cast<const float4*>(st + r * kStage + c4);
                const float4 g = *reinterpret_cast<const float4*>(gamma + c4);
                const float4 be = *reinterpret_cast<const float4*>(beta + c4);
                float4 o;
                o.x = __fadd_rn(__fmul_rn(__fmul_rn(xv.x - s->mean[c4], s->rstd[c4]), g.x), be.x);
                o.y = __fadd_rn(__fmul_rn(__fmul_rn(xv.y - s->mean[c4 + 1], s->rstd[c4 + 1]), g.y), be.y);
                o.z = __fadd_rn(__fmul_rn(__fmul_rn(xv.z - s->mean[c4 + 2], s->rstd[c4 + 2]), g.z), be.z);
                o.w = __fadd_rn(__fmul_rn(__fmul_rn(xv.w - s->mean[c4 + 3], s->rstd[c4 + 3]), g.w), be.w);
                *reinterpret_cast<float4*>(Yb + (size_t)r * 128 + c4) = o;
            }
            fence_proxy_async();   // the staging area is the next instance's TMA destination
        }
        tc_fence_before();
        __syncthreads();
    }
    if (warp == 0) tmem_dealloc(tmem, 512);
}

// ---- multi-head self-attention of one instance, fp32 on the CUDA cores (nn/attention.py:66-136) -------------------
// qkv [n, 384] = (q | k | v) x (8 heads x 16); out [n, 128] = heads concatenated (the input of out_proj).  No mask
// (the encoder attends over all nodes).  K and V of the instance sit in shared memory; a thread owns (row, head)
// tasks and runs an online softmax over the keys, four keys per maximum update.
constexpr int kAttnThreads = 256;

__global__ void __launch_bounds__(kAttnThreads) enc_attention_kernel(const float* __restrict__ qkv,
                                                                      float* __restrict__ out, int64_t B, int n) {
    extern __shared__ __align__(16) float kv[];    // K [n][128] then V [n][128]
    float* Ks = kv;
    float* Vs = kv + (size_t)n * 128;
    const float kScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16), exponentials in base 2
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        const float* base = qkv + b * (int64_t)n * 384;
        __syncthreads();
        for (int i = threadIdx.x; i < n * 32; i += kAttnThreads) {
            const int r = i >> 5, c4 = (i & 31) * 4;
            *reinterpret_cast<float4*>(Ks + r * 128 + c4) = *reinterpret_cast<const float4*>(base + (size_t)r * 384 + 128 + c4);
            *reinterpret_cast<float4*>(Vs + r * 128 + c4) = *reinterpret_cast<const float4*>(base + (size_t)r * 384 + 256 + c4);
        }
        __syncthreads();
        // lanes of a warp take consecutive rows of the SAME head: their K / V reads are warp-wide broadcasts
        for (int task = threadIdx.x; task < n * 8; task += kAttnThreads) {
            const int h = task / n, r = task - h * n;
            float q[16], acc[16];
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 t = *reinterpret_cast<const float4*>(base + (size_t)r * 384 + 16 * h + i);
                q[i] = t.x * kScale; q[i + 1] = t.y * kScale; q[i + 2] = t.z * kScale; q[i + 3] = t.w * kScale;
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = 0.f;
            float m = -INFINITY, l = 0.f;
            for (int j0 = 0; j0 < n; j0 += 4) {
                float sc[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int jj = min(j0 + u, n - 1);
                    const float* kr = Ks + jj * 128 + 16 * h;
                    float d0 = 0.f, d1 = 0.f;
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 t = *reinterpret_cast<const float4*>(kr + i);
                        d0 = fmaf(q[i], t.x, d0);
                        d1 = fmaf(q[i + 1], t.y, d1);
                        d0 = fmaf(q[i + 2], t.z, d0);
                        d1 = fmaf(q[i + 3], t.w, d1);
                    }
                    sc[u] = (j0 + u < n) ? d0 + d1 : -INFINITY;
                }
                const float mn = fmaxf(fmaxf(m, fmaxf(sc[0], sc[1])), fmaxf(sc[2], sc[3]));
                const float corr = ex2_approx(m - mn);     // exp2(-inf) = 0 on the first block
                m = mn;
                l *= corr;
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] *= corr;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float p = ex2_approx(sc[u] - mn);
                    l += p;
                    const float* vr = Vs + min(j0 + u, n - 1) * 128 + 16 * h;
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 t = *reinterpret_cast<const float4*>(vr + i);
                        acc[i] = fmaf(p, t.x, acc[i]);
                        acc[i + 1] = fmaf(p, t.y, acc[i + 1]);
                        acc[i + 2] = fmaf(p, t.z, acc[i + 2]);
                        acc[i + 3] = fmaf(p, t.w, acc[i + 3]);
                    }
                }
            }
            const float inv = 1.0f / l;
            float* o = out + (b * (int64_t)n + r) * 128 + 16 * h;
#pragma unroll
            for (int i = 0; i < 16; i += 4)
                *reinterpret_cast<float4*>(o + i) = make_float4(acc[i] * inv, acc[i + 1] * inv, acc[i + 2] * inv, acc[i + 3] * inv);
        }
    }
}

}  // namespace enc
}  // namespace r4

// ---- C ABI ---------------------------------------------------------------------------------------------------------
using namespace r4;

extern "C" int64_t rl4co_enc_packed_bytes(int32_t n_out, int32_t k) {
    if (n_out <= 0 || k <= 0 || n_out % 128 || k % 128) return 0;
    return (int64_t)(n_out / 128) * (k / 128) * (int64_t)enc::kTile;
}

extern "C" int rl4co_enc_pack_weight(const float* W, int32_t n_out, int32_t k, void* packed, void* stream) {
    R4_CHECK_ARG(W && packed && n_out > 0 && k > 0 && n_out % 128 == 0 && k % 128 == 0,
                 "enc_pack_weight: n_out and k must be positive multiples of 128 (got %d, %d)", n_out, k);
    enc::enc_pack_weight_kernel<<<(n_out / 128) * (k / 128), 256, 0, (cudaStream_t)stream>>>(W, n_out, k, (unsigned char*)packed);
    R4_CHECK_LAUNCH("enc_pack_weight");
    return RL4CO_OK;
}

template <int EPI>
static int launch_linear(const float* X, const void* Wp, const float* bias, const float* R, const float* gamma,
                         const float* beta, float eps, float* Y, int64_t B, int n, int K, int n_out,
                         cudaStream_t stream) {
    auto kern = enc::enc_linear_kernel<EPI>;
    static bool prepared = false;
    if (!prepared) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)enc::kSmemBytes) != cudaSuccess) {
            set_error("enc_linear: cannot reserve %zu bytes of shared memory", enc::kSmemBytes);
            return RL4CO_E_LAUNCH;
        }
        prepared = true;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)(B < sms ? B : sms);
    kern<<<grid, enc::kThreads, enc::kSmemBytes, stream>>>(X, (const unsigned char*)Wp, bias, R, gamma, beta, eps, Y, B,
                                                           n, K, n_out);
    R4_CHECK_LAUNCH("enc_linear");
    return RL4CO_OK;
}

extern "C" int rl4co_enc_linear(const float* X, const void* Wpacked, const float* bias, const float* residual,
                                const float* gamma, const float* beta, float eps, float* Y, int64_t B, int32_t n,
                                int32_t k, int32_t n_out, int32_t epilogue, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    R4_CHECK_ARG(X && Wpacked && bias && Y && B >= 0, "enc_linear: null pointer");
    R4_CHECK_ARG(n >= 1 && n <= enc::kRows, "enc_linear: 1 <= nodes <= 128 (got %d)", n);
    R4_CHECK_ARG(k > 0 && k % 128 == 0 && n_out > 0 && n_out % 128 == 0 && n_out <= 512,
                 "enc_linear: k and n_out must be multiples of 128, n_out <= 512 (got %d, %d)", k, n_out);
    if (B == 0) return RL4CO_OK;
    switch (epilogue) {
        case enc::EPI_PLAIN:
            return launch_linear<enc::EPI_PLAIN>(X, Wpacked, bias, nullptr, nullptr, nullptr, 0.f, Y, B, n, k, n_out, stream);
        case enc::EPI_RELU:
            return launch_linear<enc::EPI_RELU>(X, Wpacked, bias, nullptr, nullptr, nullptr, 0.f, Y, B, n, k, n_out, stream);
        case enc::EPI_RES_NORM:
            R4_CHECK_ARG(residual && gamma && beta && n_out == 128,
                         "enc_linear: the residual + instance-norm epilogue needs residual, gamma, beta and n_out = 128");
            return launch_linear<enc::EPI_RES_NORM>(X, Wpacked, bias, residual, gamma, beta, eps, Y, B, n, k, n_out, stream);
        default:
            R4_CHECK_ARG(false, "enc_linear: unknown epilogue %d", epilogue);
    }
    return RL4CO_OK;
}

extern "C" int rl4co_enc_attention(const float* qkv, float* out, int64_t B, int32_t n, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    R4_CHECK_ARG(qkv && out && B >= 0 && n >= 1 && n <= enc::kRows, "enc_attention: bad args (nodes = %d)", n);
    if (B == 0) return RL4CO_OK;
    const size_t smem = (size_t)n * 128 * 2 * sizeof(float);
    static bool prepared = false;
    if (!prepared) {
        if (cudaFuncSetAttribute(enc::enc_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 128 * 128 * 2 * (int)sizeof(float)) != cudaSuccess) {
            set_error("enc_attention: cannot reserve shared memory");
            return RL4CO_E_LAUNCH;
        }
        prepared = true;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t want = 2LL * sms;
    const unsigned grid = (unsigned)(B < want ? B : want);
    enc::enc_attention_kernel<<<grid, enc::kAttnThreads, smem, stream>>>(qkv, out, B, n);
    R4_CHECK_LAUNCH("enc_attention");
    return RL4CO_OK;
}
