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
constexpr int kRing = 36;                   // store slot row stride: 32 output columns + 16 B pad (same bank argument)

// EPI_RES_AFFINE: Y = (R + X W^T + b) * gamma + beta with per-channel gamma / beta = eval-mode BatchNorm1d folded by the
// host (nn/ops.py:46-47 with running statistics: gamma = w / sqrt(var + eps), beta = b - mean * gamma)
enum { EPI_PLAIN = 0, EPI_RELU = 1, EPI_RES_NORM = 2, EPI_RES_AFFINE = 3 };

struct Smem {
    uint64_t tma_bar[2];
    uint64_t mma_bar[2];
    uint64_t kdone_bar;    // the MMAs of a K block are complete (K > 128 only): its A tile may be overwritten
    uint64_t done_bar;     // all MMAs of the instance are complete
    uint32_t tmem_ptr;
    uint32_t pad_;
    int next_chunk[2][4];  // [instance parity][TMEM lane quadrant]: next 32-column chunk of the epilogue to drain
    alignas(16) float mean[kRows];
    alignas(16) float rstd[kRows];
    alignas(16) float part[2][16 * kRows];  // per-warp partial column sums (mean pass, variance pass), <= 16 warps
};
constexpr uint32_t kOffX = 0, kOffW = kTile, kOffS = 3 * kTile;
constexpr size_t kSmemBytes = kOffS + sizeof(Smem);

// one 64 KB W tile as R4_TMA_SPLIT bulk copies counted on the same mbarrier (expect_tx covers the whole tile)
#ifndef R4_TMA_SPLIT
#define R4_TMA_SPLIT 1
#endif
__device__ __forceinline__ void bulk_tile(void* dst, const void* src, uint64_t* bar) {
    constexpr uint32_t piece = kTile / R4_TMA_SPLIT;
#pragma unroll
    for (int i = 0; i < R4_TMA_SPLIT; ++i)
        bulk_g2s(reinterpret_cast<unsigned char*>(dst) + i * piece, reinterpret_cast<const unsigned char*>(src) + i * piece,
                 piece, bar);
}

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
// Run by the four producer warps (4..7).  A warp pass covers 8 rows x 4 chunks: lanes write 512 contiguous bytes
// (conflict-free) and read whole 32-byte sectors.
__device__ __forceinline__ void convert_x(const float* __restrict__ Xb, int n, int K, int kb, unsigned char* xs) {
    const int pw = (threadIdx.x >> 5) - 4, lane = threadIdx.x & 31;
#pragma unroll
    for (int part = 0; part < 4; ++part) {
        float4 a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int v = (part * 4 + i) * 4 + pw;            // 64 warp passes over 4 warps
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
            const int v = (part * 4 + i) * 4 + pw;
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

// The same conversion by all eight warps, split into its global loads (64 registers per thread, issued while the
// previous K block's MMAs still run) and the split + shared-memory stores (once those MMAs are done with the tile).
__device__ __forceinline__ void load_x8(const float* __restrict__ Xb, int n, int K, int kb, float4 (&a)[8], float4 (&b)[8]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int v = i * 8 + warp;                           // 64 warp passes over 8 warps
        const int r = 8 * (v >> 2) + (lane & 7), c = 4 * (v & 3) + (lane >> 3);
        a[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        b[i] = a[i];
        if (r < n) {
            const float* src = Xb + (size_t)r * K + kb * 128 + 8 * c;
            a[i] = *reinterpret_cast<const float4*>(src);
            b[i] = *reinterpret_cast<const float4*>(src + 4);
        }
    }
}
__device__ __forceinline__ void store_x8(const float4 (&a)[8], const float4 (&b)[8], unsigned char* xs) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int v = i * 8 + warp;
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


// ---- residual + norm epilogue of a 128-column accumulator (TMEM columns [0,128) of `lane_base`) -------------------
// EPI_RES_NORM: residual + InstanceNorm over the node axis (nn/ops.py:48-49; biased variance, eps inside the sqrt);
// EPI_RES_AFFINE: residual + per-channel affine (eval-mode BatchNorm folded by the host).  Run by the first NT threads
// of the CTA (NT = 128, 256 or 512: warp w drains the columns [CW (w / 4), +CW) of TMEM lane quadrant w % 4,
// CW = 128 / (NT / 128)); `stage` is a [128][kStage] fp32 area (idle operand buffers); `part` holds 2 x (NT / 32) x 128
// floats.  `lane_base` = accumulator base with the caller's quadrant lane offset.
template <int EPI, int NT>
__device__ __forceinline__ void epilogue_res_norm(uint32_t lane_base, unsigned char* stage, float* part, float* mean_s,
                                                  float* rstd_s, const float* __restrict__ bias,
                                                  const float* __restrict__ Rb, const float* __restrict__ gamma,
                                                  const float* __restrict__ beta, float eps, float* __restrict__ Yb,
                                                  int n) {
    constexpr int NW = NT / 32, CW = 128 / (NT / 128);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = 32 * (warp & 3) + lane, c0 = CW * (warp >> 2);
    float* st = reinterpret_cast<float*>(stage);
#pragma unroll 2
    for (int col = c0; col < c0 + CW; col += 16) {
        uint32_t v[16];
        tmem_ld16(lane_base + (uint32_t)col, v);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
            const float4 bb = *reinterpret_cast<const float4*>(bias + col + i);
            *reinterpret_cast<float4*>(st + row * kStage + col + i) =
                make_float4(__uint_as_float(v[i]) + bb.x, __uint_as_float(v[i + 1]) + bb.y,
                            __uint_as_float(v[i + 2]) + bb.z, __uint_as_float(v[i + 3]) + bb.w);
        }
    }
    bar_sync(1, NT);
    if (EPI == EPI_RES_NORM) {
        // column statistics, two passes (mean, then biased variance).  A warp takes every NW-th node row and a lane
        // four channels: the residual rows are read as coalesced 512-byte segments, all of a warp's loads are in
        // flight together, and the NW partial sums meet in shared memory.
        const int c4 = 4 * lane;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int r = warp; r < n; r += NW) {
            const float4 rr = *reinterpret_cast<const float4*>(Rb + (size_t)r * 128 + c4);
            float4 v = *reinterpret_cast<const float4*>(st + r * kStage + c4);
            v = make_float4(rr.x + v.x, rr.y + v.y, rr.z + v.z, rr.w + v.w);
            *reinterpret_cast<float4*>(st + r * kStage + c4) = v;
            acc = make_float4(acc.x + v.x, acc.y + v.y, acc.z + v.z, acc.w + v.w);
        }
        *reinterpret_cast<float4*>(part + warp * 128 + c4) = acc;
        bar_sync(1, NT);
        const int c = threadIdx.x;
        if (c < 128) {
            float sum = 0.f;
#pragma unroll
            for (int w = 0; w < NW; ++w) sum += part[w * 128 + c];
            mean_s[c] = sum / (float)n;
        }
        bar_sync(1, NT);
        const float4 m4 = *reinterpret_cast<const float4*>(mean_s + c4);
        acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int r = warp; r < n; r += NW) {
            const float4 v = *reinterpret_cast<const float4*>(st + r * kStage + c4);
            const float dx = v.x - m4.x, dy = v.y - m4.y, dz = v.z - m4.z, dw = v.w - m4.w;
            acc = make_float4(fmaf(dx, dx, acc.x), fmaf(dy, dy, acc.y), fmaf(dz, dz, acc.z), fmaf(dw, dw, acc.w));
        }
        *reinterpret_cast<float4*>(part + (NW + warp) * 128 + c4) = acc;
        bar_sync(1, NT);
        if (c < 128) {
            float sq = 0.f;
#pragma unroll
            for (int w = 0; w < NW; ++w) sq += part[(NW + w) * 128 + c];
            rstd_s[c] = rsqrtf(sq / (float)n + eps);
        }
    }
    bar_sync(1, NT);
    for (int idx = threadIdx.x; idx < n * 32; idx += NT) {
        const int r = idx >> 5, c4 = (idx & 31) * 4;
        const float4 xv = *reinterpret_cast<const float4*>(st + r * kStage + c4);
        const float4 g = *reinterpret_cast<const float4*>(gamma + c4);
        const float4 be = *reinterpret_cast<const float4*>(beta + c4);
        float4 o;
        if (EPI == EPI_RES_AFFINE) {
            const float4 rr = *reinterpret_cast<const float4*>(Rb + (size_t)r * 128 + c4);
            o.x = __fadd_rn(__fmul_rn(rr.x + xv.x, g.x), be.x);
            o.y = __fadd_rn(__fmul_rn(rr.y + xv.y, g.y), be.y);
            o.z = __fadd_rn(__fmul_rn(rr.z + xv.z, g.z), be.z);
            o.w = __fadd_rn(__fmul_rn(rr.w + xv.w, g.w), be.w);
        } else {
            o.x = __fadd_rn(__fmul_rn(__fmul_rn(xv.x - mean_s[c4], rstd_s[c4]), g.x), be.x);
            o.y = __fadd_rn(__fmul_rn(__fmul_rn(xv.y - mean_s[c4 + 1], rstd_s[c4 + 1]), g.y), be.y);
            o.z = __fadd_rn(__fmul_rn(__fmul_rn(xv.z - mean_s[c4 + 2], rstd_s[c4 + 2]), g.z), be.z);
            o.w = __fadd_rn(__fmul_rn(__fmul_rn(xv.w - mean_s[c4 + 3], rstd_s[c4 + 3]), g.w), be.w);
        }
        *reinterpret_cast<float4*>(Yb + (size_t)r * 128 + c4) = o;
    }
    fence_proxy_async();   // the staging area is the next instance's TMA destination
}

#ifdef R4_ENC_PROF
__device__ long long g_enc_prof[16];
#define ENC_STAMP(k) do { if (blockIdx.x == 0 && threadIdx.x == PROF_T) { const long long now__ = clock64(); g_enc_prof[k] += now__ - tprev; tprev = now__; } } while (0)
#else
#define ENC_STAMP(k) do { } while (0)
#endif

// ---- Y[b] = epilogue(X[b] W^T + bias): one CTA per instance at a time, persistent over the batch ----------------
// Tiles are walked K-block outer, N-block inner (the order enc_pack_weight_kernel stores them); the accumulators of
// all N blocks stay in TMEM (n_out <= 512 columns), so X is read once.  Warps 0..3 are the epilogue warps (one TMEM
// lane quadrant each), warps 4..7 the producers: while the epilogue of instance b drains TMEM, the producers already
// convert the A tile of instance b+1, and one elected producer lane streams the W tiles (TMA) and issues the MMAs.
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
        mbar_init(&s->kdone_bar, 1);
        mbar_init(&s->done_bar, 1);
        for (int i = 0; i < 8; ++i) s->next_chunk[i >> 2][i & 3] = 0;
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
    const bool producer = warp >= 4;
    // phase counters (the same in every thread; tma/mma ones matter to the issuing lane only)
    uint32_t n_tma0 = 0, n_tma1 = 0, n_mma0 = 0, n_mma1 = 0, n_kdone = 0, n_done = 0;
    bool tile0_in_flight = false;   // issuing lane: W tile 0 of this instance was requested during the last epilogue
    const int row = 32 * (warp & 3) + lane;                 // TMEM lane = node row (warps w and w+4 share a quadrant)
    const uint32_t lane_base = tmem + ((uint32_t)(32 * (warp & 3)) << 16);

    if (producer && (int64_t)blockIdx.x < B) convert_x(X + (int64_t)blockIdx.x * n * K, n, K, 0, xs);
    fence_proxy_async();
    __syncthreads();

#ifdef R4_ENC_PROF
    constexpr int PROF_T = 0;
    long long tprev = clock64();
#endif
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        const float* Xb = X + b * (int64_t)n * K;
        for (int kb = 0; kb < nK; ++kb) {
            if (kb > 0) {   // next K block: its loads fly while the MMAs of the previous block still read the A tile
                float4 xa[8], xb[8];
                load_x8(Xb, n, K, kb, xa, xb);
                mbar_wait(&s->kdone_bar, n_kdone & 1u);
                ++n_kdone;
                tc_fence_after();
                store_x8(xa, xb, xs);
                fence_proxy_async();
                __syncthreads();
            }
            if (warp == 4) {
                if (elect_one()) {
                    for (int nb = 0; nb < nN; ++nb) {
                        const int t = kb * nN + nb, buf = t & 1;
                        if (t == 0 && !tile0_in_flight) {
                            mbar_expect_tx(&s->tma_bar[0], kTile);
                            bulk_tile(wsm, Wp, &s->tma_bar[0]);
                        }
                        if (t + 1 < T) {   // next W tile into the other buffer once its previous reader (tile t-1) is done
                            if (t >= 1) {
                                mbar_wait(&s->mma_bar[buf ^ 1], ((buf ? n_mma0 : n_mma1) - 1u) & 1u);
                            }
                            mbar_expect_tx(&s->tma_bar[buf ^ 1], kTile);
                            bulk_tile(wsm + (size_t)(buf ^ 1) * kTile, Wp + (size_t)(t + 1) * kTile, &s->tma_bar[buf ^ 1]);
                        }
                        mbar_wait(&s->tma_bar[buf], (buf ? n_tma1 : n_tma0) & 1u);
                        tc_fence_after();
                        issue_tile(tmem, (uint32_t)nb * 128u, xs_a, ws_a + (uint32_t)buf * kTile, kb == 0);
                        tc_commit(&s->mma_bar[buf]);
                        if (buf) { ++n_tma1; ++n_mma1; } else { ++n_tma0; ++n_mma0; }
                    }
                    tc_commit(kb + 1 < nK ? &s->kdone_bar : &s->done_bar);
                }
                __syncwarp();
            }
        }
        ENC_STAMP(0);
        mbar_wait(&s->done_bar, n_done & 1u);
        ++n_done;
        tc_fence_after();
        ENC_STAMP(1);

        if (producer) {
            // A tile of the next instance while the epilogue warps drain TMEM (all MMAs that read xs are complete);
            // its first W tile too when W buffer 0 is not part of the epilogue's staging area
            const int64_t bn = b + gridDim.x;
            if (EPI < EPI_RES_NORM && warp == 4) {
                if (elect_one()) {
                    tile0_in_flight = bn < B;
                    if (tile0_in_flight) {
                        mbar_expect_tx(&s->tma_bar[0], kTile);
                        bulk_tile(wsm, Wp, &s->tma_bar[0]);
                    }
                }
                __syncwarp();
            }
            if (bn < B) convert_x(X + bn * (int64_t)n * K, n, K, 0, xs);
            fence_proxy_async();
        }
        if (EPI < EPI_RES_NORM) {
            // TMEM -> registers -> (+bias, relu) -> the warp's staging slot (thread = row) -> coalesced stores (8 lanes
            // per 128-byte row segment).  Lane-per-row stores straight from registers cost one L1 transaction per
            // 16 bytes and made this epilogue 70 % of the kernel; through the slot it is one per 128 bytes.  A warp
            // only reads rows its own lanes wrote: __syncwarp is the only synchronisation.  The phase is latency
            // bound, so both warps of a lane quadrant pull 32-column chunks from a shared counter (the producer warp
            // joins when its conversion is done).
            float* slot = reinterpret_cast<float*>(wsm + kTile) + warp * 32 * kRing;
            float* Yb = Y + b * (int64_t)n * n_out;
            const int sub = lane >> 3, piece = lane & 7, q = warp & 3;
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&s->next_chunk[n_done & 1u][q], 1);
                c = __shfl_sync(kFull, c, 0);
                const int c32 = 32 * c;
                if (c32 >= n_out) break;
                uint32_t v0[16], v1[16];
                tmem_ld16(lane_base + (uint32_t)c32, v0);
                tmem_ld16(lane_base + (uint32_t)c32 + 16u, v1);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    const float4 b0 = *reinterpret_cast<const float4*>(bias + c32 + i);
                    const float4 b1 = *reinterpret_cast<const float4*>(bias + c32 + 16 + i);
                    float4 o0 = make_float4(__uint_as_float(v0[i]) + b0.x, __uint_as_float(v0[i + 1]) + b0.y,
                                            __uint_as_float(v0[i + 2]) + b0.z, __uint_as_float(v0[i + 3]) + b0.w);
                    float4 o1 = make_float4(__uint_as_float(v1[i]) + b1.x, __uint_as_float(v1[i + 1]) + b1.y,
                                            __uint_as_float(v1[i + 2]) + b1.z, __uint_as_float(v1[i + 3]) + b1.w);
                    if (EPI == EPI_RELU) {
                        o0 = make_float4(fmaxf(o0.x, 0.f), fmaxf(o0.y, 0.f), fmaxf(o0.z, 0.f), fmaxf(o0.w, 0.f));
                        o1 = make_float4(fmaxf(o1.x, 0.f), fmaxf(o1.y, 0.f), fmaxf(o1.z, 0.f), fmaxf(o1.w, 0.f));
                    }
                    *reinterpret_cast<float4*>(slot + lane * kRing + i) = o0;
                    *reinterpret_cast<float4*>(slot + lane * kRing + 16 + i) = o1;
                }
                __syncwarp();
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int r = 4 * it + sub, grow = 32 * q + r;
                    const float4 o = *reinterpret_cast<const float4*>(slot + r * kRing + 4 * piece);
                    if (grow < n) *reinterpret_cast<float4*>(Yb + (size_t)grow * n_out + c32 + 4 * piece) = o;
                }
                __syncwarp();
            }
        } else if (!producer) {
            epilogue_res_norm<EPI, 128>(lane_base, wsm, s->part[0], s->mean, s->rstd, bias, R + b * (int64_t)n * 128, gamma, beta,
                                   eps, Y + b * (int64_t)n * 128, n);
        }
        ENC_STAMP(2);
        // the other parity's counters were last used one instance ago: reset them for the next one
        if (threadIdx.x < 4) s->next_chunk[(n_done & 1u) ^ 1u][threadIdx.x] = 0;
        tc_fence_before();
        __syncthreads();
        ENC_STAMP(3);
    }
    if (warp == 0) tmem_dealloc(tmem, 512);
}


// ---- fused feed-forward block: Y[b] = Norm(X[b] + relu(X[b] W1^T + b1) W2^T + b2) (attnnet.py:50-60, nn/mlp.py) -----
// The hidden activations (n x 512 fp32 = 205 KB per instance, written and read back by the two-kernel form) never
// leave the SM.  Per 128-wide hidden block k:  M1(k): acc1[k&1] = X W1[k]^T  (24 MMAs, operands in shared memory);
// the four epilogue warps drain acc1, add b1, apply relu, split to fp16 hi/lo and write the block back IN PLACE as a
// TS-form A operand (16-column chunk c: hi pairs in [c, c+8), lo pairs in [c+8, c+16));  M2(k): out += A W2[:, k]^T
// (24 MMAs, A from TMEM).  The MMAs are issued in the order M1(0) M1(1) M2(0) M1(2) M2(1) ... : the tensor pipe works on
// M1(k+1) while block k is converted, block k+1 is converted (other buffer) while M2(k) runs, and M1(k+2) overwrites
// acc1[k & 1] behind M2(k) in the pipe's issue order.  W tiles stream through the two 64 KB buffers in that same order.
// TMEM: acc1 / A x 2 [0,256), out [256,384).  The final residual + norm epilogue is the one of enc_linear_kernel.
//
// Measured and NOT kept (profiles/r02_enc_ffn_phases.md): eight conversion warps + a ninth issuing warp, the X tile as a
// TS-form A operand in TMEM (first GEMM without shared-memory A reads), a per-CTA rotation of the hidden-block order (L2
// hot-spot test) - each left the kernel at 0.150 ms: it is bound by the 512 KB of W tiles every instance streams from L2
// (~20 B/clk per SM with all SMs streaming, 9.5 k of 41 k cycles per instance exposed as W-tile waits).
struct FfnSmem {
    uint64_t tma_bar[2];     // W tile landed in buffer i
    uint64_t wfree_bar[2];   // the MMAs reading W buffer i are complete
    uint64_t acc_full[2];    // M1 into acc1[i] is complete
    uint64_t a_ready[2];     // the epilogue warps have converted acc1[i] in place into the A operand   (128 arrivals)
    uint64_t done_bar;       // every MMA of the instance is complete
    uint32_t tmem_ptr;
    uint32_t pad_;
    alignas(16) float mean[kRows];
    alignas(16) float rstd[kRows];
    alignas(16) float part[2][16 * kRows];
};
constexpr size_t kFfnSmemBytes = kOffS + sizeof(FfnSmem);

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// out (+)= A[tmem] W2tile^T : 8 k-steps x (hi*hi + hi*lo + lo*hi), A = 128 channels as fp16 pairs (64 + 64 columns)
// A_STEP = TMEM columns between consecutive k-steps of the A operand: 8 (hi block [0,64) then lo block [64,128)) or 16
// (k-step ks: hi at 16 ks, lo at 16 ks + 8 - the layout an accumulator converted IN PLACE has)
template <uint32_t A_STEP = 8>
__device__ __forceinline__ void issue_tile_ts(uint32_t dcol_addr, uint32_t a_hi, uint32_t a_lo, uint32_t ws, bool first) {
    const uint32_t id = idesc_f16(kRows, 128);
    uint64_t b_hi = smem_desc(ws, kLBO, kSBO), b_lo = smem_desc(ws + kHalf, kLBO, kSBO);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
        mma_ts(dcol_addr, a_hi + A_STEP * ks, b_hi, id, (first && ks == 0) ? 0u : 1u);
        mma_ts(dcol_addr, a_hi + A_STEP * ks, b_lo, id, 1u);
        mma_ts(dcol_addr, a_lo + A_STEP * ks, b_hi, id, 1u);
        b_hi += (2u * kLBO) >> 4;
        b_lo += (2u * kLBO) >> 4;
    }
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
    enc_ffn_kernel(const float* __restrict__ X, const unsigned char* __restrict__ W1p, const float* __restrict__ b1,
                   const unsigned char* __restrict__ W2p, const float* __restrict__ b2, const float* __restrict__ gamma,
                   const float* __restrict__ beta, float eps, float* __restrict__ Y, int64_t B, int n, int nH) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* xs = smem + kOffX;
    unsigned char* wsm = smem + kOffW;
    FfnSmem* s = reinterpret_cast<FfnSmem*>(smem + kOffS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s->tma_bar[i], 1);
            mbar_init(&s->wfree_bar[i], 1);
            mbar_init(&s->acc_full[i], 1);
            mbar_init(&s->a_ready[i], 128);
        }
        mbar_init(&s->done_bar, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(&s->tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s->tmem_ptr;
    const uint32_t xs_a = smem_u32(xs), ws_a = smem_u32(wsm);
    const bool producer = warp >= 4;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
    constexpr uint32_t kColOut = 256;   // [0,256): the two M1 accumulators, each converted in place into an A operand
    const int nSeq = 2 * nH;
    // completed-phase counters (identical in every thread; each role only uses its own)
    uint32_t n_tma[2] = {0, 0}, n_wfree[2] = {0, 0}, n_full[2] = {0, 0}, n_aready[2] = {0, 0};
    uint32_t n_done = 0;
#ifdef R4_ENC_PROF
    // g_enc_prof: [0] issuer: wait W tile, [1] wait acc_free, [2] wait a_ready, [3] wait wfree, [4] issue, [5] tail waits,
    // [6] wait done (all threads, thread 0), [7] final epilogue + conversion, [8] epilogue warps: wait acc_full,
    // [9] wait a_free, [10] drain/convert/store
    long long tp = clock64();
#define FFN_STAMP(k) do { if (blockIdx.x == 0) { const long long now__ = clock64(); g_enc_prof[k] += now__ - tp; tp = now__; } } while (0)
#else
#define FFN_STAMP(k) do { } while (0)
#endif

    if (producer && (int64_t)blockIdx.x < B) convert_x(X + (int64_t)blockIdx.x * n * 128, n, 128, 0, xs);
    fence_proxy_async();
    __syncthreads();

    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        if (warp == 4) {
            if (elect_one()) {
                // W tile of sequence position sq: even positions after 0 alternate M1(k) / M2(k-1)
                auto tile_src = [&](int sq) -> const unsigned char* {
                    if (sq == 0) return W1p;
                    if (sq == nSeq - 1) return W2p + (size_t)(nH - 1) * kTile;
                    return (sq & 1) ? W1p + (size_t)((sq + 1) >> 1) * kTile : W2p + (size_t)((sq >> 1) - 1) * kTile;
                };
                if (b == (int64_t)blockIdx.x) {               // later instances: requested at the end of the previous one
                    mbar_expect_tx(&s->tma_bar[0], kTile);
                    bulk_tile(wsm, tile_src(0), &s->tma_bar[0]);
                }
                FFN_STAMP(4);
                for (int sq = 0; sq < nSeq; ++sq) {
                    const int buf = sq & 1;
                    mbar_wait(&s->tma_bar[buf], n_tma[buf] & 1u);
                    ++n_tma[buf];
                    FFN_STAMP(0);
                    const bool is_m1 = (sq == 0) || ((sq & 1) && sq != nSeq - 1);
                    if (is_m1) {
                        // acc1[k & 1] held the A operand of M2(k-2), issued before this M1: the tensor pipe keeps the order
                        const int k = (sq == 0) ? 0 : (sq + 1) >> 1;
                        FFN_STAMP(1);
                        tc_fence_after();
                        issue_tile(tmem, (uint32_t)(k & 1) * 128u, xs_a, ws_a + (uint32_t)buf * kTile, true);
                        tc_commit(&s->acc_full[k & 1]);
                    } else {
                        const int k = (sq == nSeq - 1) ? nH - 1 : (sq >> 1) - 1;
                        mbar_wait(&s->a_ready[k & 1], n_aready[k & 1] & 1u);
                        ++n_aready[k & 1];
                        FFN_STAMP(2);
                        tc_fence_after();
                        issue_tile_ts<16>(tmem + kColOut, tmem + (uint32_t)(k & 1) * 128u, tmem + (uint32_t)(k & 1) * 128u + 8u,
                                          ws_a + (uint32_t)buf * kTile, k == 0);
                        FFN_STAMP(6);                    // issue of an M2 (A from TMEM); [4] then holds the M1 issues
                    }
                    tc_commit(sq + 1 < nSeq ? &s->wfree_bar[buf] : &s->done_bar);
                    FFN_STAMP(4);
                    // the next tile goes into the other buffer once ITS previous reader (position sq-1, queued ahead of
                    // the MMAs just issued) is complete; the copy then runs under the MMAs of position sq
                    if (sq + 1 < nSeq) {
                        if (sq >= 1) {
                            mbar_wait(&s->wfree_bar[buf ^ 1], n_wfree[buf ^ 1] & 1u);
                            ++n_wfree[buf ^ 1];
                        }
                        FFN_STAMP(3);
                        mbar_expect_tx(&s->tma_bar[buf ^ 1], kTile);
                        bulk_tile(wsm + (size_t)(buf ^ 1) * kTile, tile_src(sq + 1), &s->tma_bar[buf ^ 1]);
                    }
                }
                // the buffer commits nobody waited for, and the last two drains: keep waits == arrivals per barrier
                // (by now they are all complete or about to be; the instance ends at done_bar)
                mbar_wait(&s->wfree_bar[(nSeq - 2) & 1], n_wfree[(nSeq - 2) & 1] & 1u);
                ++n_wfree[(nSeq - 2) & 1];
                FFN_STAMP(5);
            }
            __syncwarp();
        } else if (!producer) {
            // the four epilogue warps: hidden block k -> + b1 -> relu -> fp16 hi/lo A operand
            for (int k = 0; k < nH; ++k) {
                mbar_wait(&s->acc_full[k & 1], n_full[k & 1] & 1u);
                ++n_full[k & 1];
                if (threadIdx.x == 0) { FFN_STAMP(8); }
                tc_fence_after();
                // in place: chunk c (16 fp32 columns) -> hi pairs in [c, c+8), lo pairs in [c+8, c+16); M2(k) then reads
                // this buffer while block k+1 is converted in the other one
                const uint32_t src = lane_base + (uint32_t)(k & 1) * 128u;
#pragma unroll 2
                for (int c = 0; c < 128; c += 16) {
                    uint32_t v[16], hi[8], lo[8];
                    tmem_ld16(src + (uint32_t)c, v);
                    tc_wait_ld();
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 bb = *reinterpret_cast<const float4*>(b1 + 128 * k + c + i);
                        const float f0 = fmaxf(__uint_as_float(v[i]) + bb.x, 0.f);
                        const float f1 = fmaxf(__uint_as_float(v[i + 1]) + bb.y, 0.f);
                        const float f2 = fmaxf(__uint_as_float(v[i + 2]) + bb.z, 0.f);
                        const float f3 = fmaxf(__uint_as_float(v[i + 3]) + bb.w, 0.f);
                        split2(f0, f1, hi[i >> 1], lo[i >> 1]);
                        split2(f2, f3, hi[(i >> 1) + 1], lo[(i >> 1) + 1]);
                    }
                    tmem_st8(src + (uint32_t)c, hi);
                    tmem_st8(src + (uint32_t)c + 8u, lo);
                }
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(&s->a_ready[k & 1]);
                if (threadIdx.x == 0) { FFN_STAMP(10); }
            }
        }
        mbar_wait(&s->done_bar, n_done & 1u);
        ++n_done;
        tc_fence_after();
#ifdef R4_ENC_PROF
        if (threadIdx.x == 0 || threadIdx.x == 128) tp = clock64();
#endif
        // residual + norm by all eight warps, then the next instance's A tile by all eight (every MMA reading xs is done);
        // the first W tile of the next instance is requested first so that its flight covers the conversion
        epilogue_res_norm<EPI, kThreads>(lane_base + kColOut, wsm, s->part[0], s->mean, s->rstd, b2,
                                         X + b * (int64_t)n * 128, gamma, beta, eps, Y + b * (int64_t)n * 128, n);
        tc_fence_before();
        __syncthreads();
        const int64_t bn = b + gridDim.x;
        if (bn < B) {
            if (warp == 4) {
                if (elect_one()) {
                    mbar_expect_tx(&s->tma_bar[0], kTile);
                    bulk_tile(wsm, W1p, &s->tma_bar[0]);
                }
                __syncwarp();
            }
            float4 xa[8], xb[8];
            load_x8(X + bn * (int64_t)n * 128, n, 128, 0, xa, xb);
            store_x8(xa, xb, xs);
            fence_proxy_async();
        }
        __syncthreads();
#ifdef R4_ENC_PROF
        if (threadIdx.x == 0) { FFN_STAMP(7); }
        if (threadIdx.x == 128 && blockIdx.x == 0) tp = clock64();
#endif
    }
    if (warp == 0) tmem_dealloc(tmem, 512);
}

// ---- multi-head self-attention of one instance, fp32 on the CUDA cores (nn/attention.py:66-136) -------------------
// qkv [n, 384] = (q | k | v) x (8 heads x 16); out [n, 128] = heads concatenated (the input of out_proj).  No mask
// (the encoder attends over all nodes).  K and V of the instance sit in shared memory; a thread owns (head, two
// rows) tasks and runs an online softmax over the keys, four keys per maximum update.  The kernel is bound by the
// shared-memory pipe (a 128-bit broadcast load still costs four passes), hence two query rows per K / V load.
constexpr int kAttnMaxThreads = 512;
constexpr int kRPT = 2;   // query rows per thread

__global__ void __launch_bounds__(kAttnMaxThreads) enc_attention_kernel(const float* __restrict__ qkv,
                                                                         float* __restrict__ out, int64_t B, int n) {
    extern __shared__ __align__(16) float kv[];    // K [n][128] then V [n][128]
    float* Ks = kv;
    float* Vs = kv + (size_t)n * 128;
    const float kScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16), exponentials in base 2
    const int npair = (n + kRPT - 1) / kRPT;
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        const float* base = qkv + b * (int64_t)n * 384;
        __syncthreads();
        for (int i = threadIdx.x; i < n * 32; i += blockDim.x) {
            const int r = i >> 5, c4 = (i & 31) * 4;
            *reinterpret_cast<float4*>(Ks + r * 128 + c4) = *reinterpret_cast<const float4*>(base + (size_t)r * 384 + 128 + c4);
            *reinterpret_cast<float4*>(Vs + r * 128 + c4) = *reinterpret_cast<const float4*>(base + (size_t)r * 384 + 256 + c4);
        }
        __syncthreads();
        // lanes of a warp take consecutive row pairs of the SAME head: their K / V reads are warp-wide broadcasts
        for (int task = threadIdx.x; task < npair * 8; task += blockDim.x) {
            const int h = task / npair, r0 = kRPT * (task - h * npair);
            float q[kRPT][16], acc[kRPT][16], m[kRPT], l[kRPT];
#pragma unroll
            for (int a = 0; a < kRPT; ++a) {
                const int r = min(r0 + a, n - 1);
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    const float4 t = *reinterpret_cast<const float4*>(base + (size_t)r * 384 + 16 * h + i);
                    q[a][i] = t.x * kScale; q[a][i + 1] = t.y * kScale; q[a][i + 2] = t.z * kScale; q[a][i + 3] = t.w * kScale;
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[a][i] = 0.f;
                m[a] = -INFINITY;
                l[a] = 0.f;
            }
            for (int j0 = 0; j0 < n; j0 += 4) {
                float sc[kRPT][4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float* kr = Ks + min(j0 + u, n - 1) * 128 + 16 * h;
                    float d[kRPT][2];
#pragma unroll
                    for (int a = 0; a < kRPT; ++a) d[a][0] = d[a][1] = 0.f;
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 t = *reinterpret_cast<const float4*>(kr + i);
#pragma unroll
                        for (int a = 0; a < kRPT; ++a) {
                            d[a][0] = fmaf(q[a][i], t.x, d[a][0]);
                            d[a][1] = fmaf(q[a][i + 1], t.y, d[a][1]);
                            d[a][0] = fmaf(q[a][i + 2], t.z, d[a][0]);
                            d[a][1] = fmaf(q[a][i + 3], t.w, d[a][1]);
                        }
                    }
#pragma unroll
                    for (int a = 0; a < kRPT; ++a) sc[a][u] = (j0 + u < n) ? d[a][0] + d[a][1] : -INFINITY;
                }
                float p[kRPT][4];
#pragma unroll
                for (int a = 0; a < kRPT; ++a) {
                    const float mn = fmaxf(fmaxf(m[a], fmaxf(sc[a][0], sc[a][1])), fmaxf(sc[a][2], sc[a][3]));
                    const float corr = ex2_approx(m[a] - mn);     // exp2(-inf) = 0 on the first block
                    m[a] = mn;
                    l[a] *= corr;
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc[a][i] *= corr;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        p[a][u] = ex2_approx(sc[a][u] - mn);
                        l[a] += p[a][u];
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float* vr = Vs + min(j0 + u, n - 1) * 128 + 16 * h;
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 t = *reinterpret_cast<const float4*>(vr + i);
#pragma unroll
                        for (int a = 0; a < kRPT; ++a) {
                            acc[a][i] = fmaf(p[a][u], t.x, acc[a][i]);
                            acc[a][i + 1] = fmaf(p[a][u], t.y, acc[a][i + 1]);
                            acc[a][i + 2] = fmaf(p[a][u], t.z, acc[a][i + 2]);
                            acc[a][i + 3] = fmaf(p[a][u], t.w, acc[a][i + 3]);
                        }
                    }
                }
            }
#pragma unroll
            for (int a = 0; a < kRPT; ++a) {
                if (r0 + a < n) {
                    const float inv = 1.0f / l[a];
                    float* o = out + (b * (int64_t)n + r0 + a) * 128 + 16 * h;
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        *reinterpret_cast<float4*>(o + i) = make_float4(acc[a][i] * inv, acc[a][i + 1] * inv,
                                                                         acc[a][i + 2] * inv, acc[a][i + 3] * inv);
                }
            }
        }
    }
}

// ---- multi-head self-attention of one instance on tcgen05 (n <= 112) ----------------------------------------------
// Same structure as the glimpse of the rollout kernel: S_h = Q_h K_h^T (M = 128 query rows, N = NPAD keys, K = 16) into
// TMEM, softmax in registers (thread = query row, warp w -> lane quadrant w % 4, head w / 4 of the current group of
// four), P written back in place as fp16 hi/lo, O_h = P_h V_h with P as the TMEM A operand.  Q, K: the X-tile layout
// of the linear kernel (head h = K-chunks 2h, 2h+1); V^T: channels x keys.  512 threads, one instance per CTA.
constexpr int kAT = 512;
constexpr uint32_t kColO = 448;

struct AttnSmem {
    uint64_t mma_bar;
    uint32_t tmem_ptr;
    uint32_t pad_;
};

__global__ void __launch_bounds__(kAT, 1) enc_attention_tc_kernel(const float* __restrict__ qkv, float* __restrict__ out,
                                                                  int64_t B, int n, int npad) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* qs = smem;                 // Q tile hi | lo (64 KB)
    unsigned char* ks = smem + kTile;         // K tile
    unsigned char* vs = smem + 2 * kTile;     // V^T tile: byte(ch, key) = (ch/8) sbo_v + (key/8) 128 + (ch%8) 16 + (key%8) 2
    AttnSmem* s = reinterpret_cast<AttnSmem*>(smem + 3 * kTile);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    const uint32_t sbo_v = (uint32_t)(npad / 8) * 128u;
    const int nch = npad / 16;
    if (threadIdx.x == 0) {
        mbar_init(&s->mma_bar, 4);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(&s->tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s->tmem_ptr;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    const uint32_t qs_a = smem_u32(qs), ks_a = smem_u32(ks), vs_a = smem_u32(vs);
    uint32_t par = 0;
    const float kLog2e = 1.4426950408889634f;

    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        const float* base = qkv + b * (int64_t)n * 384;
        // ---- operands: Q (pre-scaled by 1/sqrt(16), exact), K, V^T as fp16 hi/lo tiles -----------------------------
        for (int v = warp; v < 2 * 64; v += 16) {            // 64 warp passes (8 rows x 4 chunks) per tile, Q then K
            const int which = v >> 6, vv = v & 63;
            const int rg = vv >> 2, c = 4 * (vv & 3) + (lane >> 3), r = 8 * rg + (lane & 7);
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bq = a;
            if (r < n) {
                const float* src = base + (size_t)r * 384 + which * 128 + 8 * c;
                a = *reinterpret_cast<const float4*>(src);
                bq = *reinterpret_cast<const float4*>(src + 4);
            }
            const float sc = which == 0 ? 0.25f : 1.0f;
            uint32_t h4[4], l4[4];
            split2(a.x * sc, a.y * sc, h4[0], l4[0]);
            split2(a.z * sc, a.w * sc, h4[1], l4[1]);
            split2(bq.x * sc, bq.y * sc, h4[2], l4[2]);
            split2(bq.z * sc, bq.w * sc, h4[3], l4[3]);
            unsigned char* dst = (which == 0 ? qs : ks) + (uint32_t)rg * kSBO + (uint32_t)c * kLBO + (uint32_t)(lane & 7) * 16u;
            *reinterpret_cast<uint4*>(dst) = make_uint4(h4[0], h4[1], h4[2], h4[3]);
            *reinterpret_cast<uint4*>(dst + kHalf) = make_uint4(l4[0], l4[1], l4[2], l4[3]);
        }
        for (int it = threadIdx.x; it < (npad / 8) * 128; it += kAT) {   // (8-key chunk, channel): lanes over channels
            const int rc = it >> 7, ch = it & 127;
            float vv[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = rc * 8 + i;
                vv[i] = (r < n) ? base[(size_t)r * 384 + 256 + ch] : 0.f;
            }
            uint32_t vh[4], vl[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) split2(vv[2 * p], vv[2 * p + 1], vh[p], vl[p]);
            const uint32_t off = (uint32_t)(ch >> 3) * sbo_v + (uint32_t)rc * 128u + (uint32_t)(ch & 7) * 16u;
            *reinterpret_cast<uint4*>(vs + off) = make_uint4(vh[0], vh[1], vh[2], vh[3]);
            *reinterpret_cast<uint4*>(vs + kHalf + off) = make_uint4(vl[0], vl[1], vl[2], vl[3]);
        }
        fence_proxy_async();
        __syncthreads();

        float o_keep[16];
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int h = 4 * g + j;
            // scores of head h into columns [j NPAD, (j+1) NPAD): one k-step, hi*hi + hi*lo + lo*hi
            if (q == j) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(kRows, npad);
                    const uint64_t a_hi = smem_desc(qs_a + 2u * h * kLBO, kLBO, kSBO);
                    const uint64_t a_lo = smem_desc(qs_a + kHalf + 2u * h * kLBO, kLBO, kSBO);
                    const uint64_t b_hi = smem_desc(ks_a + 2u * h * kLBO, kLBO, kSBO);
                    const uint64_t b_lo = smem_desc(ks_a + kHalf + 2u * h * kLBO, kLBO, kSBO);
                    const uint32_t d = tmem + (uint32_t)(j * npad);
                    mma_ss(d, a_hi, b_hi, id, 0u);
                    mma_ss(d, a_hi, b_lo, id, 1u);
                    mma_ss(d, a_lo, b_hi, id, 1u);
                    tc_commit(&s->mma_bar);
                }
                __syncwarp();
            }
            mbar_wait(&s->mma_bar, par);
            par ^= 1u;
            tc_fence_after();
            // softmax over the n keys of this thread's (row, head): pass 1 maximum, pass 2 exp / split / store in place
            const uint32_t sp = lane_base + (uint32_t)(j * npad);
            float m = -INFINITY;
            for (int c = 0; c < nch; ++c) {
                uint32_t v[16];
                tmem_ld16(sp + 16u * c, v);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (16 * c + i < n) m = fmaxf(m, __uint_as_float(v[i]));
            }
            const float ml = m * kLog2e;
            float l0 = 0.f, l1 = 0.f;
            for (int c = 0; c < nch; ++c) {
                uint32_t v[16], hi[8], lo[8];
                tmem_ld16(sp + 16u * c, v);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k0 = 16 * c + 2 * i;
                    const float p0 = (k0 < n) ? ex2_approx(fmaf(__uint_as_float(v[2 * i]), kLog2e, -ml)) : 0.f;
                    const float p1 = (k0 + 1 < n) ? ex2_approx(fmaf(__uint_as_float(v[2 * i + 1]), kLog2e, -ml)) : 0.f;
                    l0 += p0;
                    l1 += p1;
                    split2(p0, p1, hi[i], lo[i]);
                }
                tmem_st8(sp + 16u * c, hi);
                tmem_st8(sp + 16u * c + 8u, lo);
            }
            tc_wait_st();
            const float inv_l = 1.0f / (l0 + l1);
            tc_fence_before();
            __syncthreads();
            if (q == j) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(kRows, 16);
                    uint64_t bv_hi = smem_desc(vs_a + 2u * h * sbo_v, 128u, sbo_v);
                    uint64_t bv_lo = smem_desc(vs_a + kHalf + 2u * h * sbo_v, 128u, sbo_v);
                    const uint32_t d = tmem + kColO + 16u * j;
                    uint32_t a = tmem + (uint32_t)(j * npad);
                    for (int jk = 0; jk < nch; ++jk) {
                        mma_ts(d, a, bv_hi, id, jk > 0 ? 1u : 0u);
                        mma_ts(d, a, bv_lo, id, 1u);
                        mma_ts(d, a + 8u, bv_hi, id, 1u);
                        a += 16u;
                        bv_hi += (2u * 128u) >> 4;
                        bv_lo += (2u * 128u) >> 4;
                    }
                    tc_commit(&s->mma_bar);
                }
                __syncwarp();
            }
            mbar_wait(&s->mma_bar, par);
            par ^= 1u;
            tc_fence_after();
            uint32_t o[16];
            tmem_ld16(lane_base + kColO + 16u * j, o);
            tc_wait_ld();
            if (g == 0) {
#pragma unroll
                for (int i = 0; i < 16; ++i) o_keep[i] = __uint_as_float(o[i]) * inv_l;
                tc_fence_before();
                __syncthreads();          // every warp has drained O and is done with P before the columns are reused
            } else {
                // all MMAs are complete: Q and K tiles are free -> fp32 staging [128][kStage] for coalesced stores
                float* st = reinterpret_cast<float*>(smem);
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    *reinterpret_cast<float4*>(st + row * kStage + 16 * j + i) =
                        make_float4(o_keep[i], o_keep[i + 1], o_keep[i + 2], o_keep[i + 3]);
                    *reinterpret_cast<float4*>(st + row * kStage + 64 + 16 * j + i) =
                        make_float4(__uint_as_float(o[i]) * inv_l, __uint_as_float(o[i + 1]) * inv_l,
                                    __uint_as_float(o[i + 2]) * inv_l, __uint_as_float(o[i + 3]) * inv_l);
                }
            }
        }
        tc_fence_before();
        __syncthreads();
        {
            const float* st = reinterpret_cast<const float*>(smem);
            float* ob = out + b * (int64_t)n * 128;
            for (int idx = threadIdx.x; idx < n * 32; idx += kAT) {
                const int r = idx >> 5, c4 = (idx & 31) * 4;
                *reinterpret_cast<float4*>(ob + (size_t)r * 128 + c4) = *reinterpret_cast<const float4*>(st + r * kStage + c4);
            }
        }
        __syncthreads();
    }
    if (warp == 0) tmem_dealloc(tmem, 512);
}


// ---- fused attention block: Y[b] = Norm(X[b] + MHA(X[b]) Wo^T + bo) (attnnet.py:43-49, attention.py:66-136) ---------
// One kernel for what was qkv-linear -> attention -> out-proj+norm (three launches, 154 KB of q|k|v and 51 KB of
// heads per instance through HBM).  Per instance, 512 threads:
//   1. X -> fp16 hi/lo tile (region R0);  Q = X Wq^T, K = X Wk^T as usual (rows = nodes) and V^T = Wv X^T with the
//      operand roles swapped (A = the packed Wv tile, B = the X tile), so the accumulator rows are channels and the
//      V^T operand tile of the PV product is written with contiguous 16-byte stores - no transpose through memory.
//      W tiles stream through R1 / R2 (TMA bulk copies).  Accumulators: TMEM [0,384).
//   2. the accumulators are drained, biased (Q also scaled by 1/sqrt(16), exact), split and stored as operand tiles:
//      Q -> R0 (X is dead), K -> R1, V^T -> R2.
//   3. attention exactly as enc_attention_tc_kernel (two groups of four heads, scores in TMEM, softmax in registers,
//      P in place as the TS-form A operand of the PV product); Wo streams into R0 once the last scores are done.
//   4. the normalised heads go back to TMEM as a TS-form A operand, out = heads Wo^T (24 MMAs), residual + norm.
struct BlockSmem {
    // every barrier but mma_bar completes exactly once per instance (shared phase bit `ph`)
    uint64_t wq_bar, wk_bar, wv_bar, wo_bar;   // W tile landed (TMA)
    uint64_t r1free_bar;     // the Q product has finished reading Wq (R1 may take Wv)
    uint64_t qkv_bar;        // the three projection products are complete
    uint64_t mma_bar;        // attention MMAs of the four teams (4 arrivals), four completions per instance
    uint64_t out_bar;        // out-proj MMAs complete
    uint32_t tmem_ptr;
    uint32_t pad_;
    alignas(16) float mean[kRows];
    alignas(16) float rstd[kRows];
    alignas(16) float part[2][16 * kRows];
};
constexpr size_t kBlockSmemBytes = 3 * kTile + sizeof(BlockSmem);

// accumulator block (128 columns at `src`, thread = row) -> + bias, * scale -> fp16 hi/lo K-major tile at `tile`
// bias_per_row = false: bias indexed by column (Q, K: channel = column); true: by row (V^T: channel = row)
__device__ __forceinline__ void drain_to_tile(uint32_t src, unsigned char* tile, const float* __restrict__ bias,
                                              bool bias_per_row, float scale, int row) {
    const float brow = bias_per_row ? bias[row] : 0.f;
#pragma unroll 2
    for (int c = 0; c < 128; c += 16) {
        uint32_t v[16];
        tmem_ld16(src + (uint32_t)c, v);
        tc_wait_ld();
        float f[16];
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
            float4 bb = make_float4(brow, brow, brow, brow);
            if (!bias_per_row) bb = *reinterpret_cast<const float4*>(bias + c + i);
            f[i] = (__uint_as_float(v[i]) + bb.x) * scale;
            f[i + 1] = (__uint_as_float(v[i + 1]) + bb.y) * scale;
            f[i + 2] = (__uint_as_float(v[i + 2]) + bb.z) * scale;
            f[i + 3] = (__uint_as_float(v[i + 3]) + bb.w) * scale;
        }
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {           // two 8-element K-chunks
            uint32_t h4[4], l4[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) split2(f[8 * hh + 2 * p], f[8 * hh + 2 * p + 1], h4[p], l4[p]);
            const uint32_t off = (uint32_t)(row >> 3) * kSBO + (uint32_t)((c >> 3) + hh) * kLBO + (uint32_t)(row & 7) * 16u;
            *reinterpret_cast<uint4*>(tile + off) = make_uint4(h4[0], h4[1], h4[2], h4[3]);
            *reinterpret_cast<uint4*>(tile + kHalf + off) = make_uint4(l4[0], l4[1], l4[2], l4[3]);
        }
    }
}

template <int EPI>
__global__ void __launch_bounds__(kAT, 1)
    enc_attn_block_kernel(const float* __restrict__ X, const unsigned char* __restrict__ Wqkvp,
                          const float* __restrict__ bqkv, const unsigned char* __restrict__ Wop,
                          const float* __restrict__ bo, const float* __restrict__ gamma, const float* __restrict__ beta,
                          float eps, float* __restrict__ Y, int64_t B, int n, int npad) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* r0 = smem;                 // X tile -> Q tile -> Wo tile
    unsigned char* r1 = smem + kTile;         // Wq, Wv -> K tile -> staging
    unsigned char* r2 = smem + 2 * kTile;     // Wk -> V^T tile -> staging
    BlockSmem* s = reinterpret_cast<BlockSmem*>(smem + 3 * kTile);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    const int nch = npad / 16;
    if (threadIdx.x == 0) {
        mbar_init(&s->wq_bar, 1);
        mbar_init(&s->wk_bar, 1);
        mbar_init(&s->wv_bar, 1);
        mbar_init(&s->wo_bar, 1);
        mbar_init(&s->r1free_bar, 1);
        mbar_init(&s->qkv_bar, 1);
        mbar_init(&s->mma_bar, 4);
        mbar_init(&s->out_bar, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(&s->tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s->tmem_ptr;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    const uint32_t r0_a = smem_u32(r0), r1_a = smem_u32(r1), r2_a = smem_u32(r2);
    uint32_t ph = 0, par = 0;   // ph: instance parity (once-per-instance barriers); par: mma_bar
    const float kLog2e = 1.4426950408889634f;

    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        // ---- 1. X tile; Wq -> R1, Wk -> R2 in flight meanwhile ---------------------------------------------------
        if (warp == 4) {
            if (elect_one()) {
                mbar_expect_tx(&s->wq_bar, kTile);
                bulk_tile(r1, Wqkvp, &s->wq_bar);
                mbar_expect_tx(&s->wk_bar, kTile);
                bulk_tile(r2, Wqkvp + kTile, &s->wk_bar);
            }
            __syncwarp();
        }
        {
            const float* Xb = X + b * (int64_t)n * 128;
            for (int v = warp; v < 64; v += 16) {                // 64 warp passes: 8 rows x 4 chunks each
                const int rg = v >> 2, c = 4 * (v & 3) + (lane >> 3), r = 8 * rg + (lane & 7);
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bq = a;
                if (r < n) {
                    const float* src = Xb + (size_t)r * 128 + 8 * c;
                    a = *reinterpret_cast<const float4*>(src);
                    bq = *reinterpret_cast<const float4*>(src + 4);
                }
                uint32_t h4[4], l4[4];
                split2(a.x, a.y, h4[0], l4[0]);
                split2(a.z, a.w, h4[1], l4[1]);
                split2(bq.x, bq.y, h4[2], l4[2]);
                split2(bq.z, bq.w, h4[3], l4[3]);
                unsigned char* dst = r0 + (uint32_t)rg * kSBO + (uint32_t)c * kLBO + (uint32_t)(lane & 7) * 16u;
                *reinterpret_cast<uint4*>(dst) = make_uint4(h4[0], h4[1], h4[2], h4[3]);
                *reinterpret_cast<uint4*>(dst + kHalf) = make_uint4(l4[0], l4[1], l4[2], l4[3]);
            }
        }
        fence_proxy_async();
        __syncthreads();
        // ---- projections: Q [0,128), K [128,256), V^T [256,384) -------------------------------------------------
        if (warp == 4) {
            if (elect_one()) {
                mbar_wait(&s->wq_bar, ph);
                tc_fence_after();
                issue_tile(tmem, 0u, r0_a, r1_a, true);            // Q = X Wq^T
                tc_commit(&s->r1free_bar);
                mbar_wait(&s->wk_bar, ph);
                tc_fence_after();
                issue_tile(tmem, 128u, r0_a, r2_a, true);          // K = X Wk^T
                mbar_wait(&s->r1free_bar, ph);                     // the Q product has finished reading Wq
                mbar_expect_tx(&s->wv_bar, kTile);
                bulk_tile(r1, Wqkvp + 2 * (size_t)kTile, &s->wv_bar);
                mbar_wait(&s->wv_bar, ph);
                tc_fence_after();
                issue_tile(tmem, 256u, r1_a, r0_a, true);          // V^T = Wv X^T (operand roles swapped)
                tc_commit(&s->qkv_bar);
            }
            __syncwarp();
        }
        mbar_wait(&s->qkv_bar, ph);
        tc_fence_after();
        // ---- 2. operand tiles: teams 0..2 drain one block each ---------------------------------------------------
        if (j == 0) drain_to_tile(lane_base + 0u, r0, bqkv, false, 0.25f, row);            // Q / sqrt(16)
        else if (j == 1) drain_to_tile(lane_base + 128u, r1, bqkv + 128, false, 1.0f, row);
        else if (j == 2) drain_to_tile(lane_base + 256u, r2, bqkv + 256, true, 1.0f, row);  // rows = channels
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        tc_fence_after();

        // ---- 3. attention (enc_attention_tc_kernel with V^T at the full 128-key stride) --------------------------
        float o_keep[16], o_last[16];
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int h = 4 * g + j;
            if (q == j) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(kRows, npad);
                    const uint64_t a_hi = smem_desc(r0_a + 2u * h * kLBO, kLBO, kSBO);
                    const uint64_t a_lo = smem_desc(r0_a + kHalf + 2u * h * kLBO, kLBO, kSBO);
                    const uint64_t b_hi = smem_desc(r1_a + 2u * h * kLBO, kLBO, kSBO);
                    const uint64_t b_lo = smem_desc(r1_a + kHalf + 2u * h * kLBO, kLBO, kSBO);
                    const uint32_t d = tmem + (uint32_t)(j * npad);
                    mma_ss(d, a_hi, b_hi, id, 0u);
                    mma_ss(d, a_hi, b_lo, id, 1u);
                    mma_ss(d, a_lo, b_hi, id, 1u);
                    tc_commit(&s->mma_bar);
                }
                __syncwarp();
            }
            mbar_wait(&s->mma_bar, par);
            par ^= 1u;
            tc_fence_after();
            if (g == 1 && warp == 4) {            // every score product is done: the Q tile is dead, Wo streams in
                if (elect_one()) {
                    fence_proxy_async();
                    mbar_expect_tx(&s->wo_bar, kTile);
                    bulk_tile(r0, Wop, &s->wo_bar);
                }
                __syncwarp();
            }
            const uint32_t sp = lane_base + (uint32_t)(j * npad);
            float m = -INFINITY;
            for (int c = 0; c < nch; ++c) {
                uint32_t v[16];
                tmem_ld16(sp + 16u * c, v);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (16 * c + i < n) m = fmaxf(m, __uint_as_float(v[i]));
            }
            const float ml = m * kLog2e;
            float l0 = 0.f, l1 = 0.f;
            for (int c = 0; c < nch; ++c) {
                uint32_t v[16], hi[8], lo[8];
                tmem_ld16(sp + 16u * c, v);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k0 = 16 * c + 2 * i;
                    const float p0 = (k0 < n) ? ex2_approx(fmaf(__uint_as_float(v[2 * i]), kLog2e, -ml)) : 0.f;
                    const float p1 = (k0 + 1 < n) ? ex2_approx(fmaf(__uint_as_float(v[2 * i + 1]), kLog2e, -ml)) : 0.f;
                    l0 += p0;
                    l1 += p1;
                    split2(p0, p1, hi[i], lo[i]);
                }
                tmem_st8(sp + 16u * c, hi);
                tmem_st8(sp + 16u * c + 8u, lo);
            }
            tc_wait_st();
            const float inv_l = 1.0f / (l0 + l1);
            tc_fence_before();
            __syncthreads();
            if (q == j) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(kRows, 16);
                    uint64_t bv_hi = smem_desc(r2_a + 2u * h * kSBO, kLBO, kSBO);
                    uint64_t bv_lo = smem_desc(r2_a + kHalf + 2u * h * kSBO, kLBO, kSBO);
                    const uint32_t d = tmem + kColO + 16u * j;
                    uint32_t a = tmem + (uint32_t)(j * npad);
                    for (int jk = 0; jk < nch; ++jk) {
                        mma_ts(d, a, bv_hi, id, jk > 0 ? 1u : 0u);
                        mma_ts(d, a, bv_lo, id, 1u);
                        mma_ts(d, a + 8u, bv_hi, id, 1u);
                        a += 16u;
                        bv_hi += (2u * kLBO) >> 4;
                        bv_lo += (2u * kLBO) >> 4;
                    }
                    tc_commit(&s->mma_bar);
                }
                __syncwarp();
            }
            mbar_wait(&s->mma_bar, par);
            par ^= 1u;
            tc_fence_after();
            uint32_t o[16];
            tmem_ld16(lane_base + kColO + 16u * j, o);
            tc_wait_ld();
            if (g == 0) {
#pragma unroll
                for (int i = 0; i < 16; ++i) o_keep[i] = __uint_as_float(o[i]) * inv_l;
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) o_last[i] = __uint_as_float(o[i]) * inv_l;
            }
            tc_fence_before();
            __syncthreads();          // every warp has drained O and is done with P before the columns are reused
            tc_fence_after();
        }
        // ---- 4. heads -> TS-form A operand (hi [0,64), lo [64,128)), out = heads Wo^T into [128,256) ---------------
        {
            uint32_t hi[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) split2(o_keep[2 * i], o_keep[2 * i + 1], hi[i], lo[i]);
            tmem_st8(lane_base + 8u * j, hi);                      // head j: channels [16j, 16j+16)
            tmem_st8(lane_base + 64u + 8u * j, lo);
#pragma unroll
            for (int i = 0; i < 8; ++i) split2(o_last[2 * i], o_last[2 * i + 1], hi[i], lo[i]);
            tmem_st8(lane_base + 8u * (4 + j), hi);                // head 4 + j
            tmem_st8(lane_base + 64u + 8u * (4 + j), lo);
            tc_wait_st();
        }
        tc_fence_before();
        __syncthreads();
        if (warp == 4) {
            if (elect_one()) {
                mbar_wait(&s->wo_bar, ph);                         // Wo has landed in R0
                tc_fence_after();
                issue_tile_ts(tmem + 128u, tmem, tmem + 64u, r0_a, true);
                tc_commit(&s->out_bar);
            }
            __syncwarp();
        }
        mbar_wait(&s->out_bar, ph);
        ph ^= 1u;
        tc_fence_after();
        epilogue_res_norm<EPI, kAT>(lane_base + 128u, r1, s->part[0], s->mean, s->rstd, bo, X + b * (int64_t)n * 128, gamma,
                                    beta, eps, Y + b * (int64_t)n * 128, n);
        tc_fence_before();
        __syncthreads();
    }
    if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace enc
}  // namespace r4

// ---- C ABI ---------------------------------------------------------------------------------------------------------
using namespace r4;

#ifdef R4_ENC_PROF
extern "C" int rl4co_enc_debug_prof(long long* out16, int reset) {
    if (reset) { long long z[16] = {0}; cudaMemcpyToSymbol(r4::enc::g_enc_prof, z, sizeof(z)); return 0; }
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out16, r4::enc::g_enc_prof, 16 * sizeof(long long));
    return 0;
}
#endif

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
        case enc::EPI_RES_AFFINE:
            R4_CHECK_ARG(residual && gamma && beta && n_out == 128,
                         "enc_linear: the residual + affine epilogue needs residual, gamma, beta and n_out = 128");
            return launch_linear<enc::EPI_RES_AFFINE>(X, Wpacked, bias, residual, gamma, beta, 0.f, Y, B, n, k, n_out, stream);
        default:
            R4_CHECK_ARG(false, "enc_linear: unknown epilogue %d", epilogue);
    }
    return RL4CO_OK;
}


template <int EPI>
static int launch_ffn(const float* X, const void* W1p, const float* b1, const void* W2p, const float* b2,
                      const float* gamma, const float* beta, float eps, float* Y, int64_t B, int n, int hidden,
                      cudaStream_t stream) {
    auto kern = enc::enc_ffn_kernel<EPI>;
    static bool prepared = false;
    if (!prepared) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)enc::kFfnSmemBytes) != cudaSuccess) {
            set_error("enc_ffn: cannot reserve %zu bytes of shared memory", enc::kFfnSmemBytes);
            return RL4CO_E_LAUNCH;
        }
        prepared = true;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)(B < sms ? B : sms);
    kern<<<grid, enc::kThreads, enc::kFfnSmemBytes, stream>>>(X, (const unsigned char*)W1p, b1, (const unsigned char*)W2p,
                                                              b2, gamma, beta, eps, Y, B, n, hidden / 128);
    R4_CHECK_LAUNCH("enc_ffn");
    return RL4CO_OK;
}

extern "C" int rl4co_enc_ffn(const float* X, const void* W1packed, const float* b1, const void* W2packed,
                             const float* b2, const float* gamma, const float* beta, float eps, float* Y, int64_t B,
                             int32_t n, int32_t hidden, int32_t epilogue, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    R4_CHECK_ARG(X && W1packed && b1 && W2packed && b2 && gamma && beta && Y && B >= 0, "enc_ffn: null pointer");
    R4_CHECK_ARG(n >= 1 && n <= enc::kRows, "enc_ffn: 1 <= nodes <= 128 (got %d)", n);
    R4_CHECK_ARG(hidden >= 128 && hidden <= 512 && hidden % 128 == 0, "enc_ffn: hidden must be 128..512 in steps of 128 (got %d)", hidden);
    if (B == 0) return RL4CO_OK;
    if (epilogue == enc::EPI_RES_NORM)
        return launch_ffn<enc::EPI_RES_NORM>(X, W1packed, b1, W2packed, b2, gamma, beta, eps, Y, B, n, hidden, stream);
    if (epilogue == enc::EPI_RES_AFFINE)
        return launch_ffn<enc::EPI_RES_AFFINE>(X, W1packed, b1, W2packed, b2, gamma, beta, 0.f, Y, B, n, hidden, stream);
    R4_CHECK_ARG(false, "enc_ffn: epilogue must be RES_NORM or RES_AFFINE (got %d)", epilogue);
    return RL4CO_OK;
}


template <int EPI>
static int launch_attn_block(const float* X, const void* Wqkvp, const float* bqkv, const void* Wop, const float* bo,
                             const float* gamma, const float* beta, float eps, float* Y, int64_t B, int n,
                             cudaStream_t stream) {
    auto kern = enc::enc_attn_block_kernel<EPI>;
    static bool prepared = false;
    if (!prepared) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)enc::kBlockSmemBytes) != cudaSuccess) {
            set_error("enc_attn_block: cannot reserve %zu bytes of shared memory", enc::kBlockSmemBytes);
            return RL4CO_E_LAUNCH;
        }
        prepared = true;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)(B < sms ? B : sms);
    kern<<<grid, enc::kAT, enc::kBlockSmemBytes, stream>>>(X, (const unsigned char*)Wqkvp, bqkv, (const unsigned char*)Wop,
                                                            bo, gamma, beta, eps, Y, B, n, (n + 15) & ~15);
    R4_CHECK_LAUNCH("enc_attn_block");
    return RL4CO_OK;
}

extern "C" int rl4co_enc_attn_block(const float* X, const void* Wqkv_packed, const float* bqkv, const void* Wo_packed,
                                    const float* bo, const float* gamma, const float* beta, float eps, float* Y,
                                    int64_t B, int32_t n, int32_t epilogue, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    R4_CHECK_ARG(X && Wqkv_packed && bqkv && Wo_packed && bo && gamma && beta && Y && B >= 0, "enc_attn_block: null pointer");
    R4_CHECK_ARG(n >= 1 && ((n + 15) & ~15) <= 112, "enc_attn_block: 1 <= nodes <= 112 (got %d)", n);
    if (B == 0) return RL4CO_OK;
    if (epilogue == enc::EPI_RES_NORM)
        return launch_attn_block<enc::EPI_RES_NORM>(X, Wqkv_packed, bqkv, Wo_packed, bo, gamma, beta, eps, Y, B, n, stream);
    if (epilogue == enc::EPI_RES_AFFINE)
        return launch_attn_block<enc::EPI_RES_AFFINE>(X, Wqkv_packed, bqkv, Wo_packed, bo, gamma, beta, 0.f, Y, B, n, stream);
    R4_CHECK_ARG(false, "enc_attn_block: epilogue must be RES_NORM or RES_AFFINE (got %d)", epilogue);
    return RL4CO_OK;
}

extern "C" int rl4co_enc_attention(const float* qkv, float* out, int64_t B, int32_t n, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    R4_CHECK_ARG(qkv && out && B >= 0 && n >= 1 && n <= enc::kRows, "enc_attention: bad args (nodes = %d)", n);
    if (B == 0) return RL4CO_OK;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int npad = (n + 15) & ~15;
    if (npad <= 112) {   // tensor-core kernel: four heads of NPAD score columns + 64 output columns fit the 512 of TMEM
        const size_t smem = 3 * (size_t)enc::kTile + sizeof(enc::AttnSmem);
        static bool prepared_tc = false;
        if (!prepared_tc) {
            if (cudaFuncSetAttribute(enc::enc_attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)smem) != cudaSuccess) {
                set_error("enc_attention: cannot reserve %zu bytes of shared memory", smem);
                return RL4CO_E_LAUNCH;
            }
            prepared_tc = true;
        }
        const unsigned grid = (unsigned)(B < sms ? B : sms);
        enc::enc_attention_tc_kernel<<<grid, enc::kAT, smem, stream>>>(qkv, out, B, n, npad);
        R4_CHECK_LAUNCH("enc_attention (tcgen05)");
        return RL4CO_OK;
    }
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
    const int64_t want = 2LL * sms;
    const unsigned grid = (unsigned)(B < want ? B : want);
    // 8 * ceil(n / 2) (head, row pair) tasks: size the CTA so that its warp-tasks split into two full rounds
    const int warp_tasks = (8 * ((n + enc::kRPT - 1) / enc::kRPT) + 31) / 32;
    const int threads = 32 * ((warp_tasks + 1) / 2);
    enc::enc_attention_kernel<<<grid, threads, smem, stream>>>(qkv, out, B, n);
    R4_CHECK_LAUNCH("enc_attention");
    return RL4CO_OK;
}
