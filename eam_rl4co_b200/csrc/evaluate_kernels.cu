// Teacher-forced evaluate (decoding.py:452-465, constructive/base.py:204-229) of ALL decoding steps at once, forward and
// backward: log p(action_t | prefix) for every (instance, start, step) row and its gradient w.r.t. the glimpse query and the
// per-instance glimpse key / glimpse value / logit key tables.  SURVEY.md §8(b): rl4co_eval_ll_fwd / rl4co_eval_ll_bwd.
//
// With the actions known, the rows (instance b, start s, step t) are independent: row r of instance b has its query q[b,r,:]
// (context.py:118-157 gathers, built by the caller with autograd), its feasibility mask and its taken action.  All rows of an
// instance share the three [n,128] tables, which a CTA keeps in shared memory as fp16 hi / lo MMA operand tiles.  Forward, both
// parts on tcgen05 over tiles of 128 rows (TMEM lane = row):
//   glimpse  (attention.py:308-322): per head S = (0.25 Q_h) K_h^T, masked softmax numerator P in place as the TS-form A operand,
//            O_h = P V_h (masked nodes have probability exactly 0 and contribute no gradient);
//   pointer  (attention.py:296-301,328; decoding.py:140-190): logits = heads . lkf[n] / sqrt(E), then tanh clip, mask,
//            temperature, log_softmax, gather at the taken action.
// The backward recomputes the probabilities from the saved log-sum-exps (flash-attention style) in two passes, both on tcgen05:
//   row pass : everything that reduces over NODES, TMEM lane = row - du = d loss / d raw logit and dheads = dU lkf (the logits are
//             recomputed as an MMA per 128-row tile, du converted in place into the A operand of the second product); then
//             S = Q K^T and dP = dO V^T per head, dS = P (dP - D) in place, dq = 0.25 dS K;
//   node pass: the reductions over ROWS, TMEM lane = node - dgk = dS^T Q, dgv = P^T dO, dlkf = dU^T O - as MMAs with split-fp16
//             operands and fp32 accumulators that stay in TMEM across a CTA's row tiles.  (fp32 shared-memory atomics are CAS
//             loops on sm_100: a first version that accumulated the table gradients in shared memory spent 25 of 31 ms of the c5
//             evaluate in ATOMS.CAST.SPIN.)
//
// exp / tanh / log on the MUFU approximations (ex2, rcp, lg2: the arithmetic of the tensor-path rollout kernels).
#include "decode_core.cuh"
#include "tc_ptx.cuh"

namespace r4 {
namespace ev {

constexpr float kInvSqrtE = 0.08838834764831845f;   // 1 / sqrt(128)

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

// ---- the kernels below: every product of the evaluate on tcgen05 -------------------------------------------------------------------
//   backward, reductions over ROWS: dlkf = dU^T O (eval_dlk_tc_kernel); dgv = P^T dO, dgk = dS^T (0.25 Q) (eval_dkv_tc_kernel:
//     transposed score tiles, P^T / dS^T as TS-form A operands in TMEM);
//   backward, row passes: du and dheads = dU lkf (eval_ptr_bwd_tc_kernel); dq = 0.25 dS K and D = dO . O (eval_dq_tc_kernel);
//   forward: glimpse (eval_glimpse_fwd_tc_kernel) and pointer (eval_ptr_fwd_tc_kernel).
// du_t [B,R,128] is d loss / d (raw logit of node n), column = node; columns >= n are zero.
// (Round-2 history, measured on c5: CUDA-core versions - a warp per four rows x eight heads with the tables in shared memory for
// the row work, a warp per two nodes with register accumulators for the reductions over rows - took 2.1 ms forward and 7.5 ms
// backward; the tensor-core kernels take 1.06 and 2.65 ms.)
constexpr int kMaxNodes = 128;   // four mask words per row, 128-node MMA tiles

// cp.async staging of row tiles (eval_dlk_tc_kernel)
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// ---- backward: power-of-two scale of the incoming gradient -------------------------------------------------------------------------
// The backward MMAs take their operands as fp16 hi + lo: ~22 bits of mantissa but fp16's EXPONENT range.  dlogp of a training step
// is advantage / (B S) ~ 1e-4, the quantities derived from it (du, dheads, dS) reach into fp16's subnormals and the gradients lose
// accuracy with the loss scale (measured: 2e-3 relative at a cotangent of 1e-4, 0.2 at 1e-6, all zero at 1e-8).  Every output is
// linear in dlogp, so the whole backward runs on dlogp * s with s = 2^-ceil(log2 max|dlogp|) (exact in fp32) and the kernels that
// write results multiply by 1 / s in their epilogues.  ws[0] = s, ws[1] = 1 / s.
__global__ void __launch_bounds__(1024)
    eval_grad_scale_kernel(const float* __restrict__ dlogp, int64_t rows, float* __restrict__ ws) {
    __shared__ float s_m[32];
    float m = 0.f;
    for (int64_t i = threadIdx.x; i < rows; i += 1024) {
        const float a = fabsf(dlogp[i]);
        m = (a <= 3.0e38f && a > m) ? a : m;      // ignores NaN / inf
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, o));
    if ((threadIdx.x & 31) == 0) s_m[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = s_m[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, o));
        if (threadIdx.x == 0) {
            int e = 0;
            if (m > 0.f) {
                frexpf(m, &e);                    // m = f 2^e, f in [0.5, 1)
                e = e > 100 ? 100 : (e < -100 ? -100 : e);
            }
            ws[0] = ldexpf(1.0f, -e);
            ws[1] = ldexpf(1.0f, e);
        }
    }
}

// ---- backward, pointer part of the node pass on the tensor cores: dlkf[b] = dU^T O (tcgen05, split fp16) -------------------------
// Per CTA a [128 x 128] fp32 accumulator in TMEM (rows = the 128 du_t columns = nodes; columns = the 128
// head channels), summed over the CTA's row tiles of 64 rows (the K dimension).  A tile of du_t and heads is staged as fp32
// with cp.async (rows padded to 132 floats), converted by the CUDA cores into two K-major fp16 hi/lo operand tiles - the
// conversion IS the transposition: a thread reads eight consecutive rows of one column and writes one 16-byte K chunk - and
// multiplied with 12 MMAs (hi*hi + hi*lo + lo*hi, four k-steps).  0.45e9 FMA warp-instructions of the register version
// become ~1.5 k conversion instructions per tile.
namespace dlk {
constexpr int kT = 64;                          // rows per tile = K of one tile product
constexpr int kStride = 132;                    // staged fp32 row stride (floats)
constexpr uint32_t kOpHalf = 16 * 1024;         // [128 rows][64 K] fp16: 16 row groups x (8 chunks x 128 B)
constexpr uint32_t kOffStage = 4 * kOpHalf;     // A hi, A lo, B hi, B lo, then 2 stages x (du tile, heads tile)
constexpr uint32_t kStageFloats = 2 * kT * kStride;
constexpr uint32_t kOffBar = kOffStage + 2 * kStageFloats * 4;
constexpr size_t kSmem = kOffBar + 64;
}  // namespace dlk

__global__ void __launch_bounds__(256, 1)
    eval_dlk_tc_kernel(const float* __restrict__ heads, const float* __restrict__ du_t, const float* __restrict__ gscale,
                       float* __restrict__ dlkf, int64_t R, int n) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smb[];
    unsigned char* op = smb;
    float* stage = reinterpret_cast<float*>(smb + dlk::kOffStage);
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(smb + dlk::kOffBar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smb + dlk::kOffBar + 16);
    const int64_t b = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int64_t r0, r1;
    row_range(R, r0, r1);
    if (r0 >= r1) return;
    const int ntiles = (int)((r1 - r0 + dlk::kT - 1) / dlk::kT);
    if (threadIdx.x == 0) {
        mbar_init(mma_bar, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(tmem_slot, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    auto issue = [&](int t) {
        const int64_t rt = r0 + (int64_t)t * dlk::kT;
        const int rows = (int)(r1 - rt < dlk::kT ? r1 - rt : dlk::kT);
        float* st = stage + (t & 1) * dlk::kStageFloats;
        for (int i = threadIdx.x; i < rows * 32; i += blockDim.x) {
            const int r = i >> 5, c4 = i & 31;
            cp_async16(st + r * dlk::kStride + 4 * c4, du_t + (b * R + rt + r) * kE + 4 * c4);
            cp_async16(st + (dlk::kT + r) * dlk::kStride + 4 * c4, heads + (b * R + rt + r) * kE + 4 * c4);
        }
        cp_async_commit();
    };
    uint32_t par = 0;
    issue(0);
    for (int t = 0; t < ntiles; ++t) {
        if (t + 1 < ntiles) {
            issue(t + 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (t > 0) {                                     // the MMAs of tile t-1 have read the operand tiles
            mbar_wait(mma_bar, par);
            par ^= 1u;
        }
        const int64_t rt = r0 + (int64_t)t * dlk::kT;
        const int rows = (int)(r1 - rt < dlk::kT ? r1 - rt : dlk::kT);
        const float* st = stage + (t & 1) * dlk::kStageFloats;
#pragma unroll 2
        for (int j = 0; j < 8; ++j) {
            const int i = threadIdx.x + 256 * j;         // item = (operand, 8-row group g, column m)
            const int which = i >> 10, g = (i >> 7) & 7, m = i & 127;
            const float* src = st + (which * dlk::kT + 8 * g) * dlk::kStride + m;
            float v[8];
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) v[kk] = (8 * g + kk < rows) ? src[kk * dlk::kStride] : 0.f;
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int w = 0; w < 4; ++w) split2(v[2 * w], v[2 * w + 1], hi[w], lo[w]);
            unsigned char* dst = op + (uint32_t)which * 2u * dlk::kOpHalf + (uint32_t)(m >> 3) * 1024u + (uint32_t)g * 128u +
                                 (uint32_t)(m & 7) * 16u;
            *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + dlk::kOpHalf) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        fence_proxy_async();
        __syncthreads();
        if (warp == 0) {
            if (elect_one()) {
                tc_fence_after();
                const uint32_t id = idesc_f16(128, 128);
                const uint32_t oa = smem_u32(op);
                uint64_t a_hi = smem_desc(oa, 128, 1024), a_lo = smem_desc(oa + dlk::kOpHalf, 128, 1024);
                uint64_t b_hi = smem_desc(oa + 2 * dlk::kOpHalf, 128, 1024), b_lo = smem_desc(oa + 3 * dlk::kOpHalf, 128, 1024);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    mma_ss(tmem, a_hi, b_hi, id, (t == 0 && ks == 0) ? 0u : 1u);
                    mma_ss(tmem, a_hi, b_lo, id, 1u);
                    mma_ss(tmem, a_lo, b_hi, id, 1u);
                    a_hi += 256u >> 4;
                    a_lo += 256u >> 4;
                    b_hi += 256u >> 4;
                    b_lo += 256u >> 4;
                }
                tc_commit(mma_bar);
            }
            __syncwarp();
        }
    }
    mbar_wait(mma_bar, par);
    tc_fence_after();
    {   // accumulator row m = du_t column m = node m; one red per element
        const int m = 32 * (warp & 3) + lane, node = m;
        const uint32_t src = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + 64u * (uint32_t)(warp >> 2);
#pragma unroll 1
        for (int c = 0; c < 64; c += 16) {
            uint32_t v[16];
            tmem_ld16(src + (uint32_t)c, v);
            tc_wait_ld();
            if (node < n) {
                const float inv_s = gscale[1];
                float* dst = dlkf + (b * n + node) * kE + 64 * (warp >> 2) + c;
#pragma unroll
                for (int i = 0; i < 16; ++i) atomicAdd(dst + i, __uint_as_float(v[i]) * inv_s);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 128);
}

// ---- backward, row pass of the pointer on the tensor cores: du_t and dheads ------------------------------------------------------
// Per tile of 128 rows: U = heads lkf^T (24 split-fp16 MMAs, A = the heads tile converted from global fp32, B = the resident
// lkf tile) lands with TMEM lane = row; thread (row, 64 nodes) computes du = g (1[n = a] - softmax(z)[n]) dz/du / sqrt(E) on
// the feasible nodes (softmax from the saved lse_z), stores it (du_t feeds eval_dlk_tc_kernel) and writes it back IN PLACE as
// a TS-form A operand; dheads = du lkf is 24 TS-form MMAs against the resident TRANSPOSED lkf tile.  The dot products that
// made up most of the CUDA-core version's instructions (1.6 ms) are gone.
// Item mapping of the (row, 8-channel chunk) conversions into a K-major operand tile ("item_rc"): consecutive lanes take the 8 rows
// of a core matrix (16 bytes apart in the tile), then the next chunk: a warp's STS.128 covers 512 contiguous bytes.  (Lanes running
// over the chunks of ONE row are 128 bytes apart in the tile - a 16-way bank conflict that cost the pointer kernels half their time.)
namespace pbw {
constexpr int kT = 128;
constexpr uint32_t kHalf = 32 * 1024;           // [128 x 128] fp16 K-major: 16 row groups x 16 chunks x 128 B
constexpr uint32_t kOffL = 0;                   // lkf   [node][ch] hi | lo   (B of U:      N = nodes, K = channels)
constexpr uint32_t kOffLT = 2 * kHalf;          // lkf^T [ch][node] hi | lo   (B of dheads: N = channels, K = nodes)
constexpr uint32_t kOffH = 4 * kHalf;           // heads tile [row][ch] hi | lo (A of U)
constexpr uint32_t kOffRow = 6 * kHalf;         // per row of the tile: g, lse_z (f32), taken (i32), 4 mask words
constexpr uint32_t kOffBar = kOffRow + kT * 28;
constexpr size_t kSmem = kOffBar + 64;
constexpr int kThreads = 512;                   // thread = (row, 32 nodes): 16 warps for the elementwise pass
}  // namespace pbw

__device__ __forceinline__ float tanh_fast_ev(float v) {   // 1 - 2 / (1 + e^(2v)) on MUFU ex2 / rcp: abs. error ~2e-7
    const float e2 = __expf(2.0f * v);
    return 1.0f - __fdividef(2.0f, e2 + 1.0f);
}

__global__ void __launch_bounds__(pbw::kThreads, 1)
    eval_ptr_bwd_tc_kernel(const float* __restrict__ lkf, const uint32_t* __restrict__ maskw, const int64_t* __restrict__ taken,
                           const float* __restrict__ heads, const float* __restrict__ lse_z, const float* __restrict__ dlogp,
                           const float* __restrict__ gscale, float* __restrict__ dheads, float* __restrict__ du_t, int64_t R, int n, float clip, float temperature) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smb[];
    float* s_g = reinterpret_cast<float*>(smb + pbw::kOffRow);
    float* s_lz = s_g + pbw::kT;
    int* s_a = reinterpret_cast<int*>(s_lz + pbw::kT);
    uint32_t* s_mw = reinterpret_cast<uint32_t*>(s_a + pbw::kT);
    uint64_t* bar1 = reinterpret_cast<uint64_t*>(smb + pbw::kOffBar);
    uint64_t* bar2 = bar1 + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smb + pbw::kOffBar + 32);
    const int64_t b = blockIdx.x;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int64_t r0, r1;
    row_range(R, r0, r1);
    if (r0 >= r1) return;
    const int ntiles = (int)((r1 - r0 + pbw::kT - 1) / pbw::kT);
    if (tid == 0) {
        mbar_init(bar1, 1);
        mbar_init(bar2, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(tmem_slot, 256);
    // resident B operands: lkf (K = channels) and its transpose (K = nodes); nodes >= n are zero
    for (int i = tid; i < 128 * 16; i += pbw::kThreads) {
        const int m = 8 * (i >> 7) + (i & 7), c = (i >> 3) & 15;   // lkf: (node m, 8-channel chunk c), see item_rc
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bb = a;
        if (m < n) {
            a = ld4(lkf + ((size_t)b * n + m) * kE + 8 * c);
            bb = ld4(lkf + ((size_t)b * n + m) * kE + 8 * c + 4);
        }
        uint32_t hi[4], lo[4];
        split2(a.x, a.y, hi[0], lo[0]);
        split2(a.z, a.w, hi[1], lo[1]);
        split2(bb.x, bb.y, hi[2], lo[2]);
        split2(bb.z, bb.w, hi[3], lo[3]);
        unsigned char* dst = smb + pbw::kOffL + (uint32_t)(m >> 3) * 2048u + (uint32_t)c * 128u + (uint32_t)(m & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + pbw::kHalf) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    for (int i = tid; i < 128 * 16; i += pbw::kThreads) {
        const int ch = i & 127, g = i >> 7;              // lkf^T: (channel ch, 8-node group g); lanes run over channels
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = (8 * g + k < n) ? lkf[((size_t)b * n + 8 * g + k) * kE + ch] : 0.f;
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) split2(v[2 * w], v[2 * w + 1], hi[w], lo[w]);
        unsigned char* dst = smb + pbw::kOffLT + (uint32_t)(ch >> 3) * 2048u + (uint32_t)g * 128u + (uint32_t)(ch & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + pbw::kHalf) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t sb = smem_u32(smb);
    const int q4 = warp & 3, cq = warp >> 2;             // TMEM lane quadrant (row = 32 q4 + lane), node quarter
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q4) << 16);
    const float sc = (clip > 0.f ? clip : 1.0f) / temperature;
    uint32_t p1 = 0, p2 = 0;
    // The fp32 heads of a tile (4 items = 8 float4 per thread) and its per-row scalars travel through registers ONE TILE AHEAD:
    // the loads of tile t+1 are issued right after tile t's first product and land during its elementwise pass.
    float4 pre[8];
    float pre_g = 0.f, pre_lz = 0.f;
    int pre_a = -1;
    uint32_t pre_mw[4] = {0u, 0u, 0u, 0u};
    auto prefetch = [&](int t) {
        const int64_t rt = r0 + (int64_t)t * pbw::kT;
        const int rows = t < ntiles ? (int)(r1 - rt < pbw::kT ? r1 - rt : pbw::kT) : 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = tid + pbw::kThreads * k;
            const int r = 8 * (i >> 7) + (i & 7), c = (i >> 3) & 15;   // conflict-free operand stores (item_rc)
            pre[2 * k] = pre[2 * k + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < rows) {
                pre[2 * k] = ld4(heads + (b * R + rt + r) * kE + 8 * c);
                pre[2 * k + 1] = ld4(heads + (b * R + rt + r) * kE + 8 * c + 4);
            }
        }
        if (tid < pbw::kT) {
            const bool ok = tid < rows;
            const int64_t row = b * R + (ok ? rt + tid : r0);
            pre_g = ok ? dlogp[row] * gscale[0] : 0.f;
            pre_lz = ok ? lse_z[row] : 0.f;
            pre_a = ok ? (int)taken[row] : -1;
#pragma unroll
            for (int k = 0; k < 4; ++k) pre_mw[k] = ok ? maskw[row * kNB + k] : 0u;
        }
    };
    prefetch(0);
    for (int t = 0; t < ntiles; ++t) {
        const int64_t rt = r0 + (int64_t)t * pbw::kT;
        const int rows = (int)(r1 - rt < pbw::kT ? r1 - rt : pbw::kT);
        // heads tile -> K-major fp16 hi | lo (item = (row, 8-channel chunk)); per-row scalars
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = tid + pbw::kThreads * k;
            const int r = 8 * (i >> 7) + (i & 7), c = (i >> 3) & 15;
            const float4 a = pre[2 * k], bb = pre[2 * k + 1];
            uint32_t hi[4], lo[4];
            split2(a.x, a.y, hi[0], lo[0]);
            split2(a.z, a.w, hi[1], lo[1]);
            split2(bb.x, bb.y, hi[2], lo[2]);
            split2(bb.z, bb.w, hi[3], lo[3]);
            unsigned char* dst = smb + pbw::kOffH + (uint32_t)(r >> 3) * 2048u + (uint32_t)c * 128u + (uint32_t)(r & 7) * 16u;
            *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + pbw::kHalf) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        if (tid < pbw::kT) {
            s_g[tid] = pre_g;
            s_lz[tid] = pre_lz;
            s_a[tid] = pre_a;
#pragma unroll
            for (int k = 0; k < 4; ++k) s_mw[tid * 4 + k] = pre_mw[k];
        }
        fence_proxy_async();
        __syncthreads();
        if (warp == 0) {
            if (elect_one()) {
                tc_fence_after();
                const uint32_t id = idesc_f16(128, 128);
                uint64_t a_hi = smem_desc(sb + pbw::kOffH, 128, 2048), a_lo = smem_desc(sb + pbw::kOffH + pbw::kHalf, 128, 2048);
                uint64_t b_hi = smem_desc(sb + pbw::kOffL, 128, 2048), b_lo = smem_desc(sb + pbw::kOffL + pbw::kHalf, 128, 2048);
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    mma_ss(tmem, a_hi, b_hi, id, ks == 0 ? 0u : 1u);     // U = heads lkf^T
                    mma_ss(tmem, a_hi, b_lo, id, 1u);
                    mma_ss(tmem, a_lo, b_hi, id, 1u);
                    a_hi += 256u >> 4;
                    a_lo += 256u >> 4;
                    b_hi += 256u >> 4;
                    b_lo += 256u >> 4;
                }
                tc_commit(bar1);
            }
            __syncwarp();
        }
        prefetch(t + 1);
        mbar_wait(bar1, p1);
        p1 ^= 1u;
        tc_fence_after();
        {
            const int r = 32 * q4 + lane;
            const bool rv = r < rows;
            const float g = s_g[r], lz = s_lz[r];
            const int a = s_a[r];
            const uint32_t mw = s_mw[r * 4 + cq];
            float* dut = du_t + (b * R + rt + (rv ? r : 0)) * kE;
#pragma unroll 1
            for (int c = 0; c < 32; c += 16) {
                const int col = 32 * cq + c;
                uint32_t uv[16];
                tmem_ld16(lane_base + (uint32_t)col, uv);
                tc_wait_ld();
                const uint32_t bits = (mw >> c) & 0xffffu;
                float du[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float u = __uint_as_float(uv[i]) * kInvSqrtE;
                    float zv, dzu;
                    if (clip > 0.f) {
                        const float th = tanh_fast_ev(u);
                        zv = th * sc;
                        dzu = sc * (1.0f - th * th);
                    } else {
                        zv = u * sc;
                        dzu = sc;
                    }
                    const float pz = __expf(zv - lz);
                    du[i] = (bits >> i) & 1u ? g * ((col + i == a ? 1.0f : 0.0f) - pz) * dzu * kInvSqrtE : 0.f;
                }
                if (rv) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        *reinterpret_cast<float4*>(dut + col + i) = make_float4(du[i], du[i + 1], du[i + 2], du[i + 3]);
                }
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int i = 0; i < 16; i += 2) split2(du[i], du[i + 1], hi[i >> 1], lo[i >> 1]);
                tmem_st8(lane_base + (uint32_t)col, hi);
                tmem_st8(lane_base + (uint32_t)col + 8u, lo);
            }
        }
        tc_wait_st();
        tc_fence_before();
        __syncthreads();
        if (warp == 0) {
            if (elect_one()) {
                tc_fence_after();
                const uint32_t id = idesc_f16(128, 128);
                uint64_t b_hi = smem_desc(sb + pbw::kOffLT, 128, 2048), b_lo = smem_desc(sb + pbw::kOffLT + pbw::kHalf, 128, 2048);
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    mma_ts(tmem + 128u, tmem + 16u * ks, b_hi, id, ks == 0 ? 0u : 1u);   // dheads = du lkf
                    mma_ts(tmem + 128u, tmem + 16u * ks, b_lo, id, 1u);
                    mma_ts(tmem + 128u, tmem + 16u * ks + 8u, b_hi, id, 1u);
                    b_hi += 256u >> 4;
                    b_lo += 256u >> 4;
                }
                tc_commit(bar2);
            }
            __syncwarp();
        }
        mbar_wait(bar2, p2);
        p2 ^= 1u;
        tc_fence_after();
        {
            const int r = 32 * q4 + lane;
            float* dh = dheads + (b * R + rt + (r < rows ? r : 0)) * kE + 32 * cq;
#pragma unroll 1
            for (int c = 0; c < 32; c += 16) {
                uint32_t v[16];
                tmem_ld16(lane_base + 128u + 32u * (uint32_t)cq + (uint32_t)c, v);
                tc_wait_ld();
                if (r < rows) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        *reinterpret_cast<float4*>(dh + c + i) = make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]),
                                                                            __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
                }
            }
        }
        tc_fence_before();
        __syncthreads();   // the heads tile, the row scalars and the TMEM columns are reused by the next tile
    }
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ---- forward, pointer part on the tensor cores: logp and lse_z from the heads the glimpse kernel wrote ------------------------------
// Per tile of 128 rows the logits U = heads lkf^T are 24 split-fp16 MMAs against the resident lkf tile (as in the backward's row
// pass above); thread (row, 32 nodes) applies scale, tanh clip, mask and temperature, and the four threads of a row combine maximum,
// sum and the taken node's value through shared memory.  Software pipeline over the tiles of a CTA: the fp32 heads of tile t+2
// arrive by cp.async while tile t+1 is converted to the fp16 operand and multiplied into the second accumulator, and tile t's
// logits are read from the first.
namespace pfw {
constexpr int kT = 128;
constexpr int kStride = 132;                    // floats per staged row (+4: the conversion's 16 rows per warp hit distinct banks)
constexpr uint32_t kHalf = 32 * 1024;
constexpr uint32_t kOffL = 0;                   // lkf [node][ch] hi | lo (B operand, resident)
constexpr uint32_t kOffH = 2 * kHalf;           // heads tile [row][ch] hi | lo (A operand)
constexpr uint32_t kOffStage = 4 * kHalf;       // fp32 heads tile as it arrives
constexpr int kThreads = 512;                   // thread = (row, 32 nodes): 16 warps hide the MUFU / tcgen05.ld latencies
constexpr uint32_t kOffRow = kOffStage + kT * kStride * 4;   // per row: taken (i32), 4 mask words; 4 x (max, sum, za) exchange
constexpr uint32_t kOffBar = kOffRow + kT * (20 + 48);
constexpr size_t kSmem = kOffBar + 64;
}  // namespace pfw

__global__ void __launch_bounds__(pfw::kThreads, 1)
    eval_ptr_fwd_tc_kernel(const float* __restrict__ lkf, const uint32_t* __restrict__ maskw, const int64_t* __restrict__ taken,
                           const float* __restrict__ heads, float* __restrict__ logp, float* __restrict__ lse_z, int64_t R, int n,
                           float clip, float temperature) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smb[];
    float* s_stage = reinterpret_cast<float*>(smb + pfw::kOffStage);
    int* s_a = reinterpret_cast<int*>(smb + pfw::kOffRow);
    uint32_t* s_mw = reinterpret_cast<uint32_t*>(s_a + pfw::kT);
    float* s_x = reinterpret_cast<float*>(s_mw + 4 * pfw::kT);       // [4 node quarters][3][128]: max, sum, za
    uint64_t* bars = reinterpret_cast<uint64_t*>(smb + pfw::kOffBar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smb + pfw::kOffBar + 32);
    const int64_t b = blockIdx.x;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int64_t r0, r1;
    row_range(R, r0, r1);
    if (r0 >= r1) return;
    const int ntiles = (int)((r1 - r0 + pfw::kT - 1) / pfw::kT);
    if (tid == 0) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(tmem_slot, 256);

    auto fetch = [&](int t) {   // fp32 heads of tile t -> staging (rows past the end keep stale values: their logits are masked)
        const int64_t rt = r0 + (int64_t)t * pfw::kT;
        const int rows = (int)(r1 - rt < pfw::kT ? r1 - rt : pfw::kT);
        for (int i = tid; i < rows * 32; i += pfw::kThreads) {
            const int r = i >> 5, c4 = i & 31;
            cp_async16(s_stage + r * pfw::kStride + 4 * c4, heads + (b * R + rt + r) * kE + 4 * c4);
        }
        cp_async_commit();
    };
    auto convert = [&]() {      // staging -> K-major fp16 hi | lo, item = (row, 8-channel chunk)
        for (int i = tid; i < pfw::kT * 16; i += pfw::kThreads) {
            const int r = 8 * (i >> 7) + (i & 7), c = (i >> 3) & 15;   // conflict-free operand stores (item_rc)
            const float4 a = *reinterpret_cast<const float4*>(s_stage + r * pfw::kStride + 8 * c);
            const float4 bb = *reinterpret_cast<const float4*>(s_stage + r * pfw::kStride + 8 * c + 4);
            uint32_t hi[4], lo[4];
            split2(a.x, a.y, hi[0], lo[0]);
            split2(a.z, a.w, hi[1], lo[1]);
            split2(bb.x, bb.y, hi[2], lo[2]);
            split2(bb.z, bb.w, hi[3], lo[3]);
            unsigned char* dst = smb + pfw::kOffH + (uint32_t)(r >> 3) * 2048u + (uint32_t)c * 128u + (uint32_t)(r & 7) * 16u;
            *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + pfw::kHalf) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
    };
    const uint32_t sb = smem_u32(smb);
    auto issue = [&](uint32_t tmem, int t) {   // U(t) = heads lkf^T into accumulator t & 1
        if (warp == 0) {
            if (elect_one()) {
                tc_fence_after();
                const uint32_t id = idesc_f16(128, 128);
                uint64_t a_hi = smem_desc(sb + pfw::kOffH, 128, 2048), a_lo = smem_desc(sb + pfw::kOffH + pfw::kHalf, 128, 2048);
                uint64_t b_hi = smem_desc(sb + pfw::kOffL, 128, 2048), b_lo = smem_desc(sb + pfw::kOffL + pfw::kHalf, 128, 2048);
                const uint32_t d = tmem + 128u * (uint32_t)(t & 1);
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    mma_ss(d, a_hi, b_hi, id, ks == 0 ? 0u : 1u);
                    mma_ss(d, a_hi, b_lo, id, 1u);
                    mma_ss(d, a_lo, b_hi, id, 1u);
                    a_hi += 256u >> 4;
                    a_lo += 256u >> 4;
                    b_hi += 256u >> 4;
                    b_lo += 256u >> 4;
                }
                tc_commit(bars + (t & 1));
            }
            __syncwarp();
        }
    };
    // per-row scalars of a tile travel through registers one tile ahead (thread = row)
    int reg_a = -1;
    uint32_t reg_mw[4] = {0u, 0u, 0u, 0u};
    auto fetch_rows = [&](int t) {
        if (tid < pfw::kT) {
            const int64_t rt = r0 + (int64_t)t * pfw::kT;
            const bool ok = t < ntiles && rt + tid < r1;
            const int64_t row = b * R + (ok ? rt + tid : r0);
            reg_a = ok ? (int)taken[row] : -1;
#pragma unroll
            for (int k = 0; k < 4; ++k) reg_mw[k] = ok ? maskw[row * kNB + k] : 0u;
        }
    };

    fetch(0);
    fetch_rows(0);
    for (int i = tid; i < 128 * 16; i += pfw::kThreads) {   // resident B operand: lkf (K = channels), nodes >= n are zero
        const int m = 8 * (i >> 7) + (i & 7), c = (i >> 3) & 15;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bb = a;
        if (m < n) {
            a = ld4(lkf + ((size_t)b * n + m) * kE + 8 * c);
            bb = ld4(lkf + ((size_t)b * n + m) * kE + 8 * c + 4);
        }
        uint32_t hi[4], lo[4];
        split2(a.x, a.y, hi[0], lo[0]);
        split2(a.z, a.w, hi[1], lo[1]);
        split2(bb.x, bb.y, hi[2], lo[2]);
        split2(bb.z, bb.w, hi[3], lo[3]);
        unsigned char* dst = smb + pfw::kOffL + (uint32_t)(m >> 3) * 2048u + (uint32_t)c * 128u + (uint32_t)(m & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + pfw::kHalf) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    cp_async_wait<0>();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    convert();
    fence_proxy_async();
    __syncthreads();
    issue(tmem, 0);
    if (ntiles > 1) fetch(1);

    const int q4 = warp & 3, cq = warp >> 2;             // TMEM lane quadrant (row = 32 q4 + lane), node quarter
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q4) << 16);
    const float sc = (clip > 0.f ? clip : 1.0f) / temperature;
    uint32_t ph[2] = {0u, 0u};
    for (int t = 0; t < ntiles; ++t) {
        const int64_t rt = r0 + (int64_t)t * pfw::kT;
        const int rows = (int)(r1 - rt < pfw::kT ? r1 - rt : pfw::kT);
        if (tid < pfw::kT) {
            s_a[tid] = reg_a;
#pragma unroll
            for (int k = 0; k < 4; ++k) s_mw[tid * 4 + k] = reg_mw[k];
        }
        fetch_rows(t + 1);
        mbar_wait(bars + (t & 1), ph[t & 1]);             // U(t) is complete: the operand tile is free
        ph[t & 1] ^= 1u;
        tc_fence_after();
        if (t + 1 < ntiles) {
            cp_async_wait<0>();
            __syncthreads();
            convert();
            fence_proxy_async();
            __syncthreads();
            issue(tmem, t + 1);
            if (t + 2 < ntiles) fetch(t + 2);
        } else {
            __syncthreads();
        }
        const int r = 32 * q4 + lane;
        const int a = s_a[r];
        const uint32_t mw = s_mw[r * 4 + cq];
        float z[32];
        float zmax = -INFINITY, za = -INFINITY;
#pragma unroll
        for (int c = 0; c < 32; c += 16) {
            uint32_t uv[16];
            tmem_ld16(lane_base + 128u * (uint32_t)(t & 1) + (uint32_t)(32 * cq + c), uv);
            tc_wait_ld();
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float u = __uint_as_float(uv[i]) * kInvSqrtE;
                const float zv = (clip > 0.f ? tanh_fast_ev(u) : u) * sc;     // decoding.py:171-172, :180
                z[c + i] = (mw >> (c + i)) & 1u ? zv : -INFINITY;              // :177
                zmax = fmaxf(zmax, z[c + i]);
                if (32 * cq + c + i == a) za = z[c + i];
            }
        }
        s_x[(cq * 3 + 0) * pfw::kT + r] = zmax;
        s_x[(cq * 3 + 2) * pfw::kT + r] = za;
        __syncthreads();
        float mx = zmax;
#pragma unroll
        for (int k = 0; k < 4; ++k) mx = fmaxf(mx, s_x[(k * 3 + 0) * pfw::kT + r]);
        const float mref = (mx == -INFINITY) ? 0.f : mx;
        float zsum = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) zsum += __expf(z[i] - mref);
        s_x[(cq * 3 + 1) * pfw::kT + r] = zsum;
        __syncthreads();
        if (cq == 0 && r < rows) {
            float tot = 0.f, zt = -INFINITY;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                tot += s_x[(k * 3 + 1) * pfw::kT + r];
                zt = fmaxf(zt, s_x[(k * 3 + 2) * pfw::kT + r]);
            }
            const bool any = tot > 0.f;
            const float lz = any ? mref + __logf(tot) : 0.f;
            logp[b * R + rt + r] = any ? zt - lz : 0.f;      // :190 log_softmax, :462 gather
            lse_z[b * R + rt + r] = lz;
        }
        tc_fence_before();
        __syncthreads();   // the row scalars, the exchange words and accumulator t & 1 are reused
    }
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ---- backward, glimpse part of the node pass on the tensor cores: dgk = dS^T (0.25 Q), dgv = P^T dO ------------------------------
// The TRANSPOSED score tile is computed so that P^T and dS^T come out with TMEM lane = node, which is the M dimension the two
// reductions over rows need (the flash-attention backward arrangement).  Per CTA (instance, row range), per tile of 128 rows,
// per head h:
//   S^T  = K_h (0.25 Q_h)^T   [nodes x rows]   3 MMAs (split fp16), A = K_h tile (resident), B = Q_h tile   -> TMEM [0,128)
//   dP^T = V_h dO_h^T          [nodes x rows]   3 MMAs,              A = V_h tile (resident), B = dO_h tile  -> TMEM [128,256)
//   thread (node, 64 of the tile's rows): P^T = exp(S^T - lse_h[row]) on the feasible (row, node) pairs, dS^T = P^T (dP^T - D_h[row]),
//     both written back IN PLACE as fp16 hi/lo TS-form A operands (16-column chunk: hi pairs [c, c+8), lo pairs [c+8, c+16))
//   dgv_h += P^T dO_h, dgk_h += dS^T (0.25 Q_h): 2 x 24 TS-form MMAs (N = 16, K = the tile's 128 rows) into TMEM [256 + 32 h, +32)
// The accumulators of the CTA's heads live in TMEM across all row tiles of the CTA.  (Column numbers above: KT = 128.)  The B operands of the second
// pair of products are the transposed head slices, produced by the same transposing conversion as in eval_dlk_tc_kernel.
// Template: KT = rows per tile (128 or 64), NH = heads per CTA (8, or 4 with blockIdx.z selecting the half).  <64, 4> needs
// 256 TMEM columns and 93 KB of shared memory: TWO CTAs per SM, whose serial stage -> convert -> MMA -> elementwise -> MMA
// chains overlap.
template <int KT, int NH>
struct Dkv {
    static constexpr uint32_t kHeadTile = 4096;              // [128 nodes x 16] fp16 K-major (hi or lo): 16 row groups x 2 chunks x 128 B
    static constexpr uint32_t kRowTile = KT * 32;            // [KT rows x 16] or [16 x KT rows] fp16 K-major (hi or lo)
    static constexpr uint32_t kOffK = 0;                     // K_h hi | lo for NH heads
    static constexpr uint32_t kOffV = NH * 2 * kHeadTile;
    static constexpr uint32_t kOffBq = 2 * kOffV;            // Q_h tile hi | lo   (B of S^T: N = rows, K = 16)
    static constexpr uint32_t kOffBdo = kOffBq + 2 * kRowTile;
    static constexpr uint32_t kOffBqT = kOffBdo + 2 * kRowTile;   // Q_h^T hi | lo: [16 dims x KT rows], 2 row groups x KT/8 chunks
    static constexpr uint32_t kOffBdoT = kOffBqT + 2 * kRowTile;
    static constexpr uint32_t kOffStage = kOffBdoT + 2 * kRowTile;   // fp32 head slices [2][KT][17]
    static constexpr uint32_t kStagePitch = 17;
    static constexpr uint32_t kOffLse = kOffStage + 2 * KT * kStagePitch * 4;   // [NH][KT] f32
    static constexpr uint32_t kOffDd = kOffLse + NH * KT * 4;
    static constexpr uint32_t kOffMaskW = kOffDd + NH * KT * 4;                 // [KT rows][4] u32
    static constexpr uint32_t kOffColMask = kOffMaskW + KT * 16;                // [128 nodes][KT / 32] u32
    static constexpr uint32_t kOffBar = kOffColMask + 128 * (KT / 32) * 4;
    static constexpr size_t kSmem = kOffBar + 64;
    static constexpr uint32_t kTmemCols = (2 * KT + 32 * NH) <= 256 ? 256 : 512;
};

template <int KT, int NH>
__global__ void __launch_bounds__(256, (KT == 64 && NH == 4) ? 2 : 1)
    eval_dkv_tc_kernel(const float* __restrict__ q, const float* __restrict__ gk, const float* __restrict__ gv,
                       const uint32_t* __restrict__ maskw, const float* __restrict__ lse_h, const float* __restrict__ dheads,
                       const float* __restrict__ dd, const float* __restrict__ gscale, float* __restrict__ dgk,
                       float* __restrict__ dgv, int64_t R, int n) {
    using namespace tc;
    using L = Dkv<KT, NH>;
    constexpr int NW = KT / 32;                          // 32-row words per tile
    constexpr int CPT = KT / 2;                          // tile rows (TMEM columns) per thread
    extern __shared__ __align__(128) unsigned char smb[];
    float* stage = reinterpret_cast<float*>(smb + L::kOffStage);
    float* s_lse = reinterpret_cast<float*>(smb + L::kOffLse);
    float* s_dd = reinterpret_cast<float*>(smb + L::kOffDd);
    uint32_t* s_mw = reinterpret_cast<uint32_t*>(smb + L::kOffMaskW);
    uint32_t* s_cm = reinterpret_cast<uint32_t*>(smb + L::kOffColMask);
    uint64_t* bar1 = reinterpret_cast<uint64_t*>(smb + L::kOffBar);
    uint64_t* bar2 = bar1 + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smb + L::kOffBar + 32);
    const int64_t b = blockIdx.x;
    const int h0 = (int)blockIdx.z * NH;                 // this CTA's heads [h0, h0 + NH)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int64_t r0, r1;
    row_range(R, r0, r1);
    if (r0 >= r1) return;
    const int ntiles = (int)((r1 - r0 + KT - 1) / KT);
    if (tid == 0) {
        mbar_init(bar1, 1);
        mbar_init(bar2, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(tmem_slot, L::kTmemCols);
    // resident A operands: K_h and V_h of the CTA's heads, [128 nodes x 16 dims] K-major hi | lo (nodes >= n are zero rows)
    for (int i = tid; i < 2 * NH * 128 * 2; i += 256) {
        const int tab = i / (NH * 256), hl = (i >> 8) % NH, m = (i >> 1) & 127, c = i & 1;
        const float* src = (tab ? gv : gk) + ((size_t)b * n + m) * kE + 16 * (h0 + hl) + 8 * c;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bb = a;
        if (m < n) {
            a = ld4(src);
            bb = ld4(src + 4);
        }
        uint32_t hi[4], lo[4];
        split2(a.x, a.y, hi[0], lo[0]);
        split2(a.z, a.w, hi[1], lo[1]);
        split2(bb.x, bb.y, hi[2], lo[2]);
        split2(bb.z, bb.w, hi[3], lo[3]);
        unsigned char* dst = smb + (tab ? L::kOffV : L::kOffK) + (uint32_t)hl * 2u * L::kHeadTile + (uint32_t)(m >> 3) * 256u +
                             (uint32_t)c * 128u + (uint32_t)(m & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + L::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t sb = smem_u32(smb);
    const int q4 = warp & 3, chalf = warp >> 2;          // TMEM lane quadrant (node = 32 q4 + lane), row half of the tile
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q4) << 16);
    uint32_t p1 = 0, p2 = 0;
    bool mma2_pending = false;
    for (int t = 0; t < ntiles; ++t) {
        const int64_t rt = r0 + (int64_t)t * KT;
        const int rows = (int)(r1 - rt < KT ? r1 - rt : KT);
        // per-row vectors of the tile: lse and D head-major, mask words (zero beyond the valid rows)
        for (int i = tid; i < KT * NH; i += 256) {
            const int r = i / NH, hl = i % NH;
            const bool ok = r < rows;
            s_lse[hl * KT + r] = ok ? lse_h[(b * R + rt + r) * kH + h0 + hl] : 0.f;
            s_dd[hl * KT + r] = ok ? dd[(b * R + rt + r) * kH + h0 + hl] : 0.f;
        }
        for (int i = tid; i < KT * 4; i += 256) s_mw[i] = (i >> 2) < rows ? maskw[(b * R + rt) * kNB + i] : 0u;
        __syncthreads();
        // feasibility transposed to (node, 32-row word): warp w takes row word w % NW and a block of 128 NW / 8 nodes
        {
            constexpr int NPB = 128 * NW / 8;            // nodes per warp
            const int j = warp % NW, nb0 = (warp / NW) * NPB;
            uint32_t mw[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) mw[k] = s_mw[(32 * j + lane) * 4 + k];
            for (int nd = nb0; nd < nb0 + NPB; ++nd) {
                const uint32_t word = (nd >> 5) == 0 ? mw[0] : ((nd >> 5) == 1 ? mw[1] : ((nd >> 5) == 2 ? mw[2] : mw[3]));
                const uint32_t bal = __ballot_sync(kFull, (word >> (nd & 31)) & 1u);
                if (lane == 0) s_cm[nd * NW + j] = bal;
            }
        }
        __syncthreads();
        const int node = 32 * q4 + lane;
        uint32_t cm[CPT / 32 > 0 ? CPT / 32 : 1];        // this thread's CPT rows: CPT = 64 -> two words, 32 -> one
#pragma unroll
        for (int k = 0; k < (CPT >= 32 ? CPT / 32 : 1); ++k) cm[k] = s_cm[node * NW + (CPT >= 32 ? (CPT / 32) * chalf + k : 0)];
        // fp32 head slices of q and dheads of the tile's rows: loaded into registers one head ahead (the loads of head h + 1
        // fly under the MMAs and the elementwise pass of head h), then staged as [KT][17]
        constexpr int NPRE = KT * 8 / 256;
        float4 pre[NPRE];
        auto prefetch = [&](int h) {
#pragma unroll
            for (int k = 0; k < NPRE; ++k) {
                const int i = tid + 256 * k;
                const int which = i / (KT * 4), r = (i >> 2) % KT, c4 = i & 3;
                pre[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < rows) pre[k] = ld4((which ? dheads : q) + (b * R + rt + r) * kE + 16 * h + 4 * c4);
            }
        };
        prefetch(h0);
        for (int hl = 0; hl < NH; ++hl) {
            const int h = h0 + hl;
#pragma unroll
            for (int k = 0; k < NPRE; ++k) {
                const int i = tid + 256 * k;
                const int which = i / (KT * 4), r = (i >> 2) % KT, c4 = i & 3;
                float* d = stage + (which * KT + r) * L::kStagePitch + 4 * c4;
                d[0] = pre[k].x, d[1] = pre[k].y, d[2] = pre[k].z, d[3] = pre[k].w;
            }
            if (hl + 1 < NH) prefetch(h + 1);
            // (double-buffering the four operand tiles below over consecutive heads was measured: 2.03 against 1.88 ms)
            if (mma2_pending) {                          // the TS MMAs of the previous head have read BqT / BdoT (and Bq / Bdo)
                mbar_wait(bar2, p2);
                p2 ^= 1u;
                mma2_pending = false;
            }
            __syncthreads();
            // K-major operand tiles: (a) [rows x 16] for the S^T / dP^T products, (b) transposed [16 x rows] for dgk / dgv
            if (tid < 2 * KT) {
                const int which = tid / KT, r = tid % KT;                        // (a): thread = (tensor, row), two 8-dim chunks
                const float sc = which ? 1.0f : 0.25f;
                const float* src = stage + (which * KT + r) * L::kStagePitch;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int w = 0; w < 4; ++w) split2(src[8 * c + 2 * w] * sc, src[8 * c + 2 * w + 1] * sc, hi[w], lo[w]);
                    unsigned char* dst = smb + (which ? L::kOffBdo : L::kOffBq) + (uint32_t)(r >> 3) * 256u + (uint32_t)c * 128u +
                                         (uint32_t)(r & 7) * 16u;
                    *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    *reinterpret_cast<uint4*>(dst + L::kRowTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                }
            }
#pragma unroll
            for (int jj = 0; jj < KT * 4 / 256; ++jj) {                          // (b): item = (tensor, 8-row group g, dim)
                const int i = tid + 256 * jj;
                const int which = i / (KT * 2), g = (i >> 4) % (KT / 8), dm = i & 15;
                const float sc = which ? 1.0f : 0.25f;
                const float* src = stage + (which * KT + 8 * g) * L::kStagePitch + dm;
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int w = 0; w < 4; ++w)
                    split2(src[(2 * w) * L::kStagePitch] * sc, src[(2 * w + 1) * L::kStagePitch] * sc, hi[w], lo[w]);
                unsigned char* dst = smb + (which ? L::kOffBdoT : L::kOffBqT) + (uint32_t)(dm >> 3) * (uint32_t)(KT * 16) +
                                     (uint32_t)g * 128u + (uint32_t)(dm & 7) * 16u;
                *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(dst + L::kRowTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
            fence_proxy_async();
            __syncthreads();
            if (warp == 0) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(128, KT);
                    const uint32_t ko = sb + L::kOffK + (uint32_t)hl * 2u * L::kHeadTile, vo = sb + L::kOffV + (uint32_t)hl * 2u * L::kHeadTile;
                    const uint64_t k_hi = smem_desc(ko, 128, 256), k_lo = smem_desc(ko + L::kHeadTile, 128, 256);
                    const uint64_t v_hi = smem_desc(vo, 128, 256), v_lo = smem_desc(vo + L::kHeadTile, 128, 256);
                    const uint64_t q_hi = smem_desc(sb + L::kOffBq, 128, 256), q_lo = smem_desc(sb + L::kOffBq + L::kRowTile, 128, 256);
                    const uint64_t o_hi = smem_desc(sb + L::kOffBdo, 128, 256), o_lo = smem_desc(sb + L::kOffBdo + L::kRowTile, 128, 256);
                    mma_ss(tmem, k_hi, q_hi, id, 0u);                            // S^T
                    mma_ss(tmem, k_hi, q_lo, id, 1u);
                    mma_ss(tmem, k_lo, q_hi, id, 1u);
                    mma_ss(tmem + (uint32_t)KT, v_hi, o_hi, id, 0u);             // dP^T
                    mma_ss(tmem + (uint32_t)KT, v_hi, o_lo, id, 1u);
                    mma_ss(tmem + (uint32_t)KT, v_lo, o_hi, id, 1u);
                    tc_commit(bar1);
                }
                __syncwarp();
            }
            mbar_wait(bar1, p1);
            p1 ^= 1u;
            tc_fence_after();
            // thread = (node, rows [CPT chalf, +CPT) of the tile): P^T and dS^T in place
#pragma unroll 1
            for (int c = 0; c < CPT; c += 16) {
                const int col = CPT * chalf + c;
                uint32_t sv[16], dv[16];
                tmem_ld16(lane_base + (uint32_t)col, sv);
                tmem_ld16(lane_base + (uint32_t)KT + (uint32_t)col, dv);
                tc_wait_ld();
                const uint32_t bits = (cm[CPT > 32 ? (c >> 5) : 0] >> (c & 31)) & 0xffffu;
                uint32_t phi[8], plo[8], dhi[8], dlo[8];
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    const float4 l4 = ld4(s_lse + hl * KT + col + i), d4 = ld4(s_dd + hl * KT + col + i);
                    const float pa = (bits >> i) & 1u ? fexp(__uint_as_float(sv[i]) - l4.x) : 0.f;
                    const float pb = (bits >> (i + 1)) & 1u ? fexp(__uint_as_float(sv[i + 1]) - l4.y) : 0.f;
                    const float pc = (bits >> (i + 2)) & 1u ? fexp(__uint_as_float(sv[i + 2]) - l4.z) : 0.f;
                    const float pd = (bits >> (i + 3)) & 1u ? fexp(__uint_as_float(sv[i + 3]) - l4.w) : 0.f;
                    split2(pa, pb, phi[i >> 1], plo[i >> 1]);
                    split2(pc, pd, phi[(i >> 1) + 1], plo[(i >> 1) + 1]);
                    split2(pa * (__uint_as_float(dv[i]) - d4.x), pb * (__uint_as_float(dv[i + 1]) - d4.y), dhi[i >> 1], dlo[i >> 1]);
                    split2(pc * (__uint_as_float(dv[i + 2]) - d4.z), pd * (__uint_as_float(dv[i + 3]) - d4.w), dhi[(i >> 1) + 1],
                           dlo[(i >> 1) + 1]);
                }
                tmem_st8(lane_base + (uint32_t)col, phi);
                tmem_st8(lane_base + (uint32_t)col + 8u, plo);
                tmem_st8(lane_base + (uint32_t)KT + (uint32_t)col, dhi);
                tmem_st8(lane_base + (uint32_t)KT + (uint32_t)col + 8u, dlo);
            }
            tc_wait_st();
            tc_fence_before();
            __syncthreads();
            if (warp == 0) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(128, 16);
                    uint64_t ot_hi = smem_desc(sb + L::kOffBdoT, 128, KT * 16), ot_lo = smem_desc(sb + L::kOffBdoT + L::kRowTile, 128, KT * 16);
                    uint64_t qt_hi = smem_desc(sb + L::kOffBqT, 128, KT * 16), qt_lo = smem_desc(sb + L::kOffBqT + L::kRowTile, 128, KT * 16);
                    const uint32_t dv_col = tmem + 2u * KT + 32u * (uint32_t)hl, dk_col = dv_col + 16u;
#pragma unroll
                    for (int ks = 0; ks < KT / 16; ++ks) {
                        const uint32_t acc = (t == 0 && ks == 0) ? 0u : 1u;
                        mma_ts(dv_col, tmem + 16u * ks, ot_hi, id, acc);          // dgv_h += P^T dO_h
                        mma_ts(dv_col, tmem + 16u * ks, ot_lo, id, 1u);
                        mma_ts(dv_col, tmem + 16u * ks + 8u, ot_hi, id, 1u);
                        mma_ts(dk_col, tmem + (uint32_t)KT + 16u * ks, qt_hi, id, acc);   // dgk_h += dS^T (0.25 Q_h)
                        mma_ts(dk_col, tmem + (uint32_t)KT + 16u * ks, qt_lo, id, 1u);
                        mma_ts(dk_col, tmem + (uint32_t)KT + 16u * ks + 8u, qt_hi, id, 1u);
                        ot_hi += 256u >> 4;
                        ot_lo += 256u >> 4;
                        qt_hi += 256u >> 4;
                        qt_lo += 256u >> 4;
                    }
                    tc_commit(bar2);
                }
                __syncwarp();
            }
            mma2_pending = true;
        }
        __syncthreads();   // s_lse / s_dd / masks of this tile are rewritten by the next one
    }
    if (mma2_pending) mbar_wait(bar2, p2);
    tc_fence_after();
    {   // accumulators: lane = node, columns 2 KT + 32 hl + (0: dgv, 16: dgk) + d; warp halves split the heads
        const int node = 32 * q4 + lane;
#pragma unroll 1
        for (int hl = (NH / 2) * chalf; hl < (NH / 2) * chalf + NH / 2; ++hl) {
            uint32_t v[16], k[16];
            tmem_ld16(lane_base + 2u * KT + 32u * (uint32_t)hl, v);
            tmem_ld16(lane_base + 2u * KT + 32u * (uint32_t)hl + 16u, k);
            tc_wait_ld();
            if (node < n) {
                const float inv_s = gscale[1];
                float* pv = dgv + ((size_t)b * n + node) * kE + 16 * (h0 + hl);
                float* pk = dgk + ((size_t)b * n + node) * kE + 16 * (h0 + hl);
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    atomicAdd(pv + i, __uint_as_float(v[i]) * inv_s);
                    atomicAdd(pk + i, __uint_as_float(k[i]) * inv_s);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, L::kTmemCols);
}

// ---- mask bytes [rows, n] -> bit words [rows, 4] (the form every evaluate kernel reads) ----------------------------------------
__global__ void __launch_bounds__(256)
    eval_pack_mask_kernel(const uint8_t* __restrict__ mask, uint32_t* __restrict__ maskw, int64_t rows, int n) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < rows; row += nw) {
        uint32_t w[kNB];
#pragma unroll
        for (int blk = 0; blk < kNB; ++blk) {
            const int i = 32 * blk + lane;
            w[blk] = __ballot_sync(kFull, i < n && mask[row * n + i] != 0);
        }
        if (lane < kNB) maskw[row * kNB + lane] = sel4u(w, lane);
    }
}

// ---- forward, glimpse on the tensor cores: heads = softmax(0.25 Q_h K_h^T + mask) V_h, lse_h ------------------------------------
// CTA = (instance, row range, four heads); per tile of 128 rows and head h:
//   S = (0.25 Q_h) K_h^T  [rows x nodes]  3 split-fp16 MMAs, A = the tile's Q_h slice, B = the resident K_h tile -> TMEM [0, npad)
//   thread (row, 64 nodes): row maximum over the feasible nodes (the two threads of a row exchange through shared memory),
//     P = exp(S - max) written back IN PLACE as the fp16 hi / lo TS-form A operand, row sum l
//   O_h = P V_h           [rows x 16]     3 x npad / 16 TS-form MMAs (N = 16), B = the resident V_h^T tile -> TMEM [128 + 16 h, +16)
//   heads[row, 16 h ..] = O_h / l, lse_h = max + log l: stored while the next head's S product runs.
// 256 TMEM columns and ~100 KB of shared memory: two CTAs per SM whose serial S -> softmax -> O chains overlap.
namespace gfw {
constexpr int kNH = 4;
constexpr uint32_t kHeadTile = 4096;                  // [128 x 16] or [16 x 128] fp16 K-major (hi or lo)
constexpr uint32_t kOffK = 0;                         // K_h   [node][dim] hi | lo per head (B of S: N = nodes, K = 16)
constexpr uint32_t kOffVT = kNH * 2 * kHeadTile;      // V_h^T [dim][node] hi | lo per head (B of O: N = 16, K = nodes)
constexpr uint32_t kOffQ = 2 * kOffVT;                // Q_h   [row][dim]  hi | lo per head (A of S), rewritten per tile
constexpr uint32_t kOffMaskW = 3 * kOffVT;            // [128 rows][4] u32
constexpr uint32_t kOffX = kOffMaskW + 128 * 16;      // exchange between the two threads of a row: max [2][128], sum [2][128]
constexpr uint32_t kOffBar = kOffX + 4 * 128 * 4;
constexpr size_t kSmem = kOffBar + 64;
}  // namespace gfw

__global__ void __launch_bounds__(256, 2)
    eval_glimpse_fwd_tc_kernel(const float* __restrict__ q, const float* __restrict__ gk, const float* __restrict__ gv,
                               const uint32_t* __restrict__ maskw, float* __restrict__ heads, float* __restrict__ lse_h,
                               int64_t R, int n) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smb[];
    uint32_t* s_mw = reinterpret_cast<uint32_t*>(smb + gfw::kOffMaskW);
    float* s_max = reinterpret_cast<float*>(smb + gfw::kOffX);
    float* s_sum = s_max + 256;
    uint64_t* bar1 = reinterpret_cast<uint64_t*>(smb + gfw::kOffBar);
    uint64_t* bar2 = bar1 + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smb + gfw::kOffBar + 32);
    const int64_t b = blockIdx.x;
    const int h0 = (int)blockIdx.z * gfw::kNH;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int64_t r0, r1;
    row_range(R, r0, r1);
    if (r0 >= r1) return;
    const int ntiles = (int)((r1 - r0 + 127) / 128);
    const int npad = (n + 15) & ~15;
    if (tid == 0) {
        mbar_init(bar1, 1);
        mbar_init(bar2, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(tmem_slot, 256);
    // resident B operands (nodes >= n are zero).  K_h: item = (head, node m, 8-dim chunk c), lanes over m & 7 first (item_rc)
    for (int i = tid; i < gfw::kNH * 256; i += 256) {
        const int hl = i >> 8, m = 8 * ((i >> 4) & 15) + (i & 7), c = (i >> 3) & 1;
        const float* src = gk + ((size_t)b * n + m) * kE + 16 * (h0 + hl) + 8 * c;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bb = a;
        if (m < n) {
            a = ld4(src);
            bb = ld4(src + 4);
        }
        uint32_t hi[4], lo[4];
        split2(a.x, a.y, hi[0], lo[0]);
        split2(a.z, a.w, hi[1], lo[1]);
        split2(bb.x, bb.y, hi[2], lo[2]);
        split2(bb.z, bb.w, hi[3], lo[3]);
        unsigned char* dst = smb + gfw::kOffK + (uint32_t)hl * 2u * gfw::kHeadTile + (uint32_t)(m >> 3) * 256u + (uint32_t)c * 128u +
                             (uint32_t)(m & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + gfw::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    // V_h^T: item = (head, 8-node group g, dim dm); the conversion is the transposition (8 nodes of one dim = one K chunk)
    for (int i = tid; i < gfw::kNH * 256; i += 256) {
        const int hl = i >> 8, g = (i >> 4) & 15, dm = i & 15;
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = (8 * g + k < n) ? gv[((size_t)b * n + 8 * g + k) * kE + 16 * (h0 + hl) + dm] : 0.f;
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) split2(v[2 * w], v[2 * w + 1], hi[w], lo[w]);
        unsigned char* dst = smb + gfw::kOffVT + (uint32_t)hl * 2u * gfw::kHeadTile + (uint32_t)(dm >> 3) * 2048u + (uint32_t)g * 128u +
                             (uint32_t)(dm & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + gfw::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    // the tile's Q slices (4 heads x 128 rows x 2 chunks = 4 items per thread, item k = head k) and mask words travel through
    // registers one tile ahead
    float4 pre[8];
    uint32_t pre_mw[4] = {0u, 0u, 0u, 0u};
    auto prefetch = [&](int t) {
        const int64_t rt = r0 + (int64_t)t * 128;
        const int rows = t < ntiles ? (int)(r1 - rt < 128 ? r1 - rt : 128) : 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = 8 * ((tid >> 4) & 15) + (tid & 7), c = (tid >> 3) & 1;
            pre[2 * k] = pre[2 * k + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < rows) {
                const float* src = q + (b * R + rt + r) * kE + 16 * (h0 + k) + 8 * c;
                pre[2 * k] = ld4(src);
                pre[2 * k + 1] = ld4(src + 4);
            }
        }
        if (tid < 128) {
            const bool ok = tid < rows;
#pragma unroll
            for (int k = 0; k < 4; ++k) pre_mw[k] = ok ? maskw[(b * R + rt + tid) * kNB + k] : 0u;
        }
    };
    prefetch(0);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t sb = smem_u32(smb);
    const int q4 = warp & 3, chalf = warp >> 2;          // TMEM lane quadrant (row = 32 q4 + lane), node half
    const int r = 32 * q4 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q4) << 16);
    uint32_t p1 = 0, p2 = 0;
    bool pv_pending = false;
    // epilogue of the previous (tile, head): valid in the threads of node half 0
    float epi_inv = 0.f, epi_lse = 0.f;
    int64_t epi_rt = 0;
    int epi_rows = 0, epi_hl = 0;
    auto epilogue = [&]() {
        if (chalf == 0) {
            uint32_t v[16];
            tmem_ld16(lane_base + 128u + 16u * (uint32_t)epi_hl, v);
            tc_wait_ld();
            if (r < epi_rows) {
                float* dst = heads + (b * R + epi_rt + r) * kE + 16 * (h0 + epi_hl);
#pragma unroll
                for (int i = 0; i < 16; i += 4)
                    *reinterpret_cast<float4*>(dst + i) =
                        make_float4(__uint_as_float(v[i]) * epi_inv, __uint_as_float(v[i + 1]) * epi_inv,
                                    __uint_as_float(v[i + 2]) * epi_inv, __uint_as_float(v[i + 3]) * epi_inv);
                lse_h[(b * R + epi_rt + r) * kH + h0 + epi_hl] = epi_lse;
            }
        }
    };
    for (int t = 0; t < ntiles; ++t) {
        const int64_t rt = r0 + (int64_t)t * 128;
        const int rows = (int)(r1 - rt < 128 ? r1 - rt : 128);
        // (every S product of the previous tile has completed: its heads' bar1 were waited for)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int rr = 8 * ((tid >> 4) & 15) + (tid & 7), c = (tid >> 3) & 1;
            const float4 a = pre[2 * k], bb = pre[2 * k + 1];
            uint32_t hi[4], lo[4];
            split2(a.x * 0.25f, a.y * 0.25f, hi[0], lo[0]);
            split2(a.z * 0.25f, a.w * 0.25f, hi[1], lo[1]);
            split2(bb.x * 0.25f, bb.y * 0.25f, hi[2], lo[2]);
            split2(bb.z * 0.25f, bb.w * 0.25f, hi[3], lo[3]);
            unsigned char* dst = smb + gfw::kOffQ + (uint32_t)k * 2u * gfw::kHeadTile + (uint32_t)(rr >> 3) * 256u + (uint32_t)c * 128u +
                                 (uint32_t)(rr & 7) * 16u;
            *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + gfw::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        if (tid < 128) {
#pragma unroll
            for (int k = 0; k < 4; ++k) s_mw[tid * 4 + k] = pre_mw[k];
        }
        fence_proxy_async();
        __syncthreads();
        prefetch(t + 1);
        const uint32_t m0 = s_mw[r * 4 + 2 * chalf], m1 = s_mw[r * 4 + 2 * chalf + 1];
        for (int hl = 0; hl < gfw::kNH; ++hl) {
            if (pv_pending) {                            // O of the previous head is complete and P has been read
                mbar_wait(bar2, p2);
                p2 ^= 1u;
                tc_fence_after();
            }
            if (warp == 0) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(128, npad);
                    const uint32_t qo = sb + gfw::kOffQ + (uint32_t)hl * 2u * gfw::kHeadTile;
                    const uint32_t ko = sb + gfw::kOffK + (uint32_t)hl * 2u * gfw::kHeadTile;
                    const uint64_t q_hi = smem_desc(qo, 128, 256), q_lo = smem_desc(qo + gfw::kHeadTile, 128, 256);
                    const uint64_t k_hi = smem_desc(ko, 128, 256), k_lo = smem_desc(ko + gfw::kHeadTile, 128, 256);
                    mma_ss(tmem, q_hi, k_hi, id, 0u);
                    mma_ss(tmem, q_hi, k_lo, id, 1u);
                    mma_ss(tmem, q_lo, k_hi, id, 1u);
                    tc_commit(bar1);
                }
                __syncwarp();
            }
            if (pv_pending) {
                epilogue();
                pv_pending = false;
            }
            mbar_wait(bar1, p1);
            p1 ^= 1u;
            tc_fence_after();
            // pass 1: maximum over this thread's feasible nodes
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < 64; c += 16) {
                const int col = 64 * chalf + c;
                if (col >= npad) break;
                uint32_t sv[16];
                tmem_ld16(lane_base + (uint32_t)col, sv);
                tc_wait_ld();
                const uint32_t bits = ((c < 32 ? m0 : m1) >> (c & 31)) & 0xffffu;
#pragma unroll
                for (int i = 0; i < 16; ++i) mx = fmaxf(mx, (bits >> i) & 1u ? __uint_as_float(sv[i]) : -INFINITY);
            }
            s_max[chalf * 128 + r] = mx;
            __syncthreads();
            mx = fmaxf(mx, s_max[(chalf ^ 1) * 128 + r]);
            const float mref = (mx == -INFINITY) ? 0.f : mx;
            // pass 2: P = exp(S - max) in place (fp16 hi | lo per 16-column chunk), row sum
            float lsum = 0.f;
#pragma unroll 1
            for (int c = 0; c < 64; c += 16) {
                const int col = 64 * chalf + c;
                if (col >= npad) break;
                uint32_t sv[16];
                tmem_ld16(lane_base + (uint32_t)col, sv);
                tc_wait_ld();
                const uint32_t bits = ((c < 32 ? m0 : m1) >> (c & 31)) & 0xffffu;
                uint32_t phi[8], plo[8];
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    const float pa = (bits >> i) & 1u ? fexp(__uint_as_float(sv[i]) - mref) : 0.f;
                    const float pb = (bits >> (i + 1)) & 1u ? fexp(__uint_as_float(sv[i + 1]) - mref) : 0.f;
                    lsum += pa + pb;
                    split2(pa, pb, phi[i >> 1], plo[i >> 1]);
                }
                tmem_st8(lane_base + (uint32_t)col, phi);
                tmem_st8(lane_base + (uint32_t)col + 8u, plo);
            }
            s_sum[chalf * 128 + r] = lsum;
            tc_wait_st();
            tc_fence_before();
            __syncthreads();
            if (chalf == 0) {
                const float l = lsum + s_sum[128 + r];
                const bool any = l > 0.f;
                epi_inv = any ? 1.0f / l : 0.f;
                epi_lse = any ? mref + __logf(l) : 0.f;
            }
            epi_rt = rt;
            epi_rows = rows;
            epi_hl = hl;
            if (warp == 0) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(128, 16);
                    const uint32_t vo = sb + gfw::kOffVT + (uint32_t)hl * 2u * gfw::kHeadTile;
                    uint64_t v_hi = smem_desc(vo, 128, 2048), v_lo = smem_desc(vo + gfw::kHeadTile, 128, 2048);
                    const uint32_t d = tmem + 128u + 16u * (uint32_t)hl;
                    for (int ks = 0; ks < npad / 16; ++ks) {
                        mma_ts(d, tmem + 16u * ks, v_hi, id, ks == 0 ? 0u : 1u);   // O_h = P V_h
                        mma_ts(d, tmem + 16u * ks, v_lo, id, 1u);
                        mma_ts(d, tmem + 16u * ks + 8u, v_hi, id, 1u);
                        v_hi += 256u >> 4;
                        v_lo += 256u >> 4;
                    }
                    tc_commit(bar2);
                }
                __syncwarp();
            }
            pv_pending = true;
        }
    }
    if (pv_pending) {
        mbar_wait(bar2, p2);
        tc_fence_after();
        epilogue();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ---- backward, glimpse part of the row pass on the tensor cores: dq = 0.25 dS K, D = dO . O ---------------------------------------
// CTA = (instance, row range, TWO heads); per tile of 128 rows and head h (TMEM lane = row):
//   S  = (0.25 Q_h) K_h^T and dP = dO_h V_h^T   [rows x nodes]   3 + 3 split-fp16 MMAs against the resident K_h / V_h tiles
//   thread (row, 64 nodes): P = exp(S - lse_h[row]) on the feasible nodes (the saved log-sum-exp: no maximum pass),
//     dS = P (dP - D_h[row]) written over S IN PLACE as the fp16 hi / lo TS-form A operand
//   dq_h = dS K_h   [rows x 16]   3 x npad / 16 TS-form MMAs (N = 16) against the resident K_h^T tile; stored (x 0.25) while the next
//     head's products run.  D_h[row] = dO_h . O_h comes from the staging of the tile (also written out: the dK / dV kernel reads it).
// TMEM: S at 0, dP at 112 (128 if n > 112), two dq slots behind them: 256 columns for n <= 112 (two CTAs per SM; ~84 KB of shared
// memory each), 512 beyond (the second CTA of an SM then waits for the first one's columns).
namespace dqk {
constexpr int kNH = 2;
constexpr uint32_t kHeadTile = 4096;                  // [128 x 16] or [16 x 128] fp16 K-major (hi or lo)
constexpr uint32_t kOffK = 0;                         // K_h   [node][dim] hi | lo per head (B of S)
constexpr uint32_t kOffV = kNH * 2 * kHeadTile;       // V_h   [node][dim] hi | lo per head (B of dP)
constexpr uint32_t kOffKT = 2 * kOffV;                // K_h^T [dim][node] hi | lo per head (B of dq)
constexpr uint32_t kOffQ = 3 * kOffV;                 // Q_h  tile [row][dim] hi | lo per head (A of S)
constexpr uint32_t kOffDO = 4 * kOffV;                // dO_h tile [row][dim] hi | lo per head (A of dP)
constexpr uint32_t kOffMaskW = 5 * kOffV;             // [128 rows][4] u32
constexpr uint32_t kOffLse = kOffMaskW + 128 * 16;    // [2 heads][128] f32
constexpr uint32_t kOffD = kOffLse + kNH * 128 * 4;   // [2 heads][128] f32
constexpr uint32_t kOffBar = kOffD + kNH * 128 * 4;
constexpr size_t kSmem = kOffBar + 64;
}  // namespace dqk

__global__ void __launch_bounds__(256, 2)
    eval_dq_tc_kernel(const float* __restrict__ q, const float* __restrict__ gk, const float* __restrict__ gv,
                      const uint32_t* __restrict__ maskw, const float* __restrict__ heads, const float* __restrict__ lse_h,
                      const float* __restrict__ dheads, const float* __restrict__ gscale, float* __restrict__ dq,
                      float* __restrict__ dd_out, int64_t R, int n) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smb[];
    uint32_t* s_mw = reinterpret_cast<uint32_t*>(smb + dqk::kOffMaskW);
    float* s_lse = reinterpret_cast<float*>(smb + dqk::kOffLse);
    float* s_d = reinterpret_cast<float*>(smb + dqk::kOffD);
    uint64_t* bar1 = reinterpret_cast<uint64_t*>(smb + dqk::kOffBar);
    uint64_t* bar2 = bar1 + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smb + dqk::kOffBar + 32);
    const int64_t b = blockIdx.x;
    const int h0 = (int)blockIdx.z * dqk::kNH;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int64_t r0, r1;
    row_range(R, r0, r1);
    if (r0 >= r1) return;
    const int ntiles = (int)((r1 - r0 + 127) / 128);
    const int npad = (n + 15) & ~15;
    const uint32_t col_dp = npad <= 112 ? 112u : 128u, col_dq = 2u * col_dp, tmem_cols = npad <= 112 ? 256u : 512u;
    if (tid == 0) {
        mbar_init(bar1, 1);
        mbar_init(bar2, 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(tmem_slot, tmem_cols);
    // resident B operands (nodes >= n are zero).  K_h, V_h: item = (table, head, node m, 8-dim chunk c), lanes over m & 7 first
    for (int i = tid; i < 2 * dqk::kNH * 256; i += 256) {
        const int tab = i >> 9, hl = (i >> 8) & 1, m = 8 * ((i >> 4) & 15) + (i & 7), c = (i >> 3) & 1;
        const float* src = (tab ? gv : gk) + ((size_t)b * n + m) * kE + 16 * (h0 + hl) + 8 * c;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bb = a;
        if (m < n) {
            a = ld4(src);
            bb = ld4(src + 4);
        }
        uint32_t hi[4], lo[4];
        split2(a.x, a.y, hi[0], lo[0]);
        split2(a.z, a.w, hi[1], lo[1]);
        split2(bb.x, bb.y, hi[2], lo[2]);
        split2(bb.z, bb.w, hi[3], lo[3]);
        unsigned char* dst = smb + (tab ? dqk::kOffV : dqk::kOffK) + (uint32_t)hl * 2u * dqk::kHeadTile + (uint32_t)(m >> 3) * 256u +
                             (uint32_t)c * 128u + (uint32_t)(m & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + dqk::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    // K_h^T: item = (head, 8-node group g, dim dm): 8 nodes of one dim = one K chunk
    for (int i = tid; i < dqk::kNH * 256; i += 256) {
        const int hl = i >> 8, g = (i >> 4) & 15, dm = i & 15;
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = (8 * g + k < n) ? gk[((size_t)b * n + 8 * g + k) * kE + 16 * (h0 + hl) + dm] : 0.f;
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) split2(v[2 * w], v[2 * w + 1], hi[w], lo[w]);
        unsigned char* dst = smb + dqk::kOffKT + (uint32_t)hl * 2u * dqk::kHeadTile + (uint32_t)(dm >> 3) * 2048u + (uint32_t)g * 128u +
                             (uint32_t)(dm & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + dqk::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    // per tile: the head slices of q, dheads and heads (2 heads x 128 rows x 2 chunks = 2 items per thread and tensor, item k =
    // head k), lse (thread = (head, row)) and the mask words travel through registers one tile ahead
    float4 pq[4], pg[4], po[4];
    float pre_lse = 0.f;
    uint32_t pre_mw[4] = {0u, 0u, 0u, 0u};
    const int sr = 8 * ((tid >> 4) & 15) + (tid & 7), sc = (tid >> 3) & 1;     // staging row and chunk of this thread
    auto prefetch = [&](int t) {
        const int64_t rt = r0 + (int64_t)t * 128;
        const int rows = t < ntiles ? (int)(r1 - rt < 128 ? r1 - rt : 128) : 0;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            pq[2 * k] = pq[2 * k + 1] = pg[2 * k] = pg[2 * k + 1] = po[2 * k] = po[2 * k + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (sr < rows) {
                const int64_t off = (b * R + rt + sr) * kE + 16 * (h0 + k) + 8 * sc;
                pq[2 * k] = ld4(q + off);
                pq[2 * k + 1] = ld4(q + off + 4);
                pg[2 * k] = ld4(dheads + off);
                pg[2 * k + 1] = ld4(dheads + off + 4);
                po[2 * k] = ld4(heads + off);
                po[2 * k + 1] = ld4(heads + off + 4);
            }
        }
        {
            const int hl = tid >> 7, rr = tid & 127;
            pre_lse = rr < rows ? lse_h[(b * R + rt + rr) * kH + h0 + hl] : 0.f;
        }
        if (tid < 128) {
            const bool ok = tid < rows;
#pragma unroll
            for (int k = 0; k < 4; ++k) pre_mw[k] = ok ? maskw[(b * R + rt + tid) * kNB + k] : 0u;
        }
    };
    prefetch(0);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t sb = smem_u32(smb);
    const int q4 = warp & 3, chalf = warp >> 2;          // TMEM lane quadrant (row = 32 q4 + lane), node half
    const int r = 32 * q4 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q4) << 16);
    uint32_t p1 = 0, p2 = 0;
    bool pending = false;
    int64_t epi_rt = 0;
    int epi_rows = 0, epi_h = 0, cnt = 0;
    const float out_sc = 0.25f * gscale[1];              // the query scale and the inverse gradient scale
    auto epilogue = [&]() {   // dq of the previous (tile, head): slot (cnt - 1) & 1, 8 of its 16 columns per node half
        uint32_t v[8];
        tmem_ld8(lane_base + col_dq + 16u * (uint32_t)((cnt - 1) & 1) + 8u * (uint32_t)chalf, v);
        tc_wait_ld();
        if (r < epi_rows) {
            float* dst = dq + (b * R + epi_rt + r) * kE + 16 * epi_h + 8 * chalf;
            *reinterpret_cast<float4*>(dst) = make_float4(out_sc * __uint_as_float(v[0]), out_sc * __uint_as_float(v[1]),
                                                          out_sc * __uint_as_float(v[2]), out_sc * __uint_as_float(v[3]));
            *reinterpret_cast<float4*>(dst + 4) = make_float4(out_sc * __uint_as_float(v[4]), out_sc * __uint_as_float(v[5]),
                                                              out_sc * __uint_as_float(v[6]), out_sc * __uint_as_float(v[7]));
        }
    };
    for (int t = 0; t < ntiles; ++t) {
        const int64_t rt = r0 + (int64_t)t * 128;
        const int rows = (int)(r1 - rt < 128 ? r1 - rt : 128);
        // (the S / dP products of the previous tile have completed: their bar1 were waited for)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float4 a = pq[2 * k], bb = pq[2 * k + 1], g0 = pg[2 * k], g1 = pg[2 * k + 1], o0 = po[2 * k], o1 = po[2 * k + 1];
            uint32_t hi[4], lo[4];
            split2(a.x * 0.25f, a.y * 0.25f, hi[0], lo[0]);
            split2(a.z * 0.25f, a.w * 0.25f, hi[1], lo[1]);
            split2(bb.x * 0.25f, bb.y * 0.25f, hi[2], lo[2]);
            split2(bb.z * 0.25f, bb.w * 0.25f, hi[3], lo[3]);
            const uint32_t off = (uint32_t)k * 2u * dqk::kHeadTile + (uint32_t)(sr >> 3) * 256u + (uint32_t)sc * 128u + (uint32_t)(sr & 7) * 16u;
            *reinterpret_cast<uint4*>(smb + dqk::kOffQ + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(smb + dqk::kOffQ + off + dqk::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            split2(g0.x, g0.y, hi[0], lo[0]);
            split2(g0.z, g0.w, hi[1], lo[1]);
            split2(g1.x, g1.y, hi[2], lo[2]);
            split2(g1.z, g1.w, hi[3], lo[3]);
            *reinterpret_cast<uint4*>(smb + dqk::kOffDO + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(smb + dqk::kOffDO + off + dqk::kHeadTile) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            // D = dO . O: this thread's 8 channels, the other chunk of the row is 8 lanes away
            float dpart = g0.x * o0.x + g0.y * o0.y + g0.z * o0.z + g0.w * o0.w + g1.x * o1.x + g1.y * o1.y + g1.z * o1.z + g1.w * o1.w;
            dpart += __shfl_xor_sync(kFull, dpart, 8);
            if (sc == 0) {
                s_d[k * 128 + sr] = dpart;
                if (sr < rows) dd_out[(b * R + rt + sr) * kH + h0 + k] = dpart;
            }
        }
        s_lse[tid] = pre_lse;                              // tid = head * 128 + row
        if (tid < 128) {
#pragma unroll
            for (int k = 0; k < 4; ++k) s_mw[tid * 4 + k] = pre_mw[k];
        }
        fence_proxy_async();
        __syncthreads();
        prefetch(t + 1);
        const uint32_t m0 = s_mw[r * 4 + 2 * chalf], m1 = s_mw[r * 4 + 2 * chalf + 1];
        for (int hl = 0; hl < dqk::kNH; ++hl) {
            if (pending) {                               // dq of the previous head is complete and dS has been read
                mbar_wait(bar2, p2);
                p2 ^= 1u;
                tc_fence_after();
            }
            if (warp == 0) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(128, npad);
                    const uint32_t ho = (uint32_t)hl * 2u * dqk::kHeadTile;
                    const uint64_t q_hi = smem_desc(sb + dqk::kOffQ + ho, 128, 256), q_lo = smem_desc(sb + dqk::kOffQ + ho + dqk::kHeadTile, 128, 256);
                    const uint64_t k_hi = smem_desc(sb + dqk::kOffK + ho, 128, 256), k_lo = smem_desc(sb + dqk::kOffK + ho + dqk::kHeadTile, 128, 256);
                    const uint64_t g_hi = smem_desc(sb + dqk::kOffDO + ho, 128, 256), g_lo = smem_desc(sb + dqk::kOffDO + ho + dqk::kHeadTile, 128, 256);
                    const uint64_t v_hi = smem_desc(sb + dqk::kOffV + ho, 128, 256), v_lo = smem_desc(sb + dqk::kOffV + ho + dqk::kHeadTile, 128, 256);
                    mma_ss(tmem, q_hi, k_hi, id, 0u);                              // S
                    mma_ss(tmem, q_hi, k_lo, id, 1u);
                    mma_ss(tmem, q_lo, k_hi, id, 1u);
                    mma_ss(tmem + col_dp, g_hi, v_hi, id, 0u);                     // dP
                    mma_ss(tmem + col_dp, g_hi, v_lo, id, 1u);
                    mma_ss(tmem + col_dp, g_lo, v_hi, id, 1u);
                    tc_commit(bar1);
                }
                __syncwarp();
            }
            if (pending) {
                epilogue();
                pending = false;
            }
            mbar_wait(bar1, p1);
            p1 ^= 1u;
            tc_fence_after();
            const float lse = s_lse[hl * 128 + r], dval = s_d[hl * 128 + r];
#pragma unroll 1
            for (int c = 0; c < 64; c += 16) {
                const int col = 64 * chalf + c;
                if (col >= npad) break;
                uint32_t sv[16], dv[16];
                tmem_ld16(lane_base + (uint32_t)col, sv);
                tmem_ld16(lane_base + col_dp + (uint32_t)col, dv);
                tc_wait_ld();
                const uint32_t bits = ((c < 32 ? m0 : m1) >> (c & 31)) & 0xffffu;
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    const float pa = (bits >> i) & 1u ? fexp(__uint_as_float(sv[i]) - lse) : 0.f;
                    const float pb = (bits >> (i + 1)) & 1u ? fexp(__uint_as_float(sv[i + 1]) - lse) : 0.f;
                    split2(pa * (__uint_as_float(dv[i]) - dval), pb * (__uint_as_float(dv[i + 1]) - dval), hi[i >> 1], lo[i >> 1]);
                }
                tmem_st8(lane_base + (uint32_t)col, hi);
                tmem_st8(lane_base + (uint32_t)col + 8u, lo);
            }
            tc_wait_st();
            tc_fence_before();
            __syncthreads();
            epi_rt = rt;
            epi_rows = rows;
            epi_h = h0 + hl;
            if (warp == 0) {
                if (elect_one()) {
                    tc_fence_after();
                    const uint32_t id = idesc_f16(128, 16);
                    const uint32_t ko = sb + dqk::kOffKT + (uint32_t)hl * 2u * dqk::kHeadTile;
                    uint64_t kt_hi = smem_desc(ko, 128, 2048), kt_lo = smem_desc(ko + dqk::kHeadTile, 128, 2048);
                    const uint32_t d = tmem + col_dq + 16u * (uint32_t)(cnt & 1);
                    for (int ks = 0; ks < npad / 16; ++ks) {
                        mma_ts(d, tmem + 16u * ks, kt_hi, id, ks == 0 ? 0u : 1u);   // dq_h = dS K_h
                        mma_ts(d, tmem + 16u * ks, kt_lo, id, 1u);
                        mma_ts(d, tmem + 16u * ks + 8u, kt_hi, id, 1u);
                        kt_hi += 256u >> 4;
                        kt_lo += 256u >> 4;
                    }
                    tc_commit(bar2);
                }
                __syncwarp();
            }
            ++cnt;
            pending = true;
        }
    }
    if (pending) {
        mbar_wait(bar2, p2);
        tc_fence_after();
        epilogue();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, tmem_cols);
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

extern "C" int32_t rl4co_eval_ll_max_nodes(void) { return ev::kMaxNodes; }

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
    {   // mask bytes -> bit words
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const int64_t want = (B * R + 7) / 8;
        ev::eval_pack_mask_kernel<<<(unsigned)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8), 256, 0, (cudaStream_t)stream>>>(
            mask, maskw, B * R, n_nodes);
        R4_CHECK_LAUNCH("eval_ll_fwd (mask words)");
    }
    if ((rc = ev::set_smem(ev::eval_glimpse_fwd_tc_kernel, ev::gfw::kSmem, "eval_ll_fwd (glimpse)"))) return rc;
    dim3 g2((unsigned)B, (unsigned)ev::grid_split(2 * B, 2), 2u);
    ev::eval_glimpse_fwd_tc_kernel<<<g2, 256, ev::gfw::kSmem, (cudaStream_t)stream>>>(q, gk, gv, maskw, heads, lse_h, R, n_nodes);
    R4_CHECK_LAUNCH("eval_ll_fwd (glimpse)");
    dim3 grid((unsigned)B, (unsigned)ev::grid_split(B, 1));
    if ((rc = ev::set_smem(ev::eval_ptr_fwd_tc_kernel, ev::pfw::kSmem, "eval_ll_fwd (pointer)"))) return rc;
    ev::eval_ptr_fwd_tc_kernel<<<grid, ev::pfw::kThreads, ev::pfw::kSmem, (cudaStream_t)stream>>>(lkf, maskw, taken, heads, logp, lse_z, R,
                                                                                   n_nodes, tanh_clipping, temperature);
    R4_CHECK_LAUNCH("eval_ll_fwd (pointer)");
    return RL4CO_OK;
}

extern "C" int rl4co_eval_ll_bwd(const float* q, const float* gk, const float* gv, const float* lkf, const uint32_t* maskw,
                                 const int64_t* taken, const float* heads, const float* lse_h, const float* lse_z,
                                 const float* dlogp, float* dheads, float* du_t, float* dd, float* scale_ws, float* dq,
                                 float* dgk, float* dgv, float* dlkf, int64_t B, int64_t R, int32_t n_nodes, float tanh_clipping,
                                 float temperature, void* stream) {
    R4_CHECK_ARG(q && gk && gv && lkf && maskw && taken && heads && lse_h && lse_z && dlogp && dheads && du_t && dd && scale_ws &&
                     dq && dgk && dgv && dlkf,
                 "eval_ll_bwd: null pointer");
    int rc;
    if ((rc = eval_check("eval_ll_bwd", B, R, n_nodes, temperature))) return rc;
    if (B == 0 || R == 0) return RL4CO_OK;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 g1((unsigned)B, (unsigned)ev::grid_split(B, 1));
    ev::eval_grad_scale_kernel<<<1, 1024, 0, st>>>(dlogp, B * R, scale_ws);
    R4_CHECK_LAUNCH("eval_ll_bwd (gradient scale)");
    if ((rc = ev::set_smem(ev::eval_ptr_bwd_tc_kernel, ev::pbw::kSmem, "eval_ll_bwd (row pass, pointer)"))) return rc;
    ev::eval_ptr_bwd_tc_kernel<<<g1, ev::pbw::kThreads, ev::pbw::kSmem, st>>>(lkf, maskw, taken, heads, lse_z, dlogp, scale_ws, dheads,
                                                                              du_t, R, n_nodes, tanh_clipping, temperature);
    R4_CHECK_LAUNCH("eval_ll_bwd (row pass, pointer)");
    if ((rc = ev::set_smem(ev::eval_dq_tc_kernel, ev::dqk::kSmem, "eval_ll_bwd (row pass, glimpse)"))) return rc;
    dim3 g2((unsigned)B, (unsigned)ev::grid_split(4 * B, 2), 4u);
    ev::eval_dq_tc_kernel<<<g2, 256, ev::dqk::kSmem, st>>>(q, gk, gv, maskw, heads, lse_h, dheads, scale_ws, dq, dd, R, n_nodes);
    R4_CHECK_LAUNCH("eval_ll_bwd (row pass)");
    // node pass (tcgen05): dlkf = dU^T O, then dgv = P^T dO and dgk = dS^T (0.25 Q)
    if ((rc = ev::set_smem(ev::eval_dlk_tc_kernel, ev::dlk::kSmem, "eval_ll_bwd (dlkf)"))) return rc;
    dim3 g3((unsigned)B, (unsigned)ev::grid_split(B, 1));
    ev::eval_dlk_tc_kernel<<<g3, 256, ev::dlk::kSmem, st>>>(heads, du_t, scale_ws, dlkf, R, n_nodes);
    R4_CHECK_LAUNCH("eval_ll_bwd (dlkf)");
    {   // two CTAs per SM: 64-row tiles, four heads each (blockIdx.z); <128, 8> - one CTA per SM - measured 1.96 against 1.47 ms
        auto kern = ev::eval_dkv_tc_kernel<64, 4>;
        if ((rc = ev::set_smem(kern, ev::Dkv<64, 4>::kSmem, "eval_ll_bwd (dgk, dgv)"))) return rc;
        dim3 g4((unsigned)B, (unsigned)ev::grid_split(2 * B, 2), 2u);
        kern<<<g4, 256, ev::Dkv<64, 4>::kSmem, st>>>(q, gk, gv, maskw, lse_h, dheads, dd, scale_ws, dgk, dgv, R, n_nodes);
    }
    R4_CHECK_LAUNCH("eval_ll_bwd (dgk, dgv)");
    return RL4CO_OK;
}
