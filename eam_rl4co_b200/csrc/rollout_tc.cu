// Tensor-core (tcgen05 + TMEM) path of the multistart rollout: one CTA = one instance, its S <= 128
// start rows are the M dimension of every MMA.
//
// Per decode step and instance (am/decoder.py:161-198, attention.py:282-328, decoding.py:140-190):
//   Q[s,:]      = pa[first_s] + pb[cur_s]                      (gather, CUDA cores -> smem A tile)
//   S_h[s,n]    = Q_h[s,:] . (0.25 K_h[n,:])                   8 x (128 x NPAD x 16)   tcgen05, A,B smem
//   P_h         = exp(S_h - max) on feasible nodes             TMEM -> regs -> TMEM (in place, fp16 hi/lo)
//   O_h[s,:]    = P_h[s,:] . V_h[:, :]                         8 x (128 x 16 x NPAD)   tcgen05, A TMEM
//   heads       = O / rowsum(P)                                TMEM -> regs -> TMEM (fp16 hi/lo)
//   logits[s,n] = heads[s,:] . (L W_out)[n,:] / sqrt(E)        128 x NPAD x 128        tcgen05, A TMEM
//   tanh clip, mask, log-softmax, argmax / Gumbel-max / given  TMEM -> regs, 4 threads per row
//   env step (tsp/env.py:62-88, cvrp/env.py:68-144)            bit-packed mask words in registers/smem
//
// Precision: every fp32 operand x is split into fp16 hi + lo (x = hi + lo to 2^-22) and every
// product is three MMAs (hi*hi + hi*lo + lo*hi) accumulated in fp32 in TMEM, which reproduces the
// reference's fp32 logits to ~1e-6 (oracle/numerics_tc.py).
//
// Shared memory: the six static operand tiles of the instance (K, L, V^T; hi and lo), brought in
// once per instance by TMA bulk copies from an image built by tc_pack_kernel, stay resident for
// all T steps; the per-step Q tile (64 channels = 4 heads at a time) is rebuilt by the CUDA cores.
#include <mutex>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace r4 {
namespace tc {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kRows = 128;      // MMA M: start rows of one instance, padded
constexpr int kMaxNPad = 112;   // nodes padded to a multiple of 16 (MMA N / K granularity)
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kColSP = 0;   // scores / probabilities of 4 heads: 4 x NPAD columns
constexpr uint32_t kColO = 448;  // glimpse accumulators of 4 heads: 4 x 16 columns
constexpr uint32_t kColHA = 0;   // heads as A operand: hi [0,64), lo [64,128)
constexpr uint32_t kColLG = 128; // logits accumulator: NPAD columns
// K-major, no-swizzle operand tiles (see tc_ptx.cuh::smem_desc)
constexpr uint32_t kLBO = 128;                 // static tiles: the K-chunks of a row group are contiguous
constexpr uint32_t kSBO_KL = 16 * kLBO;        // K / L tiles: 16 chunks (128 channels) per row group
// Q tiles: one per team (head slot) = 128 rows x 16 channels, hi then lo; a row group is its two K-chunks
constexpr uint32_t kLBO_Q = 128;
constexpr uint32_t kSBO_Q = 2 * kLBO_Q;
constexpr uint32_t kQHalfBytes = (kRows / 8) * kSBO_Q;   // 4 KB: the hi (or lo) half of a team's tile
constexpr uint32_t kQTeamBytes = 2 * kQHalfBytes;        // 8 KB = 128 threads x 16 words (also the head stash)
constexpr uint32_t kQTileBytes = 2 * kQTeamBytes;        // (x 2 = all four teams, see kOffTiles)
constexpr int kTeamBar = 5;                              // named barriers 5..8: the four warps of a team

__host__ __device__ inline int npad_of(int n) { return (n + 15) & ~15; }
__host__ __device__ inline uint32_t tile_bytes(int npad) { return (uint32_t)npad * 256u; }

// shared-memory carve-up
struct Smem {
    uint32_t maskw[4][kRows];   // feasible bits, word j = nodes [32j, 32j+32), SoA for conflict-free reads
    int cur[kRows];
    int first[kRows];
    float rem[kRows];           // CVRP: vehicle_capacity - used_capacity
    float red_a[4][kRows];      // epilogue exchanges between the 4 threads of a row
    float red_b[4][kRows];
    float red_c[4][kRows];
    float red_lp[4][kRows];
    int red_i[4][kRows];
    int newvis[kRows];
    float2 locs[kRows];
    float demand[kRows];        // indexed by node (node 0 unused)
    uint64_t tma_bar;
    uint64_t mma_bar;           // logits MMAs (one commit)
    uint64_t team_bar[4];       // scores / PV MMAs of one team (1 arrival)
    uint32_t tmem_ptr;
    uint32_t pad_;
    float lnf[4][kRows];        // ln(#feasible nodes) of the row at the previous step, one private copy per thread
    float lnf_depot[4][kRows];  // CVRP: the same at the row's last step AT THE DEPOT (< 0: none yet) ...
    float lse_depot[2][4][kRows];   // ... and the log-sum-exp of the thread's two heads at that step
};

// Shared-memory map: [Smem][Q hi][Q lo][K_hi][K_lo][L_hi][L_lo][Vt_hi][Vt_lo].  Everything but the tile size is a
// compile-time offset from the dynamic shared-memory base, so the context costs three registers.
constexpr uint32_t kOffQ = (uint32_t)((sizeof(Smem) + 127) / 128 * 128);
constexpr uint32_t kOffTiles = kOffQ + 2u * kQTileBytes;

struct Ctx {
    unsigned char* smem;
    uint32_t tmem;
    int npad;
    __device__ __forceinline__ Smem* s() const { return reinterpret_cast<Smem*>(smem); }
    __device__ __forceinline__ unsigned char* q_p(int team) const { return smem + kOffQ + (uint32_t)team * kQTeamBytes; }
    __device__ __forceinline__ uint32_t q_hi(int team) const { return base() + kOffQ + (uint32_t)team * kQTeamBytes; }
    __device__ __forceinline__ uint32_t q_lo(int team) const { return q_hi(team) + kQHalfBytes; }
    __device__ __forceinline__ unsigned char* tiles_p() const { return smem + kOffTiles; }
    __device__ __forceinline__ uint32_t base() const { return smem_u32(smem); }
    __device__ __forceinline__ uint32_t tile(int i) const { return base() + kOffTiles + (uint32_t)i * tile_bytes(npad); }
    __device__ __forceinline__ uint32_t k_hi() const { return tile(0); }
    __device__ __forceinline__ uint32_t k_lo() const { return tile(1); }
    __device__ __forceinline__ uint32_t l_hi() const { return tile(2); }
    __device__ __forceinline__ uint32_t l_lo() const { return tile(3); }
    __device__ __forceinline__ uint32_t v_hi() const { return tile(4); }
    __device__ __forceinline__ uint32_t v_lo() const { return tile(5); }
    __device__ __forceinline__ int nch() const { return npad >> 4; }
    __device__ __forceinline__ uint32_t sbo_v() const { return (uint32_t)npad * 16u; }
};

__host__ inline size_t smem_bytes(int npad) { return (size_t)kOffTiles + 6ull * tile_bytes(npad); }

// ---- operand image: fp32 cache tables -> fp16 hi/lo tiles in UMMA layout ------------------------------
// image of instance b = [K_hi][K_lo][L_hi][L_lo][Vt_hi][Vt_lo], each npad*256 bytes:
//   K, L : rows = nodes, K-dim = channels; byte(r, ch) = (r/8)*2048 + (ch/8)*128 + (r%8)*16 + (ch%8)*2
//   Vt   : rows = channels, K-dim = nodes; byte(ch, r) = (ch/8)*sbo_v + (r/8)*128 + (ch%8)*16 + (r%8)*2
// K is pre-scaled by 1/sqrt(d) = 0.25 (exact), L by 1/sqrt(E) (attention.py:299-301, 308-322).
__global__ void __launch_bounds__(256) tc_pack_kernel(rl4co_cache c, unsigned char* __restrict__ img, int npad,
                                                      int32_t* err) {
    const int64_t b = blockIdx.x;
    const int n = c.n_nodes;
    const uint32_t tb = tile_bytes(npad);
    unsigned char* im = img + b * 6ull * tb;
    const float* gk = c.gk + b * (int64_t)n * kE;
    const float* gv = c.gv + b * (int64_t)n * kE;
    const float* lk = c.lk + b * (int64_t)n * kE;
    const uint32_t sbo_v = (uint32_t)(npad / 8) * 128u;
    bool bad = false;
    // K and L: one item = (row r, 8-channel chunk)
    for (int it = threadIdx.x; it < npad * 16; it += blockDim.x) {
        const int r = it >> 4, ch = (it & 15) * 8;
        const uint32_t off = (uint32_t)(r >> 3) * kSBO_KL + (uint32_t)(it & 15) * kLBO + (uint32_t)(r & 7) * 16u;
        uint4 khi = make_uint4(0, 0, 0, 0), klo = khi, lhi = khi, llo = khi;
        if (r < n) {
            const float4 k0 = ld4(gk + r * kE + ch), k1 = ld4(gk + r * kE + ch + 4);
            const float4 l0 = ld4(lk + r * kE + ch), l1 = ld4(lk + r * kE + ch + 4);
            const float kk[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
            const float ll[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
            uint32_t kh[4], kl[4], lh[4], lw[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const float a0 = 0.25f * kk[2 * p], a1 = 0.25f * kk[2 * p + 1];
                const float b0 = ll[2 * p] / 11.313708498984761f, b1 = ll[2 * p + 1] / 11.313708498984761f;
                bad |= !(fabsf(a0) <= 65504.f) || !(fabsf(a1) <= 65504.f) || !(fabsf(b0) <= 65504.f) ||
                       !(fabsf(b1) <= 65504.f);
                split2(a0, a1, kh[p], kl[p]);
                split2(b0, b1, lh[p], lw[p]);
            }
            khi = make_uint4(kh[0], kh[1], kh[2], kh[3]);
            klo = make_uint4(kl[0], kl[1], kl[2], kl[3]);
            lhi = make_uint4(lh[0], lh[1], lh[2], lh[3]);
            llo = make_uint4(lw[0], lw[1], lw[2], lw[3]);
        }
        *reinterpret_cast<uint4*>(im + 0ull * tb + off) = khi;
        *reinterpret_cast<uint4*>(im + 1ull * tb + off) = klo;
        *reinterpret_cast<uint4*>(im + 2ull * tb + off) = lhi;
        *reinterpret_cast<uint4*>(im + 3ull * tb + off) = llo;
    }
    // Vt: one item = (8-node chunk rc, channel ch); lanes run over channels (coalesced reads)
    for (int it = threadIdx.x; it < (npad / 8) * kE; it += blockDim.x) {
        const int rc = it >> 7, ch = it & 127;
        float vv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int r = rc * 8 + i;
            vv[i] = (r < n) ? gv[r * kE + ch] : 0.f;
            bad |= !(fabsf(vv[i]) <= 65504.f);
        }
        uint32_t vh[4], vl[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) split2(vv[2 * p], vv[2 * p + 1], vh[p], vl[p]);
        const uint32_t off = (uint32_t)(ch >> 3) * sbo_v + (uint32_t)rc * kLBO + (uint32_t)(ch & 7) * 16u;
        *reinterpret_cast<uint4*>(im + 4ull * tb + off) = make_uint4(vh[0], vh[1], vh[2], vh[3]);
        *reinterpret_cast<uint4*>(im + 5ull * tb + off) = make_uint4(vl[0], vl[1], vl[2], vl[3]);
    }
    if (bad) raise_err(err, RL4CO_ERR_FP16_RANGE, b);
}

// ---- per-step phases ------------------------------------------------------------------------------
// Thread roles: warp w -> TMEM lane quadrant q = w % 4 (hardware rule for tcgen05.ld/st), j = w / 4;
// the thread owns row 32q + lane and, within the row, head j of the current group, mask word j,
// and the j-th quarter of the logit columns.

// Teams.  The four warps {4j .. 4j+3} (one per scheduler / TMEM lane quadrant) form team j and own head slot j:
// heads j and 4+j, the score / probability columns [j NPAD, (j+1) NPAD), the PV accumulator [16j, 16j+16), a
// private Q tile, a private mbarrier and named barrier 5+j.  Between the step's first instruction and the
// logits GEMM a team synchronises only with itself, so the teams drift apart and one team's MMA / L2 waits
// are covered by the softmax arithmetic of the others (every scheduler holds one warp of each team).

// Q tile of head h (channels [16h, 16h+16)) by the team's four warps: context.py:118-140 (TSP, pa[first] +
// pb[cur]) / :70-74,155-157 (CVRP, pb[cur] + (cap - used) * wcap), split to fp16 hi/lo.  Warp q covers rows
// [32q, 32q+32) in four passes of 8 rows; four lanes read the 64 contiguous bytes of a row (whole sectors).
// Rows >= S are written as zeros.
template <int ENV, class CX>
__device__ __forceinline__ void build_q(const CX& x, const rl4co_cache& c, int64_t b, int S, int h, int team,
                                        int q, int lane) {
    const int64_t base = b * (int64_t)c.n_nodes;
    const int pc = lane & 3;                           // 4-channel piece of the head's 16 channels
    const int choff = 16 * h + 4 * pc;
    float4 qa[4], qb[4];
    float rem[4];
#pragma unroll
    for (int it = 0; it < 4; ++it) {                   // every gather first: the L2 latencies overlap
        const int r = 32 * q + 8 * it + (lane >> 2);
        qa[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        qb[it] = qa[it];
        rem[it] = 0.f;
        if (r < S) {
            qb[it] = ld4(c.pb + (base + x.s()->cur[r]) * kE + choff);
            if (ENV == RL4CO_ENV_TSP) qa[it] = ld4(c.pa + (base + x.s()->first[r]) * kE + choff);
            else rem[it] = x.s()->rem[r];
        }
    }
    float4 wc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ENV == RL4CO_ENV_CVRP) wc = ld4(c.wcap + choff);
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        const int r = 32 * q + 8 * it + (lane >> 2);
        float4 v;
        if (ENV == RL4CO_ENV_TSP)
            v = make_float4(qa[it].x + qb[it].x, qa[it].y + qb[it].y, qa[it].z + qb[it].z, qa[it].w + qb[it].w);
        else
            v = make_float4(fmaf(rem[it], wc.x, qb[it].x), fmaf(rem[it], wc.y, qb[it].y),
                            fmaf(rem[it], wc.z, qb[it].z), fmaf(rem[it], wc.w, qb[it].w));
        uint32_t h0, l0, h1, l1;
        split2(v.x, v.y, h0, l0);
        split2(v.z, v.w, h1, l1);
        unsigned char* t = x.q_p(team) + (uint32_t)(r >> 3) * kSBO_Q + (uint32_t)(pc >> 1) * kLBO_Q +
                           (uint32_t)(r & 7) * 16u + (uint32_t)(pc & 1) * 8u;
        *reinterpret_cast<uint2*>(t) = make_uint2(h0, h1);
        *reinterpret_cast<uint2*>(t + kQHalfBytes) = make_uint2(l0, l1);
    }
    fence_proxy_async();
}

__device__ __forceinline__ void wait_mma(uint64_t* bar, uint32_t& par) {
    mbar_wait(bar, par);
    par ^= 1u;
    tc_fence_after();
}

// barrier of the 128 threads of a team, and the same with an AND-vote
__device__ __forceinline__ void team_sync(int team) { bar_sync(kTeamBar + team, 128); }
__device__ __forceinline__ bool team_all(int team, bool pred) {
    uint32_t r;
    asm volatile(
        "{\n\t.reg .pred p, r;\n\tsetp.ne.u32 p, %1, 0;\n\tbar.red.and.pred r, %2, 128, p;\n\t"
        "selp.u32 %0, 1, 0, r;\n\t}"
        : "=r"(r)
        : "r"((uint32_t)pred), "r"(kTeamBar + team)
        : "memory");
    return r != 0u;
}

// scores of head h in slot `team`: S = Q_h . K_h^T, three MMAs (hi*hi + hi*lo + lo*hi); issued by one thread
// of the team, completion on the team's mbarrier
__device__ __forceinline__ void issue_scores(const Ctx& x, int h, int team) {
    tc_fence_after();
    const uint32_t id = idesc_f16(kRows, x.npad);
    const uint64_t aq_hi = smem_desc(x.q_hi(team), kLBO_Q, kSBO_Q);
    const uint64_t aq_lo = smem_desc(x.q_lo(team), kLBO_Q, kSBO_Q);
    const uint64_t bk_hi = smem_desc(x.k_hi() + 2u * h * kLBO, kLBO, kSBO_KL);
    const uint64_t bk_lo = smem_desc(x.k_lo() + 2u * h * kLBO, kLBO, kSBO_KL);
    const uint32_t d = x.tmem + kColSP + (uint32_t)(team * x.npad);
    mma_ss(d, aq_hi, bk_hi, id, 0u);
    mma_ss(d, aq_hi, bk_lo, id, 1u);
    mma_ss(d, aq_lo, bk_hi, id, 1u);
    tc_commit(&x.s()->team_bar[team]);
}

// O = P . V_h : A = P from TMEM (hi/lo pairs per 16-node chunk), B = V^T tile rows [16h, 16h+16)
__device__ __forceinline__ void issue_pv(const Ctx& x, int h, int team) {
    tc_fence_after();
    const uint32_t id = idesc_f16(kRows, 16);
    uint64_t bv_hi = smem_desc(x.v_hi() + 2u * h * x.sbo_v(), kLBO, x.sbo_v());
    uint64_t bv_lo = smem_desc(x.v_lo() + 2u * h * x.sbo_v(), kLBO, x.sbo_v());
    const uint32_t d = x.tmem + kColO + 16u * team;
    uint32_t a = x.tmem + kColSP + (uint32_t)(team * x.npad);
    for (int jk = 0; jk < x.nch(); ++jk) {
        mma_ts(d, a, bv_hi, id, jk > 0 ? 1u : 0u);
        mma_ts(d, a, bv_lo, id, 1u);
        mma_ts(d, a + 8u, bv_hi, id, 1u);
        a += 16u;
        bv_hi += (2u * kLBO) >> 4;   // next 16 nodes = two 16-byte K-chunks further
        bv_lo += (2u * kLBO) >> 4;
    }
    tc_commit(&x.s()->team_bar[team]);
}

// logits = heads . (L W_out / sqrt(E))^T : A = heads from TMEM (k-step jk = head jk), B = L tile; one chain of
// 24 dependent MMAs into kColLG (profiles/ubench_mma.cu: as fast as two half-chains, and the readers load one
// accumulator instead of summing two).
__device__ __forceinline__ void issue_logits(const Ctx& x) {
    tc_fence_after();
    const uint32_t id = idesc_f16(kRows, x.npad);
    const uint32_t d = x.tmem + kColLG;
#pragma unroll
    for (int jk = 0; jk < 8; ++jk) {
        const uint64_t bl_hi = smem_desc(x.l_hi() + 2u * jk * kLBO, kLBO, kSBO_KL);
        const uint64_t bl_lo = smem_desc(x.l_lo() + 2u * jk * kLBO, kLBO, kSBO_KL);
        const uint32_t a = x.tmem + kColHA + 8u * jk;
        mma_ts(d, a, bl_hi, id, jk > 0 ? 1u : 0u);
        mma_ts(d, a, bl_lo, id, 1u);
        mma_ts(d, a + 64u, bl_hi, id, 1u);
    }
    tc_commit(&x.s()->mma_bar);
}

// Masked softmax numerator of one (row, head): S (fp32, NPAD columns of TMEM) -> P = exp(S - max) on
// feasible nodes, 0 elsewhere (attention.py:308-322: masked scores are -inf -> exactly 0), written
// back in place as fp16 hi/lo: chunk c (16 nodes) -> hi in columns [16c,16c+8), lo in [16c+8,16c+16).
// Returns 1 / rowsum (0 for a row without feasible nodes: padding rows).
__device__ __forceinline__ void shift_mask16(uint32_t& a0, uint32_t& a1, uint32_t& a2, uint32_t& a3) {
    a0 = __funnelshift_r(a0, a1, 16);
    a1 = __funnelshift_r(a1, a2, 16);
    a2 = __funnelshift_r(a2, a3, 16);
    a3 >>= 16;
}

#define R4_STAMP(k)                                       \
    do {                                                  \
        if (prof != nullptr) {                            \
            const long long now__ = clock64();            \
            prof[k] += now__ - tprev;                     \
            tprev = now__;                                \
        }                                                 \
    } while (0)

// scores of one 16-column chunk with the infeasible nodes set to -inf (attention.py:308-322): one select
// per element serves both the running maximum and the exponent (exp(-inf) = 0 exactly)
__device__ __forceinline__ void chunk_mask(const uint32_t (&v)[16], uint32_t bits, float (&sv)[16]) {
#pragma unroll
    for (int i = 0; i < 16; ++i) sv[i] = (bits & (1u << i)) ? __uint_as_float(v[i]) : -INFINITY;
}
__device__ __forceinline__ float chunk_max(const float (&sv)[16]) {
    float m0 = fmaxf(sv[0], sv[1]), m1 = fmaxf(sv[2], sv[3]);
#pragma unroll
    for (int i = 4; i < 16; i += 2) {
        m0 = fmaxf(m0, sv[i]);
        m1 = fmaxf(m1, sv[i + 1]);
    }
    return fmaxf(m0, m1);
}
// P of one 16-column chunk -> fp16 hi/lo words, row sums in two chains
__device__ __forceinline__ void chunk_exp(const float (&sv)[16], float ml, float& l0, float& l1,
                                          uint32_t (&hi)[8], uint32_t (&lo)[8]) {
    const float kLog2e = 1.4426950408889634f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float p0 = ex2_approx(fmaf(sv[2 * i], kLog2e, -ml));
        const float p1 = ex2_approx(fmaf(sv[2 * i + 1], kLog2e, -ml));
        l0 += p0;
        l1 += p1;
        split2(p0, p1, hi[i], lo[i]);
    }
}
// the last chunk when it holds <= 12 real nodes: `tq` 4-column loads instead of 16 columns
__device__ __forceinline__ void load_tail(uint32_t addr, int tq, uint32_t (&v)[16]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        uint32_t v4[4] = {0u, 0u, 0u, 0u};
        if (k < tq) tmem_ld4(addr + 4u * k, v4);
#pragma unroll
        for (int i = 0; i < 4; ++i) v[4 * k + i] = v4[i];
    }
#pragma unroll
    for (int i = 12; i < 16; ++i) v[i] = 0u;
}

// One-pass masked softmax numerator of one (row, head), in place: S (fp32, NPAD columns) -> P (fp16 hi/lo).
// TMEM reads are the scarce resource here (~128 B/clk/SM), so S is read ONCE: chunk c is exponentiated
// against the running row maximum m_run of the chunks seen so far.  When a chunk's masked maximum exceeds
// m_run by more than 10 (e^10 still fits fp16), the warp takes the rare rescale path: m_run is raised and
// the chunks already written are multiplied by e^(m_old - m_new).  Any common reference cancels in
// heads = (P V) / rowsum(P), and P >= e^(s - m_final), so nothing is lost against the two-pass form.
__device__ __forceinline__ void rescale_chunk(uint32_t addr, float f) {
    uint32_t w[16];
    tmem_ld16(addr, w);
    tc_wait_ld();
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
        const float2 l = __half22float2(*reinterpret_cast<const __half2*>(&w[8 + i]));
        split2((h.x + l.x) * f, (h.y + l.y) * f, hi[i], lo[i]);
    }
    tmem_st8(addr, hi);
    tmem_st8(addr + 8u, lo);
}

// One 16-node chunk.  spec (CTA-uniform): the reference refl2 (= ref * log2e) is the caller's and there is no
// maximum, vote or rescale between the TMEM load and the exponentials; otherwise the exact online form.
__device__ __forceinline__ void softmax_chunk(uint32_t sp, int ch, const uint32_t (&v)[16], uint32_t bits, bool spec,
                                              float& refl2, float& m_run, float& l0, float& l1) {
    float sv[16];
    chunk_mask(v, bits, sv);
    if (!spec) {
        const float cm = chunk_max(sv);
        const bool need = cm > m_run + 10.0f;          // also the first feasible chunk of the row (m_run = -inf)
        const bool resc = need && m_run != -INFINITY;  // a row with no feasible node so far wrote exact zeros only
        if (__any_sync(kFull, resc)) {
            const float f = resc ? ex2_approx((m_run - cm) * 1.4426950408889634f) : 1.0f;
            l0 *= f;
            l1 *= f;
            for (int pc = 0; pc < ch; ++pc) rescale_chunk(sp + 16u * pc, f);
            tc_wait_st();
        }
        if (need) m_run = cm;
        // a row with no feasible node so far has m_run = -inf: reference 0 keeps exp(-inf - ref) = 0, not NaN
        refl2 = (m_run == -INFINITY ? 0.f : m_run) * 1.4426950408889634f;
    }
    uint32_t hi[8], lo[8];
    chunk_exp(sv, refl2, l0, l1, hi, lo);
    tmem_st8(sp + 16u * ch, hi);
    tmem_st8(sp + 16u * ch + 8u, lo);
}

// spec = false: exact online form, returns rowsum(P) and the reference m_out it is relative to.
// spec = true : P = exp(S - ref) against the caller's reference (ref * log2e in refl2); the caller checks the
//               returned row sum and falls back to the exact form when it left the safe window.
template <class CX>
__device__ __forceinline__ float softmax_inplace(const CX& x, uint32_t sp, int row, float* dbg_scores, int n,
                                                 bool spec, float refl2, float& m_out) {
    const uint32_t w0 = x.s()->maskw[0][row], w1 = x.s()->maskw[1][row], w2 = x.s()->maskw[2][row],
                   w3 = x.s()->maskw[3][row];
    // full 16-column chunks, then an optional short tail (n = 100: 6 chunks + 4 columns)
    const int lastv = n - 16 * (x.nch() - 1);
    const int tq = lastv <= 12 ? (lastv + 3) >> 2 : 0;
    const int nfull = tq ? x.nch() - 1 : x.nch();
    float m_run = -INFINITY, l0 = 0.f, l1 = 0.f;
    uint32_t a0 = w0, a1 = w1, a2 = w2, a3 = w3;
    uint32_t va[16], vb[16];
    if (nfull > 0) tmem_ld16(sp, va);
    for (int ch = 0; ch < nfull; ch += 2) {   // the next chunk's TMEM load is in flight while one is processed
        tc_wait_ld();
        if (ch + 1 < nfull) tmem_ld16(sp + 16u * (ch + 1), vb);
        if (dbg_scores != nullptr) {
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if (16 * ch + i < n) dbg_scores[16 * ch + i] = __uint_as_float(va[i]);
        }
        softmax_chunk(sp, ch, va, a0, spec, refl2, m_run, l0, l1);
        shift_mask16(a0, a1, a2, a3);
        if (ch + 1 < nfull) {
            tc_wait_ld();
            if (ch + 2 < nfull) tmem_ld16(sp + 16u * (ch + 2), va);
            if (dbg_scores != nullptr) {
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (16 * (ch + 1) + i < n) dbg_scores[16 * (ch + 1) + i] = __uint_as_float(vb[i]);
            }
            softmax_chunk(sp, ch + 1, vb, a0, spec, refl2, m_run, l0, l1);
            shift_mask16(a0, a1, a2, a3);
        }
    }
    if (tq) {
        load_tail(sp + 16u * nfull, tq, va);
        tc_wait_ld();
        if (dbg_scores != nullptr) {
#pragma unroll
            for (int i = 0; i < 12; ++i)
                if (16 * nfull + i < n) dbg_scores[16 * nfull + i] = __uint_as_float(va[i]);
        }
        if (spec) {
            // only the tq 4-column groups that hold real nodes are exponentiated; the padding columns of the
            // chunk (K of the PV MMA) are written as exact zeros
            uint32_t hi[8], lo[8];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (k < tq) {
#pragma unroll
                    for (int i = 2 * k; i < 2 * k + 2; ++i) {
                        const float s0 = (a0 & (1u << (2 * i))) ? __uint_as_float(va[2 * i]) : -INFINITY;
                        const float s1 = (a0 & (1u << (2 * i + 1))) ? __uint_as_float(va[2 * i + 1]) : -INFINITY;
                        const float p0 = ex2_approx(fmaf(s0, 1.4426950408889634f, -refl2));
                        const float p1 = ex2_approx(fmaf(s1, 1.4426950408889634f, -refl2));
                        l0 += p0;
                        l1 += p1;
                        split2(p0, p1, hi[i], lo[i]);
                    }
                } else {
                    hi[2 * k] = hi[2 * k + 1] = 0u;
                    lo[2 * k] = lo[2 * k + 1] = 0u;
                }
            }
            tmem_st8(sp + 16u * nfull, hi);
            tmem_st8(sp + 16u * nfull + 8u, lo);
        } else {
            softmax_chunk(sp, nfull, va, a0, false, refl2, m_run, l0, l1);   // mask bits beyond n are 0 -> exact zeros
        }
    }
    tc_wait_st();
    m_out = m_run;
    return l0 + l1;
}

// All GEMM phases of one decoder step; on return the raw logits of the instance sit in TMEM
// columns [kColLG, kColLG + n) (row = lane).  dbg_* (nullable) receive scores [S?,8,n] / heads [.,128]
// of this thread's row for the parity tests.
// Speculative softmax reference.  P only has to be exp(S - ref) for SOME ref that keeps the row sum l inside
// the window where the fp16 hi/lo split is accurate relative to l (l >= 12.5: the 2^-25 absolute floor of the
// lo halves stays below 2^-22 l over 100 nodes) and finite (l < 6e4 bounds every element by fp16's maximum).
// The row's log-sum-exp moves slowly from one decode step to the next, so each (row, head) keeps lse of the
// previous step and aims the next row sum at 2^10: the window is lse_new - lse_old in (-4.4, +4.1).  A miss
// anywhere in the CTA (block-wide vote that replaces the barrier after the softmax) re-issues the score MMAs
// of the group and takes the exact online form.  The first step of an instance has no reference yet.
constexpr float kSpecTargetLn = 6.931471805599453f;   // ln 1024
constexpr float kSpecLMin = 12.5f, kSpecLMax = 6.0e4f;

template <int ENV>
__device__ __forceinline__ void decode_step_gemms(const Ctx& x, const rl4co_cache& c, int64_t b, int S,
                                                  uint32_t& mma_par, uint32_t& team_par, float* dbg_scores_row,
                                                  float* dbg_heads_row, long long* prof, long long& tprev,
                                                  float (&lse)[2], bool have_ref) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    const uint32_t lane_base = x.tmem + ((uint32_t)(32 * q) << 16);
    const int n = c.n_nodes;
    // the team's MMAs are issued by one elected lane of its warp on scheduler j: one issuing warp per scheduler
    const bool issuer_warp = (q == j);
    uint64_t* tbar = &x.s()->team_bar[j];
    const float kLog2e = 1.4426950408889634f;
    // heads of round 0 wait in the team's Q tile while round 1 is processed: the tile is idle between the end
    // of scores(1) and the next step's Q build (8 KB = 128 threads x 16 words, conflict-free)
    float* stash = reinterpret_cast<float*>(x.q_p(j)) + (threadIdx.x & 127);
    float hv[16];
    const bool at_depot = (ENV == RL4CO_ENV_CVRP) && x.s()->cur[row] == 0;
    build_q<ENV>(x, c, b, S, j, j, q, lane);
    team_sync(j);
    R4_STAMP(0);
    if (issuer_warp) { if (elect_one()) issue_scores(x, j, j); }
    // The feasible set can change by two orders of magnitude in one step (CVRP: a vehicle with no fitting customer
    // returns to the depot and every unvisited customer opens up).  With near-uniform attention the log-sum-exp
    // follows ln(#feasible), with peaked attention it does not move: the reference is shifted by half of the
    // change of ln(#feasible), which keeps both extremes inside the window.
    {
        const int nf = __popc(x.s()->maskw[0][row]) + __popc(x.s()->maskw[1][row]) + __popc(x.s()->maskw[2][row]) +
                       __popc(x.s()->maskw[3][row]);
        const float lnf = __logf((float)(nf > 0 ? nf : 1));
        // openings are abrupt (1 -> 60 nodes when a full vehicle returns), closings gradual: all of an increase
        const float dl = lnf - x.s()->lnf[j][row];
        float shift = have_ref ? (dl > 0.f ? 1.0f : 0.5f) * dl : 0.f;              // first step: no reference yet
        // CVRP at the depot: the query is the same every time the row is there (cur = depot, full capacity), only
        // the mask has shrunk, while the step before (a nearly empty vehicle at a customer) says little about it:
        // the reference comes from the row's previous depot step.
        if (ENV == RL4CO_ENV_CVRP && have_ref && at_depot && x.s()->lnf_depot[j][row] >= 0.f) {
            shift = lnf - x.s()->lnf_depot[j][row];
            lse[0] = x.s()->lse_depot[0][j][row];
            lse[1] = x.s()->lse_depot[1][j][row];
        }
        x.s()->lnf[j][row] = lnf;
        lse[0] += shift;
        lse[1] += shift;
    }
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const int h = 4 * g + j;
        wait_mma(tbar, team_par);                      // scores of round g
        R4_STAMP(1 + 5 * g);
        if (g == 1) {
#pragma unroll
            for (int i = 0; i < 16; ++i) stash[i * 128] = hv[i];
        }
        const uint32_t sp = lane_base + kColSP + (uint32_t)(j * x.npad);
        float* dbg = dbg_scores_row ? dbg_scores_row + h * n : nullptr;
        float l, m;
        bool spec = have_ref;
        const float ref = lse[g] - kSpecTargetLn;
        for (;;) {
            l = softmax_inplace(x, sp, row, dbg, n, spec, ref * kLog2e, m);
            if (spec) m = ref;
            // l == 0: row without feasible nodes (padding rows)
            const bool ok = !spec || (l == 0.f) || (l >= kSpecLMin && l < kSpecLMax);
            tc_fence_before();
            if (team_all(j, ok)) break;
            // rare: the reference missed somewhere in the team -> scores again, then the exact form
            if (prof != nullptr) prof[15] += 1;        // debug counter: fallbacks of team 0
            spec = false;
            if (g == 1) {                              // the stash sits on the Q tile the MMAs need: rebuild it
#pragma unroll
                for (int i = 0; i < 16; ++i) hv[i] = stash[i * 128];
                team_sync(j);
                build_q<ENV>(x, c, b, S, h, j, q, lane);
                team_sync(j);
            }
            if (issuer_warp) { if (elect_one()) issue_scores(x, h, j); }
            wait_mma(tbar, team_par);
            if (g == 1) {
#pragma unroll
                for (int i = 0; i < 16; ++i) stash[i * 128] = hv[i];
            }
        }
        const float inv_l = l > 0.f ? 1.0f / l : 0.f;
        lse[g] = l > 0.f ? m + __logf(l) : 0.f;
        if (ENV == RL4CO_ENV_CVRP && at_depot) {
            x.s()->lse_depot[g][j][row] = lse[g];
            if (g == 1) x.s()->lnf_depot[j][row] = x.s()->lnf[j][row];
        }
        R4_STAMP(2 + 5 * g);
        if (issuer_warp) { if (elect_one()) issue_pv(x, h, j); }
        if (g == 0) build_q<ENV>(x, c, b, S, 4 + j, j, q, lane);   // next round's Q while PV runs (scores(0) is done)
        R4_STAMP(3 + 5 * g);
        wait_mma(tbar, team_par);                      // PV of round g
        R4_STAMP(4 + 5 * g);
        if (g == 0) {                                  // scores(1) run while O(0) is drained
            team_sync(j);
            if (issuer_warp) { if (elect_one()) issue_scores(x, 4 + j, j); }
        }
        uint32_t o[16];
        tmem_ld16(lane_base + kColO + 16u * j, o);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) hv[i] = __uint_as_float(o[i]) * inv_l;
        tc_fence_before();
        if (g == 0) R4_STAMP(5);
    }
    // every team is done with its probabilities: the head operand and the logits accumulators reuse those columns
    __syncthreads();
    tc_fence_after();
    // heads -> A operand in TMEM (fp16 hi/lo), k-step = head
#pragma unroll
    for (int g = 1; g >= 0; --g) {
        const int h = 4 * g + j;
        if (g == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) hv[i] = stash[i * 128];
        }
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) split2(hv[2 * i], hv[2 * i + 1], hi[i], lo[i]);
        tmem_st8(lane_base + kColHA + 8u * h, hi);
        tmem_st8(lane_base + kColHA + 64u + 8u * h, lo);
        if (dbg_heads_row != nullptr) {
#pragma unroll
            for (int i = 0; i < 16; ++i) dbg_heads_row[16 * h + i] = hv[i];
        }
    }
    tc_wait_st();
    tc_fence_before();
    __syncthreads();
    R4_STAMP(10);
    if (warp == 0) { if (elect_one()) issue_logits(x); }
    wait_mma(&x.s()->mma_bar, mma_par);
    R4_STAMP(11);
}

// ---- selection: process_logits + Greedy / Sampling / Evaluate (decoding.py:140-190, 346-465) --------
// Four threads per row, thread j takes the j-th quarter of the logit columns (units of 4 columns);
// max / sum / argmax are combined through shared memory with a 128-thread named barrier per quadrant.
struct RowSel {
    int action;
    float logp;
};

// tanh(v) = 1 - 2 / (1 + e^(2v)) on the MUFU ex2 / rcp approximations: absolute error ~2e-7
// (+-1 exactly at the tails, NaN propagates); the reference's clipped logit 10*tanh is matched to ~2e-6.
__device__ __forceinline__ float tanh_fast(float v) {
    const float e2 = ex2_approx(v * 2.8853900817779268f);
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e2 + 1.0f));
    return fmaf(-2.0f, r, 1.0f);
}
// |v| <= 1: tanh(v) = v R(v^2), R of degree 6 (weighted least squares on Chebyshev nodes; absolute error
// 1.1e-7 evaluated in fp32).  Eight FMA-pipe instructions and no MUFU: the selection phase is MUFU-bound
// otherwise (ex2 + rcp for the tanh and ex2 for the softmax at 8 cycles per warp-instruction each).
__device__ __forceinline__ float tanh_small(float v) {
    const float u = v * v;
    float r = 9.568327223e-04f;
    r = fmaf(r, u, -5.928727333e-03f);
    r = fmaf(r, u, 1.994509064e-02f);
    r = fmaf(r, u, -5.324891955e-02f);
    r = fmaf(r, u, 1.331909001e-01f);
    r = fmaf(r, u, -3.333206773e-01f);
    r = fmaf(r, u, 9.999996424e-01f);
    return v * r;
}

// log_softmax needs log sum_e exp(x_e - ref) for ANY ref >= max_e x_e that keeps the sum a normal float.  With
// tanh clipping (decoding.py:171-172) x = clip * tanh(.) / temperature <= clip / temperature, so ref is a
// constant and the row needs no max exchange: greedy / evaluate take ONE 128-thread barrier per step.  Without
// clipping the row maximum is exchanged first.  Greedy takes the argmax over x itself (log_softmax is monotone;
// first index on exact ties as torch.argmax).
template <int MODE>
__device__ __forceinline__ RowSel select_row(const Ctx& x, const rl4co_rollout_args& a, int n, int q, int j,
                                             int row, bool valid, int64_t grow, int t, int T, uint32_t nkey,
                                             int given, int& err_bits) {
    Smem* s = x.s();
    const uint32_t lg = x.tmem + ((uint32_t)(32 * q) << 16) + kColLG;
    const int U = (n + 3) >> 2;
    const int u0 = (U * j) >> 2, u1 = (U * (j + 1)) >> 2;
    // 28 columns from 4*u0 on (columns past this thread's range are ignored below)
    float tv[28];                                                   // tanh(logit) (or the logit), -inf if infeasible
    {
        uint32_t a16[16], a8[8], a4[4];
        tmem_ld16(lg + 4u * u0, a16);
        tmem_ld8(lg + 4u * u0 + 16u, a8);
        tmem_ld4(lg + 4u * u0 + 24u, a4);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) tv[i] = __uint_as_float(a16[i]);
#pragma unroll
        for (int i = 0; i < 8; ++i) tv[16 + i] = __uint_as_float(a8[i]);
#pragma unroll
        for (int i = 0; i < 4; ++i) tv[24 + i] = __uint_as_float(a4[i]);
    }
    const uint32_t w0 = s->maskw[0][row], w1 = s->maskw[1][row], w2 = s->maskw[2][row], w3 = s->maskw[3][row];
    const int sh = 4 * u0, wi = sh >> 5;
    const uint32_t lo = wi == 0 ? w0 : (wi == 1 ? w1 : (wi == 2 ? w2 : w3));
    const uint32_t hi = wi == 0 ? w1 : (wi == 1 ? w2 : (wi == 2 ? w3 : 0u));
    const int cnt = 4 * (u1 - u0);
    const uint32_t inrange = cnt >= 32 ? 0xffffffffu : ((1u << cnt) - 1u);
    const uint32_t win = __funnelshift_r(lo, hi, sh & 31) & inrange;  // feasible bits of nodes sh .. sh+27

    // every run-time-uniform option is hoisted out of the unrolled element loops: the loops stay
    // branch-free so that the 28 independent chains of a thread overlap
    const bool clipped = a.tanh_clipping > 0.f;                     // decoding.py:171-172
    const float sc = (clipped ? a.tanh_clipping : 1.0f) / a.temperature;   // :179 folded into the clip scale
    const float kL2e = 1.4426950408889634f;
    if (clipped) {
        float am0 = 0.f, am1 = 0.f;
#pragma unroll
        for (int e = 0; e < 28; e += 2) {
            am0 = fmaxf(am0, fabsf(tv[e]));
            am1 = fmaxf(am1, fabsf(tv[e + 1]));
        }
        if (__all_sync(kFull, fmaxf(am0, am1) <= 1.0f)) {           // NaN fails the test -> MUFU path propagates it
#pragma unroll
            for (int e = 0; e < 28; ++e) tv[e] = tanh_small(tv[e]);
        } else {
#pragma unroll
            for (int e = 0; e < 28; ++e) tv[e] = tanh_fast(tv[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < 28; ++e) tv[e] = ((win >> e) & 1u) ? tv[e] : -INFINITY;   // :177

    // The constant reference is only safe while exp(x - ref) >= exp(-2 sc) stays a normal float: sc <= 40
    // (temperature 0.1 gives sc = 100 and every term would flush to zero -> lse = -inf).  Beyond that the row
    // maximum is exchanged like in the unclipped case.
    float ref;                                                      // >= every x of the row
    if (clipped && sc <= 40.0f) {
        ref = sc;
    } else {
        float lmax = -INFINITY;
#pragma unroll
        for (int e = 0; e < 28; ++e) lmax = fmaxf(lmax, tv[e]);
        s->red_a[j][row] = lmax;
        bar_sync(1 + q, 128);
        const float m = fmaxf(fmaxf(s->red_a[0][row], s->red_a[1][row]), fmaxf(s->red_a[2][row], s->red_a[3][row]));
        ref = (m == -INFINITY) ? 0.f : m * sc;
    }
    // sum_e exp(x_e - ref), x_e = sc * tv_e: one FFMA + MUFU + FADD per element (exp(-inf) = 0)
    const float c1 = sc * kL2e, c0 = -ref * kL2e;
    float ls0 = 0.f, ls1 = 0.f;
#pragma unroll
    for (int e = 0; e < 28; e += 2) {
        ls0 += ex2_approx(fmaf(tv[e], c1, c0));
        ls1 += ex2_approx(fmaf(tv[e + 1], c1, c0));
    }
    s->red_b[j][row] = ls0 + ls1;

    float best = -INFINITY, best_x = -INFINITY;
    int best_i = 0x7fffffff;
    constexpr bool sampling = MODE == RL4CO_MODE_SAMPLING;
    if (MODE == RL4CO_MODE_EVALUATE) {
#pragma unroll
        for (int e = 0; e < 28; ++e) {
            if (e < cnt && sh + e == given) {
                best = 0.f;
                best_i = given;
                best_x = tv[e];
            }
        }
    } else if (!sampling) {
#pragma unroll
        for (int e = 0; e < 28; ++e) {
            if (tv[e] > best) {   // ascending node order + strict > = first max index (torch.argmax)
                best = tv[e];
                best_i = sh + e;
            }
        }
        best_x = best;
    }
    if (!sampling) {
        s->red_c[j][row] = best;
        s->red_i[j][row] = best_i;
        s->red_lp[j][row] = best_x;
    }
    bar_sync(1 + q, 128);
    const float tot = (s->red_b[0][row] + s->red_b[1][row]) + (s->red_b[2][row] + s->red_b[3][row]);
    // attention.py:303-304 (a feasible logit is NaN); tot == 0 cannot happen with a feasible node (the maximum
    // contributes exp(0) or at least exp(-2 sc)) and is flagged as the same numerical failure
    if ((tot != tot || !(tot > 0.f)) && valid) err_bits |= RL4CO_ERR_NAN_LOGITS;
    const float lse = logf(tot);                                    // log_softmax(x)_e = (x_e - ref) - lse  :190
    if (a.logp_all != nullptr && valid) {
#pragma unroll
        for (int e = 0; e < 28; ++e)
            if (e < cnt && sh + e < n) a.logp_all[(grow * T + t) * (int64_t)n + sh + e] = (tv[e] * sc - ref) - lse;
    }
    if (sampling && a.gumbel != nullptr) {
        // Gumbel-max over logp + g (decoding.py:403-420 restated with supplied noise)
#pragma unroll
        for (int e = 0; e < 28; ++e) {
            const float lp = (tv[e] * sc - ref) - lse;              // -inf stays -inf
            float g = 0.f;
            if ((win >> e) & 1u) g = a.gumbel[(grow * T + t) * (int64_t)n + sh + e];
            const float cand = lp + g;
            if (cand > best) {
                best = cand;
                best_i = sh + e;
                best_x = tv[e];
            }
        }
    } else if (sampling) {
        // Counter noise: argmax_e (logp_e + g_e), g = -ln(-ln u), is argmax_e p_e / (-ln u_e) (the exponential race:
        // exp is monotone and the row's normaliser is common), so neither logarithm of the Gumbel transform nor the
        // log-softmax of every element is needed: one lg2, one rcp and the exp(x_e - ref) of the row sum per element.
        // -lg2 u must be accurate RELATIVE to itself: lg2.approx is (2^-22 absolute near 1), so u > 0.98 takes the
        // series of -ln(1 - d), d = 1 - u exact.  Keys agree with logp + g to ~1e-5, far inside the near-tie bar.
        best = 0.f;                                                 // an infeasible node has key 0 and never wins
#pragma unroll
        for (int e = 0; e < 28; ++e) {
            const float u = noise_u(nkey, sh + e);
            const float d = 1.0f - u;
            const float ser = d * fmaf(d, fmaf(d, fmaf(d, 0.36067376f, 0.48089835f), 0.72134752f), 1.44269504f);
            float l2;
            asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(u));
            const float big_e = u > 0.98f ? ser : -l2;             // -lg2 u > 0
            float r;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(big_e));
            const float cand = ex2_approx(fmaf(tv[e], c1, c0)) * r;   // exp(-inf) = 0 for infeasible nodes
            if (cand > best) {
                best = cand;
                best_i = sh + e;
                best_x = tv[e];
            }
        }
    }
    if (sampling) {
        s->red_c[j][row] = best;
        s->red_i[j][row] = best_i;
        s->red_lp[j][row] = best_x;
        bar_sync(1 + q, 128);
    }
    float fb = s->red_c[0][row];
    int fi = s->red_i[0][row];
    float fx = s->red_lp[0][row];
#pragma unroll
    for (int jj = 1; jj < 4; ++jj) {
        const float cb = s->red_c[jj][row];
        if (cb > fb) {
            fb = cb;
            fi = s->red_i[jj][row];
            fx = s->red_lp[jj][row];
        }
    }
    RowSel sel;
    sel.action = (fi == 0x7fffffff) ? 0 : fi;
    sel.logp = (fx * sc - ref) - lse;
    if (valid && j == 0) {
        const uint32_t ww = (sel.action >> 5) == 0 ? w0 : ((sel.action >> 5) == 1 ? w1 : ((sel.action >> 5) == 2 ? w2 : w3));
        if (MODE != RL4CO_MODE_EVALUATE && !((ww >> (sel.action & 31)) & 1u)) err_bits |= RL4CO_ERR_INFEASIBLE;
    }
    // the log-probability is consumed by thread 1 of the row only (record_action): lse and this check sink there
    if (valid && j == 1 && !(sel.logp > -1000.f)) err_bits |= RL4CO_ERR_NEG_INF_LOGP;   // decoding.py:56-58
    return sel;
}

// ---- env step (tsp/env.py:62-88, cvrp/env.py:68-100,132-144), 4 threads per row ------------------------
// The row's bookkeeping is spread over its four threads so that no warp carries a long serial chain between the
// selection and the step-end barrier: thread 0 owns what the next step needs (current / first node, done flag),
// thread 1 the log-likelihood, thread 2 the tour length, thread 3 the action record.
struct RowState {
    uint32_t word;   // TSP: feasible bits of word j; CVRP: visited bits of word j
    float used, cap; // CVRP (tracked redundantly by the 4 threads of the row)
    float ll;        // thread 1
    float len;       // thread 2
    int prev, first_action;   // thread 2
    int nvis, length;         // thread 0
    bool done;                // thread 0
};

// Outputs of the step that nothing in the next step waits for; called after the step-end barrier, where the
// three short chains (logf, sqrt, scattered stores) of a row run in different warps under the next Q gathers.
template <int ENV, class CX>
__device__ __forceinline__ void record_action(const CX& x, const rl4co_rollout_args& a, RowState& st, int j,
                                              bool valid, int64_t grow, int t, int T, int act, float lpa) {
    if (!valid) return;
    if (j == 1) {
        a.logp[grow * T + t] = lpa;
        st.ll += lpa;
    } else if (j == 2) {
        // tour length so far (ops.py:82-95): edge prev -> act; CVRP starts at the depot
        const int from = (st.prev < 0) ? ((ENV == RL4CO_ENV_CVRP) ? 0 : act) : st.prev;
        const float2 p0 = x.s()->locs[from], p1 = x.s()->locs[act];
        const float dx = p1.x - p0.x, dy = p1.y - p0.y;
        st.len += sqrtf(dx * dx + dy * dy);
        if (st.prev < 0) st.first_action = act;
        st.prev = act;
    } else if (j == 3) {
        a.actions[grow * T + t] = act;
    }
}

template <int ENV, class CX>
__device__ __forceinline__ void apply_action(const CX& x, const rl4co_rollout_args& a, RowState& st, int n, int q,
                                             int j, int row, bool valid, int t, int act) {
    auto* s = x.s();
    if (j == 0 && valid) s->cur[row] = act;
    if (ENV == RL4CO_ENV_TSP) {
        if (valid && (act >> 5) == j) st.word &= ~(1u << (act & 31));
        s->maskw[j][row] = st.word;
        if (j == 0 && valid) {
            if (t == 0) s->first[row] = act;
            if (!st.done && t + 1 >= n) {
                st.done = true;
                st.length = t + 1;
            }
        }
    } else {
        bool newly = false;
        if (valid && (act >> 5) == j) {
            newly = !((st.word >> (act & 31)) & 1u);
            st.word |= 1u << (act & 31);
        }
        int di = act - 1;
        di = di < 0 ? 0 : (di > n - 2 ? n - 2 : di);
        st.used = (st.used + s->demand[di + 1]) * ((act != 0) ? 1.0f : 0.0f);   // cvrp/env.py:73-80
        const float lim = st.cap + 1e-5f;                                       // :135
        uint32_t feas = 0u;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) {
            const int node = 32 * j + i;
            const bool ok = valid && node >= 1 && node < n && !((st.word >> i) & 1u) &&
                            !((s->demand[node < kRows ? node : 0] + st.used) > lim);
            feas |= (ok ? 1u : 0u) << i;
        }
        s->maskw[j][row] = feas;
        if (newly) s->newvis[row] = 1;
        bar_sync(1 + q, 128);
        if (j == 0 && valid) {
            const bool any_free = (s->maskw[0][row] | s->maskw[1][row] | s->maskw[2][row] | s->maskw[3][row]) != 0u;
            if (!((act == 0) && any_free)) s->maskw[0][row] = feas | 1u;         // :141-144
            st.nvis += s->newvis[row];
            s->newvis[row] = 0;
            s->rem[row] = st.cap - st.used;
            if (!st.done && st.nvis == n) {                                      // :87
                st.done = true;
                st.length = t + 1;
            }
        }
    }
}

// ---- context set-up shared by the kernels --------------------------------------------------------------
__device__ __forceinline__ Ctx make_ctx(unsigned char* smem, int npad) {
    Ctx x;
    x.smem = smem;
    x.npad = npad;
    x.tmem = 0;
    return x;
}

// barriers, TMEM, zeroed Q tiles (rows >= S are never written afterwards and must stay finite)
__device__ __forceinline__ void cta_setup(Ctx& x) {
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        mbar_init(&x.s()->tma_bar, 1);
        mbar_init(&x.s()->mma_bar, 1);   // the logits issuer's commit
        for (int i = 0; i < 4; ++i) mbar_init(&x.s()->team_bar[i], 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(&x.s()->tmem_ptr, kTmemCols);
    for (uint32_t i = threadIdx.x; i < 2u * kQTileBytes / 16u; i += kThreads)
        reinterpret_cast<uint4*>(x.q_p(0))[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < kRows; i += kThreads) {
        x.s()->newvis[i] = 0;
        x.s()->cur[i] = 0;
        x.s()->first[i] = 0;
        x.s()->rem[i] = 0.f;
        x.s()->demand[i] = 0.f;
        x.s()->locs[i] = make_float2(0.f, 0.f);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    x.tmem = x.s()->tmem_ptr;
}

// the instance's six operand tiles: one TMA bulk copy each, one mbarrier phase
__device__ __forceinline__ void load_tiles(const Ctx& x, unsigned char* smem, const unsigned char* img, int64_t b) {
    const uint32_t tb = tile_bytes(x.npad);
    fence_proxy_async();
    mbar_expect_tx(&x.s()->tma_bar, 6u * tb);
    const unsigned char* src = img + b * 6ull * tb;
#pragma unroll
    for (int i = 0; i < 6; ++i) bulk_g2s(x.tiles_p() + (size_t)i * tb, src + (size_t)i * tb, tb, &x.s()->tma_bar);
}

// ---- persistent fused rollout (constructive/base.py:221-263), tensor path ------------------------------
// One instantiation per (env, decode mode): the selection code of the other modes (noise, teacher forcing) would
// otherwise share the 128-register budget of the greedy path.
template <int ENV, int MODE>
__global__ void __launch_bounds__(kThreads, 1)
    rollout_tc_kernel(rl4co_cache c, rl4co_rollout_args a, const unsigned char* __restrict__ img, int npad) {
    extern __shared__ __align__(128) unsigned char smem[];
    Ctx x = make_ctx(smem, npad);
    cta_setup(x);
    Smem* s = x.s();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    const int n = c.n_nodes, S = a.num_starts, T = a.max_steps;
    const bool valid = row < S;
    const int ncust = (ENV == RL4CO_ENV_TSP) ? n : n - 1;
    const int smod = (a.start_mod > 0 && a.start_mod <= ncust) ? a.start_mod : ncust;   // ops.py:147-161
    uint32_t tma_par = 0, mma_par = 0, team_par = 0;
    // optional phase timing (debug): thread 0 of CTA 0 accumulates clock64 deltas per phase
#ifdef R4_PHASE_PROF   // profile builds only (profiles/phase.sh): the counters cost four registers of a full budget
    long long* prof = (a.phase_cycles != nullptr && blockIdx.x == 0 && threadIdx.x == 0) ? a.phase_cycles : nullptr;
    long long tprev = clock64();
#else
    long long* const prof = nullptr;
    long long tprev = 0;
#endif

    for (int64_t b = blockIdx.x; b < c.B; b += gridDim.x) {
        if (threadIdx.x == 0) load_tiles(x, smem, img, b);
        for (int i = threadIdx.x; i < n; i += kThreads) {
            s->locs[i] = reinterpret_cast<const float2*>(a.locs)[b * (int64_t)n + i];
            if (ENV == RL4CO_ENV_CVRP && i >= 1) s->demand[i] = a.demand[b * (int64_t)(n - 1) + i - 1];
        }
        const int64_t grow = (int64_t)row * c.B + b;  // start-major (ops.py:13-16)
        RowState st;
        st.used = 0.f;
        st.cap = (ENV == RL4CO_ENV_CVRP) ? a.vehicle_capacity[b] : 1.f;
        st.ll = 0.f;
        st.len = 0.f;
        st.prev = -1;
        st.first_action = 0;
        st.nvis = 0;
        st.length = 0;
        st.done = false;
        if (ENV == RL4CO_ENV_TSP) {  // tsp/env.py:90-115: every node feasible
            const int rem = n - 32 * j;
            st.word = !valid ? 0u : (rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u)));
        } else {
            st.word = 0u;            // cvrp/env.py:102-130: nothing visited
        }
        int err_bits = 0;
        s->lnf_depot[j][row] = -1.f;
        // forced first action (decoding.py:305-327; select_start_nodes ops.py:147-161)
        int act = 0;
        if (valid) {
            if (MODE == RL4CO_MODE_EVALUATE)
                act = (int)a.actions_in[grow * a.actions_in_stride];
            else
                act = (ENV == RL4CO_ENV_TSP) ? (row % smod) : (row % smod) + 1;
            if (act < 0 || act >= n) {
                err_bits |= RL4CO_ERR_INFEASIBLE;
                act = 0;
            }
        }
        mbar_wait(&s->tma_bar, tma_par);
        tma_par ^= 1u;
        __syncthreads();
        apply_action<ENV>(x, a, st, n, q, j, row, valid, 0, act);
        int all_done = __syncthreads_and(!valid || j != 0 || st.done);
        record_action<ENV>(x, a, st, j, valid, grow, 0, T, act, 0.f);
        R4_STAMP(14);

        float lse[2] = {0.f, 0.f};   // log-sum-exp of this thread's two (row, head) score rows, previous step
        for (int t = 1; !all_done; ++t) {
            if (t >= T) {
                if (valid && j == 0 && !st.done) err_bits |= RL4CO_ERR_MAX_STEPS;
                break;
            }
            decode_step_gemms<ENV>(x, c, b, S, mma_par, team_par, nullptr, nullptr, prof, tprev, lse, t > 1);
            int given = 0;
            if (MODE == RL4CO_MODE_EVALUATE && valid) {
                given = (t < a.actions_in_T) ? (int)a.actions_in[grow * a.actions_in_stride + t] : 0;
                if (given < 0 || given >= n) {
                    if (j == 0) err_bits |= RL4CO_ERR_INFEASIBLE;
                    given = 0;
                }
            }
            const uint32_t nkey = (MODE == RL4CO_MODE_SAMPLING) ? noise_key(a.seed, a.inst_offset + b, row, t) : 0u;
#ifdef R4_WARP_PROF   // experiment: per-warp select / env / step-end wait cycles of CTA 0 -> phase_cycles[16 + 4 warp + k]
            const long long wp0 = clock64();
#endif
            const RowSel sel = select_row<MODE>(x, a, n, q, j, row, valid, grow, t, T, nkey, given, err_bits);
            R4_STAMP(12);
#ifdef R4_WARP_PROF
            const long long wp1 = clock64();
#endif
            apply_action<ENV>(x, a, st, n, q, j, row, valid, t, sel.action);
            tc_fence_before();
#ifdef R4_WARP_PROF
            const long long wp2 = clock64();
#endif
            all_done = __syncthreads_and(!valid || j != 0 || st.done);
#ifdef R4_WARP_PROF
            if (a.phase_cycles != nullptr && blockIdx.x == 0 && lane == 0) {
                const long long wp3 = clock64();
                a.phase_cycles[16 + 4 * warp + 0] += wp1 - wp0;
                a.phase_cycles[16 + 4 * warp + 1] += wp2 - wp1;
                a.phase_cycles[16 + 4 * warp + 2] += wp3 - wp2;
            }
#endif
            R4_STAMP(13);
            record_action<ENV>(x, a, st, j, valid, grow, t, T, sel.action, sel.logp);
        }
        if (valid && j == 2) {
            // closing edge (roll(-1) wraps to the first element: TSP first action, CVRP depot)
            const int to = (ENV == RL4CO_ENV_CVRP) ? 0 : st.first_action;
            const float2 p0 = s->locs[st.prev], p1 = s->locs[to];
            const float dx = p1.x - p0.x, dy = p1.y - p0.y;
            st.len += sqrtf(dx * dx + dy * dy);
            a.reward[grow] = -st.len;
        }
        if (valid && j == 1) a.log_likelihood[grow] = st.ll;
        if (valid && j == 0) {
            const int len = st.done ? st.length : T;
            a.lengths[grow] = len;
            atomicMax(a.max_len, len);
        }
        if (valid && err_bits) raise_err(a.err, err_bits, grow);
        __syncthreads();  // locs / state of this instance are dead before the next one overwrites them
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(x.tmem, kTmemCols);
}

// ---- one decoder step for all rows (am/decoder.py:161-198): RAW logits for every node --------------------
template <int ENV>
__global__ void __launch_bounds__(kThreads, 1)
    decode_step_tc_kernel(rl4co_cache c, const unsigned char* __restrict__ img, int npad, int S,
                          const int64_t* __restrict__ first_node, const int64_t* __restrict__ current_node,
                          const float* __restrict__ used, const float* __restrict__ cap,
                          const uint8_t* __restrict__ mask, float* __restrict__ logits_out,
                          float* __restrict__ dbg_scores, float* __restrict__ dbg_heads, int32_t* err) {
    extern __shared__ __align__(128) unsigned char smem[];
    Ctx x = make_ctx(smem, npad);
    cta_setup(x);
    Smem* s = x.s();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    const int n = c.n_nodes;
    const bool valid = row < S;
    uint32_t tma_par = 0, mma_par = 0, team_par = 0;
    for (int64_t b = blockIdx.x; b < c.B; b += gridDim.x) {
        if (threadIdx.x == 0) load_tiles(x, smem, img, b);
        const int64_t grow = (int64_t)row * c.B + b;
        uint32_t word = 0u;
        if (valid) {
            for (int i = 0; i < 32; ++i) {
                const int node = 32 * j + i;
                if (node < n && mask[grow * n + node] != 0) word |= 1u << i;
            }
            if (j == 0) {
                s->cur[row] = (int)current_node[grow];
                s->first[row] = (ENV == RL4CO_ENV_TSP) ? (int)first_node[grow] : 0;
                s->rem[row] = (ENV == RL4CO_ENV_CVRP) ? cap[grow] - used[grow] : 0.f;
            }
        }
        s->maskw[j][row] = word;
        s->lnf_depot[j][row] = -1.f;
        mbar_wait(&s->tma_bar, tma_par);
        tma_par ^= 1u;
        __syncthreads();
        long long tprev = 0;
        float lse[2] = {0.f, 0.f};
        decode_step_gemms<ENV>(x, c, b, S, mma_par, team_par,
                               (dbg_scores != nullptr && valid) ? dbg_scores + grow * (int64_t)(8 * n) : nullptr,
                               (dbg_heads != nullptr && valid) ? dbg_heads + grow * (int64_t)kE : nullptr, nullptr,
                               tprev, lse, false);
        // this thread's quarter of the logit columns
        const uint32_t lg = x.tmem + ((uint32_t)(32 * q) << 16) + kColLG;
        const int U = (n + 3) >> 2;
        const int u0 = (U * j) >> 2, u1 = (U * (j + 1)) >> 2;
        bool nan_local = false;
        for (int u = u0; u < u1; ++u) {
            uint32_t v4[4];
            tmem_ld4(lg + 4u * u, v4);
            tc_wait_ld();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int node = 4 * u + i;
                if (valid && node < n) {
                    const float v = __uint_as_float(v4[i]);
                    logits_out[grow * n + node] = v;
                    nan_local |= (v != v);
                }
            }
        }
        if (nan_local) raise_err(err, RL4CO_ERR_NAN_LOGITS, grow);  // attention.py:303-304
        tc_fence_before();
        __syncthreads();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(x.tmem, kTmemCols);
}

#include "rollout_tc2.cuh"

}  // namespace tc

// ---- host-side launchers ---------------------------------------------------------------------------------
int check_cache(const rl4co_cache* c, int env, const char* who);

bool tc_supported(const rl4co_cache* c, int S) {
    return c->tc_img != nullptr && S >= 8 && S <= tc::kRows && tc::npad_of(c->n_nodes) <= tc::kMaxNPad;
}

template <typename K>
static int tc_prepare(K kern, int npad, const char* who) {
    const size_t smem = tc::smem_bytes(npad);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        set_error("%s: cudaFuncSetAttribute(%zu B smem): %s", who, smem, cudaGetErrorString(e));
        return RL4CO_E_LAUNCH;
    }
    return RL4CO_OK;
}

static unsigned tc_grid(int64_t B) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return (unsigned)(B < sms ? B : sms);
}

// streaming kernel (operand image v2): two CTAs per SM
static int launch_rollout_tc2(const rl4co_cache* cache, const rl4co_rollout_args& a, cudaStream_t st) {
    const int npad = tc::npad_of(cache->n_nodes);
    const size_t smem = tc::t2::smem_bytes2(npad);
    const unsigned char* img = reinterpret_cast<const unsigned char*>(cache->tc_img);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)(cache->B < 2 * sms ? cache->B : 2 * sms);
#define R4_LAUNCH_TC2(ENV_, MODE_)                                                                                  \
    do {                                                                                                            \
        auto kern = tc::t2::rollout_tc2_kernel<ENV_, MODE_>;                                                        \
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);         \
        if (e != cudaSuccess) {                                                                                     \
            set_error("rollout_tc2: cudaFuncSetAttribute(%zu B smem): %s", smem, cudaGetErrorString(e));            \
            return RL4CO_E_LAUNCH;                                                                                  \
        }                                                                                                           \
        kern<<<grid, tc::t2::kThreads2, smem, st>>>(*cache, a, img, npad);                                          \
    } while (0)
    const bool tsp = a.env == RL4CO_ENV_TSP;
    if (a.mode == RL4CO_MODE_GREEDY) {
        if (tsp) R4_LAUNCH_TC2(RL4CO_ENV_TSP, RL4CO_MODE_GREEDY);
        else R4_LAUNCH_TC2(RL4CO_ENV_CVRP, RL4CO_MODE_GREEDY);
    } else if (a.mode == RL4CO_MODE_SAMPLING) {
        if (tsp) R4_LAUNCH_TC2(RL4CO_ENV_TSP, RL4CO_MODE_SAMPLING);
        else R4_LAUNCH_TC2(RL4CO_ENV_CVRP, RL4CO_MODE_SAMPLING);
    } else {
        if (tsp) R4_LAUNCH_TC2(RL4CO_ENV_TSP, RL4CO_MODE_EVALUATE);
        else R4_LAUNCH_TC2(RL4CO_ENV_CVRP, RL4CO_MODE_EVALUATE);
    }
#undef R4_LAUNCH_TC2
    R4_CHECK_LAUNCH("rollout_tc2");
    return RL4CO_OK;
}

int launch_rollout_tc(const rl4co_cache* cache, const rl4co_rollout_args& a, cudaStream_t st) {
    if (cache->tc_layout == RL4CO_TC_LAYOUT_STREAM) return launch_rollout_tc2(cache, a, st);
    const int npad = tc::npad_of(cache->n_nodes);
    const size_t smem = tc::smem_bytes(npad);
    const unsigned char* img = reinterpret_cast<const unsigned char*>(cache->tc_img);
    int rc;
#define R4_LAUNCH_TC(ENV_, MODE_)                                                                                \
    do {                                                                                                         \
        if ((rc = tc_prepare(tc::rollout_tc_kernel<ENV_, MODE_>, npad, "rollout_tc"))) return rc;                \
        tc::rollout_tc_kernel<ENV_, MODE_><<<tc_grid(cache->B), tc::kThreads, smem, st>>>(*cache, a, img, npad); \
    } while (0)
    const bool tsp = a.env == RL4CO_ENV_TSP;
    if (a.mode == RL4CO_MODE_GREEDY) {
        if (tsp) R4_LAUNCH_TC(RL4CO_ENV_TSP, RL4CO_MODE_GREEDY);
        else R4_LAUNCH_TC(RL4CO_ENV_CVRP, RL4CO_MODE_GREEDY);
    } else if (a.mode == RL4CO_MODE_SAMPLING) {
        if (tsp) R4_LAUNCH_TC(RL4CO_ENV_TSP, RL4CO_MODE_SAMPLING);
        else R4_LAUNCH_TC(RL4CO_ENV_CVRP, RL4CO_MODE_SAMPLING);
    } else {
        if (tsp) R4_LAUNCH_TC(RL4CO_ENV_TSP, RL4CO_MODE_EVALUATE);
        else R4_LAUNCH_TC(RL4CO_ENV_CVRP, RL4CO_MODE_EVALUATE);
    }
#undef R4_LAUNCH_TC
    R4_CHECK_LAUNCH("rollout_tc");
    return RL4CO_OK;
}

}  // namespace r4

using namespace r4;

extern "C" int64_t rl4co_tc_image_bytes(int32_t n_nodes) {
    if (n_nodes < 2 || tc::npad_of(n_nodes) > tc::kMaxNPad) return 0;
    return 6ll * tc::tile_bytes(tc::npad_of(n_nodes));
}

extern "C" int32_t rl4co_tc_preferred_layout(int64_t B) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // two instances per SM only pay when there is more than one instance per SM to run
    return B > sms ? RL4CO_TC_LAYOUT_STREAM : RL4CO_TC_LAYOUT_RESIDENT;
}

extern "C" int rl4co_tc_pack(const rl4co_cache* cache, void* image, int32_t* err, void* stream) {
    R4_CHECK_ARG(cache && cache->gk && cache->gv && cache->lk && cache->B > 0 && image && err, "tc_pack: null pointer");
    R4_CHECK_ARG(cache->embed_dim == kE && cache->num_heads == kH, "tc_pack: only embed_dim=128, num_heads=8");
    R4_CHECK_ARG(rl4co_tc_image_bytes(cache->n_nodes) > 0, "tc_pack: n_nodes=%d outside the tensor path (<= %d)",
                 cache->n_nodes, tc::kMaxNPad);
    R4_CHECK_ARG(cache->tc_layout == RL4CO_TC_LAYOUT_RESIDENT || cache->tc_layout == RL4CO_TC_LAYOUT_STREAM,
                 "tc_pack: unknown tc_layout %d", cache->tc_layout);
    if (cache->tc_layout == RL4CO_TC_LAYOUT_STREAM)
        tc::t2::tc2_pack_kernel<<<(unsigned)cache->B, 256, 0, (cudaStream_t)stream>>>(
            *cache, reinterpret_cast<unsigned char*>(image), tc::npad_of(cache->n_nodes), err);
    else
    tc::tc_pack_kernel<<<(unsigned)cache->B, 256, 0, (cudaStream_t)stream>>>(
        *cache, reinterpret_cast<unsigned char*>(image), tc::npad_of(cache->n_nodes), err);
    R4_CHECK_LAUNCH("tc_pack");
    return RL4CO_OK;
}

extern "C" int rl4co_tc_decode_step(int32_t env, const rl4co_cache* cache, const int64_t* first_node,
                                    const int64_t* current_node, const float* used_capacity,
                                    const float* vehicle_capacity, const uint8_t* mask, float* logits_out,
                                    float* dbg_scores, float* dbg_heads, int32_t* err, int64_t rows,
                                    void* stream) {
    R4_CHECK_ARG(env == RL4CO_ENV_TSP || env == RL4CO_ENV_CVRP, "tc_decode_step: bad env %d", env);
    int rc = check_cache(cache, env, "tc_decode_step");
    if (rc) return rc;
    if (rows == 0) return RL4CO_OK;
    R4_CHECK_ARG(rows > 0 && rows % cache->B == 0, "tc_decode_step: rows=%lld not a multiple of B=%lld",
                 (long long)rows, (long long)cache->B);
    const int S = (int)(rows / cache->B);
    R4_CHECK_ARG(tc_supported(cache, S), "tc_decode_step: needs tc_img, 8 <= S=%d <= 128, n_nodes=%d <= %d", S,
                 cache->n_nodes, tc::kMaxNPad);
    R4_CHECK_ARG(current_node && mask && logits_out && err, "tc_decode_step: null pointer");
    R4_CHECK_ARG(cache->tc_layout == RL4CO_TC_LAYOUT_RESIDENT, "tc_decode_step: needs the resident operand image (tc_layout 0)");
    if (env == RL4CO_ENV_TSP) R4_CHECK_ARG(first_node, "tc_decode_step: TSP needs first_node");
    if (env == RL4CO_ENV_CVRP) R4_CHECK_ARG(used_capacity && vehicle_capacity, "tc_decode_step: CVRP needs capacities");
    const int npad = tc::npad_of(cache->n_nodes);
    const size_t smem = tc::smem_bytes(npad);
    const unsigned char* img = reinterpret_cast<const unsigned char*>(cache->tc_img);
    cudaStream_t st = (cudaStream_t)stream;
    if (env == RL4CO_ENV_TSP) {
        if ((rc = tc_prepare(tc::decode_step_tc_kernel<RL4CO_ENV_TSP>, npad, "tc_decode_step"))) return rc;
        tc::decode_step_tc_kernel<RL4CO_ENV_TSP><<<tc_grid(cache->B), tc::kThreads, smem, st>>>(
            *cache, img, npad, S, first_node, current_node, used_capacity, vehicle_capacity, mask, logits_out,
            dbg_scores, dbg_heads, err);
    } else {
        if ((rc = tc_prepare(tc::decode_step_tc_kernel<RL4CO_ENV_CVRP>, npad, "tc_decode_step"))) return rc;
        tc::decode_step_tc_kernel<RL4CO_ENV_CVRP><<<tc_grid(cache->B), tc::kThreads, smem, st>>>(
            *cache, img, npad, S, first_node, current_node, used_capacity, vehicle_capacity, mask, logits_out,
            dbg_scores, dbg_heads, err);
    }
    R4_CHECK_LAUNCH("tc_decode_step");
    return RL4CO_OK;
}

// ---- allocation-free variant of the tensor-path rollout ---------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  The only scratch the
// tensor-path rollout needs beyond the caller's inputs / outputs is the fp16 hi/lo operand image of the instances
// (rl4co_tc_pack's output, 6 * npad * 256 B per instance).  Here it lives in a file-scope device array of
// RL4CO_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  In HOST code `g_noalloc_scratch` is
// only a stand-in for the symbol, not a device address: its address is resolved once per device, with that device
// current, by cudaGetSymbolAddress, and that pointer is what the kernels receive.
// The pack kernel rewrites every byte of the image it uses (real rows and zero padding alike), so nothing has to be
// cleared between calls; all calls share the array, so calls that overlap on the GPU (different streams) race.
static __device__ __align__(256) unsigned char g_noalloc_scratch[RL4CO_NOALLOC_SCRATCH_BYTES];

namespace r4 {
static int noalloc_scratch_ptr(unsigned char** out) {
    constexpr int kMaxDev = 64;
    static unsigned char* ptrs[kMaxDev] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDev) {
        set_error("rollout_noalloc: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return RL4CO_E_LAUNCH;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);     // device `dev` is current
        if (e != cudaSuccess || p == nullptr) {
            set_error("rollout_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
            return RL4CO_E_LAUNCH;
        }
        ptrs[dev] = static_cast<unsigned char*>(p);
    }
    *out = ptrs[dev];
    return RL4CO_OK;
}
}  // namespace r4

extern "C" int64_t rl4co_noalloc_max_instances(int32_t n_nodes) {
    const int64_t per = rl4co_tc_image_bytes(n_nodes);
    return per > 0 ? (int64_t)RL4CO_NOALLOC_SCRATCH_BYTES / per : 0;
}

extern "C" int rl4co_rollout_noalloc(const rl4co_cache* cache, const rl4co_rollout_args* args, void* stream) {
    R4_CHECK_ARG(args != nullptr, "rollout_noalloc: null args");
    const int env = args->env;
    R4_CHECK_ARG(env == RL4CO_ENV_TSP || env == RL4CO_ENV_CVRP, "rollout_noalloc: bad env %d", env);
    int rc = check_cache(cache, env, "rollout_noalloc");
    if (rc) return rc;
    R4_CHECK_ARG(args->mode >= 0 && args->mode <= 2, "rollout_noalloc: bad mode %d", args->mode);
    R4_CHECK_ARG(args->max_steps > 0 && args->temperature > 0.f, "rollout_noalloc: bad max_steps / temperature");
    R4_CHECK_ARG(args->locs && args->actions && args->logp && args->reward && args->log_likelihood && args->lengths &&
                     args->max_len && args->err, "rollout_noalloc: null output/input pointer");
    if (env == RL4CO_ENV_CVRP)
        R4_CHECK_ARG(args->demand && args->vehicle_capacity, "rollout_noalloc: CVRP needs demand/capacity");
    if (args->mode == RL4CO_MODE_EVALUATE)
        R4_CHECK_ARG(args->actions_in && args->actions_in_T > 0 && args->actions_in_stride >= args->actions_in_T,
                     "rollout_noalloc: evaluate needs actions_in");
    const int S = args->num_starts;
    const int64_t per = rl4co_tc_image_bytes(cache->n_nodes);
    R4_CHECK_ARG(per > 0 && S >= 8 && S <= tc::kRows,
                 "rollout_noalloc: tensor-path shapes only (8 <= num_starts=%d <= 128, n_nodes=%d <= %d)", S,
                 cache->n_nodes, tc::kMaxNPad);
    const int max_starts = (env == RL4CO_ENV_TSP) ? cache->n_nodes : cache->n_nodes - 1;
    R4_CHECK_ARG(args->mode == RL4CO_MODE_EVALUATE || S <= max_starts, "rollout_noalloc: num_starts=%d > %d start nodes",
                 S, max_starts);
    R4_CHECK_ARG(cache->B * per <= (int64_t)RL4CO_NOALLOC_SCRATCH_BYTES,
                 "rollout_noalloc: %lld instances x %lld B of operand image exceed the static scratch of %lld B "
                 "(at most %lld instances at n_nodes=%d); use rl4co_tc_pack + rl4co_rollout with a caller-owned image",
                 (long long)cache->B, (long long)per, (long long)RL4CO_NOALLOC_SCRATCH_BYTES,
                 (long long)((int64_t)RL4CO_NOALLOC_SCRATCH_BYTES / per), cache->n_nodes);
    unsigned char* scratch = nullptr;
    if ((rc = noalloc_scratch_ptr(&scratch))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    rl4co_cache c = *cache;
    c.tc_img = scratch;                                      // any caller-supplied image is ignored
    c.tc_layout = RL4CO_TC_LAYOUT_RESIDENT;
    tc::tc_pack_kernel<<<(unsigned)c.B, 256, 0, st>>>(c, scratch, tc::npad_of(c.n_nodes), args->err);
    R4_CHECK_LAUNCH("rollout_noalloc (operand pack)");
    rl4co_rollout_args a = *args;
    a.impl = RL4CO_IMPL_TC;
    return launch_rollout_tc(&c, a, st);
}
