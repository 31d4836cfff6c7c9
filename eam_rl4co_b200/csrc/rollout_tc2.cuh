// Streaming variant of the tensor-path rollout: TWO instances per SM.  Included by rollout_tc.cu inside namespace r4::tc.
//
// The resident kernel (rollout_tc_kernel) keeps the six operand tiles of an instance (172 KB at N' = 100) in shared memory
// and all 512 TMEM columns, which pins an SM at one instance whose phases (Q gathers, MMA tails, softmax, selection) run
// one after another: 55 % of the issue slots, 32 % of the MUFU and 17 % of the tensor pipe are busy
// (profiles/r02_ncu_rollout_tc.md).  Here a CTA is HALF as wide - 256 threads = 2 teams of 4 warps, 256 TMEM columns, 4 rounds
// of 2 heads per decode step - and the operands are not resident: each team streams the K / V^T slices of the head it is
// working on, and the CTA the L slices of the logits product, from the L2-resident image with TMA bulk copies into a
// 28 KB ring.  A CTA needs 109 KB of shared memory, so two CTAs of DIFFERENT instances share an SM and one's latencies
// are covered by the other's arithmetic.  L2 -> shared traffic: 172 KB per instance-step (2.7 TB/s at the measured rate,
// a quarter of the L2's bandwidth); DRAM traffic is unchanged (the image is read from HBM once per rollout).
//
// Operand image v2 ("head-sliced", tc2_pack_kernel): instance b = 8 head blocks [K_hi][K_lo][Vt_hi][Vt_lo][L_hi][L_lo],
// each slice npad*32 bytes (the 16 channels of head h); one bulk copy fetches (K_hi,K_lo), (Vt_hi,Vt_lo) or (L_hi,L_lo).
//   K, L : rows = nodes, K-dim = the head's 16 channels; byte(r, c) = (r/8)*256 + (c/8)*128 + (r%8)*16 + (c%8)*2
//   Vt   : rows = the head's 16 channels, K-dim = nodes; byte(c, r) = (c/8)*npad*16 + (r/8)*128 + (c%8)*16 + (r%8)*2
#pragma once

namespace t2 {

constexpr int kThreads2 = 256;
constexpr int kTeams = 2;            // warps {4j .. 4j+3}
constexpr int kRounds = 4;           // round g: team j works on head 2g + j
constexpr uint32_t kTmemCols2 = 256;
constexpr uint32_t kColO2 = 224;     // PV accumulators: team j at 224 + 16 j   (S / P of team j at j * NPAD)
constexpr uint32_t kColHA2 = 0;      // end of step: heads A operand hi [0,64) lo [64,128)
constexpr uint32_t kColLG2 = 128;    // logits accumulator [128, 128 + NPAD)
constexpr uint32_t kSBO_SL = 256;    // K / L slices: a row group is its two 16-byte K-chunks
constexpr int kTeamBar2 = 5;         // named barriers 5, 6: the four warps of a team; 1..4: the two warps of a lane quadrant

__host__ __device__ inline uint32_t slice_bytes(int npad) { return (uint32_t)npad * 32u; }       // hi OR lo of one head
__host__ __device__ inline uint32_t pair_bytes(int npad) { return 2u * slice_bytes(npad); }      // hi + lo: one bulk copy

struct Smem2 {
    uint32_t maskw[4][kRows];   // feasible bits, word k = nodes [32k, 32k+32)
    int cur[kRows];
    int first[kRows];
    float rem[kRows];           // CVRP: vehicle_capacity - used_capacity
    float red_a[2][kRows];      // exchanges between the 2 threads of a row
    float red_b[2][kRows];
    float red_c[2][kRows];
    float red_lp[2][kRows];
    int red_i[2][kRows];
    int newvis[kRows];
    float2 locs[kRows];
    float demand[kRows];        // indexed by node (node 0 unused)
    uint64_t kv_bar[2][2];      // [team][0: K slot, 1: V slot]: the slot's bulk copy has landed (5 copies per decode step)
    uint64_t l_bar[4];          // L slices 4..7 landing in the stash area (1 copy per decode step)
    uint64_t mma_bar;           // logits MMAs (one commit)
    uint64_t team_bar[2];       // scores / PV MMAs of one team
    uint32_t tmem_ptr;
    uint32_t n_lload;           // L pairs requested into the stash area so far (issuing lane of warp 0)
    uint32_t n_load[2][2];      // [team][slot]: bulk copies requested so far (written by the team's issuing lane)
    float lnf[2][kRows];        // ln(#feasible nodes) of the row at the previous step (one private copy per thread)
    float lnf_depot[2][kRows];  // CVRP: the same at the row's last step AT THE DEPOT (< 0: none yet) ...
    float lse_depot[4][2][kRows];   // ... and the log-sum-exp of the thread's four heads at that step
};

// shared memory: [Smem2][Q tile team 0][Q tile team 1][stash: 3 rounds x 2 teams x 8 KB][ring: 4 slots of pair_bytes]
constexpr uint32_t kOffQ2 = (uint32_t)((sizeof(Smem2) + 127) / 128 * 128);
constexpr uint32_t kOffStash2 = kOffQ2 + 2u * kQTeamBytes;
constexpr uint32_t kStashBytes2 = 3u * 2u * kQTeamBytes;
constexpr uint32_t kOffRing2 = kOffStash2 + kStashBytes2;

struct Ctx2 {
    unsigned char* smem;
    uint32_t tmem;
    int npad;
    __device__ __forceinline__ Smem2* s() const { return reinterpret_cast<Smem2*>(smem); }
    __device__ __forceinline__ unsigned char* q_p(int team) const { return smem + kOffQ2 + (uint32_t)team * kQTeamBytes; }
    __device__ __forceinline__ uint32_t base() const { return smem_u32(smem); }
    __device__ __forceinline__ uint32_t q_hi(int team) const { return base() + kOffQ2 + (uint32_t)team * kQTeamBytes; }
    __device__ __forceinline__ uint32_t q_lo(int team) const { return q_hi(team) + kQHalfBytes; }
    __device__ __forceinline__ float* stash(int g, int team) const {
        return reinterpret_cast<float*>(smem + kOffStash2 + (uint32_t)(2 * g + team) * kQTeamBytes);
    }
    // ring slot (team, kv): a (hi, lo) pair of slices
    __device__ __forceinline__ unsigned char* slot_p(int team, int kv) const {
        return smem + kOffRing2 + (uint32_t)(2 * team + kv) * pair_bytes(npad);
    }
    __device__ __forceinline__ uint32_t slot(int team, int kv) const {
        return base() + kOffRing2 + (uint32_t)(2 * team + kv) * pair_bytes(npad);
    }
    // L slices 4..7 land in the (then idle) stash area
    __device__ __forceinline__ unsigned char* lst_p(int i) const { return smem + kOffStash2 + (uint32_t)i * pair_bytes(npad); }
    __device__ __forceinline__ uint32_t lst(int i) const { return base() + kOffStash2 + (uint32_t)i * pair_bytes(npad); }
    __device__ __forceinline__ int nch() const { return npad >> 4; }
    __device__ __forceinline__ uint32_t sbo_v() const { return (uint32_t)npad * 16u; }
};

__host__ inline size_t smem_bytes2(int npad) { return (size_t)kOffRing2 + 4ull * pair_bytes(npad); }

// ---- operand image v2 ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tc2_pack_kernel(rl4co_cache c, unsigned char* __restrict__ img, int npad,
                                                       int32_t* err) {
    const int64_t b = blockIdx.x;
    const int n = c.n_nodes;
    const uint32_t sl = slice_bytes(npad);
    unsigned char* im = img + b * 48ull * sl;
    const float* gk = c.gk + b * (int64_t)n * kE;
    const float* gv = c.gv + b * (int64_t)n * kE;
    const float* lk = c.lk + b * (int64_t)n * kE;
    const uint32_t sbo_v = (uint32_t)npad * 16u;
    bool bad = false;
    // K and L: one item = (row r, 8-channel chunk)
    for (int it = threadIdx.x; it < npad * 16; it += blockDim.x) {
        const int r = it >> 4, ck = it & 15, ch = ck * 8, h = ck >> 1;
        const uint32_t off = (uint32_t)(r >> 3) * kSBO_SL + (uint32_t)(ck & 1) * kLBO + (uint32_t)(r & 7) * 16u;
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
        unsigned char* hb = im + (size_t)(6 * h) * sl + off;
        *reinterpret_cast<uint4*>(hb + 0ull * sl) = khi;
        *reinterpret_cast<uint4*>(hb + 1ull * sl) = klo;
        *reinterpret_cast<uint4*>(hb + 4ull * sl) = lhi;
        *reinterpret_cast<uint4*>(hb + 5ull * sl) = llo;
    }
    // Vt: one item = (8-node chunk rc, channel ch); lanes run over channels (coalesced reads)
    for (int it = threadIdx.x; it < (npad / 8) * kE; it += blockDim.x) {
        const int rc = it >> 7, ch = it & 127, h = ch >> 4, c16 = ch & 15;
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
        const uint32_t off = (uint32_t)(c16 >> 3) * sbo_v + (uint32_t)rc * kLBO + (uint32_t)(c16 & 7) * 16u;
        unsigned char* hb = im + (size_t)(6 * h) * sl + off;
        *reinterpret_cast<uint4*>(hb + 2ull * sl) = make_uint4(vh[0], vh[1], vh[2], vh[3]);
        *reinterpret_cast<uint4*>(hb + 3ull * sl) = make_uint4(vl[0], vl[1], vl[2], vl[3]);
    }
    if (bad) raise_err(err, RL4CO_ERR_FP16_RANGE, b);
}

// ---- MMA issue (one elected lane of the team's issuer warp) ---------------------------------------------------
// scores of the head whose K pair sits in the team's K slot: S = Q . K_h^T, hi*hi + hi*lo + lo*hi
template <class CX>
__device__ __forceinline__ void issue_scores2(const CX& x, int team) {
    tc_fence_after();
    const uint32_t id = idesc_f16(kRows, x.npad);
    const uint64_t aq_hi = smem_desc(x.q_hi(team), kLBO_Q, kSBO_Q);
    const uint64_t aq_lo = smem_desc(x.q_lo(team), kLBO_Q, kSBO_Q);
    const uint32_t ks = x.slot(team, 0);
    const uint64_t bk_hi = smem_desc(ks, kLBO, kSBO_SL);
    const uint64_t bk_lo = smem_desc(ks + slice_bytes(x.npad), kLBO, kSBO_SL);
    const uint32_t d = x.tmem + (uint32_t)(team * x.npad);
    mma_ss(d, aq_hi, bk_hi, id, 0u);
    mma_ss(d, aq_hi, bk_lo, id, 1u);
    mma_ss(d, aq_lo, bk_hi, id, 1u);
    tc_commit(&x.s()->team_bar[team]);
}

// O = P . V_h : A = P from TMEM, B = the V^T pair in the team's V slot
template <class CX>
__device__ __forceinline__ void issue_pv2(const CX& x, int team) {
    tc_fence_after();
    const uint32_t id = idesc_f16(kRows, 16);
    const uint32_t vs = x.slot(team, 1);
    uint64_t bv_hi = smem_desc(vs, kLBO, x.sbo_v());
    uint64_t bv_lo = smem_desc(vs + slice_bytes(x.npad), kLBO, x.sbo_v());
    const uint32_t d = x.tmem + kColO2 + 16u * team;
    uint32_t a = x.tmem + (uint32_t)(team * x.npad);
    for (int jk = 0; jk < x.nch(); ++jk) {
        mma_ts(d, a, bv_hi, id, jk > 0 ? 1u : 0u);
        mma_ts(d, a, bv_lo, id, 1u);
        mma_ts(d, a + 8u, bv_hi, id, 1u);
        a += 16u;
        bv_hi += (2u * kLBO) >> 4;
        bv_lo += (2u * kLBO) >> 4;
    }
    tc_commit(&x.s()->team_bar[team]);
}

// one (hi, lo) pair of slices of instance image `img_b` -> ring slot; `which`: 0 = K, 1 = Vt, 2 = L of head h
template <class CX>
__device__ __forceinline__ void load_pair(const CX& x, const unsigned char* img_b, int h, int which, void* dst,
                                          uint64_t* bar) {
    const uint32_t sl = slice_bytes(x.npad);
    mbar_expect_tx(bar, 2u * sl);
    bulk_g2s(dst, img_b + (size_t)(6 * h + 2 * which) * sl, 2u * sl, bar);
}
// the team's K / V slot <- pair; the issuing lane counts the loads of a slot, a waiter takes the parity from the count
template <class CX>
__device__ __forceinline__ void load_slot(const CX& x, const unsigned char* img_b, int h, int which, int team, int kv) {
    x.s()->n_load[team][kv] += 1u;
    load_pair(x, img_b, h, which, x.slot_p(team, kv), &x.s()->kv_bar[team][kv]);
}
template <class CX>
__device__ __forceinline__ void wait_slot(const CX& x, int team, int kv) {
    mbar_wait(&x.s()->kv_bar[team][kv], (x.s()->n_load[team][kv] - 1u) & 1u);
}

// logits = heads . (L W_out / sqrt(E))^T : k-step jk = head jk, its L pair sits in slot (jk & 1, jk >> 1) for jk < 4 and in
// the stash area for jk >= 4 (requested here: the stash was consumed before the barrier that precedes this call)
template <class CX>
__device__ __forceinline__ void issue_logits2(const CX& x, const unsigned char* img_b) {
    auto* s = x.s();
    fence_proxy_async();   // generic-proxy reads of the stash before the async-proxy overwrite
    s->n_lload += 1u;
#pragma unroll
    for (int i = 0; i < 4; ++i) load_pair(x, img_b, 4 + i, 2, x.lst_p(i), &s->l_bar[i]);
    tc_fence_after();
    const uint32_t id = idesc_f16(kRows, x.npad);
    const uint32_t d = x.tmem + kColLG2;
    const uint32_t sl = slice_bytes(x.npad);
#pragma unroll
    for (int jk = 0; jk < 8; ++jk) {
        uint32_t ls;
        if (jk < 4) {
            wait_slot(x, jk & 1, jk >> 1);
            ls = x.slot(jk & 1, jk >> 1);
        } else {
            mbar_wait(&s->l_bar[jk - 4], (s->n_lload - 1u) & 1u);
            ls = x.lst(jk - 4);
        }
        tc_fence_after();
        const uint64_t bl_hi = smem_desc(ls, kLBO, kSBO_SL);
        const uint64_t bl_lo = smem_desc(ls + sl, kLBO, kSBO_SL);
        const uint32_t a = x.tmem + kColHA2 + 8u * jk;
        mma_ts(d, a, bl_hi, id, jk > 0 ? 1u : 0u);
        mma_ts(d, a, bl_lo, id, 1u);
        mma_ts(d, a + 64u, bl_hi, id, 1u);
    }
    tc_commit(&s->mma_bar);
}

// ---- all GEMM phases of one decoder step: on return the raw logits sit in TMEM columns [kColLG2, kColLG2 + n) ----------
// Team j (warps 4j .. 4j+3) works on head 2g + j in round g; thread = (row 32q + lane, team j).  The speculative softmax
// reference is the resident kernel's (see rollout_tc.cu), with four (row, head) log-sum-exps per thread.
template <int ENV>
__device__ __forceinline__ void decode_step_gemms2(const Ctx2& x, const rl4co_cache& c, const unsigned char* img_b,
                                                   int64_t b, int S, uint32_t& mma_par, uint32_t& team_par,
                                                   float (&lse)[4], bool have_ref) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    const uint32_t lane_base = x.tmem + ((uint32_t)(32 * q) << 16);
    const int n = c.n_nodes;
    const bool issuer_warp = (q == j);
    Smem2* s = x.s();
    uint64_t* tbar = &s->team_bar[j];
    const float kLog2e = 1.4426950408889634f;
    float hv[16];
    const bool at_depot = (ENV == RL4CO_ENV_CVRP) && s->cur[row] == 0;
    build_q<ENV>(x, c, b, S, j, j, q, lane);
    team_sync(j);
    if (issuer_warp) {
        if (elect_one()) {
            wait_slot(x, j, 0);
            issue_scores2(x, j);
        }
    }
    {
        const int nf = __popc(s->maskw[0][row]) + __popc(s->maskw[1][row]) + __popc(s->maskw[2][row]) +
                       __popc(s->maskw[3][row]);
        const float lnf = __logf((float)(nf > 0 ? nf : 1));
        const float dl = lnf - s->lnf[j][row];
        float shift = have_ref ? (dl > 0.f ? 1.0f : 0.5f) * dl : 0.f;
        if (ENV == RL4CO_ENV_CVRP && have_ref && at_depot && s->lnf_depot[j][row] >= 0.f) {
            shift = lnf - s->lnf_depot[j][row];
#pragma unroll
            for (int g = 0; g < kRounds; ++g) lse[g] = s->lse_depot[g][j][row];
        }
        s->lnf[j][row] = lnf;
#pragma unroll
        for (int g = 0; g < kRounds; ++g) lse[g] += shift;
    }
    // a real loop (not unrolled): four inlined copies of the softmax thrash the instruction cache (warp stall "no
    // instruction" 0.61 per issue against 0.17 for the two-round resident kernel)
#pragma unroll 1
    for (int g = 0; g < kRounds; ++g) {
        const int h = 2 * g + j;
        wait_mma(tbar, team_par);                      // scores of round g
        const uint32_t sp = lane_base + (uint32_t)(j * x.npad);
        float l, m;
        bool spec = have_ref;
        const float lse_g = g == 0 ? lse[0] : (g == 1 ? lse[1] : (g == 2 ? lse[2] : lse[3]));
        const float ref = lse_g - kSpecTargetLn;
        for (;;) {
            l = softmax_inplace(x, sp, row, nullptr, n, spec, ref * kLog2e, m);
            if (spec) m = ref;
            const bool ok = !spec || (l == 0.f) || (l >= kSpecLMin && l < kSpecLMax);
            tc_fence_before();
            if (team_all(j, ok)) break;
            // rare: the reference missed somewhere in the team -> scores again (K slot and Q tile are intact), exact form
            spec = false;
            if (issuer_warp) { if (elect_one()) issue_scores2(x, j); }
            wait_mma(tbar, team_par);
        }
        const float inv_l = l > 0.f ? 1.0f / l : 0.f;
        const float lse_new = l > 0.f ? m + __logf(l) : 0.f;
        if (g == 0) lse[0] = lse_new;
        else if (g == 1) lse[1] = lse_new;
        else if (g == 2) lse[2] = lse_new;
        else lse[3] = lse_new;
        if (ENV == RL4CO_ENV_CVRP && at_depot) {
            s->lse_depot[g][j][row] = lse_new;
            if (g == kRounds - 1) s->lnf_depot[j][row] = s->lnf[j][row];
        }
        if (issuer_warp) {
            if (elect_one()) {
                wait_slot(x, j, 1);                    // V^T pair of head h
                issue_pv2(x, j);
                // scores(g) are final: the K slot takes the next round's K pair, or this team's first L pair (head j)
                if (g < kRounds - 1) load_slot(x, img_b, h + 2, 0, j, 0);
                else load_slot(x, img_b, j, 2, j, 0);
            }
        }
        if (g < kRounds - 1) build_q<ENV>(x, c, b, S, h + 2, j, q, lane);   // next round's Q while PV runs
        wait_mma(tbar, team_par);                      // PV of round g
        if (issuer_warp) {
            if (elect_one()) {
                if (g < kRounds - 1) load_slot(x, img_b, h + 2, 1, j, 1);
                else load_slot(x, img_b, 2 + j, 2, j, 1);            // second L pair (head 2 + j)
            }
        }
        if (g < kRounds - 1) {                         // next scores run while O(g) is drained
            team_sync(j);
            if (issuer_warp) {
                if (elect_one()) {
                    wait_slot(x, j, 0);
                    issue_scores2(x, j);
                }
            }
        }
        uint32_t o[16];
        tmem_ld16(lane_base + kColO2 + 16u * j, o);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) hv[i] = __uint_as_float(o[i]) * inv_l;
        tc_fence_before();
        if (g < kRounds - 1) {
            float* st = x.stash(g, j) + (threadIdx.x & 127);
#pragma unroll
            for (int i = 0; i < 16; ++i) st[i * 128] = hv[i];
        }
    }
    // both teams are done with their probabilities: the head operand and the logits accumulator reuse those columns
    __syncthreads();
    tc_fence_after();
#pragma unroll
    for (int g = kRounds - 1; g >= 0; --g) {
        const int h = 2 * g + j;
        if (g < kRounds - 1) {
            const float* st = x.stash(g, j) + (threadIdx.x & 127);
#pragma unroll
            for (int i = 0; i < 16; ++i) hv[i] = st[i * 128];
        }
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) split2(hv[2 * i], hv[2 * i + 1], hi[i], lo[i]);
        tmem_st8(lane_base + kColHA2 + 8u * h, hi);
        tmem_st8(lane_base + kColHA2 + 64u + 8u * h, lo);
    }
    tc_wait_st();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { if (elect_one()) issue_logits2(x, img_b); }
    wait_mma(&s->mma_bar, mma_par);
}

// ---- selection: process_logits + Greedy / Sampling / Evaluate (decoding.py:140-190, 346-465), TWO threads per row ----
// Thread j takes the j-th half of the logit columns (units of 4 columns) in two segments of <= 28 columns; sum / argmax
// are combined through shared memory behind a 64-thread named barrier per lane quadrant.  Same arithmetic as select_row.
__device__ __forceinline__ uint32_t load_seg(uint32_t lg, int sh, int cnt, uint32_t w0, uint32_t w1, uint32_t w2,
                                             uint32_t w3, bool clipped, float (&tv)[28]) {
    {
        uint32_t a16[16], a8[8], a4[4];
        tmem_ld16(lg + (uint32_t)sh, a16);
        tmem_ld8(lg + (uint32_t)sh + 16u, a8);
        tmem_ld4(lg + (uint32_t)sh + 24u, a4);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) tv[i] = __uint_as_float(a16[i]);
#pragma unroll
        for (int i = 0; i < 8; ++i) tv[16 + i] = __uint_as_float(a8[i]);
#pragma unroll
        for (int i = 0; i < 4; ++i) tv[24 + i] = __uint_as_float(a4[i]);
    }
    const int wi = sh >> 5;
    const uint32_t lo = wi == 0 ? w0 : (wi == 1 ? w1 : (wi == 2 ? w2 : w3));
    const uint32_t hi = wi == 0 ? w1 : (wi == 1 ? w2 : (wi == 2 ? w3 : 0u));
    const uint32_t inrange = cnt >= 32 ? 0xffffffffu : ((1u << cnt) - 1u);
    const uint32_t win = __funnelshift_r(lo, hi, sh & 31) & inrange;   // feasible bits of nodes sh .. sh+27
    if (clipped) {                                                     // decoding.py:171-172
        float am0 = 0.f, am1 = 0.f;
#pragma unroll
        for (int e = 0; e < 28; e += 2) {
            am0 = fmaxf(am0, fabsf(tv[e]));
            am1 = fmaxf(am1, fabsf(tv[e + 1]));
        }
        if (__all_sync(kFull, fmaxf(am0, am1) <= 1.0f)) {
#pragma unroll
            for (int e = 0; e < 28; ++e) tv[e] = tanh_small(tv[e]);
        } else {
#pragma unroll
            for (int e = 0; e < 28; ++e) tv[e] = tanh_fast(tv[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < 28; ++e) tv[e] = ((win >> e) & 1u) ? tv[e] : -INFINITY;   // :177
    return win;
}

template <int MODE>
__device__ __forceinline__ RowSel select_row2(const Ctx2& x, const rl4co_rollout_args& a, int n, int q, int j, int row,
                                              bool valid, int64_t grow, int t, int T, uint32_t nkey, int given,
                                              int& err_bits) {
    Smem2* s = x.s();
    const uint32_t lg = x.tmem + ((uint32_t)(32 * q) << 16) + kColLG2;
    const int U = (n + 3) >> 2;
    const int ua = (U * j) >> 1, ub = (U * (j + 1)) >> 1;
    const int um = (ua + ub + 1) >> 1;                              // segments [ua, um) and [um, ub): <= 7 units each
    const uint32_t w0 = s->maskw[0][row], w1 = s->maskw[1][row], w2 = s->maskw[2][row], w3 = s->maskw[3][row];
    const bool clipped = a.tanh_clipping > 0.f;
    const float sc = (clipped ? a.tanh_clipping : 1.0f) / a.temperature;
    const float kL2e = 1.4426950408889634f;
    constexpr bool sampling = MODE == RL4CO_MODE_SAMPLING;
    float tv[28];
    float ref = sc;                                                 // constant reference: see select_row
    if (!(clipped && sc <= 40.0f)) {
        float lmax = -INFINITY;
#pragma unroll 1
        for (int sg = 0; sg < 2; ++sg) {
            const int u0 = sg ? um : ua, u1 = sg ? ub : um;
            load_seg(lg, 4 * u0, 4 * (u1 - u0), w0, w1, w2, w3, clipped, tv);
#pragma unroll
            for (int e = 0; e < 28; ++e) lmax = fmaxf(lmax, tv[e]);
        }
        s->red_a[j][row] = lmax;
        bar_sync(1 + q, 64);
        const float m = fmaxf(s->red_a[0][row], s->red_a[1][row]);
        ref = (m == -INFINITY) ? 0.f : m * sc;
    }
    const float c1 = sc * kL2e, c0 = -ref * kL2e;
    float ls0 = 0.f, ls1 = 0.f;
    float best = sampling ? 0.f : -INFINITY, best_x = -INFINITY;    // counter-noise keys are >= 0: 0 never wins
    int best_i = 0x7fffffff;
    const bool counter_noise = sampling && a.gumbel == nullptr;
#pragma unroll 1
    for (int sg = 0; sg < 2; ++sg) {
        const int u0 = sg ? um : ua, u1 = sg ? ub : um;
        const int sh = 4 * u0, cnt = 4 * (u1 - u0);
        load_seg(lg, sh, cnt, w0, w1, w2, w3, clipped, tv);
#pragma unroll
        for (int e = 0; e < 28; e += 2) {
            ls0 += ex2_approx(fmaf(tv[e], c1, c0));
            ls1 += ex2_approx(fmaf(tv[e + 1], c1, c0));
        }
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
        } else if (counter_noise) {   // exponential race, see select_row
#pragma unroll
            for (int e = 0; e < 28; ++e) {
                const float u = noise_u(nkey, sh + e);
                const float d = 1.0f - u;
                const float ser = d * fmaf(d, fmaf(d, fmaf(d, 0.36067376f, 0.48089835f), 0.72134752f), 1.44269504f);
                float l2;
                asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(u));
                const float big_e = u > 0.98f ? ser : -l2;
                float r;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(big_e));
                const float cand = ex2_approx(fmaf(tv[e], c1, c0)) * r;
                if (cand > best) {
                    best = cand;
                    best_i = sh + e;
                    best_x = tv[e];
                }
            }
        }
    }
    s->red_b[j][row] = ls0 + ls1;
    if (!(sampling && !counter_noise)) {
        s->red_c[j][row] = best;
        s->red_i[j][row] = best_i;
        s->red_lp[j][row] = best_x;
    }
    bar_sync(1 + q, 64);
    const float tot = s->red_b[0][row] + s->red_b[1][row];
    if ((tot != tot || !(tot > 0.f)) && valid) err_bits |= RL4CO_ERR_NAN_LOGITS;   // attention.py:303-304
    const float lse = logf(tot);                                    // log_softmax(x)_e = (x_e - ref) - lse  :190
    if (a.logp_all != nullptr || (sampling && !counter_noise)) {    // second pass over the logits (still in TMEM)
        best = -INFINITY;
#pragma unroll 1
        for (int sg = 0; sg < 2; ++sg) {
            const int u0 = sg ? um : ua, u1 = sg ? ub : um;
            const int sh = 4 * u0, cnt = 4 * (u1 - u0);
            const uint32_t win = load_seg(lg, sh, cnt, w0, w1, w2, w3, clipped, tv);
            if (a.logp_all != nullptr && valid) {
#pragma unroll
                for (int e = 0; e < 28; ++e)
                    if (e < cnt && sh + e < n) a.logp_all[(grow * T + t) * (int64_t)n + sh + e] = (tv[e] * sc - ref) - lse;
            }
            if (sampling && !counter_noise) {                       // Gumbel-max over logp + g with supplied noise
#pragma unroll
                for (int e = 0; e < 28; ++e) {
                    const float lp = (tv[e] * sc - ref) - lse;
                    float g = 0.f;
                    if ((win >> e) & 1u) g = a.gumbel[(grow * T + t) * (int64_t)n + sh + e];
                    const float cand = lp + g;
                    if (cand > best) {
                        best = cand;
                        best_i = sh + e;
                        best_x = tv[e];
                    }
                }
            }
        }
        if (sampling && !counter_noise) {
            s->red_c[j][row] = best;
            s->red_i[j][row] = best_i;
            s->red_lp[j][row] = best_x;
            bar_sync(1 + q, 64);
        }
    }
    float fb = s->red_c[0][row];
    int fi = s->red_i[0][row];
    float fx = s->red_lp[0][row];
    {
        const float cb = s->red_c[1][row];
        if (cb > fb) {
            fb = cb;
            fi = s->red_i[1][row];
            fx = s->red_lp[1][row];
        }
    }
    RowSel sel;
    sel.action = (fi == 0x7fffffff) ? 0 : fi;
    sel.logp = (fx * sc - ref) - lse;
    if (valid && j == 0) {
        const int wa = sel.action >> 5;
        const uint32_t ww = wa == 0 ? w0 : (wa == 1 ? w1 : (wa == 2 ? w2 : w3));
        if (MODE != RL4CO_MODE_EVALUATE && !((ww >> (sel.action & 31)) & 1u)) err_bits |= RL4CO_ERR_INFEASIBLE;
    }
    if (valid && j == 1 && !(sel.logp > -1000.f)) err_bits |= RL4CO_ERR_NEG_INF_LOGP;   // decoding.py:56-58
    return sel;
}

// ---- env step (tsp/env.py:62-88, cvrp/env.py:68-100,132-144), 2 threads per row -------------------------------------
// thread 0: current / first node, done flag, visit count, the action record; thread 1: log-likelihood, tour length.
// Thread j owns the mask words j and j + 2.
struct RowState2 {
    uint32_t word[2];   // TSP: feasible bits; CVRP: visited bits
    float used, cap;
    float ll, len;      // thread 1
    int prev, first_action;   // thread 1
    int nvis, length;         // thread 0
    bool done;                // thread 0
};

template <int ENV>
__device__ __forceinline__ void record_action2(const Ctx2& x, const rl4co_rollout_args& a, RowState2& st, int j, bool valid,
                                               int64_t grow, int t, int T, int act, float lpa) {
    if (!valid) return;
    if (j == 1) {
        a.logp[grow * T + t] = lpa;
        st.ll += lpa;
        const int from = (st.prev < 0) ? ((ENV == RL4CO_ENV_CVRP) ? 0 : act) : st.prev;   // ops.py:82-95
        const float2 p0 = x.s()->locs[from], p1 = x.s()->locs[act];
        const float dx = p1.x - p0.x, dy = p1.y - p0.y;
        st.len += sqrtf(dx * dx + dy * dy);
        if (st.prev < 0) st.first_action = act;
        st.prev = act;
    } else {
        a.actions[grow * T + t] = act;
    }
}

template <int ENV>
__device__ __forceinline__ void apply_action2(const Ctx2& x, RowState2& st, int n, int q, int j, int row, bool valid,
                                              int t, int act) {
    Smem2* s = x.s();
    if (j == 0 && valid) s->cur[row] = act;
    if (ENV == RL4CO_ENV_TSP) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int wi = j + 2 * k;
            if (valid && (act >> 5) == wi) st.word[k] &= ~(1u << (act & 31));
            s->maskw[wi][row] = st.word[k];
        }
        if (j == 0 && valid) {
            if (t == 0) s->first[row] = act;
            if (!st.done && t + 1 >= n) {
                st.done = true;
                st.length = t + 1;
            }
        }
    } else {
        int di = act - 1;
        di = di < 0 ? 0 : (di > n - 2 ? n - 2 : di);
        st.used = (st.used + s->demand[di + 1]) * ((act != 0) ? 1.0f : 0.0f);   // cvrp/env.py:73-80
        const float lim = st.cap + 1e-5f;                                       // :135
        bool newly = false;
        uint32_t feas0 = 0u;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int wi = j + 2 * k;
            if (valid && (act >> 5) == wi) {
                newly = !((st.word[k] >> (act & 31)) & 1u);
                st.word[k] |= 1u << (act & 31);
            }
            uint32_t feas = 0u;
#pragma unroll 8
            for (int i = 0; i < 32; ++i) {
                const int node = 32 * wi + i;
                const bool ok = valid && node >= 1 && node < n && !((st.word[k] >> i) & 1u) &&
                                !((s->demand[node < kRows ? node : 0] + st.used) > lim);
                feas |= (ok ? 1u : 0u) << i;
            }
            s->maskw[wi][row] = feas;
            if (k == 0) feas0 = feas;
        }
        if (newly) s->newvis[row] = 1;
        bar_sync(1 + q, 64);
        if (j == 0 && valid) {
            const bool any_free = (s->maskw[0][row] | s->maskw[1][row] | s->maskw[2][row] | s->maskw[3][row]) != 0u;
            if (!((act == 0) && any_free)) s->maskw[0][row] = feas0 | 1u;        // :141-144
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

// barriers, TMEM, zeroed Q tiles (rows >= S are never written afterwards and must stay finite)
__device__ __forceinline__ void cta_setup2(Ctx2& x) {
    const int warp = threadIdx.x >> 5;
    Smem2* s = x.s();
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) {
            mbar_init(&s->kv_bar[i >> 1][i & 1], 1);
            mbar_init(&s->l_bar[i], 1);
            s->n_load[i >> 1][i & 1] = 0u;
        }
        s->n_lload = 0u;
        mbar_init(&s->mma_bar, 1);
        mbar_init(&s->team_bar[0], 1);
        mbar_init(&s->team_bar[1], 1);
        fence_mbar_init();
    }
    __syncwarp();
    if (warp == 0) tmem_alloc(&s->tmem_ptr, kTmemCols2);
    for (uint32_t i = threadIdx.x; i < 2u * kQTeamBytes / 16u; i += kThreads2)
        reinterpret_cast<uint4*>(x.q_p(0))[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < kRows; i += kThreads2) {
        s->newvis[i] = 0;
        s->cur[i] = 0;
        s->first[i] = 0;
        s->rem[i] = 0.f;
        s->demand[i] = 0.f;
        s->locs[i] = make_float2(0.f, 0.f);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    x.tmem = s->tmem_ptr;
}

// ---- persistent fused rollout, streaming tensor path: two CTAs (instances) per SM ---------------------------------------
#ifndef R4_TC2_MINBLOCKS
#define R4_TC2_MINBLOCKS 2
#endif
template <int ENV, int MODE>
__global__ void __launch_bounds__(kThreads2, R4_TC2_MINBLOCKS)
    rollout_tc2_kernel(rl4co_cache c, rl4co_rollout_args a, const unsigned char* __restrict__ img, int npad) {
    extern __shared__ __align__(128) unsigned char smem[];
    Ctx2 x;
    x.smem = smem;
    x.npad = npad;
    x.tmem = 0;
    cta_setup2(x);
    Smem2* s = x.s();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, j = warp >> 2, row = 32 * q + lane;
    const int n = c.n_nodes, S = a.num_starts, T = a.max_steps;
    const bool valid = row < S;
    const bool issuer_warp = (q == j);
    const int ncust = (ENV == RL4CO_ENV_TSP) ? n : n - 1;
    const int smod = (a.start_mod > 0 && a.start_mod <= ncust) ? a.start_mod : ncust;   // ops.py:147-161
    const size_t img_stride = 48ull * slice_bytes(npad);
    uint32_t mma_par = 0, team_par = 0;

    // The K / V^T pairs of the FIRST round of a decode step are requested ahead of it: here for this CTA's first instance,
    // later at the end of the previous decode step.  `pend` = the instance the outstanding request is for (-1: none).
    int64_t pend = -1;
    if ((int64_t)blockIdx.x < c.B) {
        pend = blockIdx.x;
        if (issuer_warp) {
            if (elect_one()) {
                const unsigned char* ib = img + (size_t)blockIdx.x * img_stride;
                load_slot(x, ib, j, 0, j, 0);
                load_slot(x, ib, j, 1, j, 1);
            }
        }
    }
    for (int64_t b = blockIdx.x; b < c.B; b += gridDim.x) {
        const unsigned char* img_b = img + (size_t)b * img_stride;
        for (int i = threadIdx.x; i < n; i += kThreads2) {
            s->locs[i] = reinterpret_cast<const float2*>(a.locs)[b * (int64_t)n + i];
            if (ENV == RL4CO_ENV_CVRP && i >= 1) s->demand[i] = a.demand[b * (int64_t)(n - 1) + i - 1];
        }
        const int64_t grow = (int64_t)row * c.B + b;  // start-major (ops.py:13-16)
        RowState2 st;
        st.used = 0.f;
        st.cap = (ENV == RL4CO_ENV_CVRP) ? a.vehicle_capacity[b] : 1.f;
        st.ll = 0.f;
        st.len = 0.f;
        st.prev = -1;
        st.first_action = 0;
        st.nvis = 0;
        st.length = 0;
        st.done = false;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            if (ENV == RL4CO_ENV_TSP) {  // tsp/env.py:90-115: every node feasible
                const int rem = n - 32 * (j + 2 * k);
                st.word[k] = !valid ? 0u : (rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u)));
            } else {
                st.word[k] = 0u;         // cvrp/env.py:102-130: nothing visited
            }
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
        __syncthreads();
        apply_action2<ENV>(x, st, n, q, j, row, valid, 0, act);
        int all_done = __syncthreads_and(!valid || j != 0 || st.done);
        record_action2<ENV>(x, a, st, j, valid, grow, 0, T, act, 0.f);

        float lse[4] = {0.f, 0.f, 0.f, 0.f};   // log-sum-exp of this thread's four (row, head) score rows, previous step
        for (int t = 1; !all_done; ++t) {
            if (t >= T) {
                if (valid && j == 0 && !st.done) err_bits |= RL4CO_ERR_MAX_STEPS;
                break;
            }
            decode_step_gemms2<ENV>(x, c, img_b, b, S, mma_par, team_par, lse, t > 1);
            pend = -1;
            int given = 0;
            if (MODE == RL4CO_MODE_EVALUATE && valid) {
                given = (t < a.actions_in_T) ? (int)a.actions_in[grow * a.actions_in_stride + t] : 0;
                if (given < 0 || given >= n) {
                    if (j == 0) err_bits |= RL4CO_ERR_INFEASIBLE;
                    given = 0;
                }
            }
            const uint32_t nkey = (MODE == RL4CO_MODE_SAMPLING) ? noise_key(a.seed, a.inst_offset + b, row, t) : 0u;
            const RowSel sel = select_row2<MODE>(x, a, n, q, j, row, valid, grow, t, T, nkey, given, err_bits);
            apply_action2<ENV>(x, st, n, q, j, row, valid, t, sel.action);
            tc_fence_before();
            all_done = __syncthreads_and(!valid || j != 0 || st.done);
            // first-round operands of the next decode step: this instance's, or the next instance's of this CTA
            {
                const bool same = !all_done && (t + 1 < T);
                const int64_t nb = same ? b : b + gridDim.x;
                if (nb < c.B) {
                    pend = nb;
                    if (issuer_warp) {
                        if (elect_one()) {
                            const unsigned char* ib = img + (size_t)nb * img_stride;
                            load_slot(x, ib, j, 0, j, 0);
                            load_slot(x, ib, j, 1, j, 1);
                        }
                    }
                }
            }
            record_action2<ENV>(x, a, st, j, valid, grow, t, T, sel.action, sel.logp);
        }
        // an instance that ends without consuming the pairs requested for it (done after its forced action, or max_steps
        // hit before the first decode step): wait for them to land, then request the next instance's
        if (pend == b) {
            const int64_t nb = b + gridDim.x;
            if (issuer_warp) {
                if (elect_one()) {
                    wait_slot(x, j, 0);
                    wait_slot(x, j, 1);
                    if (nb < c.B) {
                        const unsigned char* ib = img + (size_t)nb * img_stride;
                        load_slot(x, ib, j, 0, j, 0);
                        load_slot(x, ib, j, 1, j, 1);
                    }
                }
            }
            pend = nb < c.B ? nb : -1;
        }
        if (valid && j == 1) {
            // closing edge (roll(-1) wraps to the first element: TSP first action, CVRP depot)
            const int to = (ENV == RL4CO_ENV_CVRP) ? 0 : st.first_action;
            const float2 p0 = s->locs[st.prev], p1 = s->locs[to];
            const float dx = p1.x - p0.x, dy = p1.y - p0.y;
            st.len += sqrtf(dx * dx + dy * dy);
            a.reward[grow] = -st.len;
            a.log_likelihood[grow] = st.ll;
        }
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
    if (warp == 0) tmem_dealloc(x.tmem, kTmemCols2);
}

}  // namespace t2
