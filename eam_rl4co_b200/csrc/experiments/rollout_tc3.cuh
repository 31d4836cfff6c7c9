// EXPERIMENT, NOT BUILT (measured and rejected, round 2): streaming tensor-path rollout with 32 warps per SM.
// To build it: #include this file in rollout_tc.cu after rollout_tc2.cuh (inside namespace r4::tc) and launch
// t3::rollout_tc3_kernel<ENV, MODE><<<min(B, 2 * SMs), t3::kThreads3, t3::smem_bytes3(npad)>>> on a STREAM-layout image.
//
// rollout_tc2_kernel runs 16 warps per SM (2 CTAs x 8 warps at 128 registers): 62 % of the issue slots find an eligible
// warp, no pipe is saturated.  This variant doubles the warps: a CTA is 512 threads at 64 registers, two CTAs (instances)
// per SM.  Two heads are in flight per CTA as before (256 TMEM columns), but the columns (nodes) of a head's score row
// are split over TWO threads - thread = (row, head slot j, column half c) - and everything is cut to 8-column pieces so that
// the hot loops fit the halved register budget:
//   warp w: TMEM lane quadrant q = w & 3, role r = w >> 2: head slot j = r & 1, column half c = r >> 1;
//   team j = the 8 warps of head slot j (named barrier 5 + j, 256 threads);
//   selection / env step: the 4 threads of a row (roles 0..3) as in the resident kernel (named barrier 1 + q, 128 threads).
//
// Result on B200 (TSP100 POMO greedy, profiles/r02_tc3_experiment.md): parity-green on the streaming-kernel tests, but
// 9.89 ms at B = 1024 against 8.50 ms for rollout_tc2_kernel.  Issue-slot utilisation rises only from 62 % to 68 % (the
// warps of a team wait on the same MMA commits and barriers, so more of them are eligible or stalled TOGETHER), while the
// halved register budget and the finer column split cost 35 % more executed instructions (bit tests instead of predicate
// registers, 4-column TMEM stores, twice the pollers on every mbarrier, 4-way instead of 2-way selection exchanges).
#pragma once

namespace t3 {

using t2::kSBO_SL;
using t2::pair_bytes;
using t2::slice_bytes;

constexpr int kThreads3 = 512;
constexpr int kRounds = 4;
constexpr uint32_t kTmemCols3 = 256;
constexpr uint32_t kColO3 = 224;     // PV accumulators: slot j at 224 + 16 j   (S / P of slot j at j * NPAD)
constexpr uint32_t kColHA3 = 0;      // end of step: heads A operand hi [0,64) lo [64,128)
constexpr uint32_t kColLG3 = 128;    // logits accumulator [128, 128 + NPAD)

struct Smem3 {
    uint32_t maskw[4][kRows];
    int cur[kRows];
    int first[kRows];
    float rem[kRows];
    float red_a[4][kRows];      // exchanges between the 4 threads of a row (selection); red_a doubles as the row-max exchange
    float red_b[4][kRows];
    float red_c[4][kRows];
    float red_lp[4][kRows];
    int red_i[4][kRows];
    float red_l[2][2][kRows];   // [head slot][column half]: partial row sums of the softmax
    float red_m[2][2][kRows];   // ... and partial row maxima (exact fallback)
    int newvis[kRows];
    float2 locs[kRows];
    float demand[kRows];
    uint64_t kv_bar[2][2];      // [slot][0: K, 1: V]: the ring slot's bulk copy has landed
    uint64_t l_bar[4];          // L pairs 4..7 landing in the stash area
    uint64_t mma_bar;           // logits MMAs
    uint64_t team_bar[2];       // scores / PV MMAs of one head slot
    uint32_t tmem_ptr;
    uint32_t n_lload;
    uint32_t n_load[2][2];
    float lnf[4][kRows];        // ln(#feasible) at the previous step, one private copy per thread of the row
    float lnf_depot[4][kRows];
    float lse_depot[4][2][kRows];   // [round][head slot]: both column halves write the same value
};

// shared memory: [Smem3][Q tile slot 0][Q tile slot 1][stash: rounds 0, 1 x 2 slots x 8 KB][ring: 4 slots of pair_bytes];
// round 2's heads wait in the slot's Q tile (idle once the scores of round 3 are done), round 3's in registers
constexpr uint32_t kOffQ3 = (uint32_t)((sizeof(Smem3) + 127) / 128 * 128);
constexpr uint32_t kOffStash3 = kOffQ3 + 2u * kQTeamBytes;
constexpr uint32_t kStashBytes3 = 2u * 2u * kQTeamBytes;
constexpr uint32_t kOffRing3 = kOffStash3 + kStashBytes3;

struct Ctx3 {
    unsigned char* smem;
    uint32_t tmem;
    int npad;
    __device__ __forceinline__ Smem3* s() const { return reinterpret_cast<Smem3*>(smem); }
    __device__ __forceinline__ unsigned char* q_p(int slot) const { return smem + kOffQ3 + (uint32_t)slot * kQTeamBytes; }
    __device__ __forceinline__ uint32_t base() const { return smem_u32(smem); }
    __device__ __forceinline__ uint32_t q_hi(int slot) const { return base() + kOffQ3 + (uint32_t)slot * kQTeamBytes; }
    __device__ __forceinline__ uint32_t q_lo(int slot) const { return q_hi(slot) + kQHalfBytes; }
    // stash of round g (0, 1: stash area; 2: the slot's Q tile), 8 KB = 256 threads x 8 words
    __device__ __forceinline__ float* stash(int g, int slot) const {
        return reinterpret_cast<float*>(g < 2 ? smem + kOffStash3 + (uint32_t)(2 * g + slot) * kQTeamBytes : q_p(slot));
    }
    __device__ __forceinline__ unsigned char* slot_p(int slot, int kv) const {
        return smem + kOffRing3 + (uint32_t)(2 * slot + kv) * pair_bytes(npad);
    }
    __device__ __forceinline__ uint32_t slot(int slot_, int kv) const {
        return base() + kOffRing3 + (uint32_t)(2 * slot_ + kv) * pair_bytes(npad);
    }
    __device__ __forceinline__ unsigned char* lst_p(int i) const { return smem + kOffStash3 + (uint32_t)i * pair_bytes(npad); }
    __device__ __forceinline__ uint32_t lst(int i) const { return base() + kOffStash3 + (uint32_t)i * pair_bytes(npad); }
    __device__ __forceinline__ int nch() const { return npad >> 4; }
    __device__ __forceinline__ uint32_t sbo_v() const { return (uint32_t)npad * 16u; }
};

__host__ inline size_t smem_bytes3(int npad) { return (size_t)kOffRing3 + 4ull * pair_bytes(npad); }

__device__ __forceinline__ void team_sync3(int slot) { bar_sync(kTeamBar + slot, 256); }
__device__ __forceinline__ bool team_all3(int slot, bool pred) {
    uint32_t r;
    asm volatile(
        "{\n\t.reg .pred p, r;\n\tsetp.ne.u32 p, %1, 0;\n\tbar.red.and.pred r, %2, 256, p;\n\t"
        "selp.u32 %0, 1, 0, r;\n\t}"
        : "=r"(r)
        : "r"((uint32_t)pred), "r"(kTeamBar + slot)
        : "memory");
    return r != 0u;
}

// Q tile of head h by the 8 warps of the head slot: warp (q, c) covers rows [32q, 32q+32) in passes 2c, 2c+1 of 8 rows;
// four lanes read the 64 contiguous bytes of a row (context.py:118-140 / :70-74,155-157).  Rows >= S stay zero.
template <int ENV>
__device__ __forceinline__ void build_q3(const Ctx3& x, const rl4co_cache& c, int64_t b, int S, int h, int slot, int q,
                                         int ch, int lane) {
    const int64_t base = b * (int64_t)c.n_nodes;
    const int pc = lane & 3;
    const int choff = 16 * h + 4 * pc;
    float4 qa[2], qb[2];
    float rem[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int r = 32 * q + 8 * (2 * ch + k) + (lane >> 2);
        qa[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        qb[k] = qa[k];
        rem[k] = 0.f;
        if (r < S) {
            qb[k] = ld4(c.pb + (base + x.s()->cur[r]) * kE + choff);
            if (ENV == RL4CO_ENV_TSP) qa[k] = ld4(c.pa + (base + x.s()->first[r]) * kE + choff);
            else rem[k] = x.s()->rem[r];
        }
    }
    float4 wc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ENV == RL4CO_ENV_CVRP) wc = ld4(c.wcap + choff);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int r = 32 * q + 8 * (2 * ch + k) + (lane >> 2);
        float4 v;
        if (ENV == RL4CO_ENV_TSP)
            v = make_float4(qa[k].x + qb[k].x, qa[k].y + qb[k].y, qa[k].z + qb[k].z, qa[k].w + qb[k].w);
        else
            v = make_float4(fmaf(rem[k], wc.x, qb[k].x), fmaf(rem[k], wc.y, qb[k].y), fmaf(rem[k], wc.z, qb[k].z),
                            fmaf(rem[k], wc.w, qb[k].w));
        uint32_t h0, l0, h1, l1;
        split2(v.x, v.y, h0, l0);
        split2(v.z, v.w, h1, l1);
        unsigned char* t = x.q_p(slot) + (uint32_t)(r >> 3) * kSBO_Q + (uint32_t)(pc >> 1) * kLBO_Q +
                           (uint32_t)(r & 7) * 16u + (uint32_t)(pc & 1) * 8u;
        *reinterpret_cast<uint2*>(t) = make_uint2(h0, h1);
        *reinterpret_cast<uint2*>(t + kQHalfBytes) = make_uint2(l0, l1);
    }
    fence_proxy_async();
}

// This thread's share of one (row, head) score row: the 16-node chunks [k0, k1).
//   pass_max : returns the masked maximum (exact fallback, first pass);
//   otherwise: P = exp(S - ref) on feasible nodes (0 elsewhere, padding included) written in place as fp16 hi/lo - chunk k:
//              hi words in columns [16k, 16k+8), lo words in [16k+8, 16k+16) - and returns the partial row sum.
// A chunk's 16 scores are loaded before anything is stored (the lo words of its first half land on the second half's
// scores); the two 8-node halves are then processed one after the other to keep the live registers at ~30.
__device__ __forceinline__ float softmax_part3(uint32_t sp, const uint32_t (&w)[4], int k0, int k1, int n, float refl2,
                                               bool pass_max) {
    const float kLog2e = 1.4426950408889634f;
    float l0 = 0.f, l1 = 0.f, mx = -INFINITY;
#pragma unroll 1
    for (int k = k0; k < k1; ++k) {
        const int cb = 16 * k;
        const uint32_t word = (k >> 1) == 0 ? w[0] : ((k >> 1) == 1 ? w[1] : ((k >> 1) == 2 ? w[2] : w[3]));
        uint32_t bits = (word >> (cb & 31)) & 0xffffu;
        if (cb + 16 > n) bits &= (cb < n) ? ((1u << (n - cb)) - 1u) : 0u;
        uint32_t v[16];
        tmem_ld16(sp + (uint32_t)cb, v);
        tc_wait_ld();
        if (pass_max) {
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if (bits & (1u << i)) mx = fmaxf(mx, __uint_as_float(v[i]));
            continue;
        }
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = 8 * hf + 2 * i;
                const float s0 = (bits & (1u << e)) ? __uint_as_float(v[e]) : -INFINITY;
                const float s1 = (bits & (1u << (e + 1))) ? __uint_as_float(v[e + 1]) : -INFINITY;
                const float p0 = ex2_approx(fmaf(s0, kLog2e, -refl2));
                const float p1 = ex2_approx(fmaf(s1, kLog2e, -refl2));
                l0 += p0;
                l1 += p1;
                split2(p0, p1, hi[i], lo[i]);
            }
            tmem_st4(sp + (uint32_t)(cb + 4 * hf), hi);
            tmem_st4(sp + (uint32_t)(cb + 8 + 4 * hf), lo);
        }
    }
    return pass_max ? mx : l0 + l1;
}

// ---- all GEMM phases of one decoder step (see decode_step_gemms2); thread = (row, head slot j, column half ch) ------
template <int ENV>
__device__ __forceinline__ void decode_step_gemms3(const Ctx3& x, const rl4co_cache& c, const unsigned char* img_b,
                                                   int64_t b, int S, uint32_t& mma_par, uint32_t& team_par,
                                                   float (&lse)[4], bool have_ref) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, r = warp >> 2, j = r & 1, ch = r >> 1, row = 32 * q + lane;
    const uint32_t lane_base = x.tmem + ((uint32_t)(32 * q) << 16);
    const int n = c.n_nodes;
    const bool issuer_warp = (ch == 0) && (q == j);
    Smem3* s = x.s();
    uint64_t* tbar = &s->team_bar[j];
    const float kLog2e = 1.4426950408889634f;
    const int tid_team = 128 * ch + row;
    const int nch = x.nch(), k0 = ch ? (nch >> 1) : 0, k1 = ch ? nch : (nch >> 1);
    const uint32_t w[4] = {s->maskw[0][row], s->maskw[1][row], s->maskw[2][row], s->maskw[3][row]};
    const bool at_depot = (ENV == RL4CO_ENV_CVRP) && s->cur[row] == 0;
    float hv[8];
    build_q3<ENV>(x, c, b, S, j, j, q, ch, lane);
    team_sync3(j);
    if (issuer_warp) {
        if (elect_one()) {
            t2::wait_slot(x, j, 0);
            t2::issue_scores2(x, j);
        }
    }
    {
        const int nf = __popc(w[0]) + __popc(w[1]) + __popc(w[2]) + __popc(w[3]);
        const float lnf = __logf((float)(nf > 0 ? nf : 1));
        const float dl = lnf - s->lnf[r][row];
        float shift = have_ref ? (dl > 0.f ? 1.0f : 0.5f) * dl : 0.f;
        if (ENV == RL4CO_ENV_CVRP && have_ref && at_depot && s->lnf_depot[r][row] >= 0.f) {
            shift = lnf - s->lnf_depot[r][row];
#pragma unroll
            for (int g = 0; g < kRounds; ++g) lse[g] = s->lse_depot[g][j][row];
        }
        s->lnf[r][row] = lnf;
#pragma unroll
        for (int g = 0; g < kRounds; ++g) lse[g] += shift;
    }
#pragma unroll 1
    for (int g = 0; g < kRounds; ++g) {
        const int h = 2 * g + j;
        wait_mma(tbar, team_par);                      // scores of round g
        const uint32_t sp = lane_base + (uint32_t)(j * x.npad);
        float l;
        bool spec = have_ref;
        const float lse_g = g == 0 ? lse[0] : (g == 1 ? lse[1] : (g == 2 ? lse[2] : lse[3]));
        float ref = lse_g - kSpecTargetLn;
        for (;;) {
            if (!spec) {   // exact reference: the row maximum (first pass over the scores), aimed at a row sum >= e^3
                s->red_m[j][ch][row] = softmax_part3(sp, w, k0, k1, n, 0.f, true);
                team_sync3(j);
                const float m = fmaxf(s->red_m[j][0][row], s->red_m[j][1][row]);
                ref = (m == -INFINITY ? 0.f : m) - 3.0f;
            }
            s->red_l[j][ch][row] = softmax_part3(sp, w, k0, k1, n, ref * kLog2e, false);
            tc_wait_st();
            tc_fence_before();
            team_sync3(j);                             // P of the head is complete, the partial sums are visible
            l = s->red_l[j][0][row] + s->red_l[j][1][row];
            const bool ok = !spec || (l == 0.f) || (l >= kSpecLMin && l < kSpecLMax);
            if (team_all3(j, ok)) break;
            spec = false;                              // rare: scores again (K slot and Q tile are intact), exact form
            if (issuer_warp) { if (elect_one()) t2::issue_scores2(x, j); }
            wait_mma(tbar, team_par);
        }
        const float inv_l = l > 0.f ? 1.0f / l : 0.f;
        const float lse_new = l > 0.f ? ref + __logf(l) : 0.f;
        if (g == 0) lse[0] = lse_new;
        else if (g == 1) lse[1] = lse_new;
        else if (g == 2) lse[2] = lse_new;
        else lse[3] = lse_new;
        if (ENV == RL4CO_ENV_CVRP && at_depot) {
            s->lse_depot[g][j][row] = lse_new;
            if (g == kRounds - 1) s->lnf_depot[r][row] = s->lnf[r][row];
        }
        if (g == kRounds - 1) {                        // round 2's heads -> the slot's Q tile (no scores read it any more)
            float* st = x.stash(2, j) + tid_team;
#pragma unroll
            for (int i = 0; i < 8; ++i) st[i * 256] = hv[i];
        }
        if (issuer_warp) {
            if (elect_one()) {
                t2::wait_slot(x, j, 1);                // V^T pair of head h
                t2::issue_pv2(x, j);
                if (g < kRounds - 1) t2::load_slot(x, img_b, h + 2, 0, j, 0);
                else t2::load_slot(x, img_b, j, 2, j, 0);
            }
        }
        if (g < kRounds - 1) build_q3<ENV>(x, c, b, S, h + 2, j, q, ch, lane);   // next round's Q while PV runs
        wait_mma(tbar, team_par);                      // PV of round g
        if (issuer_warp) {
            if (elect_one()) {
                if (g < kRounds - 1) t2::load_slot(x, img_b, h + 2, 1, j, 1);
                else t2::load_slot(x, img_b, 2 + j, 2, j, 1);
            }
        }
        if (g < kRounds - 1) {                         // next scores run while O(g) is drained
            team_sync3(j);
            if (issuer_warp) {
                if (elect_one()) {
                    t2::wait_slot(x, j, 0);
                    t2::issue_scores2(x, j);
                }
            }
        }
        uint32_t o[8];
        tmem_ld8(lane_base + kColO3 + 16u * j + 8u * ch, o);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 8; ++i) hv[i] = __uint_as_float(o[i]) * inv_l;
        tc_fence_before();
        if (g < 2) {
            float* st = x.stash(g, j) + tid_team;
#pragma unroll
            for (int i = 0; i < 8; ++i) st[i * 256] = hv[i];
        }
    }
    __syncthreads();
    tc_fence_after();
#pragma unroll 1
    for (int g = kRounds - 1; g >= 0; --g) {
        const int h = 2 * g + j;
        if (g < kRounds - 1) {
            const float* st = x.stash(g, j) + tid_team;
#pragma unroll
            for (int i = 0; i < 8; ++i) hv[i] = st[i * 256];
        }
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split2(hv[2 * i], hv[2 * i + 1], hi[i], lo[i]);
        tmem_st4(lane_base + kColHA3 + 8u * h + 4u * ch, hi);
        tmem_st4(lane_base + kColHA3 + 64u + 8u * h + 4u * ch, lo);
    }
    tc_wait_st();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { if (elect_one()) t2::issue_logits2(x, img_b); }
    wait_mma(&s->mma_bar, mma_par);
}

// ---- selection: 4 threads per row (role r takes the r-th quarter of the logit columns), 16-column segments ---------------
__device__ __forceinline__ uint32_t load_seg16(uint32_t lg, int sh, int cnt, const uint32_t (&w)[4], bool clipped,
                                               float (&tv)[16]) {
    {
        uint32_t a16[16];
        tmem_ld16(lg + (uint32_t)sh, a16);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) tv[i] = __uint_as_float(a16[i]);
    }
    const int wi = sh >> 5;
    const uint32_t lo = wi == 0 ? w[0] : (wi == 1 ? w[1] : (wi == 2 ? w[2] : w[3]));
    const uint32_t hi = wi == 0 ? w[1] : (wi == 1 ? w[2] : (wi == 2 ? w[3] : 0u));
    const uint32_t inrange = (1u << cnt) - 1u;                         // cnt <= 16
    const uint32_t win = __funnelshift_r(lo, hi, sh & 31) & inrange;   // feasible bits of nodes sh .. sh+15
    if (clipped) {                                                     // decoding.py:171-172
        float am0 = 0.f, am1 = 0.f;
#pragma unroll
        for (int e = 0; e < 16; e += 2) {
            am0 = fmaxf(am0, fabsf(tv[e]));
            am1 = fmaxf(am1, fabsf(tv[e + 1]));
        }
        if (__all_sync(kFull, fmaxf(am0, am1) <= 1.0f)) {
#pragma unroll
            for (int e = 0; e < 16; ++e) tv[e] = tanh_small(tv[e]);
        } else {
#pragma unroll
            for (int e = 0; e < 16; ++e) tv[e] = tanh_fast(tv[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < 16; ++e) tv[e] = ((win >> e) & 1u) ? tv[e] : -INFINITY;   // :177
    return win;
}

template <int MODE>
__device__ __forceinline__ RowSel select_row3(const Ctx3& x, const rl4co_rollout_args& a, int n, int q, int r, int row,
                                              bool valid, int64_t grow, int t, int T, uint32_t nkey, int given,
                                              int& err_bits) {
    Smem3* s = x.s();
    const uint32_t lg = x.tmem + ((uint32_t)(32 * q) << 16) + kColLG3;
    const int U = (n + 3) >> 2;
    const int ua = (U * r) >> 2, ub = (U * (r + 1)) >> 2;           // <= 7 units of 4 columns
    const int um = (ub - ua > 4) ? ua + 4 : ub;                      // segments [ua, um) (<= 16 columns) and [um, ub)
    const uint32_t w[4] = {s->maskw[0][row], s->maskw[1][row], s->maskw[2][row], s->maskw[3][row]};
    const bool clipped = a.tanh_clipping > 0.f;
    const float sc = (clipped ? a.tanh_clipping : 1.0f) / a.temperature;
    const float kL2e = 1.4426950408889634f;
    constexpr bool sampling = MODE == RL4CO_MODE_SAMPLING;
    float tv[16];
    float ref = sc;                                                 // constant reference: see select_row
    if (!(clipped && sc <= 40.0f)) {
        float lmax = -INFINITY;
#pragma unroll 1
        for (int sg = 0; sg < 2; ++sg) {
            const int u0 = sg ? um : ua, u1 = sg ? ub : um;
            if (u1 <= u0) continue;
            load_seg16(lg, 4 * u0, 4 * (u1 - u0), w, clipped, tv);
#pragma unroll
            for (int e = 0; e < 16; ++e) lmax = fmaxf(lmax, tv[e]);
        }
        s->red_a[r][row] = lmax;
        bar_sync(1 + q, 128);
        const float m = fmaxf(fmaxf(s->red_a[0][row], s->red_a[1][row]), fmaxf(s->red_a[2][row], s->red_a[3][row]));
        ref = (m == -INFINITY) ? 0.f : m * sc;
    }
    const float c1 = sc * kL2e, c0 = -ref * kL2e;
    float ls0 = 0.f, ls1 = 0.f;
    float best = sampling ? 0.f : -INFINITY, best_x = -INFINITY;
    int best_i = 0x7fffffff;
    const bool counter_noise = sampling && a.gumbel == nullptr;
#pragma unroll 1
    for (int sg = 0; sg < 2; ++sg) {
        const int u0 = sg ? um : ua, u1 = sg ? ub : um;
        if (u1 <= u0) continue;
        const int sh = 4 * u0, cnt = 4 * (u1 - u0);
        load_seg16(lg, sh, cnt, w, clipped, tv);
#pragma unroll
        for (int e = 0; e < 16; e += 2) {
            ls0 += ex2_approx(fmaf(tv[e], c1, c0));
            ls1 += ex2_approx(fmaf(tv[e + 1], c1, c0));
        }
        if (MODE == RL4CO_MODE_EVALUATE) {
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                if (e < cnt && sh + e == given) {
                    best = 0.f;
                    best_i = given;
                    best_x = tv[e];
                }
            }
        } else if (!sampling) {
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                if (tv[e] > best) {   // ascending node order + strict > = first max index (torch.argmax)
                    best = tv[e];
                    best_i = sh + e;
                }
            }
            best_x = best;
        } else if (counter_noise) {   // exponential race, see select_row
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                const float u = noise_u(nkey, sh + e);
                const float d = 1.0f - u;
                const float ser = d * fmaf(d, fmaf(d, fmaf(d, 0.36067376f, 0.48089835f), 0.72134752f), 1.44269504f);
                float l2;
                asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(u));
                const float big_e = u > 0.98f ? ser : -l2;
                float rc;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(big_e));
                const float cand = ex2_approx(fmaf(tv[e], c1, c0)) * rc;
                if (cand > best) {
                    best = cand;
                    best_i = sh + e;
                    best_x = tv[e];
                }
            }
        }
    }
    s->red_b[r][row] = ls0 + ls1;
    if (!(sampling && !counter_noise)) {
        s->red_c[r][row] = best;
        s->red_i[r][row] = best_i;
        s->red_lp[r][row] = best_x;
    }
    bar_sync(1 + q, 128);
    const float tot = (s->red_b[0][row] + s->red_b[1][row]) + (s->red_b[2][row] + s->red_b[3][row]);
    if ((tot != tot || !(tot > 0.f)) && valid) err_bits |= RL4CO_ERR_NAN_LOGITS;   // attention.py:303-304
    const float lse = logf(tot);                                    // log_softmax(x)_e = (x_e - ref) - lse  :190
    if (a.logp_all != nullptr || (sampling && !counter_noise)) {    // second pass over the logits (still in TMEM)
        best = -INFINITY;
#pragma unroll 1
        for (int sg = 0; sg < 2; ++sg) {
            const int u0 = sg ? um : ua, u1 = sg ? ub : um;
            if (u1 <= u0) continue;
            const int sh = 4 * u0, cnt = 4 * (u1 - u0);
            const uint32_t win = load_seg16(lg, sh, cnt, w, clipped, tv);
            if (a.logp_all != nullptr && valid) {
#pragma unroll
                for (int e = 0; e < 16; ++e)
                    if (e < cnt && sh + e < n) a.logp_all[(grow * T + t) * (int64_t)n + sh + e] = (tv[e] * sc - ref) - lse;
            }
            if (sampling && !counter_noise) {                       // Gumbel-max over logp + g with supplied noise
#pragma unroll
                for (int e = 0; e < 16; ++e) {
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
            s->red_c[r][row] = best;
            s->red_i[r][row] = best_i;
            s->red_lp[r][row] = best_x;
            bar_sync(1 + q, 128);
        }
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
    if (valid && r == 0) {
        const int wa = sel.action >> 5;
        const uint32_t ww = wa == 0 ? w[0] : (wa == 1 ? w[1] : (wa == 2 ? w[2] : w[3]));
        if (MODE != RL4CO_MODE_EVALUATE && !((ww >> (sel.action & 31)) & 1u)) err_bits |= RL4CO_ERR_INFEASIBLE;
    }
    if (valid && r == 1 && !(sel.logp > -1000.f)) err_bits |= RL4CO_ERR_NEG_INF_LOGP;   // decoding.py:56-58
    return sel;
}

__device__ __forceinline__ void cta_setup3(Ctx3& x) {
    const int warp = threadIdx.x >> 5;
    Smem3* s = x.s();
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
    if (warp == 0) tmem_alloc(&s->tmem_ptr, kTmemCols3);
    for (uint32_t i = threadIdx.x; i < 2u * kQTeamBytes / 16u; i += kThreads3)
        reinterpret_cast<uint4*>(x.q_p(0))[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < kRows; i += kThreads3) {
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

// ---- persistent fused rollout, streaming tensor path, 2 x 16 warps per SM -------------------------------------------------
template <int ENV, int MODE>
__global__ void __launch_bounds__(kThreads3, 2)
    rollout_tc3_kernel(rl4co_cache c, rl4co_rollout_args a, const unsigned char* __restrict__ img, int npad) {
    extern __shared__ __align__(128) unsigned char smem[];
    Ctx3 x;
    x.smem = smem;
    x.npad = npad;
    x.tmem = 0;
    cta_setup3(x);
    Smem3* s = x.s();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, r = warp >> 2, row = 32 * q + lane;
    const int n = c.n_nodes, S = a.num_starts, T = a.max_steps;
    const bool valid = row < S;
    const bool issuer_warp = (r < 2) && (q == r);     // head slot j = r, column half 0
    const int ncust = (ENV == RL4CO_ENV_TSP) ? n : n - 1;
    const int smod = (a.start_mod > 0 && a.start_mod <= ncust) ? a.start_mod : ncust;   // ops.py:147-161
    const size_t img_stride = 48ull * slice_bytes(npad);
    uint32_t mma_par = 0, team_par = 0;

    int64_t pend = -1;   // see rollout_tc2_kernel
    if ((int64_t)blockIdx.x < c.B) {
        pend = blockIdx.x;
        if (issuer_warp) {
            if (elect_one()) {
                const unsigned char* ib = img + (size_t)blockIdx.x * img_stride;
                t2::load_slot(x, ib, r, 0, r, 0);
                t2::load_slot(x, ib, r, 1, r, 1);
            }
        }
    }
    for (int64_t b = blockIdx.x; b < c.B; b += gridDim.x) {
        const unsigned char* img_b = img + (size_t)b * img_stride;
        for (int i = threadIdx.x; i < n; i += kThreads3) {
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
            const int rem = n - 32 * r;
            st.word = !valid ? 0u : (rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u)));
        } else {
            st.word = 0u;            // cvrp/env.py:102-130: nothing visited
        }
        int err_bits = 0;
        s->lnf_depot[r][row] = -1.f;
        int act = 0;                 // forced first action (decoding.py:305-327; select_start_nodes ops.py:147-161)
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
        apply_action<ENV>(x, a, st, n, q, r, row, valid, 0, act);
        int all_done = __syncthreads_and(!valid || r != 0 || st.done);
        record_action<ENV>(x, a, st, r, valid, grow, 0, T, act, 0.f);

        float lse[4] = {0.f, 0.f, 0.f, 0.f};   // log-sum-exp of the (row, head 2g + j) score rows at the previous step
        for (int t = 1; !all_done; ++t) {
            if (t >= T) {
                if (valid && r == 0 && !st.done) err_bits |= RL4CO_ERR_MAX_STEPS;
                break;
            }
            decode_step_gemms3<ENV>(x, c, img_b, b, S, mma_par, team_par, lse, t > 1);
            pend = -1;
            int given = 0;
            if (MODE == RL4CO_MODE_EVALUATE && valid) {
                given = (t < a.actions_in_T) ? (int)a.actions_in[grow * a.actions_in_stride + t] : 0;
                if (given < 0 || given >= n) {
                    if (r == 0) err_bits |= RL4CO_ERR_INFEASIBLE;
                    given = 0;
                }
            }
            const uint32_t nkey = (MODE == RL4CO_MODE_SAMPLING) ? noise_key(a.seed, a.inst_offset + b, row, t) : 0u;
            const RowSel sel = select_row3<MODE>(x, a, n, q, r, row, valid, grow, t, T, nkey, given, err_bits);
            apply_action<ENV>(x, a, st, n, q, r, row, valid, t, sel.action);
            tc_fence_before();
            all_done = __syncthreads_and(!valid || r != 0 || st.done);
            {   // first-round operands of the next decode step: this instance's, or the next instance's of this CTA
                const bool same = !all_done && (t + 1 < T);
                const int64_t nb = same ? b : b + gridDim.x;
                if (nb < c.B) {
                    pend = nb;
                    if (issuer_warp) {
                        if (elect_one()) {
                            const unsigned char* ib = img + (size_t)nb * img_stride;
                            t2::load_slot(x, ib, r, 0, r, 0);
                            t2::load_slot(x, ib, r, 1, r, 1);
                        }
                    }
                }
            }
            record_action<ENV>(x, a, st, r, valid, grow, t, T, sel.action, sel.logp);
        }
        if (pend == b) {   // ended without consuming the requested pairs: see rollout_tc2_kernel
            const int64_t nb = b + gridDim.x;
            if (issuer_warp) {
                if (elect_one()) {
                    t2::wait_slot(x, r, 0);
                    t2::wait_slot(x, r, 1);
                    if (nb < c.B) {
                        const unsigned char* ib = img + (size_t)nb * img_stride;
                        t2::load_slot(x, ib, r, 0, r, 0);
                        t2::load_slot(x, ib, r, 1, r, 1);
                    }
                }
            }
            pend = nb < c.B ? nb : -1;
        }
        if (valid && r == 2) {
            // closing edge (roll(-1) wraps to the first element: TSP first action, CVRP depot)
            const int to = (ENV == RL4CO_ENV_CVRP) ? 0 : st.first_action;
            const float2 p0 = s->locs[st.prev], p1 = s->locs[to];
            const float dx = p1.x - p0.x, dy = p1.y - p0.y;
            st.len += sqrtf(dx * dx + dy * dy);
            a.reward[grow] = -st.len;
        }
        if (valid && r == 1) a.log_likelihood[grow] = st.ll;
        if (valid && r == 0) {
            const int len = st.done ? st.length : T;
            a.lengths[grow] = len;
            atomicMax(a.max_len, len);
        }
        if (valid && err_bits) raise_err(a.err, err_bits, grow);
        __syncthreads();  // locs / state of this instance are dead before the next one overwrites them
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(x.tmem, kTmemCols3);
}

}  // namespace t3
