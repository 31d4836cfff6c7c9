/*
 * rl4co_b200 — C ABI of the B200-native construction-rollout path.
 *
 * One shared library (librl4co_b200.so), `extern "C"`, plain device pointers and sizes,
 * no C++/torch types.  Every function launches on the CUDA stream passed as `stream`
 * (a cudaStream_t cast to void*), never allocates or frees, and returns 0 on success or a
 * negative RL4CO_E_* code for argument / launch errors (text via rl4co_last_error()).
 * Semantic failures the reference raises as AssertionError (NaN logits, infeasible action,
 * -inf log-prob, invalid tour, capacity exceeded) are reported through a caller-owned
 * device `int32 err[2]` word: err[0] = OR of RL4CO_ERR_* bits, err[1] = first offending row+1.
 *
 * All citations are file:line in the reference (Tarseus/eam-rl4co, RL4CO v0.6 fork).
 * Row order at this boundary is the reference's start-major order: row = s*B + b
 * (rl4co/utils/ops.py:13-16, rl4co/models/zoo/am/decoder.py:195-197).
 * dtypes follow the reference's TensorDict: actions/current_node/first_node/i int64,
 * action_mask/done bool (1 byte), visited uint8, everything else float32.
 */
#ifndef RL4CO_B200_H
#define RL4CO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RL4CO_B200_ABI_VERSION 3

/* return codes */
#define RL4CO_OK 0
#define RL4CO_E_ARG (-1)      /* bad size / null pointer / unsupported shape (e.g. N' > limit) */
#define RL4CO_E_LAUNCH (-2)   /* CUDA launch or runtime error */
#define RL4CO_E_UNSUPPORTED (-3)

/* err[0] bits — reference assertion they stand for */
#define RL4CO_ERR_NAN_LOGITS 1      /* "Logits contain NaNs"            attention.py:303-304 */
#define RL4CO_ERR_INFEASIBLE 2      /* "infeasible action selected"     decoding.py:397-399,413-415 */
#define RL4CO_ERR_NEG_INF_LOGP 4    /* "Logprobs should not be -inf"    decoding.py:56-58 */
#define RL4CO_ERR_INVALID_TOUR 8    /* "Invalid tour"                   tsp/env.py:163-168, cvrp/env.py:165-170 */
#define RL4CO_ERR_CAPACITY 16       /* "Used more than capacity"        cvrp/env.py:176-185 */
#define RL4CO_ERR_MAX_STEPS 32      /* max_steps exceeded               constructive/base.py:246-250 */
#define RL4CO_ERR_FP16_RANGE 64     /* a cache table value is outside the fp16 range of the tensor path */

/* env ids, decode modes */
#define RL4CO_ENV_TSP 0
#define RL4CO_ENV_CVRP 1
#define RL4CO_MODE_GREEDY 0         /* decoding.py:391-401 */
#define RL4CO_MODE_SAMPLING 1       /* decoding.py:403-417 with supplied / counter noise */
#define RL4CO_MODE_EVALUATE 2       /* decoding.py:452-465 */

/* rollout kernel selection */
#define RL4CO_IMPL_AUTO 0           /* tensor path when num_starts >= 8 and cache->tc_img is set */
#define RL4CO_IMPL_FP32 1           /* fp32 CUDA-core kernels                                     */
#define RL4CO_IMPL_TC 2             /* tcgen05 kernels; error if the shape is unsupported         */
#define RL4CO_IMPL_STREAM 3         /* single-start rows: warp per row, K/V/L streamed from global memory / L2 every
                                       step (AUTO takes it when the batch gives >= 8 rows per SM)  */
#define RL4CO_IMPL_COOP 4           /* single-start rows: CTA per instance, K/V/L resident in shared memory, the warps
                                       of the CTA share one row's step (AUTO takes it for small batches)            */

int rl4co_b200_abi_version(void);
const char* rl4co_last_error(void);

/* ---- TSPEnv._step — rl4co/envs/routing/tsp/env.py:62-88 -----------------------------
 * first_out = action if any(i_in == 0) else first_in   (batch-global test, :65)
 * mask_out = mask_in with bit `action` cleared (:68-70); done = no bit left (:73);
 * i_out = i_in + 1; reward_out = 0 (bool, :76).  cur_out = action.                     */
int rl4co_tsp_step(const int64_t* action, const uint8_t* mask_in, const int64_t* i_in,
                   const int64_t* first_in, uint8_t* mask_out, int64_t* cur_out,
                   int64_t* first_out, int64_t* i_out, uint8_t* done_out, uint8_t* reward_out,
                   int32_t* scratch_flag /* device int32, clobbered */,
                   int64_t rows, int32_t n_nodes, void* stream);

/* ---- CVRPEnv._step + get_action_mask — rl4co/envs/routing/cvrp/env.py:68-100,132-144 --
 * demand is [demand_rows, N] (customers only); row r uses demand row r % demand_rows so a
 * multistart caller need not batchify it.  n_nodes = N+1 (depot = node 0).               */
int rl4co_cvrp_step_mask(const int64_t* action, const float* demand, int64_t demand_rows,
                         const float* used_in, const float* vehicle_capacity,
                         const uint8_t* visited_in, float* used_out, uint8_t* visited_out,
                         int64_t* cur_out, uint8_t* done_out, uint8_t* reward_out,
                         uint8_t* mask_out, int64_t rows, int32_t n_nodes, void* stream);

/* ---- CVRPEnv.get_action_mask alone — cvrp/env.py:132-144 (used by _reset :129) -------- */
int rl4co_cvrp_action_mask(const float* demand, int64_t demand_rows, const float* used,
                           const float* vehicle_capacity, const uint8_t* visited,
                           const int64_t* current_node, uint8_t* mask_out, int64_t rows,
                           int32_t n_nodes, void* stream);

/* ---- teacher-forced state replay for the differentiable evaluate (constructive/base.py:221-229 with actions=) ------
 * actions [rows, T] start-major (row = s*B + b).  Applies _step / get_action_mask (tsp/env.py:62-88,
 * cvrp/env.py:68-100,132-144) T-1 times in one launch and writes the decoder's inputs before step t = 1..T-1 in
 * instance-major order: cur_out [B,S,T-1] i64, mask_out [B,S,T-1,n_nodes] bool, rem_out [B,S,T-1] f32
 * (vehicle_capacity - used_capacity, CVRP only; null for TSP).  demand [B, n_nodes-1], vehicle_capacity [B].     */
int rl4co_replay_states(int32_t env, const int64_t* actions, int64_t rows, int32_t T, int64_t B,
                        int32_t n_nodes, const float* demand, const float* vehicle_capacity,
                        int64_t* cur_out, uint8_t* mask_out, float* rem_out, void* stream);

/* ---- _get_reward (+ check_solution_validity) ------------------------------------------
 * tsp/env.py:152-168, cvrp/env.py:146-185, ops.py:59-95.
 * locs [loc_rows, n_nodes, 2] (CVRP: depot at node 0); row r uses locs row r % loc_rows.
 * reward[r] = -(closed tour length); CVRP prepends the depot (cvrp/env.py:148-154).
 * check != 0 sets RL4CO_ERR_INVALID_TOUR / RL4CO_ERR_CAPACITY in err (demand/capacity
 * then required for CVRP).                                                               */
int rl4co_tour_reward(int32_t env, const float* locs, int64_t loc_rows, const int64_t* actions,
                      int64_t actions_row_stride, int32_t T, const float* demand,
                      int64_t demand_rows, const float* vehicle_capacity, float* reward,
                      int32_t check, int32_t* err, int64_t rows, int32_t n_nodes, void* stream);

/* ---- process_logits + DecodingStrategy.step — decoding.py:140-190, 346-465 -------------
 * logits [rows,n] (read-only; the reference mutates a tanh temp, not the input),
 * mask bool [rows,n].  logp_all (nullable) receives log_softmax [rows,n].
 * mode GREEDY: first max index; SAMPLING: argmax(logp + gumbel) with gumbel [rows,n] given,
 * or generated from (seed, inst_offset + r % B, r / B, t, node) when gumbel == NULL;
 * EVALUATE: action_in.  Writes action_out, logp_sel.                                      */
int rl4co_select(const float* logits, const uint8_t* mask, float tanh_clipping,
                 float temperature, int32_t mode, const float* gumbel, uint32_t seed,
                 int64_t inst_offset, int64_t B, int32_t t, const int64_t* action_in,
                 float* logp_all, int64_t* action_out, float* logp_sel, int32_t* err,
                 int64_t rows, int32_t n_nodes, void* stream);

/* ---- Precomputed per-instance tables consumed by the decoder kernels --------------------
 * Built by the host with ordinary GEMMs once per rollout (am/decoder.py:206-235 stays in
 * torch).  All [B, n_nodes, E] fp32 contiguous, E = 128, H = 8:
 *   gk, gv : glimpse key / value                      (decoder.py:216-220)
 *   lk     : logit key folded with project_out:  L @ W_out     (attention.py:299,328)
 *   pa     : TSP  h @ W_ctx[:, :E]^T  (first-node half, context.py:118-140); CVRP NULL
 *   pb     : TSP  h @ W_ctx[:, E:]^T + graph_context ; CVRP h @ W_ctx[:, :E]^T + graph_context
 *   q0     : TSP  [B,E] W_ctx @ W_placeholder + graph_context (context.py:124-132); CVRP NULL
 *   wcap   : CVRP [E] W_ctx[:, E]  (remaining-capacity column, context.py:155-157); TSP NULL
 *   tc_img : [B, rl4co_tc_image_bytes(n_nodes)] bytes written by rl4co_tc_pack from gk/gv/lk: the
 *            fp16 hi/lo operand tiles of the tcgen05 kernels (K scaled by 1/sqrt(d), L by 1/sqrt(E),
 *            V transposed), laid out exactly as they sit in shared memory.  NULL = fp32 path only.  */
typedef struct rl4co_cache {
    const float* gk;
    const float* gv;
    const float* lk;
    const float* pa;
    const float* pb;
    const float* q0;
    const float* wcap;
    const void* tc_img; /* nullable: operand image built by rl4co_tc_pack (tensor path, multistart) */
    int64_t B;
    int32_t n_nodes;
    int32_t embed_dim; /* must be 128 */
    int32_t num_heads; /* must be 8   */
    int32_t tc_layout; /* layout of tc_img: RL4CO_TC_LAYOUT_* (occupies the struct's former tail padding) */
} rl4co_cache;

/* ---- AttentionModelDecoder.forward — am/decoder.py:161-198 (+ context.py, attention.py) --
 * One decoder step for all rows: returns logits [rows, n_nodes] in row order.
 * TSP state: first_node, current_node int64 [rows], i int64 [rows] (i==0 -> placeholder);
 * CVRP state: current_node int64 [rows], used_capacity, vehicle_capacity float [rows].
 * mask bool [rows, n_nodes].  Row r belongs to instance r % cache->B.                      */
int rl4co_decode_step(int32_t env, const rl4co_cache* cache, const int64_t* first_node,
                      const int64_t* current_node, const int64_t* i_step,
                      const float* used_capacity, const float* vehicle_capacity,
                      const uint8_t* mask, float* logits_out, int32_t* err, int64_t rows,
                      void* stream);

/* ---- ConstructivePolicy.forward main loop, fused — constructive/base.py:221-263 ----------
 * Persistent kernel: multistart pre-hook (decoding.py:284-332), decoder step, process_logits,
 * selection, env step and reward for ALL steps of every row, no host sync.                  */
typedef struct rl4co_rollout_args {
    int32_t env;               /* RL4CO_ENV_*                                               */
    int32_t mode;              /* RL4CO_MODE_*                                              */
    int32_t num_starts;        /* S; 0 or 1 = single start (no forced first action)         */
    int32_t max_steps;         /* T_alloc: columns available in actions / logp              */
    float tanh_clipping;       /* 10.0 (am/policy.py:70)                                    */
    float temperature;         /* 1.0                                                       */
    uint32_t seed;             /* counter-noise seed (SAMPLING, gumbel == NULL)             */
    int64_t inst_offset;       /* global index of instance 0 of this shard (noise key)      */
    const float* locs;         /* [B, n_nodes, 2] (CVRP: depot first)                       */
    const float* demand;       /* CVRP [B, n_nodes-1] (already / capacity); TSP NULL        */
    const float* vehicle_capacity; /* CVRP [B]; TSP NULL                                    */
    const float* gumbel;       /* nullable [rows, max_steps, n_nodes] supplied noise        */
    const int64_t* actions_in; /* EVALUATE: [rows, actions_in_stride]; multistart: col 0 = start */
    int64_t actions_in_stride;
    int32_t actions_in_T;
    int64_t* actions;          /* out [rows, max_steps] (pre-zeroed by caller)              */
    float* logp;               /* out [rows, max_steps] (pre-zeroed by caller)              */
    float* logp_all;           /* nullable out [rows, max_steps, n_nodes] (store_all_logp)  */
    float* reward;             /* out [rows]  = -tour length                                */
    float* log_likelihood;     /* out [rows]  = sum_t logp                                  */
    int32_t* lengths;          /* out [rows]  actions emitted before the row was done       */
    int32_t* max_len;          /* out device int32: max over rows of lengths (pre-zeroed)   */
    int32_t* err;              /* device int32[2] (pre-zeroed)                              */
    int32_t impl;              /* RL4CO_IMPL_AUTO / _FP32 (CUDA cores) / _TC (tcgen05, needs tc_img) */
    int32_t start_mod;         /* multistart: forced first action = start % start_mod (+1 for CVRP), the
                                  reference's `% env.generator.num_loc` (ops.py:147-161); 0 = number of
                                  customers of the instance (n_nodes for TSP, n_nodes-1 for CVRP)       */
    long long* phase_cycles;   /* nullable debug out, device int64[16]: clock64 cycles CTA 0 spent per phase
                                  of the tensor-path kernel (accumulated; caller zeroes)               */
} rl4co_rollout_args;

int rl4co_rollout(const rl4co_cache* cache, const rl4co_rollout_args* args, void* stream);

/* ---- Allocation-free rollout.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same contract and results as rl4co_rollout on the tensor path (8 <= num_starts <= 128, n_nodes <= 112), but the
 * fp16 operand image that rl4co_tc_pack would write into caller memory (cache->tc_img, ignored here) is packed into a
 * static device array of RL4CO_NOALLOC_SCRATCH_BYTES inside the library: the call allocates and frees nothing.
 * Limit: B * rl4co_tc_image_bytes(n_nodes) <= RL4CO_NOALLOC_SCRATCH_BYTES (rl4co_noalloc_max_instances: 48 instances
 * at n_nodes = 100, 170 at 20); larger batches return RL4CO_E_ARG.  The pack's fp16-range flag is raised in args->err.
 * Every call on a device shares the one array (fully rewritten by each call): calls that overlap on the GPU race -
 * serialise them on one stream.                                                                                      */
#define RL4CO_NOALLOC_SCRATCH_BYTES (8u << 20)
int64_t rl4co_noalloc_max_instances(int32_t n_nodes);
int rl4co_rollout_noalloc(const rl4co_cache* cache, const rl4co_rollout_args* args, void* stream);

/* ---- Tensor-core (tcgen05/TMEM) path for multistart rows: S = rows/B starts of an instance form the
 * M dimension of the MMAs (8 <= S <= 128, n_nodes <= 112).  Same arithmetic as above with every fp32
 * operand split into fp16 hi + lo and three MMAs per product (fp32 accumulate).                      */
int64_t rl4co_tc_image_bytes(int32_t n_nodes); /* bytes per instance of cache->tc_img; 0 = unsupported */
/* Two operand-image layouts (same size), selected by cache->tc_layout at pack time and read back by rl4co_rollout:
 *   RESIDENT: six whole tiles; rollout_tc_kernel keeps them in shared memory (1 instance per SM, 512 threads);
 *   STREAM  : head-sliced; rollout_tc2_kernel streams the slices from L2 with TMA (2 instances per SM, 256 threads each).
 * rl4co_tc_preferred_layout(B): STREAM when the batch has more instances than the GPU has SMs.                      */
#define RL4CO_TC_LAYOUT_RESIDENT 0
#define RL4CO_TC_LAYOUT_STREAM 1
int32_t rl4co_tc_preferred_layout(int64_t B);
int rl4co_tc_pack(const rl4co_cache* cache, void* image, int32_t* err, void* stream);
/* AttentionModelDecoder.forward (am/decoder.py:161-198) for multistart rows on the tensor path: RAW
 * logits [rows, n_nodes].  TSP callers must be past the first step (i >= 1, always true after the
 * multistart pre-hook).  dbg_scores [rows,8,n_nodes] / dbg_heads [rows,128] (nullable) receive the
 * pre-mask attention scores and the normalised glimpse heads for the parity tests.                   */
int rl4co_tc_decode_step(int32_t env, const rl4co_cache* cache, const int64_t* first_node,
                         const int64_t* current_node, const float* used_capacity,
                         const float* vehicle_capacity, const uint8_t* mask, float* logits_out,
                         float* dbg_scores, float* dbg_heads, int32_t* err, int64_t rows, void* stream);

/* ---- Graph-attention encoder (the caller of the rollout path) on tcgen05 -----------------
 * rl/zoo/am/encoder.py:70-91 -> nn/graph/attnnet.py:16-103: per layer
 *   qkv = h Wqkv^T + b (attention.py:113) ; a = MHA(qkv) (:114-131) ; h = IN(h + a Wo^T + bo) (attnnet.py:43-49,
 *   ops.py:32-56) ; h = IN(h + W2 relu(W1 h + b1) + b2) (attnnet.py:50-60).
 * All tensors fp32, contiguous, [B, n, channels] with n <= 128 nodes; one CTA owns an instance, so the instance
 * norm (statistics over the node axis) is part of the GEMM epilogue.                                              */
#define RL4CO_ENC_EPI_PLAIN 0    /* Y = X W^T + bias                                                   */
#define RL4CO_ENC_EPI_RELU 1     /* Y = relu(X W^T + bias)                                             */
#define RL4CO_ENC_EPI_RES_NORM 2 /* Y = InstanceNorm(residual + X W^T + bias) * gamma + beta, n_out=128 */
#define RL4CO_ENC_EPI_RES_AFFINE 3 /* Y = (residual + X W^T + bias) * gamma + beta, n_out=128: eval-mode BatchNorm1d
                                      (nn/ops.py:46-47) with gamma = w / sqrt(running_var + eps),
                                      beta = b - running_mean * gamma folded by the caller                       */
int64_t rl4co_enc_packed_bytes(int32_t n_out, int32_t k); /* 0 = unsupported (both must be multiples of 128) */
/* nn.Linear weight W [n_out, k] -> fp16 hi/lo MMA tiles (once per forward / weight version) */
int rl4co_enc_pack_weight(const float* W, int32_t n_out, int32_t k, void* packed, void* stream);
int rl4co_enc_linear(const float* X, const void* Wpacked, const float* bias, const float* residual,
                     const float* gamma, const float* beta, float eps, float* Y, int64_t B, int32_t n,
                     int32_t k, int32_t n_out, int32_t epilogue, void* stream);
/* Fused feed-forward block (attnnet.py:50-60, nn/mlp.py): Y = Norm(X + relu(X W1^T + b1) W2^T + b2) with the hidden
 * activations kept on chip (TMEM); W1 [hidden,128], W2 [128,hidden] packed by rl4co_enc_pack_weight, hidden in
 * {128,256,384,512}; epilogue = RL4CO_ENC_EPI_RES_NORM (gamma/beta/eps of the InstanceNorm) or _RES_AFFINE.        */
int rl4co_enc_ffn(const float* X, const void* W1packed, const float* b1, const void* W2packed, const float* b2,
                  const float* gamma, const float* beta, float eps, float* Y, int64_t B, int32_t n, int32_t hidden,
                  int32_t epilogue, void* stream);
/* Fused attention block (attnnet.py:43-49, attention.py:66-136): Y = Norm(X + MHA(X) Wo^T + bo), 8 heads x 16, no
 * mask, n <= 112; Wqkv [384,128] and Wo [128,128] packed by rl4co_enc_pack_weight; q, k, v and the heads stay on chip. */
int rl4co_enc_attn_block(const float* X, const void* Wqkv_packed, const float* bqkv, const void* Wo_packed,
                         const float* bo, const float* gamma, const float* beta, float eps, float* Y, int64_t B,
                         int32_t n, int32_t epilogue, void* stream);
/* qkv [B, n, 384] = (q | k | v) x (8 heads x 16) -> heads [B, n, 128] (input of out_proj), no mask */
int rl4co_enc_attention(const float* qkv, float* out, int64_t B, int32_t n, void* stream);

/* ---- POMO evaluation reduction after the rollout - pomo/model.py:115-145 (unbatchify ops.py:35-56, gather :59-71) ----
 * reward [S*A*B] / actions [S*A*B, T] (nullable) of a multistart rollout over an A-fold augmented batch, rows start-major
 * (row = s*A*B + a*B + b).  Outputs: max_reward [B,A], best_start [B,A] (first maximum over starts), the winning tours
 * best_multistart_actions [B,A,T]; then over augmentations (nullable): max_aug_reward [B], best_aug [B],
 * best_aug_actions [B,T].  n_aug = 1 for an un-augmented batch.                                                       */
int rl4co_best_of_starts(const float* reward, const int64_t* actions, int64_t B, int32_t n_aug, int32_t n_start,
                         int32_t T, float* max_reward, int64_t* best_start, int64_t* best_multistart_actions,
                         float* max_aug_reward, int64_t* best_aug, int64_t* best_aug_actions, void* stream);

/* ---- teacher-forced evaluate of ALL decoding steps at once, forward and backward ---------------------------------------
 * decoding.py:452-465 (Evaluate), constructive/base.py:204-229, am/decoder.py:161-198, attention.py:282-328,
 * decoding.py:140-190; the differentiable log-likelihood that reinforce.py:73-115 / pomo/model.py:103-112 backpropagate.
 * Row (b, r), r = start * (T-1) + step, is independent of every other row once the actions are known:
 *   q     [B,R,128]  glimpse query of the row (step context + fixed context, context.py:70-74,118-157; caller-built),
 *   gk, gv, lkf [B,n,128]  glimpse key / value and logit key folded with project_out (lkf = lk @ W_out),
 *   mask  [B,R,n] bool (feasible = 1; rl4co_replay_states writes it), taken [B,R] i64 (the action of the row).
 * fwd: logp [B,R] = log_softmax(clip * tanh(heads . lkf / sqrt(128)) masked / temperature)[taken], and the values the
 *      backward reuses: heads [B,R,128], lse_h [B,R,8] (log-sum-exp of each head's scores), lse_z [B,R], maskw [B,R,4]
 *      (the mask as bit words).
 * bwd: given dlogp [B,R]: dq [B,R,128] (written), dgk / dgv / dlkf [B,n,128] (ACCUMULATED: zero them first);
 *      dheads, du_t [B,R,128], dd [B,R,8] and scale_ws [2] are scratch.  The backward runs on dlogp * s, s the power of two that
 *      brings max|dlogp| into [0.5, 1) (scale_ws = {s, 1/s}), and unscales its outputs: the fp16 hi / lo MMA operands keep their
 *      accuracy at any loss scale.  n_nodes <= rl4co_eval_ll_max_nodes() (128).                                              */
int32_t rl4co_eval_ll_max_nodes(void);
int rl4co_eval_ll_fwd(const float* q, const float* gk, const float* gv, const float* lkf, const uint8_t* mask,
                      const int64_t* taken, float* logp, float* heads, float* lse_h, float* lse_z, uint32_t* maskw,
                      int64_t B, int64_t R, int32_t n_nodes, float tanh_clipping, float temperature, void* stream);
int rl4co_eval_ll_bwd(const float* q, const float* gk, const float* gv, const float* lkf, const uint32_t* maskw,
                      const int64_t* taken, const float* heads, const float* lse_h, const float* lse_z, const float* dlogp,
                      float* dheads, float* du_t, float* dd, float* scale_ws, float* dq, float* dgk, float* dgv, float* dlkf,
                      int64_t B, int64_t R, int32_t n_nodes, float tanh_clipping, float temperature, void* stream);

/* ---- the glimpse query of all teacher-forced rows and its gradient (context.py:118-157, am/decoder.py:223-224) ----------
 * q[b,s,t,:] = pb[b,cur[b,s,t],:] (+ pa[b,first[b,s],:]) (+ rem[b,s,t] * wcap[:]) (+ ctx[b,:]), 128 channels, where
 * pb / pa [B,n,128] are the node embeddings under the two halves of project_context (TSP: [first, current]; CVRP: current
 * node only and wcap = the remaining-capacity column) and ctx [B,128] the projected graph mean; optional terms are NULL
 * (pa with first, wcap with rem).  cur [B,S,Tm] and first [B,S] int64, rem [B,S,Tm] f32; node ids outside [0, n) are
 * clamped.  bwd: dpb (and dpa, dwcap [128]) are ACCUMULATED from dq [B,S*Tm,128] (zero them first); the gradient of ctx
 * is the sum of dpb over its node axis.  n_nodes <= 128.                                                              */
int rl4co_eval_query_fwd(const float* pb, const float* pa, const float* wcap, const float* ctx, const int64_t* cur,
                         const int64_t* first, const float* rem, float* q, int64_t B, int64_t S, int64_t Tm,
                         int32_t n_nodes, void* stream);
int rl4co_eval_query_bwd(const float* dq, const int64_t* cur, const int64_t* first, const float* rem, float* dpb,
                         float* dpa, float* dwcap, int64_t B, int64_t S, int64_t Tm, int32_t n_nodes, void* stream);

/* ---- instance normalisation of the encoder's TRAINING path and its backward (nn/ops.py:48-49: nn.InstanceNorm1d over the node
 * axis of x [B,N,128], affine w / bias [128], biased variance, eps inside the square root); fp32.  fwd writes y and the
 * statistics mean / rstd [B,128] the backward reads; bwd writes dx and ACCUMULATES dw / db [128] (zero them first).
 * N <= rl4co_instnorm_max_nodes() (128).  (Inference runs the fused encoder kernels, rl4co_enc_*.)                          */
int32_t rl4co_instnorm_max_nodes(void);
int rl4co_instnorm_fwd(const float* x, const float* w, const float* bias, float* y, float* mean, float* rstd, int64_t B,
                       int32_t N, float eps, void* stream);
int rl4co_instnorm_bwd(const float* x, const float* w, const float* mean, const float* rstd, const float* dy, float* dx,
                       float* dw, float* db, int64_t B, int32_t N, void* stream);

/* ---- on-device instance generation: the input spec of tsp/generator.py:53-62 and cvrp/generator.py:120-147 -----------
 * locs [B, n_loc, 2] uniform in [0,1); CVRP also depot [B,2] and demand [B, n_loc] = integer 1..9 / capacity.  Values come
 * from the counter hash keyed (seed, inst_offset + b, element), so shards of a global batch agree with the whole.   */
int rl4co_generate_instances(int32_t env, uint32_t seed, int64_t inst_offset, int64_t B, int32_t n_loc,
                             float capacity, float* locs, float* depot, float* demand, void* stream);

/* counter-noise reference values, for tests: gumbel [rows, n_nodes] at step t */
int rl4co_noise_gumbel(uint32_t seed, int64_t inst_offset, int64_t B, int32_t t, float* out,
                       int64_t rows, int32_t n_nodes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RL4CO_B200_H */
