/*
 * rl4co_slap_b200 - C ABI of the B200-native (sm_100a) rollout path.
 *
 * Drop-in boundary for the batched environment + Attention-Model decode rollout of
 * j4n1k/rl4co-slap.  The reference is pure Python, so "the reference's FFI" for this path
 * is a ctypes binding (INTEGRATION.md shows the stub a maintainer adds); each entry point
 * below names the reference interface (file:line under rl4co/) it replaces.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller (torch caching allocator);
 *    the library allocates nothing and frees nothing;
 *  - `stream` is a cudaStream_t passed as void*; calls are asynchronous and stream ordered,
 *    no host synchronisation inside;
 *  - return value: 0 = launched, non-zero = host-side argument / launch error, message via
 *    rl4co_last_error() (thread local);
 *  - device-side conditions the reference raises as AssertionError ("Invalid tour",
 *    "Used more than capacity", "infeasible action selected", "Logits contain NaNs",
 *    "Logprobs should not be -inf") are OR-ed into the caller's int32 `status` word; the
 *    host wrapper reads it once per call and raises with the reference's message;
 *  - dtypes/shapes are the reference's: int64 actions and node ids, bool (1 byte) masks,
 *    uint8 `visited`, float32 coordinates / demands / capacities.
 */
#ifndef RL4CO_SLAP_B200_H
#define RL4CO_SLAP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RL4CO_ABI_VERSION 4

/* status bits (device int32 word) */
#define RL4CO_ST_INVALID_TOUR      1   /* tsp/env.py:173, cvrp/env.py:175 */
#define RL4CO_ST_OVER_CAPACITY     2   /* cvrp/env.py:190 */
#define RL4CO_ST_INFEASIBLE_ACTION 4   /* utils/decoding.py:379,395 */
#define RL4CO_ST_NAN_LOGITS        8   /* models/nn/attention.py:285 */
#define RL4CO_ST_NEG_INF_LOGP      16  /* utils/decoding.py:59 */
#define RL4CO_ST_ACTION_RANGE      32  /* action outside [0,N): torch gather/scatter would raise */
#define RL4CO_ST_DUPLICATES        64  /* op/env.py:196-199 */
#define RL4CO_ST_OVER_LENGTH       128 /* op/env.py:211-215 */
#define RL4CO_ST_PRIZE_SHORT       256 /* pctsp/env.py:193-200 */
#define RL4CO_ST_DEPOT_TWICE       512 /* sdvrp/env.py:155-157 */
#define RL4CO_ST_DEMAND_LEFT       1024 /* sdvrp/env.py:164 */

typedef void *rl4co_stream_t;

int rl4co_abi_version(void);
const char *rl4co_last_error(void);

/* ------------------------------------------------------------------ TSP (K1) ------- */

/* TSPEnv._reset, envs/routing/tsp/env.py:95-120 (+ done/terminated of EnvBase.reset,
 * envs/common/base.py:143).  Any output pointer may be NULL. */
int rl4co_tsp_reset(int64_t B, int32_t N, int64_t *first_node, int64_t *current_node,
                    int64_t *i, uint8_t *action_mask, float *reward, uint8_t *done,
                    uint8_t *terminated, rl4co_stream_t stream);

/* TSPEnv._step, envs/routing/tsp/env.py:67-93.  In place on i / first_node / action_mask.
 * `flag_ws` is a caller-provided int32 scratch word used for the reference's batch-global
 * first-step test (`td["i"].all() == 0`, :70).  done is [B]. */
int rl4co_tsp_step(int64_t B, int32_t N, const int64_t *action, int64_t *i,
                   int64_t *first_node, int64_t *current_node, uint8_t *action_mask,
                   uint8_t *done, int32_t *flag_ws, int32_t *status, rl4co_stream_t stream);

/* Fused reset + T x _step + get_reward (tsp/env.py:67-93,157-173; utils/ops.py:65-101).
 * actions [B,T] int64.  Outputs (each may be NULL): final first_node/current_node [B],
 * i [B], action_mask [B,N], done [B], reward [B] (= -closed tour length of the T actions).
 * check_validity != 0 ORs RL4CO_ST_INVALID_TOUR when sort(actions) != arange(N). */
int rl4co_tsp_replay(int64_t B, int32_t N, int32_t T, const float *locs,
                     const int64_t *actions, int64_t *first_node, int64_t *current_node,
                     int64_t *i, uint8_t *action_mask, uint8_t *done, float *reward,
                     int32_t check_validity, int32_t *status, rl4co_stream_t stream);

/* ------------------------------------------------------------------ ATSP (K1) ------ */
/* Asymmetric TSP, envs/routing/atsp/env.py: the state machine is the TSP one (rl4co_tsp_reset / rl4co_tsp_step);
 * the fused replay takes the cost matrix [B,N,N] instead of coordinates and gathers the T tour edges
 * (reward = -sum cost[a_t, a_{t+1}], closed; atsp/env.py:152-162).  Outputs as rl4co_tsp_replay. */
int rl4co_atsp_replay(int64_t B, int32_t N, int32_t T, const float *cost_matrix, const int64_t *actions,
                      int64_t *first_node, int64_t *current_node, int64_t *i, uint8_t *action_mask,
                      uint8_t *done, float *reward, int32_t check_validity, int32_t *status,
                      rl4co_stream_t stream);

/* ------------------------------------------------------------------ CVRP (K1) ------ */
/* N counts the depot (node 0).  demand [B,N-1] is already divided by capacity. */

/* CVRPEnv.get_action_mask, envs/routing/cvrp/env.py:137-149. */
int rl4co_cvrp_action_mask(int64_t B, int32_t N, const float *demand,
                           const float *used_capacity, const float *vehicle_capacity,
                           const uint8_t *visited, const int64_t *current_node,
                           uint8_t *action_mask, rl4co_stream_t stream);

/* CVRPEnv._step, envs/routing/cvrp/env.py:73-105 (includes the mask update). */
int rl4co_cvrp_step(int64_t B, int32_t N, const int64_t *action, const float *demand,
                    const float *vehicle_capacity, float *used_capacity,
                    int64_t *current_node, uint8_t *visited, uint8_t *action_mask,
                    uint8_t *done, int32_t *status, rl4co_stream_t stream);

/* Fused reset + T x _step + get_reward + check_solution_validity
 * (cvrp/env.py:73-190).  locs [B,N,2] includes the depot at row 0. */
int rl4co_cvrp_replay(int64_t B, int32_t N, int32_t T, const float *locs,
                      const float *demand, const float *vehicle_capacity,
                      const int64_t *actions, int64_t *current_node, float *used_capacity,
                      uint8_t *visited, uint8_t *action_mask, uint8_t *done, float *reward,
                      int32_t check_validity, int32_t *status, rl4co_stream_t stream);

/* ------------------------------------------------------------------ OP (K1) -------- */
/* Orienteering problem, the sibling environment of the routing family.  N counts the depot (node 0);
 * locs [B,N,2], prize [B,N] (0 for the depot) and max_length [B,N] are the tensors of the reset state
 * (op/env.py:110-152): max_length[b,n] is the budget when ARRIVING at n (return leg and 1e-6 already
 * subtracted).  visited is the reference's bool tensor (1 byte per node), tour_length /
 * current_total_prize / i are [B], current_node [B,1]. */

/* OPEnv.get_action_mask, envs/routing/op/env.py:154-170. */
int rl4co_op_action_mask(int64_t B, int32_t N, const float *locs, const float *max_length,
                         const float *tour_length, const uint8_t *visited, const int64_t *current_node,
                         uint8_t *action_mask, rl4co_stream_t stream);

/* OPEnv._step, envs/routing/op/env.py:74-108 (transition + mask in one launch).  done is [B]. */
int rl4co_op_step(int64_t B, int32_t N, const int64_t *action, const float *locs, const float *prize,
                  const float *max_length, float *tour_length, float *current_total_prize,
                  int64_t *current_node, uint8_t *visited, int64_t *i, uint8_t *action_mask,
                  uint8_t *done, int32_t *status, rl4co_stream_t stream);

/* Fused reset-state + T x _step + _get_reward + check_solution_validity (op/env.py:74-215).  Outputs may be
 * NULL.  check_validity != 0 ORs RL4CO_ST_DUPLICATES / RL4CO_ST_OVER_LENGTH. */
int rl4co_op_replay(int64_t B, int32_t N, int32_t T, const float *locs, const float *prize,
                    const float *max_length, const int64_t *actions, int64_t *current_node,
                    float *tour_length, float *current_total_prize, uint8_t *visited,
                    uint8_t *action_mask, int64_t *i, uint8_t *done, float *reward,
                    int32_t check_validity, int32_t *status, rl4co_stream_t stream);

/* ------------------------------------------------------------------ PCTSP (K1) ----- */
/* Prize-collecting TSP and its stochastic variant (spctsp: real_prize is the stochastic prize).  N counts the
 * depot; real_prize / penalty [B,N] carry 0 for the depot (pctsp/env.py:114-118); current_node, cur_total_prize,
 * cur_total_penalty, i are [B]; visited is the reference's bool tensor. */

/* PCTSPEnv.get_action_mask, envs/routing/pctsp/env.py:141-148. */
int rl4co_pctsp_action_mask(int64_t B, int32_t N, const float *cur_total_prize, const uint8_t *visited,
                            uint8_t *action_mask, rl4co_stream_t stream);

/* PCTSPEnv._step, envs/routing/pctsp/env.py:69-101 (transition + mask in one launch).  done is [B]. */
int rl4co_pctsp_step(int64_t B, int32_t N, const int64_t *action, const float *real_prize, const float *penalty,
                     float *cur_total_prize, float *cur_total_penalty, int64_t *current_node, uint8_t *visited,
                     int64_t *i, uint8_t *action_mask, uint8_t *done, int32_t *status, rl4co_stream_t stream);

/* Fused reset-state + T x _step + _get_reward + check_solution_validity (pctsp/env.py:69-200).  Outputs may be
 * NULL; cur_total_penalty is in/out and enters holding the reset state's value (penalty.sum(-1), :123).
 * check_validity != 0 ORs RL4CO_ST_DUPLICATES / RL4CO_ST_PRIZE_SHORT. */
int rl4co_pctsp_replay(int64_t B, int32_t N, int32_t T, const float *locs, const float *real_prize,
                       const float *penalty, const int64_t *actions, int64_t *current_node,
                       float *cur_total_prize, float *cur_total_penalty, uint8_t *visited, uint8_t *action_mask,
                       int64_t *i, uint8_t *done, float *reward, int32_t check_validity, int32_t *status,
                       rl4co_stream_t stream);

/* ------------------------------------------------------------------ SDVRP (K1) ----- */
/* Split-delivery VRP, envs/routing/sdvrp/env.py.  N counts the depot; demand_with_depot [B,N] (slot 0 = depot = 0)
 * is the remaining demand, updated in place by the step; demand [B,N-1] is the original one (already / capacity). */

/* SDVRPEnv.get_action_mask, sdvrp/env.py:132-140. */
int rl4co_sdvrp_action_mask(int64_t B, int32_t N, const float *demand_with_depot, const float *used_capacity,
                            const float *vehicle_capacity, const int64_t *current_node, uint8_t *action_mask,
                            rl4co_stream_t stream);

/* SDVRPEnv._step, sdvrp/env.py:63-101 (transition + mask in one launch).  done is [B]. */
int rl4co_sdvrp_step(int64_t B, int32_t N, const int64_t *action, const float *vehicle_capacity,
                     float *demand_with_depot, float *used_capacity, int64_t *current_node,
                     uint8_t *action_mask, uint8_t *done, int32_t *status, rl4co_stream_t stream);

/* Fused reset + T x _step + get_reward (cvrp/env.py:151-160, inherited) + check_solution_validity
 * (sdvrp/env.py:142-164).  check_validity != 0 ORs RL4CO_ST_DEPOT_TWICE / RL4CO_ST_DEMAND_LEFT. */
int rl4co_sdvrp_replay(int64_t B, int32_t N, int32_t T, const float *locs, const float *demand,
                       const float *vehicle_capacity, const int64_t *actions, int64_t *current_node,
                       float *used_capacity, float *demand_with_depot, uint8_t *action_mask, uint8_t *done,
                       float *reward, int32_t check_validity, int32_t *status, rl4co_stream_t stream);

/* ------------------------------------------------------------------ SLAP (K1) ------ */
/* N storage locations (node 0 = depot, never assignable), P products. */

/* SLAPEnv._step, envs/warehousing/slap/env.py:38-93.  `product` = to_choose[:,0] as the
 * reference stores it (float32, element stride `product_stride`); the to_choose slice
 * itself is a view operation done by the host wrapper.  done is [B,1] = (i == P-1). */
int rl4co_slap_step(int64_t B, int32_t N, int32_t P, const int64_t *action,
                    const float *product, int64_t product_stride, int64_t *i,
                    int32_t *assignment, uint8_t *action_mask, uint8_t *done,
                    int32_t *status, rl4co_stream_t stream);

/* SLAPEnv._get_reward, slap/env.py:131-143: sum over O orders of minus the closed
 * Euclidean tour through locs[assignment[picklist[b,o,:]]].  picklist [B,O,K] int64. */
int rl4co_slap_reward(int64_t B, int32_t N, int32_t P, int32_t O, int32_t K,
                      const float *locs, const int32_t *assignment, const int64_t *picklist,
                      float *reward, rl4co_stream_t stream);

/* Fused reset + P x _step + _get_reward.  actions [B,P] int64 (storage location of
 * product t at step t).  Outputs may be NULL. */
int rl4co_slap_replay(int64_t B, int32_t N, int32_t P, int32_t O, int32_t K,
                      const float *locs, const int64_t *picklist, const int64_t *actions,
                      int32_t *assignment, uint8_t *action_mask, int64_t *i, uint8_t *done,
                      float *reward, int32_t *status, rl4co_stream_t stream);

/* ------------------------------------------------------- AM decoder (K3, K2) ------- */

#define RL4CO_ENV_TSP  0
#define RL4CO_ENV_CVRP 1
#define RL4CO_ENV_SLAP 2
#define RL4CO_ENV_OP   3   /* envs/routing/op/env.py; register-file kernel only (N <= 128, no logp_all); multistart forces
                            * start node s % num_loc + 1 - the reference's multinomial resampling of infeasible starts
                            * (utils/ops.py:166-176) is the caller's business */
#define RL4CO_ENV_PCTSP 4  /* envs/routing/pctsp/env.py and spctsp/env.py (prize = the real prize); same limits as OP */
#define RL4CO_ENV_SDVRP 5  /* envs/routing/sdvrp/env.py with SDVRPDynamicEmbedding; single start, register-file kernel only */

#define RL4CO_MODE_GREEDY   0  /* utils/decoding.py:371-381,410-418 */
#define RL4CO_MODE_SAMPLING 1  /* utils/decoding.py:383-397,475-483 (Philox instead of torch.multinomial) */
#define RL4CO_MODE_EVALUATE 2  /* utils/decoding.py:486-499 */

#define RL4CO_EMBED_DIM 128
#define RL4CO_NUM_HEADS 8

/* K3 - one-off projection of the encoder's node embeddings into the per-node tables the
 * rollout kernel consumes.  Replaces AttentionModelDecoder._precompute_cache
 * (models/zoo/am/decoder.py:208-237) plus the per-node halves of the context projection
 * (models/nn/env_embeddings/context.py:67-71).
 *   tables[t][m][j] = sum_e h[m][e] * W[t*E + j][e]      m in [0,M), M = B*N, E = 128
 * W is [n_tables*E, E] row-major (nn.Linear layout).  fp32 in, fp32 out, fp32-accurate. */
int rl4co_am_project(int64_t M, int32_t n_tables, const float *h, const float *W,
                     float *tables, rl4co_stream_t stream);

/* K3 on the tensor cores (tcgen05.mma, 3-term hi/lo split a w = a_hi w_hi + a_hi w_lo + a_lo w_hi, fp32 accumulation in
 * TMEM; same result as rl4co_am_project to fp32 rounding level: max error 1.5e-5 on values of size 18 against 1.0e-5 for the
 * fp32 SIMT kernel).  K = 128 products - this entry point, and rl4co_linear_tc with n_in = 128 - use FP16 halves
 * (kind::f16) with exact power-of-two range scales per 128-row tile of the input and per weight table, so that inputs of
 * any magnitude are safe; K > 128 uses TF32 halves (kind::tf32).  The weights are first packed once per weight update into
 * the shared-memory image the MMA reads; the image is OPAQUE (its format follows the rule above) and must be consumed by the
 * same library version that packed it:
 *   image: rl4co_am_weight_image_floats(n_tables) floats (a 128 KB slot per table), caller owned. */
int64_t rl4co_am_weight_image_floats(int32_t n_tables);
int rl4co_am_pack_weights(int32_t n_tables, const float *W, float *image, rl4co_stream_t stream);
int rl4co_am_project_tc(int64_t M, int32_t n_tables, const float *h, const float *image,
                        float *tables, rl4co_stream_t stream);

/* The same tensor-core kernel with a linear-layer epilogue, for the linears of the PyTorch encoder that
 * feeds the path (models/nn/graph/attnnet.py:16-63: Wqkv, out_proj, feed-forward):
 *   y[m][j] = act(sum_e x[m][e] * W[j][e] + bias[j]),  act = ReLU if relu != 0,
 * n_in and n_out multiples of 128 (<= 2048); n_in > 128 runs as 128-wide K chunks that accumulate in TMEM.
 * x [M,n_in], W [n_out,n_in], y [M,n_out] row-major; bias may be NULL.
 * image = rl4co_linear_pack(W): (n_in/128) * rl4co_am_weight_image_floats(n_out/128) floats.
 * rl4co_linear_tc_ex adds the skip connection and the eval-mode batch norm that follow out_proj and the
 * feed-forward in the encoder layer (attnnet.py:44-58, SkipConnection + Normalization) to the epilogue:
 *   y = (act(x W^T + bias) + residual) * scale + shift
 * residual [M,n_out] (must not alias y), scale / shift [n_out]; each may be NULL (0, 1, 0). */
int rl4co_linear_pack(int32_t n_out, int32_t n_in, const float *W, float *image, rl4co_stream_t stream);
/* The encoder's init embedding (models/nn/env_embeddings/init.py:51-64 TSP, :111-134 VRP; examples/slap.py:11-47): a Linear
 * with n_in = 1..8 input features and 128 outputs, y[m][j] = sum_k x[m][k] W[j][k] + bias[j] (k ascending, fp32 FMA).
 * x [M,n_in], W [128,n_in], bias [128] or NULL, y [M,128], all contiguous fp32. */
int rl4co_linear_narrow(int64_t M, int32_t n_in, const float *x, const float *W, const float *bias, float *y,
                        rl4co_stream_t stream);
int rl4co_linear_tc(int64_t M, int32_t n_out, int32_t n_in, const float *x, const float *image,
                    const float *bias, int32_t relu, float *y, rl4co_stream_t stream);
int rl4co_linear_tc_ex(int64_t M, int32_t n_out, int32_t n_in, const float *x, const float *image,
                       const float *bias, int32_t relu, const float *residual, const float *scale,
                       const float *shift, float *y, rl4co_stream_t stream);

/* Self-attention of the PyTorch encoder's layers for the path's graph sizes (N <= 128, 8 heads x 16,
 * no mask, fp32): replaces the F.scaled_dot_product_attention call of MultiHeadAttention
 * (models/nn/attention.py:65-135).  qkv [B,N,3,8,16] = the fused Wqkv output, out [B,N,128]. */
int rl4co_encoder_attention(int64_t B, int32_t N, const float *qkv, float *out, rl4co_stream_t stream);
/* Same contract with S = Q K^T on the tensor cores (tcgen05.mma kind::tf32, 3-term split, scores in TMEM);
 * softmax and P V stay fp32 on the CUDA cores.  The kernel the encoder uses; the one above is the SIMT cross-check. */
int rl4co_encoder_attention_tc(int64_t B, int32_t N, const float *qkv, float *out, rl4co_stream_t stream);
/* Kernel behind rl4co_encoder_attention_tc: 0 = auto (N <= 112: P V on the tensor cores as well, P kept in TMEM as fp16
 * hi / lo pairs, V as a range-scaled fp16 operand, four consumer groups - kernel 4; above that kernel 1), 1 = scores only,
 * 2 = both GEMMs as 3xTF32, 3 = both GEMMs with the fp16 P V and a one-pass softmax, 4 = the same with a two-pass softmax and
 * four units in flight (2-4: error for N > 112).  Process-wide; for cross-checks and timing. */
int rl4co_set_attention_impl(int32_t impl);

/* Graph context: gctx[b] = Wfixed . mean_n h[b,n,:]  (am/decoder.py:225-228). */
int rl4co_am_graph_context(int64_t B, int32_t N, const float *h, const float *Wfixed,
                           float *gctx, rl4co_stream_t stream);

/* K2 - persistent rollout: encoder-output caches -> tours, log-likelihoods, rewards in ONE
 * launch.  Replaces the hot loop ConstructivePolicy.forward
 * (models/common/constructive/base.py:229-265): AttentionModelDecoder.forward
 * (am/decoder.py:162-200), PointerAttention.forward (nn/attention.py:263-309),
 * DecodingStrategy.step / process_logits (utils/decoding.py:141-191,327-369), the multistart
 * pre_decoder_hook (utils/decoding.py:265-313), Env._step and Env.get_reward.
 *
 * Rollout r in [0,R), R = max(S,1)*B, is start s = r / B of instance b = r % B (the
 * reference's batchify order, utils/ops.py:11-34).
 *
 * tables (from rl4co_am_project, [n_tables][B*N][E]):
 *   0: glimpse keys  h.Wk^T (the 1/sqrt(16) score scale is applied in the kernel)
 *   1: glimpse values h.Wv^T
 *   2: folded logit keys h.(Wout^T Wl)^T  (project_out folded: <Wout o, l> = <o, Wout^T l>)
 *   3: (tsp, cvrp) per-node context h.Wctx[:, cur-half]^T
 *   4: (tsp) first-node context h.Wctx[:, first-half]^T - a trajectory reads ONE row of it (its first node's), so a
 *      single-start caller may leave table 4 out (n_tables = 4) and pass h / w_first instead (end of the struct): the
 *      kernel then forms that row itself at step 1, one 128x128 mat-vec per rollout, and the projection writes 20 % less
 */
typedef struct rl4co_am_rollout_args {
    int32_t env;            /* RL4CO_ENV_* */
    int32_t mode;           /* RL4CO_MODE_* */
    int64_t B;              /* instances */
    int32_t S;              /* starts per instance; <=1 means single start */
    int32_t multistart;     /* !=0: first action forced to the POMO start node, log-prob 0 */
    int32_t num_loc;        /* generator.num_loc for select_start_nodes (utils/ops.py:158-163) */
    int32_t N;              /* nodes (incl. depot for cvrp / slap) */
    int32_t T;              /* second dim of actions / logp (step capacity) */
    int32_t P;              /* slap: products (= steps) */
    int32_t O, K;           /* slap: orders, SKUs per order */
    int32_t n_tables;
    float temperature;
    float tanh_clipping;    /* <=0: no clipping */
    int32_t mask_logits;
    const float *tables;
    const float *gctx;      /* [B,E] or NULL (use_graph_context=False) */
    const float *q0;        /* [E] tsp: Wctx . W_placeholder (step-0 context), else NULL */
    const float *wcap;      /* [E] cvrp: Wctx[:, E] (remaining-capacity column), else NULL */
    const float *locs;      /* [B,N,2] */
    const float *demand;    /* cvrp [B,N-1] */
    const float *vehicle_capacity; /* cvrp [B] */
    const int64_t *picklist;/* slap [B,O,K] */
    const int64_t *actions_in; /* evaluate: [R,T] */
    uint64_t seed, offset;  /* Philox key / counter base for sampling */
    int64_t *actions;       /* [R,T] out (padded with 0 after done) */
    float *logp;            /* [R,T] chosen log-probs, may be NULL */
    float *logp_all;        /* [R,T,N] all log-probs (store_all_logp / entropy), may be NULL */
    float *ll;              /* [R] sum of chosen log-probs */
    float *reward;          /* [R] */
    int32_t *steps;         /* [R] steps until done, may be NULL */
    int32_t *assignment;    /* slap [R,P] out, may be NULL */
    int32_t *status;
    /* op (appended, ABI 2): prize [B,N] with 0 for the depot (op/env.py:124-126) and the per-node length
     * budget max_length [B,N] of the reset state (op/env.py:129-133); wcap = the remaining-length column of
     * project_context (context.py:385-397); reward = collected prize */
    const float *prize;
    const float *max_length;
    /* pctsp / spctsp (appended, ABI 2): prize = real_prize [B,N] (0 for the depot; the stochastic prize for spctsp),
     * penalty [B,N] (0 for the depot), prize_required [B]; wcap = the remaining-prize column of project_context
     * (context.py:368-382); reward = saved penalties - (tour length + all penalties) (pctsp/env.py:150-174) */
    const float *penalty;
    const float *prize_required;
    /* sdvrp (appended, ABI 2): demand / vehicle_capacity / wcap as for cvrp, plus the dynamic embedding
     * (nn/env_embeddings/dynamic.py:61-79) as wdyn [3][E] = its Linear(1, 3E) weight split into the glimpse-key and
     * glimpse-value columns and Wout^T times the logit-key column (project_out folded as for table 2) */
    const float *wdyn;
    /* multistart with caller-chosen start nodes (appended, ABI 3): start_nodes [R] int32, rollout r = s*B + b starts at
     * start_nodes[r] instead of s % num_loc (+1) - what a select_start_nodes_fn returned (utils/decoding.py:287-291), e.g.
     * the op branch of utils/ops.py:166-176 that resamples the starts from the feasible nodes.  NULL: the default rule. */
    const int32_t *start_nodes;
    /* tsp without table 4 (appended, ABI 4): h [B,N,E] = the node embeddings the tables were projected from and
     * w_first [E][E] = rows 4E..5E of the folded table weights (project_context.weight[:, :E], context.py:123-131; row =
     * output dim).  Both or neither; with them n_tables may be 4.  Taken by the compaction kernel (greedy / sampling, N <= 112,
     * standard logit post-processing), which is where it pays, and by the shared-memory kernel; the register-file kernel
     * is skipped for such calls.  The multistart tensor-core kernel reads table 4 for every start, so multistart callers
     * keep passing five tables. */
    const float *h;
    const float *w_first;
} rl4co_am_rollout_args;

int rl4co_am_rollout(const rl4co_am_rollout_args *args, rl4co_stream_t stream);

/* Backward of the teacher-forced log-likelihood: what REINFORCE.calculate_loss differentiates
 * (rl4co/models/rl/reinforce/reinforce.py:73-115 through ConstructivePolicy.forward with decode_type="evaluate",
 * models/common/constructive/base.py:206-207,253-265).  Recompute based - nothing is saved by the rollout:
 *   args      as for rl4co_am_rollout in RL4CO_MODE_EVALUATE (tables, gctx, q0 / wcap, demand, vehicle_capacity,
 *             actions_in [R,T], multistart / S / num_loc); tsp and cvrp, standard logit post-processing,
 *             N <= rl4co_am_evaluate_backward_max_nodes() (the instance's gradient accumulators live in shared memory)
 *   grad_ll   [R] dL / d log_likelihood of every rollout
 * Outputs, all into CALLER-ZEROED buffers (small batches spread the starts of an instance over several CTAs that meet
 * through atomics): g_tables [n_tables][B*N][128], g_gctx [B,128] (may be NULL), g_q0 [128] (tsp) and g_wcap [128]
 * (cvrp) (may be NULL).  The gradients with respect to the encoder output and the decoder weights follow from the table
 * projection, which stays with the caller (training.py keeps it as differentiable torch ops). */
int rl4co_am_evaluate_backward_max_nodes(void);
int rl4co_am_evaluate_backward(const rl4co_am_rollout_args *args, const float *grad_ll, float *g_tables,
                               float *g_gctx, float *g_q0, float *g_wcap, rl4co_stream_t stream);

/* rl4co_am_rollout with its scratch memory held by the library.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory around the call (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  The scratch of the rollout path is the table block
 * [n_tables][B*N][128] fp32 that rl4co_am_project_tc writes and only the rollout kernel reads; here it is a static
 * __device__ array of RL4CO_STATIC_TABLE_BYTES inside the library (held by every GPU that loads the library until the
 * process exits).  The call runs rl4co_am_project_tc(h, weight_image) into that array and rl4co_am_rollout out of it on
 * `stream`; it calls no cudaMalloc / cudaFree / allocating routine.
 *   args          as for rl4co_am_rollout, except args->tables, which must be NULL
 *   h             [B*N,128] fp32 encoder output, weight_image = rl4co_am_pack_weights(n_tables, W)
 * Limit: n_tables * B * N * 512 bytes <= RL4CO_STATIC_TABLE_BYTES (131 TSP100 instances with 5 tables, 162 CVRP100
 * instances - 16 200 POMO rollouts - with 4); larger requests return non-zero with a message naming the limit.
 * All calls on a device share the array; nothing in it needs clearing (the projection overwrites every row the rollout
 * reads).  Calls that OVERLAP on the GPU (different streams or host threads, same device) race on it - serialise them
 * on one stream or use rl4co_am_rollout with caller-owned tables. */
#define RL4CO_STATIC_TABLE_BYTES (32u << 20)
int rl4co_am_rollout_static(const rl4co_am_rollout_args *args, const float *h, const float *weight_image,
                            rl4co_stream_t stream);

/* rl4co_am_rollout picks between three kernels with identical semantics: register-file
 * resident tables with node compaction (TSP, greedy, single start, N <= 112), register-file
 * resident tables (any single start, N <= 128) and shared-memory resident tables (multistart,
 * or anything the others cannot hold).  impl = 1 forces the last one, impl = 2 skips the
 * TSP-greedy specialisation (parity tests run all of them against the oracle); impl = 0
 * restores the automatic choice. */
int rl4co_set_rollout_impl(int impl);

/* Standalone decode step on given logits: process_logits + greedy / sampling / evaluate +
 * chosen log-prob (utils/decoding.py:141-191,327-397).  logits [B,N] f32 (not modified),
 * mask [B,N] bool.  Outputs: action [B] i64, logp [B] f32, logp_all [B,N] (may be NULL). */
int rl4co_decode_step(int64_t B, int32_t N, const float *logits, const uint8_t *mask,
                      int32_t mode, float temperature, float tanh_clipping,
                      int32_t mask_logits, const int64_t *action_in, uint64_t seed,
                      uint64_t offset, int64_t *action, float *logp, float *logp_all,
                      int32_t *status, rl4co_stream_t stream);

/* ------------------------------------------------------- reductions (K4) ----------- */

/* Local partial of the reward statistics that are all-reduced across ranks (NCCL sum):
 * stats[0] += sum(reward), stats[1] += sum(reward^2), stats[2] += n   (float64).
 * Feeds RolloutBaseline mean / t-test and the logged reward mean
 * (models/rl/reinforce/baselines.py:186-224, models/rl/common/base.py:224-240). */
int rl4co_reward_stats(int64_t n, const float *reward, double *stats, rl4co_stream_t stream);

/* POMO shared baseline + REINFORCE loss partials (baselines.py:57-61, reinforce.py:103-106).
 * reward / ll are [S*B] start-major.  baseline[b] = mean_s reward; advantage [S*B];
 * loss_partial[0] += sum((reward-baseline)*ll), loss_partial[1] += S*B  (float64);
 * max_reward[b] = max_s reward (pomo/model.py:117-141).  Outputs may be NULL. */
int rl4co_shared_baseline(int64_t B, int32_t S, const float *reward, const float *ll,
                          float *baseline, float *advantage, float *max_reward,
                          int64_t *best_start, double *loss_partial, rl4co_stream_t stream);

/* ------------------------------------------- POMO val / test epilogue ---- */

/* dihedral_8_augmentation, rl4co/data/transforms.py:15-37: locs [B,N,2] -> out [8*B,N,2], copy a of instance b at
 * row a*B + b, copies in the reference's order (x,y) (1-x,y) (x,1-y) (1-x,1-y) (y,x) (1-y,x) (y,1-x) (1-y,1-x).
 * Bit exact (one fp32 subtraction per coordinate). */
int rl4co_dihedral8_augment(int64_t B, int32_t N, const float *locs, float *out, rl4co_stream_t stream);

/* symmetric_transform, rl4co/data/transforms.py:50-71: xy [rows,N,2], phi [rows] -> out [rows,N,2]: rotation by phi about
 * (0.5, 0.5); phi > 2*pi additionally swaps the two coordinates.  (The caller draws phi, transforms.py:83-87.) */
int rl4co_symmetric_augment(int64_t rows, int32_t N, const float *xy, const float *phi, float *out,
                            rl4co_stream_t stream);

/* Best of starts and of augmentations with the tours that achieved them: POMO.shared_step's val / test branch
 * (rl4co/models/zoo/pomo/model.py:113-141), DecodingStrategy._select_best (rl4co/utils/decoding.py:399-407) and the
 * max / gather of the eval drivers (rl4co/tasks/eval.py:137-147,293-304).
 * reward [S*A*B] and actions [S*A*B,T] in the reference's batchify order for (n_aug, n_start): rollout
 * r = (s*A + a)*B + b.  A = 1 or S = 1 when that dimension is absent.  Outputs (each may be NULL):
 *   max_reward [B,A], best_start [B,A] (lowest index on ties, like torch.max), best_multistart_actions [B,A,T],
 *   max_aug_reward [B], best_aug [B], best_aug_actions [B,T] (tour of the best (copy, start) pair). */
int rl4co_pomo_best(int64_t B, int32_t A, int32_t S, int32_t T, const float *reward, const int64_t *actions,
                    float *max_reward, int64_t *best_start, int64_t *best_multistart_actions,
                    float *max_aug_reward, int64_t *best_aug, int64_t *best_aug_actions, rl4co_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RL4CO_SLAP_B200_H */
