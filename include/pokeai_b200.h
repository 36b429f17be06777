/* pokeai_b200 -- C ABI of the B200-native decision hot path.
 *
 * Everything select766/pokeai does between "an observation arrives" and "an action
 * index leaves" (plus the DQN minibatch update of the same network) is exposed here as
 * plain-C entry points: extern "C", raw pointers and sizes, integer status returns.
 * There are no torch / C++ types in any signature, so the reference's Python can bind
 * it with ctypes (INTEGRATION.md shows the stub) and any other host with its FFI.
 *
 * The reference has no native interface (it is pure Python + numpy + torch-CPU); each
 * entry point therefore cites the Python function(s) it replaces, relative to the
 * reference checkout:
 *
 *   pk_encode        pokeai/ai/generic_move_model/feature_extractor.py:40-52
 *                    (-> state_feature_extractor.py:92-197, choice_to_vec.py:40-61)
 *   pk_mlp_forward   pokeai/ai/generic_move_model/mlp_model.py:21-30 called from
 *                    agent.py:23-30 (_calc_q_vector / _calc_q_vector_batch)
 *   pk_select        agent.py:32-45 (_act_by_model / _act_random), agent_train.py:30-33
 *   pk_decide        agent_val.py:8-11 / agent_train.py:24-37: transform + forward + select
 *   pk_dqn_*         trainer.py:152-198 (_update, _target_update)
 *
 * Conventions
 *   - Return value 0 == PK_OK; anything else is an error and pk_last_error() (thread
 *     local) describes it.  No C++ exception crosses this boundary.
 *   - `stream` arguments are a cudaStream_t passed as void* (NULL = legacy default
 *     stream).  Kernels are enqueued, never synchronised, unless the name ends in _host.
 *   - Device buffers are owned by the caller.  The library never allocates per call and
 *     never frees caller memory; the only allocations are inside pk_session_create.
 *   - All multi-byte fields are little-endian; all float data is IEEE fp32.
 */
#ifndef POKEAI_B200_H
#define POKEAI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PK_ABI_VERSION 1

#if defined(__GNUC__)
#define PK_API __attribute__((visibility("default")))
#else
#define PK_API
#endif

#define PK_OK 0
#define PK_ERR_INVALID 1      /* bad argument (null pointer, size, alignment, unsupported shape) */
#define PK_ERR_CUDA 2         /* a CUDA runtime call failed; message holds cudaGetErrorString */
#define PK_ERR_NOT_INIT 3     /* pk_init has not been called on the current device */
#define PK_ERR_UNSUPPORTED 4  /* valid request that this build has no kernel for */

#define PK_MAX_ACTIONS 9   /* 4 moves + 5 switches (party of 6): feature_extractor.py:30-38 */
#define PK_MAX_MOVES 4
#define PK_N_STATE 50      /* state_feature_extractor.py:74-90 */
#define PK_N_TYPES 17      /* state_feature_extractor.py:9-27 */
#define PK_MAX_LAYERS 8
#define PK_OBS_BYTES 144

/* action flags */
#define PK_FLAG_SWITCH 1u
#define PK_FLAG_FORCE 2u

/* One legal action (PossibleAction, pokeai/ai/common.py:7-14) with every string already
 * replaced by its index in the reference's JSON tables. */
typedef struct pk_action {
  uint8_t flags;            /* PK_FLAG_SWITCH | PK_FLAG_FORCE */
  uint8_t species;          /* index into all_pokemons.json */
  uint8_t item;             /* index into all_items.json ('' is item 0) */
  uint8_t n_moves;          /* 1 for a move action, <=4 (all moves of the target) for a switch */
  uint16_t move[PK_MAX_MOVES]; /* indices into all_moves_with_hiddenpower_type.json */
} pk_action;

/* Packed observation: the fields of RLPolicyObservation that the encoder reads
 * (rl_policy_observation.py:7-10, battle_status.py:41-188).  Index 0 of each pair is the
 * friend side, 1 the opponent.  144 bytes, 16-byte aligned so that one battle is nine
 * 16-byte loads. */
typedef struct pk_obs {
  uint8_t remaining[2];     /* SideStatus.remaining_pokes */
  uint8_t total[2];         /* SideStatus.total_pokes (must be non-zero) */
  uint16_t hp_cur[2];       /* ActivePokeStatus.hp_current */
  uint16_t hp_max[2];       /* ActivePokeStatus.hp_max (must be non-zero) */
  uint8_t status[2];        /* 0 none, 1..6 = psn,tox,par,brn,slp,frz, 7 = fnt/other (encodes to zeros) */
  int8_t boost[2][7];       /* atk,def,spa,spd,spe,accuracy,evasion in -6..6 */
  uint8_t opp_species;      /* opponent active species (its types come from the table) */
  uint8_t weather;          /* 0 none, 1 SunnyDay, 2 RainDance, 3 Sandstorm */
  uint8_t n_act;            /* number of legal actions; mask = first n_act slots */
  uint8_t side_cond;        /* carried, never encoded (the reference does not encode it) */
  pk_action act[PK_MAX_ACTIONS];
  uint32_t user_tag;        /* opaque to the kernels */
} pk_obs;

/* Static tables (host pointers), copied to the device once by pk_init. */
typedef struct pk_tables {
  int32_t n_species;                 /* 251 */
  int32_t n_moves;                   /* 267 */
  int32_t n_items;                   /* 47  */
  const uint32_t* species_type_mask; /* [n_species], bit t set = type t of the 17-type list */
} pk_tables;

/* MLPModel parameters in the reference's own (torch state_dict) layout, device pointers:
 *   w[i]  = layers.{i}.weight  (n_channels, in_i, 1) row-major, in_0 = in_channels
 *   b[i]  = layers.{i}.bias    (n_channels,)
 *   w_out = output.weight      (1, n_channels, 1);  b_out = output.bias (1,)
 * mlp_model.py:6-19 (bn=False; eval-mode BatchNorm must be folded by the caller). */
typedef struct pk_mlp {
  int32_t n_layers;
  int32_t n_channels;
  int32_t in_channels;
  int32_t reserved;
  const float* w[PK_MAX_LAYERS];
  const float* b[PK_MAX_LAYERS];
  const float* w_out;
  const float* b_out;
} pk_mlp;

/* Action selection mode (agent.py:32-45, agent_train.py:30-33; softmax is a north-star
 * extension with no reference counterpart). */
#define PK_SELECT_ARGMAX 0   /* q[mask==0] = -inf; argmax, first maximum wins */
#define PK_SELECT_EPSGREEDY 1 /* u_explore[b] < eps ? k-th legal slot, k = floor(u_pick[b]*n_legal) : argmax */
#define PK_SELECT_SOFTMAX 2  /* inverse-CDF sample of softmax(q/temperature) over legal slots with u_pick[b] */

typedef struct pk_select_params {
  int32_t mode;
  float eps;                 /* PK_SELECT_EPSGREEDY */
  float temperature;         /* PK_SELECT_SOFTMAX */
  int32_t reserved;
  const float* u_explore;    /* [B] uniforms in [0,1) (device), EPSGREEDY only */
  const float* u_pick;       /* [B] uniforms in [0,1) (device), EPSGREEDY and SOFTMAX */
  float* probs;              /* optional [B*A] output of the softmax probabilities, or NULL */
} pk_select_params;

/* ---- library ------------------------------------------------------------------------ */
PK_API int pk_abi_version(void);
PK_API const char* pk_last_error(void);
/* Upload the tables to `device` (cudaSetDevice is called).  Idempotent. */
PK_API int pk_init(const pk_tables* tables, int device);
/* Number of feature channels C = 50 + 2 + n_species + n_moves + n_items (617). */
PK_API int pk_n_channels(void);
/* Debugging aid: when non-NULL, pk_encode / pk_decide record %globaltimer (ns) of lane 0 at 10 phase points of
 * the first battle of every warp into device_buffer[(cta * warps + warp) * 16 + phase].  NULL switches it off. */
PK_API int pk_debug_set_trace(unsigned long long* device_buffer);
/* Programmatic dependent launch for the device-pointer entry points pk_decide / pk_encode: OFF by default, opt-in per
 * process with pk_set_overlap_launches(1) (or PK_PDL=1).  When on, the part of the kernel that only reads the PACKED
 * RECORDS and fills shared memory may overlap the tail of the previous kernel of the stream; weights are read and all
 * outputs are written after griddepcontrol.wait, i.e. after that kernel has completed.  The caller who opts in asserts
 * that the packed records are NOT produced by the kernel launched immediately before on the same stream (records that
 * arrive by cudaMemcpy, or that were written earlier, are fine).  Worth 2 us per 4 096-state launch; bench.py opts in
 * (its records are uploaded once, before the timed region).  The host-buffer sessions (pk_session_*), whose records
 * always come from host memory, use it on their own (PK_PDL_SESSION=0 turns that off); the update's next-state
 * forwards never do. */
PK_API int pk_set_overlap_launches(int on);
/* Tuning / test aid: which pk_decide kernel runs for networks of <= 64 channels.  0 = one battle per warp,
 * 2 = two battles per warp (one per half-warp in the MLP), 1 (default) = paired only when features are not
 * materialised (the measured optimum, profiles/r01_pair_vs_single.txt).  Results are identical up to fp32
 * summation order. */
PK_API int pk_debug_set_pairing(int mode);
/* Tuning / test aid: pk_decide for 64-channel networks with one or two hidden layers (the default MLP 2 x 64) has a
 * second kernel (pk_decide_tc.cuh) in which the hidden layers of the eight battles of a warp group run as ONE
 * tcgen05.mma.kind::tf32 contraction with a 3-term hi/lo split (fp32-accurate: Q stays within the 1e-5 tolerance),
 * accumulators in TMEM, operand tiles staged by TMA from the prepared weights.  0 (default) = never, 1 = when features
 * are materialised, 2 = always; PK_DECIDE_TC in the environment.  Same results up to fp32 summation order, covered by
 * the same parity tests.  Measured on B200: 28.0 us against 22.9 us per 4 096-state launch -- 24
 * dependent K = 8 MMAs per layer are latency-bound at this size and the shared memory they need takes the L1 the
 * layer-0 gathers live in -- so it is off by default. */
PK_API int pk_debug_set_decide_tc(int mode);
/* Tuning / test aid for pk_mlp_forward (dense input): 1 (default; PK_DENSE_TC in the environment) runs 64-channel
 * networks with two or more layers on the tensor cores (pk_dense_tc.cuh: TMA-streamed blocks, operands in tensor
 * memory, every layer a tcgen05.mma.kind::tf32 contraction with a 3-term hi/lo split, fp32-accurate); 0 keeps the
 * FP32-pipe kernel that other shapes use.  Same results up to fp32 summation order; covered by the same parity tests. */
PK_API int pk_debug_set_dense_tc(int on);
/* Tuning / test aid: 1 runs the 64-channel hidden layers of pk_decide's one-battle-per-warp kernel on the warp-level
 * tensor path (mma.sync tf32 with a 3-term hi/lo split: fp32-accurate, Q stays within the 1e-5 tolerance); 0 (default)
 * keeps them on the FP32 pipe (FFMA2).  Measured on B200 (profiles/r01_hidden_tc.txt): the legacy mma.sync path is no
 * faster than FFMA2 here (4.2 vs 4.7 us per 16 battles, lost again in the longer prologue), so it is off by default. */
PK_API int pk_debug_set_hidden_tc(int on);
/* Number of SMs of the initialised device (grid sizing information for callers). */
PK_API int pk_sm_count(void);

/* ---- weights -------------------------------------------------------------------------
 * The kernels read a "prepared" copy of the network: every matrix transposed to
 * (in, out) with the output width padded to a multiple of 32, all tensors in one
 * contiguous fp32 buffer.  pk_mlp_prepared_floats gives its length; pk_mlp_prepare fills
 * it from the state_dict layout (one small kernel; call again whenever weights change). */
PK_API size_t pk_mlp_prepared_floats(int n_layers, int n_channels, int in_channels);
PK_API int pk_mlp_prepare(const pk_mlp* mlp, float* prepared, void* stream);

/* ---- the decision path -------------------------------------------------------------- */
/* packed[B] -> feat (B, C, A) fp32, mask (B, A) int32.  feat must be 4-byte aligned.
 * Replaces FeatureExtractor.transform, feature_extractor.py:40-52. */
PK_API int pk_encode(const pk_obs* packed, int B, int A, float* feat, int32_t* mask, void* stream);

/* Dense features (B, C, A) -> q (B, A).  Replaces MLPModel.forward as called by
 * Agent._calc_q_vector[_batch], agent.py:23-30.  C must equal in_channels. */
PK_API int pk_mlp_forward(const float* prepared, int n_layers, int n_channels, int in_channels,
                   const float* feat, int B, int A, float* q, void* stream);

/* q (B, A), mask (B, A) -> action (B,) int32.  Replaces agent.py:32-45. */
PK_API int pk_select(const float* q, const int32_t* mask, int B, int A, const pk_select_params* sel,
              int32_t* action, void* stream);

/* One fused pass: packed[B] -> [feat (B,C,A)] [mask (B,A)] q (B,A) action (B,).
 * feat and mask may be NULL (features are then never materialised).  Replaces
 * AgentVal.act / AgentTrain.act's transform + _act_by_model / _act_random for a batch. */
PK_API int pk_decide(const pk_obs* packed, int B, int A, const float* prepared, int n_layers, int n_channels,
              float* feat, int32_t* mask, float* q, int32_t* action, const pk_select_params* sel,
              void* stream);

/* ---- DQN update (trainer.py:152-198) --------------------------------------------------
 * Transitions are held packed (two pk_obs per transition) and re-encoded on the fly.
 * All pointers are device pointers.  Parameter / gradient / Adam-moment vectors use one
 * flat fp32 layout in state_dict order (layers.0.weight, layers.0.bias, ...,
 * output.weight, output.bias); pk_mlp_param_floats gives its length. */
typedef struct pk_batch {
  int32_t n;                 /* transitions in this (local) batch */
  int32_t n_global;          /* divisor of the Huber sum: the global batch size (== n on one GPU) */
  const pk_obs* state;       /* [n] */
  const pk_obs* next_state;  /* [n]; ignored where terminal */
  const int32_t* action;     /* [n] chosen slot */
  const float* reward;       /* [n] */
  const uint8_t* terminal;   /* [n] 1 = next_state is None */
} pk_batch;

PK_API size_t pk_mlp_param_floats(int n_layers, int n_channels, int in_channels);
/* scratch floats pk_dqn_grad needs for batches of up to n_max transitions (next-state Q values, per-CTA partial
 * gradients).  Call after pk_init. */
PK_API size_t pk_dqn_scratch_floats(int n_layers, int n_channels, int in_channels, int n_max);
/* Forwards over next_state (online argmax + target Q, or target max when double_dqn == 0), Q(s)[a], TD target,
 * Huber, backward.  Writes grad (flat, pre-clamp, = d/dparams of sum_i huber_i / n_global) and
 * *loss = sum_i huber_i / n_global (device pointer, may be NULL).  prepared / prepared_target are the prepared
 * copies of the online and target networks.  Deterministic (fixed summation order) when the one-hot
 * accumulator fits in shared memory (<= 64 channels, <= 3 layers). */
PK_API int pk_dqn_grad(const float* prepared, const float* prepared_target, int n_layers, int n_channels, int A,
                const pk_batch* batch, float gamma, int double_dqn, float* grad, float* loss,
                float* scratch, void* stream);
/* pk_dqn_grad for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has no counterpart in the reference.
 * Same arguments, arithmetic and results as pk_dqn_grad minus `scratch`: the scratch is a file-scope device
 * array inside the library of PK_DQN_STATIC_SCRATCH_FLOATS floats (32 MiB, held by every GPU that loads the library
 * until the process exits), and the call performs no cudaMalloc / cudaFree and uses no allocating library routine.
 * Limit: pk_dqn_scratch_floats(n_layers, n_channels, in_channels, batch->n) must not exceed
 * PK_DQN_STATIC_SCRATCH_FLOATS (the default 2 x 64 network on a 148-SM GPU: up to ~185 000 transitions); larger
 * requests fail with PK_ERR_INVALID and a message, before anything is launched.  Every call on a device shares the
 * one array: calls whose kernels may overlap on the GPU (different streams, no ordering between them) RACE -- keep
 * them on one stream or order them with events.  The array keeps the previous call's contents; nothing in it needs
 * to start at zero (each call writes what it reads). */
#define PK_DQN_STATIC_SCRATCH_FLOATS (8u << 20)
PK_API int pk_dqn_grad_static(const float* prepared, const float* prepared_target, int n_layers, int n_channels, int A,
                       const pk_batch* batch, float gamma, int double_dqn, float* grad, float* loss, void* stream);
/* clamp(grad, -1, 1) then Adam (torch defaults, no weight decay) on the flat parameter
 * vector; also writes the new values into `prepared` (may be NULL).  step is 1-based. */
PK_API int pk_clamp_adam(float* params, const float* grad, float* exp_avg, float* exp_avg_sq, int n_layers,
                  int n_channels, int in_channels, float lr, float beta1, float beta2, float eps, int step,
                  float* prepared, void* stream);

/* Graph-replayable form of the optimiser step.  A captured CUDA graph freezes launch arguments, and two of them change
 * every step: Adam's step count (bias corrections) and the epoch of the peer exchange.  With a control block in DEVICE
 * memory they advance on the device instead: capture [pk_update_ctl_advance, pk_dqn_grad, pk_clamp_adam_ctl or
 * pk_peer_clamp_adam_ctl] once (for the peer route: once per gradient-buffer parity, replayed alternately) and replay.
 * Initialise {epoch = 0, step = 0, lr, beta1, beta2, eps} with a copy from the host; same arithmetic as pk_clamp_adam
 * (the bias corrections are computed in double, by one thread per block). */
typedef struct pk_update_ctl {
  uint32_t epoch; /* peer exchange epoch of the step being taken (after pk_update_ctl_advance: 1, 2, 3, ...) */
  int32_t step;   /* Adam step count of the step being taken (1-based) */
  float lr, beta1, beta2, eps;
  uint32_t reserved[2];
} pk_update_ctl;
PK_API int pk_update_ctl_advance(pk_update_ctl* ctl_dev, void* stream); /* epoch += 1, step += 1 */
PK_API int pk_clamp_adam_ctl(float* params, const float* grad, float* exp_avg, float* exp_avg_sq, int n_layers,
                      int n_channels, int in_channels, const pk_update_ctl* ctl_dev, float* prepared, void* stream);

/* Data-parallel form of pk_clamp_adam: gradient sum over ranks + clamp + Adam in ONE kernel over NVLink peer memory
 * (replaces dist.all_reduce(grad) + the clamp / optimizer.step of trainer.py:193-195 at world size > 1).
 * peer_grads[r] (r < world): device pointer, valid in THIS process, to rank r's gradient buffer for this epoch:
 * n_params floats (pre-clamp, as pk_dqn_grad writes them) followed by one float, the rank's loss share.
 * peer_flags[r]: rank r's flag words (>= world uint32, zero before the first call), mapped the same way.
 * epoch: 1, 2, 3, ... the same on every rank; buffers must alternate between two sets by epoch parity (a rank may
 * overwrite the set of epoch e only after its own call for epoch e + 1).  Every rank of the job must make the call.
 * grad_sum (n_params floats, may be NULL) receives the reduced pre-clamp gradient; loss_status (2 floats, may be
 * NULL): [0] = sum of the loss shares, [1] = 1.0 if a peer did not arrive within the time limit (30 s;
 * PK_PEER_TIMEOUT_MS in the environment).  A thread block that gives up returns WITHOUT writing params, exp_avg,
 * exp_avg_sq or prepared, so a timed-out step leaves the state as it was (block-granular: the blocks of one launch
 * start within microseconds of each other and wait for the same flags); the caller must treat the step as not taken.
 * The sum runs in rank order on every rank: identical bits everywhere. */
#define PK_MAX_PEERS 8
PK_API int pk_peer_clamp_adam(float* params, const float* const* peer_grads, uint32_t* const* peer_flags, int rank,
                       int world, uint32_t epoch, float* exp_avg, float* exp_avg_sq, int n_layers, int n_channels,
                       int in_channels, float lr, float beta1, float beta2, float eps, int step, float* prepared,
                       float* grad_sum, float* loss_status, void* stream);
/* Sum of a LARGE gradient vector over the ranks of one NVSwitch box through peer memory, in place: the wide network's
 * 55 MB gradient (replaces dist.all_reduce(grad) of trainer.py:192-195 at world size > 1; clamp + Adam follow with
 * pk_clamp_adam).  Buffers, flags, epochs and the parity rule are pk_peer_clamp_adam's (flag words 0..7 "written",
 * 8..15 "reduced", 16 a counter: >= 17 zeroed uint32 per rank).  Rank r sums slice r of the vector over all ranks in rank
 * order and the ranks then copy each other's slices: afterwards every rank's buffer of this epoch holds the same bits.
 * n_floats (a multiple of 4) covers everything that is to be summed (the loss share behind the gradient included).
 * status (2 floats, may be NULL): [1] = 1.0 if a peer did not arrive within PK_PEER_TIMEOUT_MS -- the exchange object
 * is then unusable.  Every rank must make the call, with the same grid (the library uses one CTA per SM). */
PK_API int pk_peer_allreduce(float* const* peer_grads, uint32_t* const* peer_flags, int rank, int world, uint32_t epoch,
                      size_t n_floats, float* status, void* stream);
/* The same with epoch, step and the Adam constants read from a device control block (see pk_update_ctl): peer_grads
 * must be the buffer set of the parity of the epoch the graph will run at. */
PK_API int pk_peer_clamp_adam_ctl(float* params, const float* const* peer_grads, uint32_t* const* peer_flags, int rank,
                           int world, const pk_update_ctl* ctl_dev, float* exp_avg, float* exp_avg_sq, int n_layers,
                           int n_channels, int in_channels, float* prepared, float* grad_sum, float* loss_status,
                           void* stream);

/* ---- train-mode BatchNorm update: MLPModel(bn=True) under Trainer._update -----------------------------------------
 * Replaces torch autograd through Conv1d(k=1, bias=False) -> BatchNorm1d (training) -> ReLU, n_layers times, then
 * Conv1d to one channel (pokeai/ai/generic_move_model/mlp_model.py:12-15,21-30) as the reference's Trainer runs it:
 * Trainer.model and Trainer.target_model are never put in eval mode (trainer.py:41-43), so every forward of the update
 * (trainer.py:163,174-175,181-182) normalises with the statistics of ITS batch -- per channel over batch x action columns,
 * biased variance, eps 1e-5 -- and moves the running statistics (momentum 0.1, unbiased variance).
 * `params`: the flat vector in MLPModel.parameters() order (layers.*.weight, bn_layers.*.weight / .bias per layer,
 * output.weight, output.bias; pk_bn_param_floats floats).  `running`: [n_layers][2][n_channels] (running_mean,
 * running_var), updated in place; NULL leaves it alone.  `feat`: dense (B, in_channels, A) features (pk_encode).
 * `ws`: pk_bn_workspace_floats(...) floats, 16-byte aligned; it keeps every layer's pre-norm and post-ReLU activations
 * for pk_bn_backward.  q: (B, A).  Plain FP32 kernels (every recorded run of the reference has bn = false). */
PK_API size_t pk_bn_param_floats(int n_layers, int n_channels, int in_channels);
PK_API size_t pk_bn_workspace_floats(int n_layers, int n_channels, int B, int A);
PK_API size_t pk_bn_backward_scratch_floats(int n_layers, int n_channels, int B, int A);
PK_API int pk_bn_forward_train(const float* params, float* running, int n_layers, int n_channels, int in_channels,
                        const float* feat, int B, int A, float* ws, float* q, void* stream);
/* The same forward straight from PACKED records (what Trainer._update_bn uses): layer 0 is gathered -- a feature column
 * is the battle's 50 state values plus at most eight one-hot rows (state_feature_extractor.py:92-197,
 * choice_to_vec.py:40-61), so 58 rows of W0^T are touched instead of 617 and no dense feature block exists.  in_channels
 * must equal pk_n_channels().  ws: pk_bn_packed_workspace_floats(...) floats.  mask (B, A) int32 may be NULL. */
PK_API size_t pk_bn_packed_workspace_floats(int n_layers, int n_channels, int in_channels, int B, int A);
PK_API int pk_bn_forward_train_packed(const float* params, float* running, int n_layers, int n_channels, int in_channels,
                               const pk_obs* packed, int B, int A, float* ws, float* q, int32_t* mask, void* stream);
/* Double-DQN target, smooth-L1 loss (mean over n_global) and d loss / d q (trainer.py:163-187).  slot[i] = row of
 * transition i's next state in the COMPACTED non-final arrays q_next_online / q_next_target / mask_next (the reference
 * forwards only the non-final next states, trainer.py:157-158), -1 when terminal.  dq: (n, A), zero except at action[i].
 * q_next_online may be NULL when double_dqn == 0. */
PK_API int pk_bn_dqn_dq(const float* q, const int32_t* action, const float* reward, const int32_t* slot,
                 const float* q_next_online, const float* q_next_target, const int32_t* mask_next, int n, int A,
                 int n_global, float gamma, int double_dqn, float* dq, float* loss, void* stream);
/* loss.backward() (trainer.py:192) through the network pk_bn_forward_train just ran on `feat` with `ws`: the flat gradient
 * in `params` order (not clamped).  scratch: pk_bn_backward_scratch_floats(...) floats, 16-byte aligned. */
PK_API int pk_bn_backward(const float* params, int n_layers, int n_channels, int in_channels, const float* feat, int B,
                   int A, const float* ws, const float* dq, float* scratch, float* grad, void* stream);
/* pk_bn_backward after pk_bn_forward_train_packed on the same records: the one-hot rows of dW0 are accumulated with
 * coalesced float atomics (their order, hence the last bit, varies from run to run), the 50 state rows by one small GEMM
 * over the batch; weights of features that never occur get a gradient of exactly 0.  scratch:
 * pk_bn_packed_backward_scratch_floats(...) floats. */
PK_API size_t pk_bn_packed_backward_scratch_floats(int n_layers, int n_channels, int in_channels, int B, int A);
PK_API int pk_bn_backward_packed(const float* params, int n_layers, int n_channels, int in_channels, const pk_obs* packed,
                          int B, int A, const float* ws, const float* dq, float* scratch, float* grad, void* stream);
/* pk_clamp_adam over a plain vector of n floats (no prepared copies): clamp to [-1, 1], Adam (trainer.py:193-195). */
PK_API int pk_clamp_adam_flat(float* params, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr,
                       float beta1, float beta2, float eps, int step, void* stream);

/* ---- host-buffer session (for non-torch callers, and the end-to-end measurement) ------
 * Owns device buffers for up to max_batch observations and one stream.  *_host calls copy
 * packed observations host->device, run pk_decide, copy q / action back and synchronise. */
typedef struct pk_session pk_session;
PK_API int pk_session_create(pk_session** out, int max_batch, int A, int materialise_features);
PK_API void pk_session_destroy(pk_session* s);
/* weights given as HOST pointers in state_dict layout */
PK_API int pk_session_load_weights(pk_session* s, const pk_mlp* host_mlp);
/* use caller-owned prepared weights on the device instead (no copy; must outlive the calls) */
PK_API int pk_session_use_prepared(pk_session* s, const float* prepared_dev, int n_layers, int n_channels);
PK_API int pk_session_decide_host(pk_session* s, const pk_obs* packed_host, int B, int mode, float eps,
                           float temperature, const float* u_explore_host, const float* u_pick_host,
                           float* q_host, int32_t* action_host);
/* The same without the final synchronise: work is queued on the session's stream and the call returns; q_host /
 * action_host are valid after pk_session_wait.  Two sessions used alternately keep two batches in flight (the
 * lock-step driver of SURVEY.md 8 f2: one group of battles is stepped on the host while the other is decided).
 * Host buffers must stay untouched between submit and wait; pageable buffers are accepted but make the copies
 * synchronous. */
PK_API int pk_session_submit_host(pk_session* s, const pk_obs* packed_host, int B, int mode, float eps,
                           float temperature, const float* u_explore_host, const float* u_pick_host,
                           float* q_host, int32_t* action_host);
PK_API int pk_session_wait(pk_session* s);
/* packed_host[B] -> feat_host (B, C, A), mask_host (B, A): host-buffer FeatureExtractor.transform
 * (feature_extractor.py:40-52).  Needs a session created with materialise_features != 0. */
PK_API int pk_session_encode_host(pk_session* s, const pk_obs* packed_host, int B, float* feat_host,
                           int32_t* mask_host);
/* copy the features / masks the last pk_session_decide_host materialised back to the host */
PK_API int pk_session_fetch(pk_session* s, int B, float* feat_host, int32_t* mask_host);
/* device pointers of the session's last results (valid until the next call) */
PK_API const float* pk_session_features(const pk_session* s);
PK_API const int32_t* pk_session_mask(const pk_session* s);

/* ---- wide network: the tensor-core path (BASELINE configs[4], 4 x 2048) -------------------------------
 * Same MLPModel (mlp_model.py:5-30) evaluated as bf16 x bf16 -> fp32 tcgen05 GEMMs with fused bias + ReLU.
 * n_channels must be a multiple of 256.  Tolerance: bf16 operands, fp32 accumulation (|dq| <= 2e-2 on O(1)
 * logits against the fp32 reference). */
/* Y (M,N) = act(A (M,K) . W (N,K)^T + bias): A, W bf16 row-major with K contiguous (W is exactly a state_dict
 * weight cast to bf16); out_bf16 and/or out_f32 row-major; K % 64 == 0, N % 256 == 0; relu != 0 applies max(.,0). */
PK_API int pk_gemm_bf16(const void* a_bf16, const void* w_bf16, const float* bias, void* out_bf16, float* out_f32,
                        int M, int N, int K, int relu, void* stream);
/* Bytes of the prepared wide network (bf16 weights, K of layer 0 padded 617 -> 640; fp32 biases and head); 0 if the
 * shape is unsupported. */
PK_API size_t pk_wide_prepared_bytes(int n_layers, int n_channels, int in_channels);
PK_API int pk_wide_prepare(const pk_mlp* m, void* wide_prepared, void* stream);
/* Bytes of caller-owned scratch for pk_wide_forward at batch B: the bf16 operand X (B*A, 640) and two
 * (B*A, n_channels) bf16 activation buffers. */
PK_API size_t pk_wide_workspace_bytes(int n_layers, int n_channels, int in_channels, int B, int A);
/* transform + forward + select for the wide network (agent_val.py:8-11 with a 4 x 2048 MLPModel): mask (B,A) int32 or
 * NULL, q (B,A) fp32, action (B,) int32 or NULL (needs mask). */
PK_API int pk_wide_forward(const pk_obs* packed, int B, int A, const void* wide_prepared, int n_layers, int n_channels,
                           void* workspace, int32_t* mask, float* q, int32_t* action, const pk_select_params* sel,
                           void* stream);

/* Y (M,N) fp32 = A^T . W with A (K,M) and W (K,N) bf16 row-major: the contraction index is the ROW of both operands
 * (weight gradient dW = dZ^T H).  M % 8 == 0, N % 32 == 0; any K (zero-filled tail). */
PK_API int pk_gemm_bf16_tn(const void* a_km_bf16, const void* w_kn_bf16, float* out_f32, int M, int N, int K, void* stream);
/* Forward + backward of the wide network: loss.backward() of Trainer._update (trainer.py:192) for MLPModel at 4 x 2048.
 * pk_wide_train_forward keeps every layer's bf16 activations in the workspace and returns q (B,A); the caller
 * forms dq = d loss / d q (B*A fp32, zero for columns that carry no loss -- trainer.py:163 gathers the chosen one);
 * pk_wide_train_backward returns the flat fp32 gradient in state_dict order (pk_mlp_param_floats floats).
 * Gradients are bit-reproducible (no atomics). */
PK_API size_t pk_wide_train_workspace_bytes(int n_layers, int n_channels, int in_channels, int B, int A);
/* col_select == NULL: every action column is a row (q is (B,A), dq is (B*A,)).  col_select (B,) int32: only that
 * column of each battle is evaluated (q and dq are (B,)) -- the gather(1, action) of trainer.py:163, 9x less work. */
PK_API int pk_wide_train_forward(const pk_obs* packed, int B, int A, const int32_t* col_select, const void* wide_prepared,
                                 int n_layers, int n_channels, void* workspace, int32_t* mask, float* q, void* stream);
PK_API int pk_wide_train_backward(const void* wide_prepared, int n_layers, int n_channels, int B, int A, int chosen_only,
                                  void* workspace, const float* dq, float* grad, void* stream);
/* The same; layer_events: n_layers CUDA events (cudaEvent_t) or NULL.  layer_events[l] is recorded on `stream` once the
 * gradient slices of layers >= l and of the head are complete: the caller can exchange them (NCCL, one bucket per
 * layer, last layer first) while the backward of the earlier layers runs. */
PK_API int pk_wide_train_backward_ev(const void* wprep, int n_layers, int n_channels, int B, int A, int chosen_only,
                              void* workspace, const float* dq, float* grad, void* const* layer_events,
                              void* stream);
/* Double-DQN target y = r + gamma * Q_target(s', argmax_legal Q_online(s')) (0 for terminal), Huber loss (mean over
 * n_global) and dq = clamp(q_sa - y, -1, 1) / n_global: trainer.py:163-187 for the wide path.  Fixed-order sums. */
PK_API int pk_wide_dqn_dq(const float* q_sa, const float* q_next_online, const float* q_next_target,
                          const int32_t* mask_next, const float* reward, const uint8_t* terminal, int n, int A,
                          int n_global, float gamma, int double_dqn, float* dq, float* loss, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* POKEAI_B200_H */
