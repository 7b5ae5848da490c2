/* tsm_b200.h -- C ABI of libtsm_b200.so: the B200 (sm_100a) kernels behind the tianshou_marl
 * on-policy multi-agent hot path (rollout collection -> replay-buffer add/sample -> GAE ->
 * PPO minibatch update).
 *
 * The reference (eric-downes/tianshou_marl) is pure Python and has no FFI; the entry points
 * below are what a ctypes binding for this path binds.  Each one cites
 * the reference function it replaces (file:line relative to the reference checkout).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no torch / C++ types.
 *  - Every pointer is a DEVICE pointer unless its name starts with h_ .
 *  - The caller owns every buffer; the library allocates nothing.  Scratch space is passed in
 *    explicitly and sized by the matching *_workspace_bytes() query.
 *  - Every call is asynchronous on `stream` (a cudaStream_t passed as void*; NULL = default
 *    stream).  Calls on one stream are ordered; a handle/struct is not thread-safe.
 *  - Return 0 on success, a negative tsm_status otherwise; tsm_last_error() gives the message
 *    of the calling thread's last failure.  No C++ exception crosses the boundary.
 *  - Row flattening: row r = env * n_agents + agent.  One sub-buffer per row.
 *    Indices exchanged with the caller are numbered ENV-MAJOR like the reference
 *    (flat = r * size + t, vecbuf.py:33-37 / manager.py:26-51); storage inside the device
 *    buffer is TIME-MAJOR ([t][r]) so that one vector step is one contiguous write.
 */
#ifndef TSM_B200_H
#define TSM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TSM_ABI_VERSION 1

typedef void* tsm_stream_t; /* cudaStream_t */

typedef enum {
    TSM_OK = 0,
    TSM_ERR_INVALID = -1,   /* bad argument (maps to ValueError) */
    TSM_ERR_CUDA = -2,      /* CUDA runtime error (maps to RuntimeError) */
    TSM_ERR_WORKSPACE = -3, /* workspace too small */
    TSM_ERR_UNSUPPORTED = -4
} tsm_status;

int tsm_abi_version(void);
const char* tsm_last_error(void);
/* sm count / compute capability of `device`; any out pointer may be NULL. */
int tsm_device_info(int device, int* h_sm_count, int* h_cc_major, int* h_cc_minor);

/* =========================================================================================
 * K3  GAE / discounted returns
 * replaces  Algorithm.compute_episodic_return   tianshou/algorithm/algorithm_base.py:651-717
 *           Algorithm.value_mask                tianshou/algorithm/algorithm_base.py:631-649
 *           _gae (numba)                        tianshou/algorithm/algorithm_base.py:1079-1134
 *           episode_mc_return_to_go (numba)     tianshou/algorithm/algorithm_base.py:1137-1151
 *
 *   v_next_i <- v_next_i * !terminated_i                      (value mask, :709)
 *   end_i     = terminated_i | truncated_i | force_end_i | ((i+1) % seg_len == 0 if seg_len>0)
 *   delta_i   = rew_i + gamma * v_next_i - v_s_i              (float64, :1128)
 *   g_i       = delta_i + (1-end_i) * gamma*lam * g_{i+1}     (float64, :1129-1133), g_n = 0
 *   adv_i = g_i ; ret_i = g_i + v_s_i                         (:715)
 * The recurrence runs over the whole flat vector exactly like the reference; segment breaks
 * come only from `end`.  State is carried in float64; I/O dtypes are selected by the flags.
 * ========================================================================================= */

/* Flat 1-D vector in the reference's batch order (one-pass segmented reverse scan with
 * decoupled look-back).  io_is_f64 selects the dtype of v_s, v_next, rew, adv_out and ret_out:
 * 0 = float32 (device-resident rollouts), 1 = float64 (the reference's numpy path, where numba
 * promotes everything to float64).  v_s may be NULL (treated as 0; MC return-to-go), v_next
 * may be NULL (0).  force_end may be NULL.  seg_len <= 0 disables the periodic break. */
size_t tsm_gae_flat_workspace_bytes(int64_t n);
int tsm_gae_flat(const void* v_s, const void* v_next, const void* rew, int io_is_f64,
                 const uint8_t* terminated, const uint8_t* truncated, const uint8_t* force_end,
                 int64_t n, int64_t seg_len, double gamma, double lam,
                 void* adv_out, void* ret_out,
                 void* workspace, size_t workspace_bytes, tsm_stream_t stream);

/* tsm_gae_flat without a workspace argument and without any allocation: the scratch (tile states) is a
 * file-scope device array inside the library.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  Limit:
 * n <= TSM_GAE_FLAT_STATIC_MAX_N elements (the scratch is 64 + ceil(n_max / 1024) * 32 bytes = 524 352 bytes, held by every
 * GPU that loads the library until exit); larger n returns TSM_ERR_INVALID with a message naming the limit.  All calls on
 * one GPU share the array: it is re-initialised on `stream` at the start of every call, and calls that overlap on the GPU
 * (different streams) race -- serialise them.  Same arguments, arithmetic and results as tsm_gae_flat otherwise. */
#define TSM_GAE_FLAT_STATIC_MAX_N 16777216
int64_t tsm_gae_flat_static_max_n(void);
int tsm_gae_flat_static(const void* v_s, const void* v_next, const void* rew, int io_is_f64,
                        const uint8_t* terminated, const uint8_t* truncated, const uint8_t* force_end,
                        int64_t n, int64_t seg_len, double gamma, double lam,
                        void* adv_out, void* ret_out, tsm_stream_t stream);

/* Time-major device-buffer layout: arrays are [size][n_rows]; row r holds `len` valid slots
 * starting at ring slot `start` (time order (start + j) % size, j < len); the last one is
 * forced to `end` (= buffer.unfinished_index(), algorithm_base.py:712-713).  One thread per
 * row, serial in t like the reference, coalesced across rows.  Outputs are [size][n_rows]. */
int tsm_gae_tm(const float* v_s, const float* v_next, const float* rew,
               const uint8_t* terminated, const uint8_t* truncated,
               int64_t size, int64_t n_rows, int64_t start, int64_t len,
               double gamma, double lam, float* adv_out, float* ret_out, tsm_stream_t stream);

/* tsm_gae_tm with return scaling (A2C._add_returns_and_advantages with return_scaling=True, a2c.py:132-146): v_s and v_next
 * are multiplied by v_scale = sqrt(ret_rms.var + eps) on load, ret_out = (unnormalised return - ret_shift) / v_scale (ret_shift
 * = 0 for return_scaling, = ret_rms.mean for Reinforce's return_standardization, reinforce.py:305-309), adv_out is the
 * (unnormalised) advantage, and ret_sums[0..1] (device doubles, zeroed by the call) receive the sum and the sum of squares
 * of the unnormalised returns over all valid slots -- the batch moments RunningMeanStd.update needs
 * (tianshou/utils/statistics.py:96-114). */
int tsm_gae_tm_scaled(const float* v_s, const float* v_next, const float* rew,
                      const uint8_t* terminated, const uint8_t* truncated,
                      int64_t size, int64_t n_rows, int64_t start, int64_t len,
                      double gamma, double lam, double v_scale, double ret_shift, float* adv_out, float* ret_out,
                      double* ret_sums, tsm_stream_t stream);

/* critic(obs_next) from the values the policy forward stored at collect time (time-major [size][n_rows], valid slots
 * (start + j) % size, j < len): v_next[slot][row] = value[next slot][row] where obs_next IS the next stored obs (row not
 * done at that slot and not the last collected slot); the other (slot * n_rows + row) positions are appended to
 * need_list (capacity len * n_rows, any order) and counted in *need_count (device int64) -- evaluate the critic on
 * obs_next for exactly those rows (tsm_net_forward with rows = need_list, scatter = 1).
 * replaces  the second critic pass of A2C._add_returns_and_advantages  tianshou/algorithm/modelfree/a2c.py:120-127 */
int tsm_values_next_prepare(const float* value, const uint8_t* done, int64_t size, int64_t n_rows, int64_t start, int64_t len,
                            float* v_next, int64_t* need_list, int64_t* need_count, tsm_stream_t stream);

/* =========================================================================================
 * K2  Device ring buffer (VectorReplayBuffer)
 * replaces  ReplayBufferManager.add             tianshou/data/buffer/manager.py:131-193
 *           ReplayBuffer._update_state_pre_add  tianshou/data/buffer/buffer_base.py:355-410
 *           ReplayBufferManager.sample_indices  tianshou/data/buffer/manager.py:195-229
 *           ReplayBuffer.sample_indices         tianshou/data/buffer/buffer_base.py:495-531
 *           _prev_index / _next_index (numba)   tianshou/data/buffer/manager.py:306-358
 *           ReplayBufferManager.unfinished_index tianshou/data/buffer/manager.py:85-91
 *           ReplayBufferManager.reset           tianshou/data/buffer/manager.py:70-75
 *           ReplayBuffer.__getitem__ (gather)   tianshou/data/buffer/buffer_base.py:591-635
 * All integer results are bit-exact with the reference.
 * ========================================================================================= */
typedef struct tsm_buf {
    int64_t n_bufs;  /* R sub-buffers (rows) */
    int64_t size;    /* slots per sub-buffer = ceil(total_size / n_bufs), vecbuf.py:35 */
    int64_t obs_dim; /* O (flattened) */
    /* per-sub-buffer counters, length n_bufs (buffer_base.py:103-106, manager.py:48-50) */
    int64_t* ins;        /* _insertion_idx */
    int64_t* sz;         /* _size  (== manager._lengths) */
    int64_t* last_index; /* manager.last_index, env-major flat */
    int64_t* ep_len;     /* _ep_len */
    int64_t* ep_start;   /* _ep_start_idx (local to the sub-buffer) */
    double* ep_ret;      /* _ep_return */
    /* payload, time-major: scalars [size][n_bufs], obs [size][n_bufs][obs_dim] */
    float* obs;
    float* obs_next;
    int32_t* act;
    float* rew;
    uint8_t* terminated;
    uint8_t* truncated;
    uint8_t* done;
    float* logp;  /* optional policy output stored per step (may be NULL) */
    float* value; /* optional critic output stored per step (may be NULL) */
} tsm_buf;

/* zero the counters (keep_statistics != 0 keeps ep_ret / ep_len, buffer_base.py:292-298). */
int tsm_buf_reset(const tsm_buf* buf, int keep_statistics, tsm_stream_t stream);

/* add n rows. buffer_ids NULL = arange(n_bufs) (then n must equal n_bufs).  Payload inputs are
 * [n] / [n][obs_dim] row-major; obs_next / logp / value may be NULL (then not stored).
 * Outputs (each may be NULL): idx_out = env-major flat slot written, ep_ret_out / ep_len_out =
 * episode return / length if the row finished an episode else 0, ep_start_out = env-major
 * flat start slot of the row's current episode.  buffer_ids must be unique within one call. */
int tsm_buf_add(const tsm_buf* buf, const float* obs, const int32_t* act, const float* rew,
                const uint8_t* terminated, const uint8_t* truncated, const float* obs_next,
                const float* logp, const float* value,
                const int64_t* buffer_ids, int64_t n,
                int64_t* idx_out, double* ep_ret_out, int64_t* ep_len_out, int64_t* ep_start_out,
                tsm_stream_t stream);

/* sample_indices(0): every valid index, sub-buffer by sub-buffer, each in time order
 * ([ins..sz) ++ [0..ins)).  `out` must hold n_bufs*size entries; the number written is stored
 * to *count_out (device int64).  workspace: tsm_buf_indices_workspace_bytes(n_bufs). */
size_t tsm_buf_indices_workspace_bytes(int64_t n_bufs);
int tsm_buf_indices_all(const tsm_buf* buf, int64_t* out, int64_t* count_out,
                        void* workspace, size_t workspace_bytes, tsm_stream_t stream);

/* prev / next for n env-major indices (taken modulo maxsize like manager.py:313,340). */
int tsm_buf_prev(const tsm_buf* buf, const int64_t* idx, int64_t n, int64_t* out, tsm_stream_t stream);
int tsm_buf_next(const tsm_buf* buf, const int64_t* idx, int64_t n, int64_t* out, tsm_stream_t stream);

/* unfinished_index(): last written slot of every sub-buffer whose done flag is false, in
 * sub-buffer order.  out holds up to n_bufs entries; *count_out (device) = number written. */
int tsm_buf_unfinished(const tsm_buf* buf, int64_t* out, int64_t* count_out,
                       void* workspace, size_t workspace_bytes, tsm_stream_t stream);

/* gather rows by env-major index into contiguous [n] / [n][obs_dim] outputs (any out may be
 * NULL).  `extra_in` are optional extra time-major float arrays ([size][n_bufs], e.g. adv,
 * returns) gathered into extra_out[k]. */
int tsm_buf_gather(const tsm_buf* buf, const int64_t* idx, int64_t n,
                   float* obs_out, float* obs_next_out, int32_t* act_out, float* rew_out,
                   uint8_t* term_out, uint8_t* trunc_out, uint8_t* done_out,
                   float* logp_out, float* value_out,
                   const float* const* h_extra_in, float* const* h_extra_out, int n_extra,
                   tsm_stream_t stream);

/* Episode statistics of one vector step (collector.py:983-1024, CollectStats :116-254): rows with
 * ep_len > 0 (the outputs of tsm_buf_add) finished an episode.  acc (device double[9], caller
 * initialises to {0,0,0,+inf,-inf,0,0,+inf,-inf}) accumulates {count, sum ret, sum ret^2, min ret,
 * max ret, sum len, sum len^2, min len, max len}.  Optionally appends (ret, len) of each episode
 * to list_ret / list_len through the device counter list_count (order within a step is not
 * defined; entries beyond list_capacity are counted but dropped). */
int tsm_episode_stats(const double* ep_ret, const int64_t* ep_len, int64_t n, double* acc, double* list_ret,
                      int64_t* list_len, int64_t* list_count, int64_t list_capacity, tsm_stream_t stream);

/* =========================================================================================
 * K1  Batched MPE (simple_spread / simple_tag) step, reset, observe, reward
 * replaces  BaseVectorEnv.step / reset          tianshou/env/venvs.py:237-322, :195-235
 *           DummyEnvWorker.send                 tianshou/env/worker/dummy.py:37-41
 *           (wrappers) PettingZooEnv.step/reset tianshou/env/pettingzoo_env.py:38-120
 *                      EnhancedPettingZooEnv    tianshou/env/enhanced_pettingzoo_env.py:90-222
 *           and the third-party simulator they drive: pettingzoo 1.24.2
 *           mpe/_mpe_utils/core.py (World.step), _mpe_utils/simple_env.py, scenarios/
 *           simple_spread.py, simple_tag.py  (not in the reference tree; parity unpinned).
 * One group of G lanes (G = 4..32, chosen from n_agents) simulates one environment; lanes
 * exchange positions with warp shuffles.  State is SoA per field, float2-vectorised:
 * pos/vel are [n_envs * n_agents] (row order), lm is [n_envs * n_landmarks].
 * ========================================================================================= */
#define TSM_MPE_SPREAD 0
#define TSM_MPE_TAG 1
#define TSM_MPE_ADVERSARY 2 /* simple_adversary_v3: 1 adversary + N good agents, N landmarks one of which is the goal; nothing
                               collides; obs = [goal - p (good agents only), landmarks - p, others - p], the adversary zero-padded */
typedef struct tsm_mpe {
    int32_t scenario;      /* TSM_MPE_SPREAD / TSM_MPE_TAG / TSM_MPE_ADVERSARY */
    int32_t n_agents;      /* A (tag: adversaries first, then good agents) */
    int32_t n_landmarks;   /* L */
    int32_t n_adversaries; /* tag only */
    int32_t obs_dim;       /* spread 6A; tag 4 + 2L + 2(A-1) + 2*n_good (good agents zero-padded); adversary 2 + 2L + 2(A-1) */
    int32_t max_cycles;    /* truncation horizon (25) */
    int32_t pos_first;     /* 1: p += v*dt before the velocity update (pettingzoo 1.24) ; 0: after */
    float local_ratio;     /* spread: reward = global*(1-lr) + local*lr */
    int64_t n_envs;
    uint64_t seed;         /* Philox key for reset draws */
    float* pos;            /* float2 [n_envs*A] */
    float* vel;            /* float2 [n_envs*A] */
    float* lm;             /* float2 [n_envs*L] */
    int32_t* steps;        /* [n_envs] cycles since reset */
    int64_t* episode;      /* [n_envs] number of resets so far (Philox counter) */
} tsm_mpe;

/* reset n envs (env_ids NULL = envs 0..n-1 in order; then n must be n_envs).  init_pos NULL =
 * draw U(-1,1)^2 agents / landmarks (tag landmarks U(-0.9,0.9)^2) from Philox keyed by
 * (seed, env, episode); else float [n][A+L][2].  obs_out (may be NULL): [n*A][obs_dim]. */
int tsm_mpe_reset(const tsm_mpe* env, const int64_t* env_ids, int64_t n, const float* init_pos,
                  float* obs_out, tsm_stream_t stream);

/* one world step for all envs.  act [n_envs*A] in 0..4.  Outputs are row-ordered:
 * obs_next [R][obs_dim], rew [R], terminated [R] (always 0), truncated [R].
 * autoreset != 0: envs that truncate are re-drawn (Philox) inside the same launch and
 * obs_carry [R][obs_dim] receives the observation the NEXT step must act on (reset obs for
 * done envs, obs_next otherwise) -- collector.py:935,1028.  obs_carry may be NULL. */
int tsm_mpe_step(const tsm_mpe* env, const int32_t* act, float* obs_next, float* rew,
                 uint8_t* terminated, uint8_t* truncated, int autoreset, float* obs_carry,
                 tsm_stream_t stream);

/* The same step for continuous_actions=True (pettingzoo mpe simple_env._set_action): act [n_envs * n_agents][5] float32, the
 * Box(0, 1, (5,)) action of every row; values outside the box are clipped first (the ClipOutOfBoundsWrapper that mpe's make_env
 * puts around continuous environments), then u = (act[2] - act[1], act[4] - act[3]) * sensitivity.  A one-hot act equals the
 * discrete action of that index bit for bit.  Everything else as tsm_mpe_step. */
int tsm_mpe_step_continuous(const tsm_mpe* e, const float* act, float* obs_next, float* rew, uint8_t* terminated,
                            uint8_t* truncated, int autoreset, float* obs_carry, tsm_stream_t stream);

/* =========================================================================================
 * Policy side of the rollout
 * ========================================================================================= */
/* uniform random actions in [0, n_act) for n rows: the collector's `random=True` branch
 * (tianshou/data/collector.py:711-726).  Philox4x32-10, counter (row, step), key (seed, 'ACT!');
 * act = mulhi(x0, n_act).  Restated by oracle/philox.py. */
int tsm_random_actions(int32_t* act, int64_t n, int n_act, uint64_t seed, uint64_t step, tsm_stream_t stream);

/* =========================================================================================
 * K4/K5/K6  actor-critic MLP, fused PPO minibatch, fused clip + Adam
 * replaces  MLP.forward / Net.forward            tianshou/utils/net/common.py:173-181, :343-369
 *           DiscreteActor / DiscreteCritic       tianshou/utils/net/discrete.py:67-89, :116-124
 *           ProbabilisticActorPolicy.forward     tianshou/algorithm/modelfree/reinforce.py:167-192
 *           _add_returns_and_advantages (critic) tianshou/algorithm/modelfree/a2c.py:113-127
 *           PPO._preprocess_batch (logp_old)     tianshou/algorithm/modelfree/ppo.py:146-162
 *           PPO._update_with_batch               tianshou/algorithm/modelfree/ppo.py:164-224
 *           Batch.split (minibatch order)        tianshou/data/batch.py:1209-1226
 *           Algorithm.Optimizer.step             tianshou/algorithm/algorithm_base.py:485-498
 *           AdamOptimizerFactory defaults        tianshou/algorithm/optim.py:89-110
 * Net shape: in_dim -> hidden -> hidden -> out_dim, ReLU, hidden in {64,128}.  Flat float32
 * parameter layout (torch.nn.Linear convention weight[out][in]):
 *     W1[H][in] b1[H] W2[H][H] b2[H] W3[out][H] b3[out]
 * Head outputs 0..n_act-1 are action logits (Categorical(logits=...)); output `value_index`
 * (>= n_act, or -1 for none) is the state value.  The shared actor-critic of the reference's PPO
 * setup (one Net trunk, DiscreteActor.last + DiscreteCritic.last) is out_dim = n_act + 1.
 * ========================================================================================= */
typedef struct tsm_net {
    int32_t in_dim, hidden, out_dim, n_act, value_index;
    int32_t engine;     /* 0: fp32 CUDA-core kernels (exact parity path); 1: tcgen05 kernels (fp16 operands, fp32 accumulate in
                           TMEM; needs hidden == 64, in_dim <= 31, out_dim <= 8); 2: the wide tcgen05 engine (tsm_tcw_*:
                           hidden == 128, in_dim <= 1023, out_dim <= 16) */
    int64_t n_params;   /* == tsm_net_param_count(in_dim, hidden, out_dim) */
    float* params;      /* [n_params] */
    float* grads;       /* [n_params] accumulated by tsm_ppo_minibatch, zeroed by tsm_adam_step */
    float* adam_m;      /* [n_params] exp_avg */
    float* adam_v;      /* [n_params] exp_avg_sq */
    int64_t* adam_step; /* [1] device step counter */
    void* tc_image;     /* engine 1 only: TSM_TC_IMAGE_BYTES device bytes, zero-initialised by the caller: the fp16 operand
                           image of params.  Refreshed by tsm_adam_step / tsm_ppo_epoch; call tsm_net_pack
                           after writing params any other way.  engine 2: tsm_tcw_image_bytes(in_dim) bytes. */
    void* x_scratch;    /* engine 2 only, optional: device scratch for the packed fp16 image of the rows tsm_net_forward /
                           tsm_net_forward_range / tsm_collect_lockstep evaluate (they take float32 rows like engines 0 / 1) */
    int64_t x_scratch_bytes; /* >= tsm_tcw_x_bytes(in_dim, rows of the largest call) */
    int32_t probs_param; /* 1: the actor is DiscreteActor(softmax_output=True) feeding Categorical(probs=...) (utils/net/discrete.py:41,85-86):
                            log-probabilities / entropy are log(clamp(softmax(logits), eps, 1 - eps)) as torch computes them */
    int32_t n_sigma;     /* 0, or n_act: the n_act head outputs are the MEANS of a diagonal Gaussian whose log standard deviations are
                            n_sigma more parameters stored behind b3 (n_params = tsm_net_param_count + n_sigma) --
                            ContinuousActorProbabilistic(unbounded=True, conditioned_sigma=False) with Independent(Normal, 1)
                            (utils/net/continuous.py:222,229-247).  Engine 0 only; tsm_ppo_minibatch then reads tsm_ppo_batch.act_f,
                            and tsm_net_forward serves mode 0 only (sampling / log-probabilities are elementwise host-side ops) */
    float mu_bound;      /* Gaussian nets: 0 = the head outputs are the means (unbounded=True); > 0: mean = mu_bound * tanh(output),
                            the reference's default unbounded=False with max_action = mu_bound (continuous.py:239-241).
                            tsm_net_forward (mode 0) returns the raw outputs either way */
    int32_t cond_sigma;  /* 1: Gaussian net with a state-dependent scale (conditioned_sigma=True, continuous.py:228-233,242): head
                            outputs [0, n_act) are the means, [n_act, 2 n_act) the log standard deviations before
                            clamp(-20, 2) (SIGMA_MIN / SIGMA_MAX), the value head (if any) follows at value_index = 2 n_act;
                            n_sigma stays 0.  Engine 0 only, 2 n_act + 1 <= 8 */
} tsm_net;
#define TSM_TC_IMAGE_BYTES 123648 /* 16 896 image + tsm_ppo_epoch's three more gradient-sized buffers of 7 168 floats + 256 bytes of barrier words + a 20 KB index table */

int64_t tsm_net_param_count(int in_dim, int hidden, int out_dim);

/* engine 1: rebuild net->tc_image from net->params (no-op for engine 0) */
int tsm_net_pack(const tsm_net* net, tsm_stream_t stream);

/* Forward over n rows of x ([.][in_dim] float32; `rows` optionally gathers x rows, and with
 * scatter != 0 results are written at those row ids instead of 0..n-1).
 *   mode 0: raw head outputs -> out [n][out_dim]            (critic(obs), a2c.py:122-124)
 *   mode 1: sample act ~ Categorical(logits) by inverse CDF on one Philox uniform keyed by
 *           (seed, 'ACT!'), counter (row id, counter)        (reinforce.py:185-189, dist.sample())
 *   mode 2: act = argmax logits                              (dist.mode, deterministic_eval)
 *   mode 3: evaluate the given act                           (dist.log_prob(act), ppo.py:157-161)
 * modes 1-3 write logp [n] (log-prob of act), optional value [n] (needs value_index >= 0),
 * optional entropy [n], and optional logits -> out [n][n_act]. */
int tsm_net_forward(const tsm_net* net, const float* x, const int64_t* rows, int64_t n, int mode, int scatter,
                    uint64_t seed, uint64_t counter, float* out, int32_t* act, float* logp, float* value,
                    float* entropy, tsm_stream_t stream);

/* tsm_net_forward (mode 1 sample / 2 argmax) over rows [row0, row0 + n) of a larger batch: x, act, logp, value are the
 * FULL arrays; the sampling key uses the global row index, so splitting a batch into ranges (e.g. to run them on
 * different streams) draws exactly the actions of one launch over the whole batch. */
int tsm_net_forward_range(const tsm_net* net, const float* x, int64_t row0, int64_t n, int mode, uint64_t seed, uint64_t counter,
                          int32_t* act, float* logp, float* value, tsm_stream_t stream);

/* The update batch, addressed in place inside the time-major device buffer ([size][n_rows]
 * arrays, obs [size][n_rows][in_dim]).  Batch position p (0 <= p < B, the order of
 * buffer.sample(0), manager.py:212-229) maps to a storage slot either through an explicit list
 * of env-major buffer indices (`indices`, parity mode: indices[order[p]] drawn by numpy) or in
 * closed form for a lock-step rollout: q = perm(p), row = q / len, slot = (start + q % len) % size
 * with perm the keyed Feistel permutation of [0, perm_n) (perm_n = 0: identity). */
typedef struct tsm_ppo_batch {
    const float* obs;
    const int32_t* act;
    const float* adv;
    const float* ret;
    const float* logp_old;
    const float* v_old;      /* only read when value_clip */
    const int64_t* indices;  /* optional [B] env-major buffer indices */
    int64_t size, n_rows, start, len;
    uint64_t perm_key;
    int64_t perm_n;
    const void* records;     /* engine 1: [size][n_rows] packed rows built by tsm_ppo_pack_records (required by engine 1's
                                tsm_ppo_minibatch / tsm_ppo_epoch; ignored by engine 0) */
    int32_t* slots;          /* optional device scratch [>= number of batch positions, len * n_rows or the length of indices]:
                                the SLOT TABLE of this view, slots[p] = storage slot (t * n_rows + row) of batch position p.
                                Filled by tsm_ppo_slot_table, or as a by-product by tsm_ppo_adv_stats for the positions it
                                visits; engine 1's tsm_ppo_minibatch / tsm_ppo_epoch then read it instead of evaluating the
                                permutation per row.  The caller refills it whenever perm_key / perm_n / start / len /
                                indices change.  NULL: computed in the kernels.  Needs size * n_rows < 2^31. */
    const float* act_f;      /* Gaussian nets (tsm_net.n_sigma > 0): [size][n_rows][n_act] float actions as stored (before map_action) */
} tsm_ppo_batch;

/* Engine 1 gathers each minibatch row as ONE 96-byte record (3 sectors of one 128-byte line) instead of 7-8 sectors spread
 * over five arrays: halves x[32] (obs as fp16, zero padded, x[31] = 1) | float adv, ret, logp_old | int32 act | float v_old |
 * 12 bytes pad; records sit at a 128-byte pitch (TSM_PPO_RECORD_BYTES per slot, the last 32 bytes unused) in a 128-byte
 * aligned array.  Packs every slot of the buffer ([size][n_rows], storage order); NULL source arrays pack as zeros. */
#define TSM_PPO_RECORD_BYTES 128
int tsm_ppo_pack_records(const tsm_net* net, const tsm_ppo_batch* batch, void* records, tsm_stream_t stream);

/* The same for the sub-buffers listed in rows [n_sel] (device int64, each in [0, n_rows)) only -- the rows of one policy group
 * (FlexibleMultiAgentPolicyManager: one update per group over its strided row set): records of every other row are left as they
 * are.  The record of slot (t, row) sits where tsm_ppo_pack_records puts it. */
int tsm_ppo_pack_records_rows(const tsm_net* net, const tsm_ppo_batch* batch, void* records, const int64_t* rows, int64_t n_sel,
                              tsm_stream_t stream);

typedef struct tsm_ppo_cfg {
    float eps_clip;   /* ppo.py:25 (0.2) */
    float dual_clip;  /* <= 0: off (ppo.py:26) */
    float vf_coef;    /* a2c.py (0.5) */
    float ent_coef;   /* a2c.py (0.01) */
    int32_t value_clip;
    int32_t adv_norm; /* advantage_normalization (ppo.py:184-186) */
    int32_t plain_pg; /* 1: policy term = -logp(act) * adv (one-step actor-critic) instead of the clipped surrogate;
                         2: CTDEPolicy.learn as written (multiagent/ctde.py:181-183): logp [n] * adv [n,1] broadcasts to
                            [n,n], i.e. every row is weighted by the minibatch mean advantage = mb_stats[0] */
    int32_t value_heads; /* > 1: value = mean of that many head outputs from value_index on (ctde.py:153-156) */
} tsm_ppo_cfg;

/* batch->slots[p] = storage slot of batch position p for p in [p0, p0 + n) (see tsm_ppo_batch.slots). */
int tsm_ppo_slot_table(const tsm_ppo_batch* batch, int64_t p0, int64_t n, tsm_stream_t stream);

/* per-minibatch {mean, 1/(std_unbiased + eps)} of adv: mb_bounds (device) [n_minibatches+1]
 * batch positions; stats_out [n_minibatches][2]. */
int tsm_ppo_adv_stats(const tsm_ppo_batch* batch, const int64_t* mb_bounds, int64_t n_minibatches, float eps,
                      float* stats_out, tsm_stream_t stream);

/* forward + clipped-surrogate / value / entropy loss + backward over batch positions
 * [mb_start, mb_start + mb_len); parameter gradients are ADDED into net->grads.
 * loss_sums (device, may be NULL) [3] += {sum_i clip_loss_i, sum_i vf_i, sum_i entropy_i}. */
int tsm_ppo_minibatch(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, int64_t mb_start,
                      int64_t mb_len, const float* mb_stats, double* loss_sums, tsm_stream_t stream);

/* g <- grads * grad_scale; optional global-norm clip (max_grad_norm > 0: g *= min(1, max/(|g|+1e-6)));
 * torch.optim.Adam step (no weight decay, no amsgrad); grads zeroed; *adam_step += 1.
 * grad_norm_out (device, may be NULL) receives |g| before clipping. */
int tsm_adam_step(const tsm_net* net, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                  float grad_scale, float* grad_norm_out, tsm_stream_t stream);

/* one epoch on one GPU: for each minibatch m: tsm_ppo_minibatch then tsm_adam_step (engine 1 runs the
 * whole epoch as ONE persistent cooperative launch when the minibatches are uniform -- optimizer steps spread over the
 * grid between grid-wide barriers -- and otherwise as a chain of programmatic dependent launches).
 * h_mb_bounds is a HOST array [n_minibatches+1]; mb_stats [n_minibatches][2] and loss_sums
 * [n_minibatches][4] are device arrays (either may be NULL). */
int tsm_ppo_epoch(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, const int64_t* h_mb_bounds,
                  int64_t n_minibatches, const float* mb_stats, float lr, float beta1, float beta2, float eps,
                  float max_grad_norm, double* loss_sums, tsm_stream_t stream);

/* =========================================================================================
 * CTDE pieces (centralized training, decentralized execution)
 * replaces  GlobalStateConstructor.build         tianshou/algorithm/multiagent/ctde.py:282-343 (modes concatenate, mean)
 *           TD target / advantage of CTDEPolicy.learn    ctde.py:149-178
 * ========================================================================================= */
/* obs [n_envs][n_agents][obs_dim] (rows of one world are contiguous: row = env * A + agent) ->
 * out [n_envs * n_agents][global_dim]: every agent row of a world receives that world's global state,
 * mode 0 "concatenate": global_dim = n_agents * obs_dim, cat(obs_0 .. obs_{A-1}) in agent order (ctde.py:291-294);
 * mode 1 "mean":        global_dim = obs_dim, mean over the agents (ctde.py:296-300). */
int tsm_global_state(const float* obs, int64_t n_envs, int n_agents, int obs_dim, int mode, float* out, tsm_stream_t stream);

/* v_cur, v_next [n][heads] critic outputs; value = mean over heads (ctde.py:153-156);
 * td[i] = rew[i] + gamma * value_next[i] * (1 - terminated[i]) (ctde.py:169); adv[i] = td[i] - value[i] (ctde.py:178).
 * value_out (optional) receives value[i]. */
int tsm_td_target(const float* v_cur, const float* v_next, int heads, const float* rew, const uint8_t* terminated, float gamma,
                  int64_t n, float* td_out, float* adv_out, float* value_out, tsm_stream_t stream);

/* =========================================================================================
 * Data-parallel update on one node: gradient all-reduce fused into the optimizer kernel
 * replaces  torch DistributedDataParallel-style gradient averaging around Algorithm.Optimizer.step
 *           (tianshou/algorithm/algorithm_base.py:485-498); SURVEY 8(e)
 * Every rank maps every rank's receive area and flag pad into its address space (NVLink peer memory; torch
 * symmetric memory provides the pointers).  Step `seq` (1, 2, ...): tsm_ppo_minibatch accumulates into the local
 * net->grads as on one GPU; the optimizer kernel then PUSHES that gradient into slot [seq & 1][rank] of every
 * rank's receive area (plain stores over NVLink, no round trip), signals and waits for all ranks (bounded spin:
 * *status = 1 on time-out instead of a hang), sums the `world` slots of its own receive area in rank order --
 * the same order on every rank, so parameters stay bit-identical -- applies grad_scale = 1 / world, clip and
 * Adam, and clears net->grads.  One launch per minibatch does gradient exchange and optimizer step; there is no
 * separate collective.
 * ========================================================================================= */
#define TSM_MAX_RANKS 8
typedef struct tsm_peers {
    int32_t world, rank;
    float* recv[TSM_MAX_RANKS];     /* per rank: [2][world][n_stride] floats, n_stride = n_params rounded up to 4 */
    uint32_t* flags[TSM_MAX_RANKS]; /* per rank: >= TSM_MAX_RANKS uint32, zero-initialised */
    int32_t* status;                /* local device int, zero-initialised */
    void* ll[TSM_MAX_RANKS];        /* per rank: [2][world][n_stride] 8-byte (value, step) pairs, zero-initialised: the receive area
                                       of the persistent epoch kernel (flag-in-data: an entry is valid once its step word equals
                                       the step waited for, so no fence or separate flag follows the peer stores).  NULL: the
                                       epoch runs as one launch pair per minibatch */
} tsm_peers;

int tsm_adam_step_dist(const tsm_net* net, const tsm_peers* peers, uint32_t seq, float lr, float beta1, float beta2, float eps,
                       float max_grad_norm, float* grad_norm_out, tsm_stream_t stream);

/* tsm_ppo_epoch across ranks: minibatch m is step seq0 + m (seq0 >= 1). */
int tsm_ppo_epoch_dist(const tsm_net* net, const tsm_ppo_batch* batch, const tsm_ppo_cfg* cfg, const int64_t* h_mb_bounds,
                       int64_t n_minibatches, const float* mb_stats, float lr, float beta1, float beta2, float eps,
                       float max_grad_norm, double* loss_sums, const tsm_peers* peers, uint32_t seq0, tsm_stream_t stream);

/* =========================================================================================
 * The lock-step rollout loop (launches only)
 * replaces  Collector._collect                   tianshou/data/collector.py:770-1098
 * for the case where every environment advances every step: n_steps x { tsm_net_forward (mode 1 sample /
 * 2 argmax; net == NULL: tsm_random_actions over n_random_act actions), tsm_mpe_step with auto-reset,
 * tsm_buf_add (ring counters, episode return / length), tsm_episode_stats }, each writing straight into the
 * time-major payload of the device buffer.  obs[slot] must hold the observation to act on.  On return
 * *slot_out is the next slot, *lockstep_len_out the new number of valid slots and *obs_in_carry_out says
 * whether the next observation is in `carry` (last step / ring wrap) or already in obs[*slot_out].
 * ========================================================================================= */
typedef struct tsm_rollout {
    float* obs;           /* [size][n_rows][obs_dim] */
    int32_t* act;         /* [size][n_rows] */
    float* rew;
    uint8_t* terminated;
    uint8_t* truncated;
    float* obs_next;      /* [size][n_rows][obs_dim] */
    float* logp;          /* optional [size][n_rows] */
    float* value;         /* optional */
    float* carry;         /* [n_rows][obs_dim] */
    int64_t size, n_rows, obs_dim, slot;
    double* ep_ret;       /* [n_rows] scratch */
    int64_t* ep_len;      /* [n_rows] scratch */
    double* acc;          /* [9] episode statistics accumulator of tsm_episode_stats */
    double* list_ret;     /* optional per-episode lists */
    int64_t* list_len;
    int64_t* list_count;
    int64_t list_cap;
} tsm_rollout;

/* One vector step of a lock-step rollout in ONE launch: tsm_mpe_step with auto-reset, then for every row the ring
 * counters of tsm_buf_add (buffer_ids = arange, act / rew / flags written to the row's current slot of buf's time-major
 * payload, ep_ret_out / ep_len_out as tsm_buf_add's outputs) and tsm_episode_stats (acc, optional episode list).
 * Needs one sub-buffer per row (buf->n_bufs == n_envs * n_agents).  The launch covers environments [env_begin, env_begin +
 * env_count); every array stays indexed by the global row, so disjoint ranges may run concurrently on different streams.
 * replaces  BaseVectorEnv.step + ReplayBufferManager.add + the done-env loop of Collector._collect
 *           (venvs.py:237-322, manager.py:131-193, collector.py:983-1024) */
int tsm_rollout_step(const tsm_mpe* env, const tsm_buf* buf, const int32_t* act, float* obs_next, float* rew,
                     uint8_t* terminated, uint8_t* truncated, float* obs_carry, double* ep_ret_out, int64_t* ep_len_out,
                     double* acc, double* list_ret, int64_t* list_len, int64_t* list_count, int64_t list_capacity,
                     int64_t env_begin, int64_t env_count /* <= 0: all environments */, tsm_stream_t stream);

int tsm_collect_lockstep(const tsm_net* net, const tsm_mpe* env, const tsm_buf* buf, const tsm_rollout* ro, int64_t n_steps,
                         int mode, uint64_t seed, uint64_t counter0, int64_t lockstep_len, int n_random_act,
                         int64_t* slot_out, int64_t* lockstep_len_out, int* obs_in_carry_out, tsm_stream_t stream);

/* ---- engine 2: the "wide" tcgen05 engine (csrc/mlp_tcw.cu): in(<= 1023) -> 128 -> 128 -> out(<= 16) ---------------------
 * The nets of the reference's CTDE path: DecentralizedActor (multiagent/ctde.py:346-379, obs -> 128 -> 128 -> n_act) and
 * CentralizedCritic (ctde.py:382-414, global state -> 128 -> 128 -> one value per agent; 600 -> ... -> 10 at BASELINE.json
 * configs[2]).  tsm_net.engine = 2, hidden = 128; tsm_net.tc_image holds tsm_tcw_image_bytes(in_dim) bytes (the fp16 operand
 * image of the parameters: call tsm_net_pack after writing params from the host; tsm_tcw_adam keeps it current).
 * Inputs are consumed as PACKED fp16 tile images (tsm_tcw_pack_x), 128 rows per tile; a minibatch is a set of whole tiles. */
int64_t tsm_tcw_image_bytes(int in_dim);
int64_t tsm_tcw_x_bytes(int in_dim, int64_t n_rows);
/* x: [*][ld] float32 rows (ld >= in_dim; the centralized critic's global state of one world is A * obs_dim consecutive floats of
 * the env-major obs block, GlobalStateConstructor "concatenate", ctde.py:291-294, so ld = in_dim = A * obs_dim needs no copy);
 * idx: optional source row per packed row (a permuted / gathered batch, data/batch.py:1209-1226); ximg: tsm_tcw_x_bytes bytes. */
int tsm_tcw_pack_x(const float* x, int64_t ld, int in_dim, const int64_t* idx, int64_t n_rows, void* ximg, tsm_stream_t stream);
/* Forward over all rows of a packed image.  mode 0: out[n_rows][out_dim] raw head outputs (critic values / actor logits,
 * discrete.py:67-89); 1 sample / 2 argmax / 3 evaluate `act`: act, logp, entropy (+ out = logits [n_rows][n_act]); the sampling
 * key of row r is (seed, counter, rng_row0 + r) exactly as in tsm_net_forward. */
int tsm_tcw_forward(const tsm_net* net, const void* ximg, int64_t n_rows, int mode, uint64_t seed, uint64_t counter, int64_t rng_row0,
                    float* out, int32_t* act, float* logp, float* entropy, tsm_stream_t stream);
typedef struct tsm_tcw_tiles {
    const void* ximg;          /* packed input image */
    int64_t n_rows;            /* rows behind ximg (the last tile may be ragged) */
    const int32_t* super_ids;  /* optional device list: logical tile q -> physical tile super_ids[q / sub] * sub + q % sub
                                  (a minibatch of "super-tiles": one critic tile of 128 worlds <-> `sub` = A actor tiles) */
    int32_t sub;
    int64_t tile0, n_tiles;    /* logical tiles [tile0, tile0 + n_tiles) of this launch */
} tsm_tcw_tiles;
typedef struct tsm_tcw_loss {
    int32_t kind;              /* 0: PPO / policy-gradient term on the logits (modelfree/ppo.py:181-197, cfg.plain_pg as in
                                     tsm_ppo_cfg) with entropy bonus; row scalars adv / logp_old / act indexed by physical row
                                  1: value loss per head (ppo.py:199-208): target / v_old / weight are [n_rows][heads]
                                  2: value = mean of the heads against one target per row (ctde.py:153-166) */
    const int32_t* act;
    const float* adv;
    const float* logp_old;
    const float* target;
    const float* v_old;        /* kind 1 with cfg.value_clip */
    const float* weight;       /* kind 1, optional 0/1 mask per head */
    int32_t heads;
} tsm_tcw_loss;
/* Forward + loss + backward over the tiles of one minibatch; gradients are ADDED to net.grads, loss sums (clip, vf, ent, rows)
 * to loss_sums[4].  inv_n = 1 / (rows of the minibatch), gscale = power of two >= rows (fp16 range of the output gradient). */
int tsm_tcw_train(const tsm_net* net, const tsm_tcw_tiles* tiles, const tsm_tcw_loss* loss, const tsm_ppo_cfg* cfg, float inv_n,
                  float gscale, const float* mb_stats, double* loss_sums, tsm_stream_t stream);
/* clip_grad_norm_ over the parameters of ALL n_nets nets jointly (ActorCritic, a2c.py:102-105; algorithm_base.py:485-498) +
 * Adam (optim.py:89-110) + fp16 image refresh + zeroed gradients.  step: 1-based number of this update; workspace:
 * TSM_TCW_ADAM_WS_BYTES zeroed device bytes kept by the caller, seq: number of earlier CLIPPING launches on it. */
#define TSM_TCW_ADAM_WS_BYTES 2048
int tsm_tcw_adam(const tsm_net* nets, int n_nets, int64_t step, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                 float grad_scale, void* workspace, uint32_t seq, float* grad_norm_out, tsm_stream_t stream);
/* advantage mean / 1 / (unbiased std + eps) of tile-ordered minibatches (ppo.py:184-186): minibatch m = super-tiles
 * [m * mb_super, (m + 1) * mb_super) of super_ids (NULL: identity), one super-tile = rows_per_super consecutive rows. */
int tsm_tcw_adv_stats(const float* adv, int64_t n_rows, const int32_t* super_ids, int64_t rows_per_super, int64_t n_super,
                      int64_t mb_super, float eps, float* stats, tsm_stream_t stream);

/* Data-parallel ranks of one node for engine 2: every rank's gradient arena and flag pad live in symmetric
 * memory (torch.distributed._symmetric_memory: rendezvous only).  tsm_tcw_adam_peers = tsm_tcw_adam with the all-reduce inside:
 * one flag store per peer ("my gradient of exchange xseq is complete"), wait for all ranks, coalesced NVLink LOADS of every
 * rank's gradient summed in rank order (bit-identical parameters on all ranks), clip, Adam; grad_scale = 1 / world.  Backward
 * launches of exchange xseq accumulate into parity (xseq & 1) of the arena; the other parity is zeroed by the optimizer. */
typedef struct tsm_tcw_peers {
    int32_t world, rank;
    float* arena[8];      /* rank q's arena: [2 parities][stride] floats, zero-initialised */
    uint32_t* flags[8];   /* rank q's flag pad: >= world uint32, zero-initialised; slot r is written by rank r */
    int64_t stride;       /* floats per parity block */
    int64_t net_off[2];   /* offset of each net's flat gradient inside a parity block (multiples of 4) */
    float* gsum;          /* local scratch, stride floats */
    int32_t* status;      /* local: set to 1 if a peer did not arrive within ~3 s */
} tsm_tcw_peers;
int tsm_tcw_adam_peers(const tsm_net* nets, int n_nets, int64_t step, float lr, float beta1, float beta2, float eps,
                       float max_grad_norm, void* workspace, uint32_t seq, const tsm_tcw_peers* peers, uint32_t xseq,
                       tsm_stream_t stream);

/* One PPO epoch over CTDE nets (the build's "MAPPO": reference PPO arithmetic, modelfree/ppo.py:164-224, with the
 * DecentralizedActor on local observations and the CentralizedCritic on the concatenated global state, one value head per
 * agent) as ONE C call: per minibatch m = super-tiles [m * mb_super, (m + 1) * mb_super) of super_ids -- critic
 * forward/value-loss/backward over those tiles of ximg_critic (one row per world and step), actor forward/policy-loss/backward
 * over the matching n_agents tiles each of ximg_actor (row = world * n_agents + agent), then tsm_tcw_adam over both nets with
 * one gradient norm (tsm_tcw_adam_peers when `peers` is set). */
typedef struct tsm_tcw_epoch {
    const tsm_net* actor;
    const tsm_net* critic;
    const void* ximg_actor;     /* packed local observations, n_rows rows */
    const void* ximg_critic;    /* packed global states, n_rows / n_agents rows */
    int64_t n_rows;
    int32_t n_agents;
    const int32_t* super_ids;   /* device, [n_super] or NULL (identity) */
    int64_t n_super, mb_super;
    const int32_t* act;         /* [n_rows] */
    const float* adv;           /* [n_rows] */
    const float* logp_old;      /* [n_rows] */
    const float* ret;           /* [n_rows] == [n_rows / n_agents][n_agents] */
    const float* v_old;         /* [n_rows], read with cfg.value_clip */
    const int64_t* h_mb_rows;   /* HOST: agent rows of each minibatch */
    const float* mb_stats;      /* device [n_mb][2] (tsm_tcw_adv_stats) or NULL */
    double* sums_actor;         /* device [n_mb][4] */
    double* sums_critic;        /* device [n_mb][4] */
    float lr, beta1, beta2, eps, max_grad_norm;
    int64_t step0;              /* optimizer steps taken before this call */
    void* adam_ws;              /* TSM_TCW_ADAM_WS_BYTES zeroed device bytes */
    uint32_t adam_seq0;         /* clipping launches on adam_ws before this call */
    const tsm_tcw_peers* peers; /* NULL on one rank; else the exchange runs inside the optimizer launch (every rank passes the
                                   same number of minibatches) */
    uint32_t xseq0;             /* exchanges done before this call */
} tsm_tcw_epoch;
int tsm_tcw_ppo_epoch(const tsm_tcw_epoch* e, const tsm_ppo_cfg* cfg, tsm_stream_t stream);

/* Policy.map_action for Box action spaces (algorithm_base.py:255-288): bound_method 0 none / 1 clip to [-1, 1] / 2 tanh, then (scaling)
 * low + (high - low) * (a + 1) / 2 per action dimension; act, out: [n][dim]; *out_of_range (optional) is set to 1 where the
 * reference's range assertion would fire. */
int tsm_map_action(const float* act, int64_t n, int dim, int bound_method, int scaling, const float* low, const float* high,
                   float* out, int* out_of_range, tsm_stream_t stream);

/* tsm_collect_lockstep for GROUPED policies (FlexibleMultiAgentPolicyManager.forward inside the rollout loop,
 * multiagent/flexible_policy.py:189-231 -> marl.py:108-185): group g = (nets[g], rows[g][0..n_rows[g]) device row list, policy
 * seed seeds[g], first sampling counter counters0[g]); one fused forward launch per group and vector step (row gather, act /
 * logp / value scattered into buffer slot t, sampling key = global row) + one tsm_rollout_step over all worlds. */
int tsm_collect_lockstep_groups(const tsm_net* const* nets, const int64_t* const* rows, const int64_t* n_rows, const uint64_t* seeds,
                                const uint64_t* counters0, int n_groups, const tsm_mpe* env, const tsm_buf* buf, const tsm_rollout* ro,
                                int64_t n_steps, int mode, int64_t lockstep_len, int64_t* slot_out, int64_t* lockstep_len_out,
                                int* obs_in_carry_out, tsm_stream_t stream);

/* tcgen05 self test (tests only): D[128][n] = A * B with A / B given as shared-memory images
 * (float32 words, copied verbatim to smem) and explicit descriptor fields
 * h_desc = {a_lbo, a_sbo, a_kstep_bytes, b_lbo, b_sbo, b_kstep_bytes}; kind::tf32, M = 128,
 * `ksteps` MMAs of K = 8 (tf32) or 16 (f16/bf16).  Pins the operand-layout semantics csrc/mlp_tc.cu relies on.
 * kind 0: tf32, 1: f16 / bf16, 3: the f16 GEMM accumulated 16 times by each of FOUR concurrently issuing threads into one
 * accumulator (D = 64 A B): accumulating tcgen05.mma from different threads of a CTA do not race (tools/tc_shared_acc_probe.py). */
int tsm_tc_selftest(const float* a_img, int a_floats, const float* b_img, int b_floats, const uint32_t* h_desc,
                    int a_mn_major, int b_mn_major, int n, int ksteps, int kind /*0 tf32 (K=8), 1 f16 (K=16)*/, int a_bf16,
                    int b_bf16, float* d_out, tsm_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TSM_B200_H */
