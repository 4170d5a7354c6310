/* mopo_b200 -- C ABI of the B200-native MOPO model-rollout hot path.
 *
 * Plain pointers and sizes only (no torch types).  Every entry point returns 0 on
 * success or a negative MOPO_E* code; mopo_last_error() gives the message for the
 * calling thread.  All device work is enqueued on the caller's `stream`
 * (a cudaStream_t passed as void*; NULL = legacy default stream); calls on one handle
 * are stream-ordered and non-reentrant (SURVEY.md 8(b) "Threading").
 *
 * Reference interfaces replaced (paths relative to xionghuichen/mopo):
 *   mopo_ens_create/set_params/set_elites  <- construct_model, mopo/models/constructor.py:7-43;
 *                                             BNN.finalize/load_params, mopo/models/bnn.py:166-281;
 *                                             BNN._end_train (_model_inds), bnn.py:337-340
 *   mopo_ens_predict                       <- BNN.predict(inputs, factored=True), bnn.py:508-546
 *                                             (math: BNN._compile_outputs, bnn.py:631-675)
 *   mopo_ens_train_*, mopo_ens_get_params  <- BNN.train's inner loop and state handling, bnn.py:369-503 (train_op / mse_loss:
 *                                             bnn.py:242-249, 677-700; _save_state / _set_state: bnn.py:287-307)
 *   mopo_env_step                          <- FakeEnv.step, mopo/models/fake_env.py:37-131
 *                                             + StaticFns.termination_fn, mopo/static/<env>.py
 *   mopo_pool_*                            <- SimpleReplayPool / FlexibleReplayPool,
 *                                             softlearning/replay_pools/flexible_replay_pool.py:39-161
 *   mopo_rollout, mopo_rollout_host        <- MOPO._rollout_model, mopo/algorithms/mopo.py:723-765 (device / host inputs)
 *   mopo_sac_*                             <- the SAC update (actor, twin critics, alpha, Polyak targets): graph mopo.py:271-457,
 *                                             MOPO._do_training / _training_batch, mopo.py:801-843
 *   mopo_sac_get/set_opt_state             <- the Adam slots + step inside the reference's TF-session checkpoint (mopo.py:417-441)
 *   mopo_policy_*                          <- the SAC actor graph (mlp_gaussian_policy + apply_squashing_func,
 *                                             mopo.py:292-312) and MOPO.get_action_meta, mopo.py:468-485
 *   mopo_ens_predict_noalloc               <- no counterpart in the reference
 */
#ifndef MOPO_B200_H
#define MOPO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MOPO_OK          0
#define MOPO_EINVAL     -1   /* bad argument (reference: AssertionError / ValueError) */
#define MOPO_ESTATE     -2   /* call order (e.g. params/elites not set; reference: RuntimeError) */
#define MOPO_ECUDA      -3   /* CUDA runtime error */
#define MOPO_EUNSUPPORTED -4 /* shape outside what the sm_100a kernels are built for */

/* termination_fn kinds (mopo/static/<env>.py) */
enum {
  MOPO_TERM_NEVER    = 0,  /* halfcheetah*.py, point2d*.py : all False */
  MOPO_TERM_HOPPER   = 1,  /* hopper.py:9-18 (incl. the abs-of-bool quirk at :12) */
  MOPO_TERM_WALKER2D = 2,  /* walker2d.py:9-17 */
  MOPO_TERM_ANT      = 3,  /* ant.py:9-16, antangle.py */
  MOPO_TERM_HUMANOID = 4   /* humanoid.py:10-14 */
};

/* penalty kinds (fake_env.py:97-115) */
enum {
  MOPO_PEN_NONE         = 0,  /* penalty_coeff == 0 */
  MOPO_PEN_LEARNED_VAR  = 1,  /* max_e ||std_e||_2 over all D1 dims (fake_env.py:110) */
  MOPO_PEN_DISAGREEMENT = 2   /* max_e ||mean_e[1:] - mean_bar[1:]||_2 (fake_env.py:99-108) */
};

/* arithmetic of the ensemble contraction (the reference contracts in fp32: fc.py:96-106) */
enum {
  MOPO_PREC_FP32 = 0,     /* fp32-accurate (the rtol-1e-3 parity path): tcgen05 kind::f16 with every operand split into a
                             bf16 (hi, lo) pair, three MMAs per K step, fp32 accumulation in TMEM (ens_tc3.cu); shapes the
                             tensor-core kernel does not cover (hidden_dim != 200, ...) run on the FFMA kernel */
  MOPO_PREC_BF16 = 1,     /* tcgen05 kind::f16, plain bf16 operands, fp32 accumulation in TMEM */
  MOPO_PREC_FP32_REF = 2  /* fp32 FFMA on the CUDA cores (ens_fp32.cu): kept as the on-device checker of the two above */
};

typedef struct mopo_ens  mopo_ens;
typedef struct mopo_pool mopo_pool;
typedef struct mopo_policy mopo_policy;

typedef struct {
  int32_t obs_dim, act_dim, rew_dim;   /* D1 = obs_dim + rew_dim ; in = obs_dim + act_dim */
  int32_t hidden_dim;                  /* 200 */
  int32_t n_hidden;                    /* 4 */
  int32_t num_networks;                /* 7 */
  int32_t num_elites;                  /* 5 */
  int32_t separate_mean_var;           /* bnn.py:657-667 (1) vs :646-655 (0) */
  int32_t device;                      /* CUDA device ordinal */
} mopo_ens_cfg;

const char* mopo_last_error(void);
int mopo_version(void);
/* SHA-256 prefix of the sources and flags this binary was built from (mopo_b200/build.py: tree_hash) */
const char* mopo_build_hash(void);
/* number of kernels this library has launched in this process (bench.py reports the delta) */
int64_t mopo_launch_count(void);

/* ---- ensemble (construct_model / BNN) ------------------------------------------------ */
int mopo_ens_create(const mopo_ens_cfg* cfg, mopo_ens** out);
int mopo_ens_destroy(mopo_ens* h);

/* Parameters in the reference's own layout (fc.py:145-154): weights[l] = [E,in_l,out_l] row-major
 * fp32, biases[l] = [E,1,out_l]; l = 0..n_hidden (last = mean head, or the 2*D1-wide combined head
 * when separate_mean_var == 0); var_w/var_b = the separate log-variance head (NULL in combined mode).
 * mu/sigma [1,in]; max_logvar/min_logvar [1,D1].  Pointers are HOST pointers when on_device == 0,
 * device pointers otherwise.  Builds the fp32 and bf16 (UMMA-ready) device images. */
int mopo_ens_set_params(mopo_ens* h, const float* const* weights, const float* const* biases,
                        const float* var_w, const float* var_b,
                        const float* mu, const float* sigma,
                        const float* max_logvar, const float* min_logvar,
                        int on_device, void* stream);

/* Device address + element count of one fp32 master image (reference layout), for in-place device-to-device updates such as
 * the NCCL broadcast of the ensemble after model training; call mopo_ens_refresh_images afterwards (rebuilds the tensor-core
 * operand images from the fp32 masters). */
enum { MOPO_IMG_W = 0, MOPO_IMG_B = 1, MOPO_IMG_MU = 2, MOPO_IMG_SIGMA = 3, MOPO_IMG_MAX_LOGVAR = 4, MOPO_IMG_MIN_LOGVAR = 5 };
int mopo_ens_param_image(mopo_ens* h, int32_t which, int32_t layer, void** ptr, int64_t* count);
int mopo_ens_refresh_images(mopo_ens* h, void* stream);

/* Elite set = BNN._model_inds (host int32[n], each in [0,E)). */
int mopo_ens_set_elites(mopo_ens* h, const int32_t* elites, int32_t n);

/* BNN.predict(x, factored=True): x [B,in] device fp32 -> mean [E,B,D1], var [E,B,D1] device fp32. */
int mopo_ens_predict(mopo_ens* h, const float* x, int64_t B, float* mean, float* var,
                     int precision, void* stream);

/* BNN.predict(x, factored=True) like mopo_ens_predict with MOPO_PREC_FP32_REF (the FFMA kernels), but allocating NOTHING: no
 * cudaMalloc / cudaFree / allocating library call on any path.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The scratch (scaled inputs, two [E,B,hidden] activation buffers, log-variances) is a static device array of
 * MOPO_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; a B that needs more is rejected with
 * MOPO_EINVAL (HalfCheetah shapes: B <= 1422).  x [B,in], mean / var [E,B,D1]: device fp32, caller-owned.  ALL calls share the
 * one array (it keeps the last call's contents; nothing in it must start at zero): calls that overlap on the GPU race. */
#define MOPO_NOALLOC_SCRATCH_BYTES (16u << 20)
int mopo_ens_predict_noalloc(mopo_ens* h, const float* x, int64_t B, float* mean, float* var, void* stream);

/* BNN.predict with 3-D inputs (bnn.py:541-545; fc.py:100-101: net e sees x[e]): x [E,B,in] device fp32 -> mean, var [E,B,D1]. */
int mopo_ens_predict3d(mopo_ens* h, const float* x, int64_t B, float* mean, float* var,
                       int precision, void* stream);

/* The contraction alone (BNN._compile_outputs(ret_log_var=True), bnn.py:631-673) on ALREADY SCALED inputs
 * xn = (x - mu) / sigma [B,in]: -> mean [E,B,D1], clamped logvar [E,B,D1].  This is the kernel the roofline
 * is quoted on (bench.py times it alone). */
int mopo_ens_forward(mopo_ens* h, const float* xn, int64_t B, float* mean, float* logvar,
                     int precision, void* stream);

/* ---- FakeEnv.step --------------------------------------------------------------------- */
typedef struct {
  const float* obs;        /* [B,obs_dim] device */
  const float* act;        /* [B,act_dim] device */
  int64_t B;
  int32_t term_kind, penalty_kind;
  float   penalty_coeff;
  int32_t deterministic;   /* fake_env.py:69-70,84-86 */
  int32_t precision;
  /* randomness: either supplied (parity runs) or Philox4x32-10 keyed by (seed, step, origin row) */
  const float*   noise;       /* [B,D1] device: N(0,1) draw of the chosen elite for each row, or NULL */
  const int32_t* model_inds;  /* [B] device: chosen net per row (values in the elite set), or NULL */
  uint64_t seed;
  uint64_t step_id;           /* Philox stream id of this step */
  const int32_t* origin;      /* [B] device: row id fed to Philox (NULL = 0..B-1) */
  /* outputs (device) */
  float*   next_obs;       /* [B,obs_dim] */
  float*   rewards;        /* [B,1] penalised */
  uint8_t* terminals;      /* [B,1] 0/1 */
  float*   penalty;        /* [B,1] or NULL */
  float*   unpenalized;    /* [B,1] or NULL */
  float*   model_means;    /* [B,D1] or NULL (info['mean'] minus the terminal column) */
  float*   model_stds;     /* [B,D1] or NULL */
  float*   log_prob;       /* [B] or NULL (fake_env.py:20-35) */
  float*   dev;            /* [B] or NULL */
  int32_t* model_inds_out; /* [B] or NULL: the net actually used per row */
} mopo_step_args;

int mopo_env_step(mopo_ens* h, const mopo_step_args* a, void* stream);

/* StaticFns.termination_fn (mopo/static/<env>.py): next_obs [B,obs_dim] device fp32 -> out [B,1] uint8. */
int mopo_termination_fn(int32_t term_kind, const float* next_obs, int64_t B, int32_t obs_dim,
                        uint8_t* out, void* stream);

/* ---- model_pool ring buffer ------------------------------------------------------------ */
enum { MOPO_F_ACTIONS = 0, MOPO_F_REWARDS = 1, MOPO_F_TERMINALS = 2, MOPO_F_OBS = 3, MOPO_F_NEXT_OBS = 4 };

int mopo_pool_create(int32_t obs_dim, int32_t act_dim, int64_t max_size, int32_t device, mopo_pool** out);
int mopo_pool_destroy(mopo_pool* p);
int64_t mopo_pool_size(const mopo_pool* p);
int64_t mopo_pool_pointer(const mopo_pool* p);
int64_t mopo_pool_max_size(const mopo_pool* p);
/* Device pointer of a field array ([max_size, width]; fp32 except terminals = uint8). */
void* mopo_pool_field(const mopo_pool* p, int32_t field);
/* add_samples (flexible_replay_pool.py:57-83): device arrays of n rows, modular index, last write wins. */
int mopo_pool_add(mopo_pool* p, const float* obs, const float* act, const float* next_obs,
                  const float* rew, const uint8_t* term, int64_t n, void* stream);
/* Rows [pointer, pointer+n) of the field arrays were filled in place (a collective writing straight into mopo_pool_field
 * memory): advance pointer / size as add_samples does (:45-48).  Requires pointer + n <= max_size. */
int mopo_pool_advance(mopo_pool* p, int64_t n);
/* batch_by_indices (:121-135): idx device int64[n]; outputs device arrays of n rows. */
int mopo_pool_gather(const mopo_pool* p, const int64_t* idx, int64_t n, float* obs, float* act,
                     float* next_obs, float* rew, uint8_t* term, void* stream);
/* random_batch (:85-92) with Philox indices in [0,size): index i of the draw uses counter (i, offset) -- `offset` is the id of
 * the draw, two draws with different offsets are independent (np.random.randint per batch). */
int mopo_pool_sample(const mopo_pool* p, int64_t n, uint64_t seed, uint64_t offset, int64_t* idx_out,
                     float* obs, float* act, float* next_obs, float* rew, uint8_t* term, void* stream);

/* ---- policy (SAC actor) inside the rollout: MOPO.get_action_meta -------------------------------------- */
/* mlp_gaussian_policy (mopo.py:292-303) with hidden_sizes (H, H), ReLU on both hidden layers, then
 *   mu = dense(net, act_dim), log_std = clip(dense(net, act_dim), -20, 2), pi = mu + N(0,1) * exp(log_std),
 * squashed by apply_squashing_func (mopo.py:281-290): action = tanh(pi) (tanh(mu) when deterministic). */
int mopo_policy_create(int32_t obs_dim, int32_t act_dim, int32_t hidden_dim, int32_t device, mopo_policy** out);
int mopo_policy_destroy(mopo_policy* p);
/* tf.layers.dense layout: kernel [in,out] row-major fp32, bias [out].  weights / biases = {hidden 0 [obs,H],
 * hidden 1 [H,H], mu [H,act], log_std [H,act]}.  Host pointers when on_device == 0. */
int mopo_policy_set_params(mopo_policy* p, const float* const* weights, const float* const* biases,
                           int on_device, void* stream);
typedef struct {
  const float* obs;          /* [B,obs_dim] device */
  int64_t B;
  int32_t deterministic;     /* get_action_meta(deterministic=True): tanh(mu) */
  int32_t random_uniform;    /* MOPO._rollout_random (mopo.py:738): uniform(-1,1) actions, the network is not run */
  int32_t precision;         /* MOPO_PREC_FP32 (FFMA, the rtol-1e-3 parity path) | MOPO_PREC_BF16 (tcgen05) */
  const float* noise;        /* [B,act_dim] device N(0,1) draw (parity runs), or NULL: Philox4x32-10 */
  uint64_t seed, step_id;    /* Philox key / stream id of this call */
  const int32_t* origin;     /* [B] device row ids fed to Philox, or NULL = origin_base + 0..B-1 */
  int64_t origin_base;
  float* actions;            /* [B,act_dim] device out */
  float* mu_out;             /* [B,act_dim] or NULL: pre-squash mean */
  float* std_out;            /* [B,act_dim] or NULL: exp(clipped log_std) */
} mopo_policy_args;
int mopo_policy_act(mopo_policy* p, const mopo_policy_args* a, void* stream);

/* ---- BNN.train (bnn.py:369-503): the optimiser loop on the device -------------------------- */
/* Start a training run: allocates (first call) or keeps the Adam state (tf.train.AdamOptimizer defaults beta1 .9, beta2 .999, eps 1e-8,
 * constructor.py:40 learning_rate 1e-3) and the per-net snapshot store.  weight_decays: HOST array of n_hidden + 2
 * rates -- hidden layers, mean head, log-variance head (constructor.py:31-36; combined head: pass the same rate twice). */
int mopo_ens_train_begin(mopo_ens* h, float lr, const float* weight_decays);
/* The optimiser state lives on the handle from the first mopo_ens_train_begin on (the reference builds ONE AdamOptimizer in
 * finalize, bnn.py:180, 252: moments and step count carry over between BNN.train() calls); mopo_ens_train_reset drops it.
 * mopo_ens_train_set_deterministic(h, 1): the deterministic training loss of bnn.py:246-247 (MSE + decays). */
int mopo_ens_train_reset(mopo_ens* h);
int mopo_ens_train_set_deterministic(mopo_ens* h, int32_t deterministic);
/* ceil(n_cols / batch_size) Adam updates (bnn.py:421-427): batch k trains net e on rows idx[e*ld_idx + k*bs .. +bs) of
 * inputs [n_rows, obs+act] (raw; the handle's scaler is applied) / targets [n_rows, obs+rew]; all device pointers.
 * loss_out: optional device float[ceil(n_cols/bs)], train_loss (bnn.py:242-244) of each batch before its update.
 * Row indices are not range-checked on the device.  Enqueues only; does not synchronise. */
int mopo_ens_train_batches(mopo_ens* h, const float* inputs, const float* targets, int64_t n_rows, const int32_t* idx,
                           int64_t ld_idx, int64_t n_cols, int32_t batch_size, float* loss_out, void* stream);
/* mse_loss (bnn.py:248): mse_out[e] = mean (mean_e - target)^2 over n rows; idx == NULL: every net sees rows 0..n-1
 * (the tiled holdout set, bnn.py:397-398), else net e sees rows idx[e*ld_idx + 0..n). */
int mopo_ens_train_mse(mopo_ens* h, const float* inputs, const float* targets, const int32_t* idx, int64_t ld_idx,
                       int64_t n, float* mse_out, void* stream);
/* _save_state(net) (bnn.py:287-290) / _set_state() (bnn.py:292-307; MOPO_ESTATE if a net was never saved, the
 * reference's KeyError).  max/min_logvar are not part of the saved state, as in the reference. */
int mopo_ens_train_snapshot(mopo_ens* h, int32_t net, void* stream);
int mopo_ens_train_restore(mopo_ens* h, void* stream);
/* Finish: rebuild the tensor-core operand images from the trained fp32 weights (synchronises); the optimiser state stays. */
int mopo_ens_train_end(mopo_ens* h, void* stream);
/* Device parameters -> HOST arrays in the reference layout (the inverse of mopo_ens_set_params; synchronises). */
int mopo_ens_get_params(mopo_ens* h, float* const* weights, float* const* biases, float* var_w, float* var_b,
                        float* max_logvar, float* min_logvar, void* stream);

/* ---- the SAC update: MOPO._do_training (mopo.py:829-843) on the device ---------------------- */
typedef struct mopo_sac mopo_sac;
typedef struct {
  float reward_scale, discount, tau;       /* mopo.py:60-66 defaults 1.0, 0.99, 5e-3 */
  float policy_lr, q_lr;                   /* Adam learning rates (mopo.py:417-441), default 3e-4 */
  float target_entropy;                    /* 'auto' = -act_dim (mopo.py:165-168) */
  int32_t target_update_interval;          /* mopo.py:839: Polyak update every k-th update (default 1) */
  int32_t reserved;
} mopo_sac_hparams;
int mopo_sac_create(int32_t obs_dim, int32_t act_dim, int32_t hidden, int32_t device, const mopo_sac_hparams* hp, mopo_sac** out);
int mopo_sac_destroy(mopo_sac* s);
/* HOST arrays, tf.layers.dense layout (kernel [in,out], bias [out]): pi_W/pi_b = {hidden0, hidden1, mu, log_std} of scope
 * main/pi; q*_W/q*_b = {hidden0, hidden1, out} of main/q1, main/q2 (input = concat(obs, act)).  Runs target_init
 * (target critics <- main, mopo.py:451-452) and resets the four Adam states. */
int mopo_sac_set_params(mopo_sac* s, const float* const* pi_W, const float* const* pi_b, const float* const* q1_W,
                        const float* const* q1_b, const float* const* q2_W, const float* const* q2_b, float log_alpha, void* stream);
/* Overwrite the target critics (HOST arrays, layout as q1_W / q1_b): restoring a checkpoint taken mid-training. */
int mopo_sac_set_targets(mopo_sac* s, const float* const* q1t_W, const float* const* q1t_b, const float* const* q2t_W,
                         const float* const* q2t_b, void* stream);
/* The inverse (HOST outputs); q1t / q2t (target critics) may be NULL. */
int mopo_sac_get_params(mopo_sac* s, float* const* pi_W, float* const* pi_b, float* const* q1_W, float* const* q1_b,
                        float* const* q2_W, float* const* q2_b, float* const* q1t_W, float* const* q1t_b,
                        float* const* q2t_W, float* const* q2t_b, float* log_alpha, void* stream);
typedef struct {
  const float *obs, *act, *rew;            /* device [n,obs_dim] [n,act_dim] [n] (rewards[..., 0]) */
  const uint8_t* term;                     /* device [n] */
  const float* next_obs;                   /* device [n,obs_dim] */
  int64_t n;
  const float* eps;                        /* optional device [2n,act_dim]: tf.random_normal of pi(s) (rows 0..n-1) and of
                                              pi(s') (rows n..2n-1); NULL = Philox stream 4 keyed (seed, update index, row) */
  uint64_t seed;
  float* losses_out;                       /* optional device float[4]: q1_loss, q2_loss, policy_loss, alpha_loss */
} mopo_sac_batch;
/* One training step: all gradients at the pre-update variables, four Adam steps, Polyak update.  Enqueues only. */
int mopo_sac_update(mopo_sac* s, const mopo_sac_batch* b, void* stream);
/* n_updates x { _training_batch (mopo.py:801-821): n_env rows of env_pool then n_model rows of model_pool, indices drawn on
 * the device; one update }.  losses_out: optional device float[n_updates][4]. */
int mopo_sac_update_from_pools(mopo_sac* s, const mopo_pool* env_pool, const mopo_pool* model_pool, int32_t n_env,
                               int32_t n_model, uint64_t seed, int32_t n_updates, float* losses_out, void* stream);
/* Copy the actor (scope main/pi) into a rollout policy handle, device to device (rebuilds its tensor-core image). */
int mopo_sac_export_policy(mopo_sac* s, mopo_policy* p, void* stream);
/* Optimiser state for checkpoint / resume (the reference checkpoints its TF session, which holds the four Adam optimisers'
 * slots, mopo.py:417-441): the 13 tensors in the order pi {W0 b0 W1 b1 Whead bhead}, q {W0 b0 W1 b1 W2 b2} (twins stacked),
 * log_alpha, with Whead / bhead the fused [mu | log_std] head.  sizes: element counts; m / v: HOST arrays of those sizes;
 * t: Adam step; draw: update index that keys the minibatch draws and the policy noise. */
#define MOPO_SAC_TENSORS 13
int mopo_sac_opt_sizes(mopo_sac* s, int64_t* sizes);
int mopo_sac_get_opt_state(mopo_sac* s, float* const* m, float* const* v, int64_t* t, uint64_t* draw, void* stream);
int mopo_sac_set_opt_state(mopo_sac* s, const float* const* m, const float* const* v, int64_t t, uint64_t draw, void* stream);

/* ---- MOPO._rollout_model ------------------------------------------------------------------ */
typedef struct {
  const float* start_obs;        /* [B,obs_dim] device */
  int64_t B;
  int32_t rollout_length;        /* h */
  const float* const* actions;   /* host array of h device pointers, each [>=n_i, act_dim] (row j = j-th
                                    surviving row of step i, as act = policy(obs) at mopo.py:734) */
  int32_t term_kind, penalty_kind;
  float   penalty_coeff;
  int32_t deterministic, precision;
  const float* const*   noise;      /* optional host array of h device ptrs [n_i,D1], or NULL */
  const int32_t* const* model_inds; /* optional host array of h device ptrs [n_i], or NULL */
  uint64_t seed;
  uint64_t rollout_id;           /* Philox stream: (rollout_id, step) */
  int64_t  origin_base;          /* global row id of start_obs[0] (batch sharding) */
  int64_t* steps_added;          /* host int64[h] out (len(obs) per step, mopo.py:748) */
  void* const* action_events;    /* optional host array of h cudaEvent_t (or NULL): actions[i] is valid once
                                    action_events[i] has completed (e.g. an H2D copy on another stream); the rollout
                                    waits for each on `stream` just before the first kernel that reads actions[i], so
                                    uploads of later steps' actions overlap the earlier steps' compute */
  mopo_policy* policy;           /* optional: act = policy(obs) on the device at every step (mopo.py:734); then
                                    `actions` may be NULL (or hold NULL entries) and nothing crosses PCIe in the loop */
  int32_t rollout_random;        /* with `policy`: uniform(-1,1) actions instead (mopo.py:736-738) */
  int32_t policy_deterministic;  /* with `policy`: tanh(mu) */
} mopo_rollout_args;

/* Runs h fused steps device-resident (no host round trip between steps), appending every step's
 * rows to the pool in the reference's order, then synchronises once to read the counts back
 * (termination kinds that never terminate need no read-back: the call returns stream-ordered, without synchronising). */
int mopo_rollout(mopo_ens* h, mopo_pool* pool, const mopo_rollout_args* a, void* stream);
/* The same with HOST inputs, as MOPO calls it (NumPy arrays): a->start_obs [B,obs_dim] and every a->actions[i] [B,act_dim]
 * are host pointers (pinned memory makes the copies asynchronous; a->action_events is ignored).  The call stages them on
 * the device itself -- start states and step 0's actions on `stream`, the later steps' actions on an internal copy stream,
 * each behind an event the rollout waits for only ahead of the first kernel that reads it.  noise / model_inds stay device. */
int mopo_rollout_host(mopo_ens* h, mopo_pool* pool, const mopo_rollout_args* a, void* stream);

/* ---- measurement ------------------------------------------------------------------------------- */
/* CUDA-event timing of the ensemble-forward kernel (the dominant kernel of a step) as it runs inside
 * mopo_env_step / mopo_rollout: events are recorded on the launching stream around each launch.
 * mopo_kernel_timing(h, 1) starts and clears collection (h, 2: the step-tail kernel of mopo_rollout instead; h, 0: off), mopo_kernel_time returns the summed milliseconds and the
 * number of launches since then (synchronising on the events) and clears. bench.py's roofline uses it. */
int mopo_kernel_timing(mopo_ens* h, int enable);
int mopo_kernel_time(mopo_ens* h, double* ms_total, int64_t* launches);

#ifdef __cplusplus
}
#endif
#endif
