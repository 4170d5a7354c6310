// The SAC update step on the device: MOPO._do_training (reference mopo/algorithms/mopo.py:829-843)
// over the graph of mopo.py:271-457 -- 2x256 ReLU Gaussian actor, twin 2x256 ReLU critics, their targets, log_alpha:
//   q_target = reward_scale r + discount (1 - done) (min(q1', q2')(s', pi(s')) - alpha logp(pi(s')))   (mopo.py:382-394)
//   q_i loss = 0.5 mean (q_i(s, a) - q_target)^2                                                       (mopo.py:413-415)
//   pi loss  = mean (alpha logp(pi(s)) - min(q1, q2)(s, pi(s)))        (reparameterised, mopo.py:371-379)
//   alpha loss = -mean log_alpha (logp(pi(s)) + target_entropy)                                        (mopo.py:437-441)
// four Adam optimisers (mopo.py:417-441), then target <- (1 - tau) target + tau main (mopo.py:448-449).
// Every gradient is taken at the pre-update variables (one session.run, mopo.py:838).  Batches are 256 rows, so the
// step is a chain of small launches: the policy runs once on [s; s'] (2n rows), the twin critics are one batched GEMM
// stack (E = 2) on [(s, a); (s, pi(s))], their targets on (s', pi(s')); fp32 FFMA GEMMs with bias / ReLU / ReLU' epilogues
// (gemm_fp32.cuh), fixed-order reductions, one multi-tensor Adam launch and one Polyak launch.  The batch can be drawn
// from the device replay pools inside the call (mopo_sac_update_from_pools: _training_batch, mopo.py:801-821).
#include "gemm_fp32.cuh"
#include "philox.cuh"
#include <string.h>

namespace {

constexpr float LOG_STD_MAX = 2.0f, LOG_STD_MIN = -20.0f, PI_EPS = 1e-8f;
constexpr int SAC_MAX_ACT = 32;

// t += 1; lr_t[0] = policy lr, lr_t[1] = Q lr for the step being applied (tf.train.AdamOptimizer)
__global__ void sac_tick_kernel(int64_t* __restrict__ t, uint64_t* __restrict__ step, float* __restrict__ lr_t, float lr_pi, float lr_q) {
  const int64_t k = ++t[0];
  ++step[0];                                   // Philox step id of this update's policy noise
  const double c = sqrt(1.0 - pow((double)ADAM_B2, (double)k)) / (1.0 - pow((double)ADAM_B1, (double)k));
  lr_t[0] = (float)((double)lr_pi * c);
  lr_t[1] = (float)((double)lr_q * c);
}

// One thread per policy row r of [s; s'] (2n rows): head [mu | log_std] -> u = mu + eps std, a = tanh(u), logp
// (gaussian_likelihood + squashing correction, mopo.py:281-296); writes the critic inputs:
//   r <  n: Xq[r] = (s_r, a_data_r),  Xq[n + r] = (s_r, a)      r >= n: Xt[r - n] = (s'_r, a)
__global__ void sac_pi_head_kernel(const float* __restrict__ HD, const float* __restrict__ eps_in, uint64_t seed,
                                   const uint64_t* __restrict__ d_step,
                                   const float* __restrict__ obs, const float* __restrict__ act, const float* __restrict__ next_obs,
                                   int n, int od, int ad, float* __restrict__ eps_out, float* __restrict__ logp,
                                   float* __restrict__ Xq, float* __restrict__ Xt) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= 2 * n) return;
  const int in_q = od + ad;
  float z[SAC_MAX_ACT];
  if (eps_in) {
    for (int d = 0; d < ad; ++d) z[d] = eps_in[(int64_t)r * ad + d];
  } else {
    const uint32_t step = (uint32_t)*d_step;
    for (int d0 = 0; d0 < ad; d0 += 4) {
      float gg[4];
      philox_normal4((uint32_t)r, step, (uint32_t)(d0 >> 2), seed, gg, MOPO_STREAM_SAC);
      for (int j = 0; j < 4 && d0 + j < ad; ++j) z[d0 + j] = gg[j];
    }
  }
  const float* s = r < n ? obs + (int64_t)r * od : next_obs + (int64_t)(r - n) * od;
  float* xo = r < n ? Xq + (int64_t)(n + r) * in_q : Xt + (int64_t)(r - n) * in_q;
  for (int k = 0; k < od; ++k) xo[k] = s[k];
  if (r < n) {
    float* xd = Xq + (int64_t)r * in_q;
    for (int k = 0; k < od; ++k) xd[k] = s[k];
    for (int d = 0; d < ad; ++d) xd[od + d] = act[(int64_t)r * ad + d];
  }
  float lp = 0.f;
  for (int d = 0; d < ad; ++d) {
    const float mu = HD[(int64_t)r * 2 * ad + d];
    const float ls = fminf(fmaxf(HD[(int64_t)r * 2 * ad + ad + d], LOG_STD_MIN), LOG_STD_MAX);
    const float sd = expf(ls);
    const float u = mu + z[d] * sd;
    const float q = z[d] * sd / (sd + PI_EPS);        // (u - mu) / (std + EPS) with u - mu = eps std exactly
    lp += -0.5f * (q * q + 2.0f * ls + 1.8378770664093453f);
    lp -= 2.0f * (0.6931471805599453f - u - softplus_f(-2.0f * u));
    xo[od + d] = tanhf(u);
    eps_out[(int64_t)r * ad + d] = z[d];
  }
  logp[r] = lp;
}

// One block: q targets, dQ for the stacked critic rows, the four losses and d alpha_loss / d log_alpha.
//   Qo [2, 2n]: twin e, rows [0,n) = q_e(s, a), [n,2n) = q_e(s, pi(s));  To [2, n] = target twins on (s', pi(s'))
//   out[0..3] = q1_loss, q2_loss, policy_loss, alpha_loss;  out[4] = d alpha_loss / d log_alpha
__global__ void __launch_bounds__(256) sac_loss_kernel(const float* __restrict__ Qo, const float* __restrict__ To,
                                                       const float* __restrict__ logp, const float* __restrict__ rew,
                                                       const uint8_t* __restrict__ term, const float* __restrict__ log_alpha,
                                                       int n, float reward_scale, float discount, float target_entropy,
                                                       float* __restrict__ dQo, float* __restrict__ out, float* __restrict__ g_alpha) {
  __shared__ float sh[8];
  const float la = *log_alpha, alpha = expf(la), inv_n = 1.0f / (float)n;
  float s1 = 0.f, s2 = 0.f, sp = 0.f, sa = 0.f;
  for (int b = threadIdx.x; b < n; b += blockDim.x) {
    const float next_value = fminf(To[b], To[n + b]) - alpha * logp[n + b];
    const float y = reward_scale * rew[b] + discount * (term[b] ? 0.0f : 1.0f) * next_value;
    const float q1 = Qo[b], q2 = Qo[2 * n + b], q1p = Qo[n + b], q2p = Qo[3 * n + b];
    dQo[b] = (q1 - y) * inv_n;
    dQo[2 * n + b] = (q2 - y) * inv_n;
    const bool first = q1p <= q2p;                 // tf.minimum sends a tie's gradient to its first argument
    dQo[n + b] = first ? -inv_n : 0.0f;
    dQo[3 * n + b] = first ? 0.0f : -inv_n;
    s1 += (q1 - y) * (q1 - y);
    s2 += (q2 - y) * (q2 - y);
    sp += alpha * logp[b] - fminf(q1p, q2p);
    sa += logp[b] + target_entropy;
  }
  s1 = block_sum(s1, sh); s2 = block_sum(s2, sh); sp = block_sum(sp, sh); sa = block_sum(sa, sh);
  if (threadIdx.x == 0) {
    out[0] = 0.5f * s1 * inv_n; out[1] = 0.5f * s2 * inv_n; out[2] = sp * inv_n; out[3] = -la * sa * inv_n; out[4] = -sa * inv_n;
    *g_alpha = -sa * inv_n;
  }
}

// d policy_loss / d [mu | log_std_raw] for the n rows of s (thread per row); dXa [2, n, ad] = d(-min q)/d a per twin.
__global__ void sac_pi_head_grad_kernel(const float* __restrict__ HD, const float* __restrict__ eps, const float* __restrict__ dXa,
                                        const float* __restrict__ log_alpha, int n, int ad, float* __restrict__ dHD) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n) return;
  const float c = expf(*log_alpha) / (float)n;
  for (int d = 0; d < ad; ++d) {
    const float mu = HD[(int64_t)b * 2 * ad + d], ls_raw = HD[(int64_t)b * 2 * ad + ad + d];
    const float ls = fminf(fmaxf(ls_raw, LOG_STD_MIN), LOG_STD_MAX);
    const float sd = expf(ls), z = eps[(int64_t)b * ad + d];
    const float u = mu + z * sd, a = tanhf(u);
    const float den = sd + PI_EPS;
    const float dev = z * sd;                          // u - mu
    const float r2 = dev / (den * den);
    const float d_a = dXa[(int64_t)b * ad + d] + dXa[((int64_t)n + b) * ad + d];
    const float g_u = c * (-r2 + 2.0f * a) + d_a * (1.0f - a * a);
    const float d_mu = g_u + c * r2;
    float d_ls = g_u * (z * sd) + c * (dev * dev * sd / (den * den * den) - 1.0f);
    if (ls_raw < LOG_STD_MIN || ls_raw > LOG_STD_MAX) d_ls = 0.0f;      // tf.clip_by_value passes no gradient outside
    dHD[(int64_t)b * 2 * ad + d] = d_mu;
    dHD[(int64_t)b * 2 * ad + ad + d] = d_ls;
  }
}

// _training_batch (mopo.py:801-821) in one launch: rows [0, n_env) from the env pool, [n_env, n) from the model pool, indices as
// mopo_pool_sample draws them (Philox stream 2, counter = (row, offset), offset = 2 (update index) + pool: every update and
// each pool has its own counter plane, so consecutive minibatches are independent draws, mopo.py:801-821) -- all five fields.
struct SacSample {
  const float *obs[2], *act[2], *nobs[2], *rew[2]; const uint8_t* term[2];
  int64_t size[2]; int n_env, n, od, ad;
};
__global__ void __launch_bounds__(256) sac_sample_kernel(SacSample a, uint64_t seed, const uint64_t* __restrict__ d_step,
                                                         float* __restrict__ bo, float* __restrict__ ba, float* __restrict__ bno,
                                                         float* __restrict__ br, uint8_t* __restrict__ bt, int64_t* __restrict__ bidx) {
  const int W = 2 * a.od + a.ad + 2;
  const uint64_t base = 2 * (*d_step - 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n * W; i += gridDim.x * blockDim.x) {
    const int j = i / W, c = i % W;
    const int w = j >= a.n_env;
    Philox4 r = philox_pool_draw((uint32_t)(w ? j - a.n_env : j), base + (uint64_t)w, seed);
    const int64_t idx = (int64_t)__umul64hi(((uint64_t)r.x << 32) | r.y, (uint64_t)a.size[w]);
    if (c < a.od) bo[(int64_t)j * a.od + c] = a.obs[w][idx * a.od + c];
    else if (c < 2 * a.od) bno[(int64_t)j * a.od + (c - a.od)] = a.nobs[w][idx * a.od + (c - a.od)];
    else if (c < 2 * a.od + a.ad) ba[(int64_t)j * a.ad + (c - 2 * a.od)] = a.act[w][idx * a.ad + (c - 2 * a.od)];
    else if (c == W - 2) { br[j] = a.rew[w][idx]; bidx[j] = idx; }
    else bt[j] = a.term[w][idx];
  }
}

constexpr int SAC_TENSORS = 13;   // pi 3 W + 3 b, q 3 W + 3 b (twins stacked), log_alpha
struct SacAdam {
  float* p[SAC_TENSORS]; const float* g[SAC_TENSORS]; float* m[SAC_TENSORS]; float* v[SAC_TENSORS];
  int64_t n[SAC_TENSORS]; int which_lr[SAC_TENSORS];
  const float* lr_t;
};
__global__ void __launch_bounds__(256) sac_adam_kernel(SacAdam a) {
  const int t = blockIdx.y;
  const float lr_t = a.lr_t[a.which_lr[t]];
  float* __restrict__ p = a.p[t]; const float* __restrict__ g = a.g[t];
  float* __restrict__ m = a.m[t]; float* __restrict__ v = a.v[t];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n[t]; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float mi = ADAM_B1 * m[i] + (1.0f - ADAM_B1) * gi;
    const float vi = ADAM_B2 * v[i] + (1.0f - ADAM_B2) * gi * gi;
    m[i] = mi; v[i] = vi;
    p[i] = p[i] - lr_t * mi / (sqrtf(vi) + ADAM_EPS);
  }
}

struct SacPolyak { float* targ[6]; const float* main[6]; int64_t n[6]; float tau; };
__global__ void __launch_bounds__(256) sac_polyak_kernel(SacPolyak a) {
  const int t = blockIdx.y;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n[t]; i += (int64_t)gridDim.x * blockDim.x)
    a.targ[t][i] = (1.0f - a.tau) * a.targ[t][i] + a.tau * a.main[t][i];
}

// dst[r, c0 + c] = src[r, c] / dst[r, c] = src[r, c0 + c]: the fused [mu | log_std] head <-> the reference's two dense layers
__global__ void sac_cols_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t rows, int n_src, int n_dst, int c0,
                                int scatter) {
  const int w = scatter ? n_src : n_dst;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * w; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / w; const int c = (int)(i % w);
    if (scatter) dst[r * n_dst + c0 + c] = src[r * n_src + c];
    else dst[r * n_dst + c] = src[r * n_src + c0 + c];
  }
}

}  // namespace

struct mopo_sac {
  int od = 0, ad = 0, H = 0, device = 0;
  bool params_set = false;
  mopo_sac_hparams hp;
  // main/pi: W0 [od,H], W1 [H,H], Wh [H,2 ad] (mu | log_std); main/q and target/q: twins stacked, W0 [2,od+ad,H], W1 [2,H,H], W2 [2,H,1]
  float *piW[3] = {}, *pib[3] = {}, *qW[3] = {}, *qb[3] = {}, *tW[3] = {}, *tb[3] = {};
  float* log_alpha = nullptr;
  // gradients and Adam moments in the order of SacAdam: pi W0 b0 W1 b1 Wh bh, q W0 b0 W1 b1 W2 b2, log_alpha
  float *g[SAC_TENSORS] = {}, *m[SAC_TENSORS] = {}, *v[SAC_TENSORS] = {};
  int64_t* t = nullptr; float* lr_t = nullptr; float* out = nullptr;   // device: Adam step, {lr_pi_t, lr_q_t}, losses[5]
  int64_t updates = 0;
  cudaStream_t side = nullptr, side2 = nullptr; cudaEvent_t ev_pi[4] = {}; cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_fork2 = nullptr, ev_join2 = nullptr;
  DevBuf ba, br, bt, bidx;                   // batch drawn from the pools
  DevBuf Xpi, H0p, H1p, HD, eps, logp, Xq, Xt, H0q, H1q, Qo, T0, T1, To, dQo, dZ1q, dZ0q, dXa, dHD, dZ1p, dZ0p;
  uint64_t* d_step = nullptr;                         // device Philox step id of the policy noise (advanced by sac_tick_kernel)
  // the captured pools iteration (tick -> draw -> update), kept across calls: MOPO calls one update per environment step
  // (_do_training_repeats, mopo.py:780-799), so the same pools / sizes recur until the next rollout grows the model pool
  cudaGraphExec_t graph = nullptr; int g_nodes = 0; uint64_t g_seed = 0;
  SacSample g_smp, last_smp; uint64_t last_seed = 0; bool have_last = false;
  void drop_graph() { if (graph) { cudaGraphExecDestroy(graph); graph = nullptr; } }
  size_t psize(int i) const {                         // elements of tensor i in SacAdam order
    const int in_q = od + ad;
    const size_t s[SAC_TENSORS] = {(size_t)od * H, (size_t)H, (size_t)H * H, (size_t)H, (size_t)H * 2 * ad, (size_t)2 * ad,
                                   (size_t)2 * in_q * H, (size_t)2 * H, (size_t)2 * H * H, (size_t)2 * H, (size_t)2 * H, 2, 1};
    return s[i];
  }
  float* param(int i) const {
    float* const ps[SAC_TENSORS] = {piW[0], pib[0], piW[1], pib[1], piW[2], pib[2], qW[0], qb[0], qW[1], qb[1], qW[2], qb[2], log_alpha};
    return ps[i];
  }
};

MOPO_API int mopo_sac_create(int32_t obs_dim, int32_t act_dim, int32_t hidden, int32_t device, const mopo_sac_hparams* hp, mopo_sac** out) {
  REQUIRE(out && hp, MOPO_EINVAL, "mopo_sac_create: null argument");
  REQUIRE(obs_dim > 0 && act_dim > 0 && act_dim <= SAC_MAX_ACT && hidden > 0, MOPO_EINVAL, "mopo_sac_create: bad dims obs=%d act=%d hidden=%d",
          obs_dim, act_dim, hidden);
  REQUIRE(hp->policy_lr > 0.f && hp->q_lr > 0.f && hp->tau >= 0.f && hp->tau <= 1.f, MOPO_EINVAL, "mopo_sac_create: bad hyper-parameters");
  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  REQUIRE(device >= 0 && device < ndev, MOPO_EINVAL, "mopo_sac_create: device %d of %d", device, ndev);
  CUDA_TRY(cudaSetDevice(device));
  mopo_sac* s = new mopo_sac();
  s->od = obs_dim; s->ad = act_dim; s->H = hidden; s->device = device; s->hp = *hp;
  const int in_q = obs_dim + act_dim, H = hidden;
  const size_t pw[3] = {(size_t)obs_dim * H, (size_t)H * H, (size_t)H * 2 * act_dim}, pb[3] = {(size_t)H, (size_t)H, (size_t)2 * act_dim};
  const size_t qw[3] = {(size_t)2 * in_q * H, (size_t)2 * H * H, (size_t)2 * H}, qb[3] = {(size_t)2 * H, (size_t)2 * H, 2};
  for (int i = 0; i < 3; ++i) {
    CUDA_TRY(cudaMalloc(&s->piW[i], pw[i] * sizeof(float))); CUDA_TRY(cudaMalloc(&s->pib[i], pb[i] * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s->qW[i], qw[i] * sizeof(float))); CUDA_TRY(cudaMalloc(&s->qb[i], qb[i] * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s->tW[i], qw[i] * sizeof(float))); CUDA_TRY(cudaMalloc(&s->tb[i], qb[i] * sizeof(float)));
  }
  CUDA_TRY(cudaMalloc(&s->log_alpha, sizeof(float)));
  for (int i = 0; i < SAC_TENSORS; ++i) {
    float** arrs[3] = {&s->g[i], &s->m[i], &s->v[i]};
    for (float** a : arrs) { CUDA_TRY(cudaMalloc(a, s->psize(i) * sizeof(float))); CUDA_TRY(cudaMemset(*a, 0, s->psize(i) * sizeof(float))); }
  }
  CUDA_TRY(cudaMalloc(&s->t, sizeof(int64_t))); CUDA_TRY(cudaMemset(s->t, 0, sizeof(int64_t)));
  CUDA_TRY(cudaMalloc(&s->d_step, sizeof(uint64_t))); CUDA_TRY(cudaMemset(s->d_step, 0, sizeof(uint64_t)));
  CUDA_TRY(cudaMalloc(&s->lr_t, 2 * sizeof(float)));
  CUDA_TRY(cudaMalloc(&s->out, 8 * sizeof(float)));
  CUDA_TRY(cudaStreamCreateWithFlags(&s->side, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
  CUDA_TRY(cudaStreamCreateWithFlags(&s->side2, cudaStreamNonBlocking));
  for (cudaEvent_t& e : s->ev_pi) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&s->ev_fork2, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&s->ev_join2, cudaEventDisableTiming));
  *out = s;
  return MOPO_OK;
}

MOPO_API int mopo_sac_destroy(mopo_sac* s) {
  if (!s) return MOPO_OK;
  cudaSetDevice(s->device);
  for (int i = 0; i < 3; ++i) { cudaFree(s->piW[i]); cudaFree(s->pib[i]); cudaFree(s->qW[i]); cudaFree(s->qb[i]); cudaFree(s->tW[i]); cudaFree(s->tb[i]); }
  for (int i = 0; i < SAC_TENSORS; ++i) { cudaFree(s->g[i]); cudaFree(s->m[i]); cudaFree(s->v[i]); }
  s->drop_graph();
  cudaFree(s->log_alpha); cudaFree(s->t); cudaFree(s->d_step); cudaFree(s->lr_t); cudaFree(s->out);
  cudaEventDestroy(s->ev_fork); cudaEventDestroy(s->ev_join); cudaEventDestroy(s->ev_fork2); cudaEventDestroy(s->ev_join2); cudaStreamDestroy(s->side);
  for (cudaEvent_t e : s->ev_pi) cudaEventDestroy(e);
  cudaStreamDestroy(s->side2);
  DevBuf* bufs[] = {&s->ba, &s->br, &s->bt, &s->bidx, &s->Xpi, &s->H0p, &s->H1p, &s->HD, &s->eps, &s->logp, &s->Xq, &s->Xt,
                    &s->H0q, &s->H1q, &s->Qo, &s->T0, &s->T1, &s->To, &s->dQo, &s->dZ1q, &s->dZ0q, &s->dXa, &s->dHD, &s->dZ1p, &s->dZ0p};
  for (DevBuf* b : bufs) b->release();
  delete s;
  return MOPO_OK;
}

static int sac_copy(float* dst, const float* src, size_t n, cudaMemcpyKind kind, cudaStream_t st) {
  CUDA_TRY(cudaMemcpyAsync(dst, src, n * sizeof(float), kind, st));
  return MOPO_OK;
}

// Host arrays in the reference's variable layout (tf.layers.dense kernels [in, out], biases [out]):
//   pi: W0, W1, Wmu, Wls / b0, b1, bmu, bls (scope main/pi);  q1, q2: W0, W1, W2 / b0, b1, b2 (main/q1, main/q2).
// Also runs target_init (mopo.py:451-452: target <- main) and resets the optimiser state.
MOPO_API int mopo_sac_set_params(mopo_sac* s, const float* const* pi_W, const float* const* pi_b, const float* const* q1_W,
                                   const float* const* q1_b, const float* const* q2_W, const float* const* q2_b, float log_alpha,
                                   void* stream) {
  REQUIRE(s && pi_W && pi_b && q1_W && q1_b && q2_W && q2_b, MOPO_EINVAL, "mopo_sac_set_params: null argument");
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int od = s->od, ad = s->ad, H = s->H, in_q = od + ad;
  const cudaMemcpyKind h2d = cudaMemcpyHostToDevice;
  TRY(sac_copy(s->piW[0], pi_W[0], (size_t)od * H, h2d, st)); TRY(sac_copy(s->pib[0], pi_b[0], H, h2d, st));
  TRY(sac_copy(s->piW[1], pi_W[1], (size_t)H * H, h2d, st)); TRY(sac_copy(s->pib[1], pi_b[1], H, h2d, st));
  TRY(s->dZ1p.reserve(((size_t)H * ad + ad) * sizeof(float)));
  float* tmp = s->dZ1p.as<float>();
  for (int k = 0; k < 2; ++k) {                                    // mu | log_std heads side by side
    TRY(sac_copy(tmp, pi_W[2 + k], (size_t)H * ad, h2d, st)); TRY(sac_copy(tmp + (size_t)H * ad, pi_b[2 + k], ad, h2d, st));
    MOPO_COUNT_LAUNCH; sac_cols_kernel<<<32, 256, 0, st>>>(tmp, s->piW[2], H, ad, 2 * ad, k * ad, 1);
    MOPO_COUNT_LAUNCH; sac_cols_kernel<<<1, 64, 0, st>>>(tmp + (size_t)H * ad, s->pib[2], 1, ad, 2 * ad, k * ad, 1);
    CUDA_TRY(cudaStreamSynchronize(st));
  }
  const float* const* qWs[2] = {q1_W, q2_W};
  const float* const* qbs[2] = {q1_b, q2_b};
  const size_t wn[3] = {(size_t)in_q * H, (size_t)H * H, (size_t)H}, bn[3] = {(size_t)H, (size_t)H, 1};
  for (int e = 0; e < 2; ++e)
    for (int l = 0; l < 3; ++l) {
      TRY(sac_copy(s->qW[l] + e * wn[l], qWs[e][l], wn[l], h2d, st)); TRY(sac_copy(s->qb[l] + e * bn[l], qbs[e][l], bn[l], h2d, st));
    }
  for (int l = 0; l < 3; ++l) {
    TRY(sac_copy(s->tW[l], s->qW[l], 2 * wn[l], cudaMemcpyDeviceToDevice, st)); TRY(sac_copy(s->tb[l], s->qb[l], 2 * bn[l], cudaMemcpyDeviceToDevice, st));
  }
  TRY(sac_copy(s->log_alpha, &log_alpha, 1, h2d, st));
  for (int i = 0; i < SAC_TENSORS; ++i) {
    CUDA_TRY(cudaMemsetAsync(s->m[i], 0, s->psize(i) * sizeof(float), st)); CUDA_TRY(cudaMemsetAsync(s->v[i], 0, s->psize(i) * sizeof(float), st));
  }
  CUDA_TRY(cudaMemsetAsync(s->t, 0, sizeof(int64_t), st));
  CUDA_TRY(cudaMemsetAsync(s->d_step, 0, sizeof(uint64_t), st));
  CUDA_TRY(cudaStreamSynchronize(st));
  s->updates = 0;
  s->params_set = true;
  return MOPO_OK;
}

MOPO_API int mopo_sac_set_targets(mopo_sac* s, const float* const* q1t_W, const float* const* q1t_b, const float* const* q2t_W,
                                    const float* const* q2t_b, void* stream) {
  REQUIRE(s && q1t_W && q1t_b && q2t_W && q2t_b, MOPO_EINVAL, "mopo_sac_set_targets: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_set_targets: parameters not set");
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int H = s->H, in_q = s->od + s->ad;
  const size_t wn[3] = {(size_t)in_q * H, (size_t)H * H, (size_t)H}, bn[3] = {(size_t)H, (size_t)H, 1};
  const float* const* Ws[2] = {q1t_W, q2t_W};
  const float* const* bs[2] = {q1t_b, q2t_b};
  for (int e = 0; e < 2; ++e)
    for (int l = 0; l < 3; ++l) {
      REQUIRE(Ws[e][l] && bs[e][l], MOPO_EINVAL, "mopo_sac_set_targets: layer %d null", l);
      TRY(sac_copy(s->tW[l] + e * wn[l], Ws[e][l], wn[l], cudaMemcpyHostToDevice, st));
      TRY(sac_copy(s->tb[l] + e * bn[l], bs[e][l], bn[l], cudaMemcpyHostToDevice, st));
    }
  CUDA_TRY(cudaStreamSynchronize(st));
  return MOPO_OK;
}

// Inverse of mopo_sac_set_params, plus the target critics (q1t / q2t, may be NULL) and log_alpha; HOST outputs.
MOPO_API int mopo_sac_get_params(mopo_sac* s, float* const* pi_W, float* const* pi_b, float* const* q1_W, float* const* q1_b,
                                   float* const* q2_W, float* const* q2_b, float* const* q1t_W, float* const* q1t_b,
                                   float* const* q2t_W, float* const* q2t_b, float* log_alpha, void* stream) {
  REQUIRE(s && pi_W && pi_b && q1_W && q1_b && q2_W && q2_b && log_alpha, MOPO_EINVAL, "mopo_sac_get_params: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_get_params: parameters not set");
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int od = s->od, ad = s->ad, H = s->H, in_q = od + ad;
  const cudaMemcpyKind d2h = cudaMemcpyDeviceToHost;
  TRY(sac_copy(pi_W[0], s->piW[0], (size_t)od * H, d2h, st)); TRY(sac_copy(pi_b[0], s->pib[0], H, d2h, st));
  TRY(sac_copy(pi_W[1], s->piW[1], (size_t)H * H, d2h, st)); TRY(sac_copy(pi_b[1], s->pib[1], H, d2h, st));
  TRY(s->dZ1p.reserve(((size_t)H * ad + ad) * sizeof(float)));
  float* tmp = s->dZ1p.as<float>();
  for (int k = 0; k < 2; ++k) {
    MOPO_COUNT_LAUNCH; sac_cols_kernel<<<32, 256, 0, st>>>(s->piW[2], tmp, H, 2 * ad, ad, k * ad, 0);
    MOPO_COUNT_LAUNCH; sac_cols_kernel<<<1, 64, 0, st>>>(s->pib[2], tmp + (size_t)H * ad, 1, 2 * ad, ad, k * ad, 0);
    TRY(sac_copy(pi_W[2 + k], tmp, (size_t)H * ad, d2h, st)); TRY(sac_copy(pi_b[2 + k], tmp + (size_t)H * ad, ad, d2h, st));
    CUDA_TRY(cudaStreamSynchronize(st));
  }
  const size_t wn[3] = {(size_t)in_q * H, (size_t)H * H, (size_t)H}, bn[3] = {(size_t)H, (size_t)H, 1};
  float* const* outW[4] = {q1_W, q2_W, q1t_W, q2t_W};
  float* const* outb[4] = {q1_b, q2_b, q1t_b, q2t_b};
  for (int k = 0; k < 4; ++k) {
    if (!outW[k] || !outb[k]) continue;
    float* const* srcW = k < 2 ? s->qW : s->tW;
    float* const* srcb = k < 2 ? s->qb : s->tb;
    const int e = k & 1;
    for (int l = 0; l < 3; ++l) {
      TRY(sac_copy(outW[k][l], srcW[l] + e * wn[l], wn[l], d2h, st)); TRY(sac_copy(outb[k][l], srcb[l] + e * bn[l], bn[l], d2h, st));
    }
  }
  TRY(sac_copy(log_alpha, s->log_alpha, 1, d2h, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return MOPO_OK;
}

// Push the actor into the rollout's policy handle (what pulling tf variables of scope main/pi would do): device to device.
MOPO_API int mopo_sac_export_policy(mopo_sac* s, mopo_policy* p, void* stream) {
  REQUIRE(s && p, MOPO_EINVAL, "mopo_sac_export_policy: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_export_policy: parameters not set");
  REQUIRE(p->obs_dim == s->od && p->act_dim == s->ad && p->H == s->H && p->device == s->device, MOPO_EINVAL,
          "mopo_sac_export_policy: policy handle is obs %d act %d hidden %d, SAC is %d / %d / %d", p->obs_dim, p->act_dim, p->H, s->od, s->ad, s->H);
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int ad = s->ad, H = s->H;
  TRY(s->dZ0p.reserve(2 * ((size_t)H * ad + ad) * sizeof(float)));
  float* tmp = s->dZ0p.as<float>();
  const size_t blk = (size_t)H * ad + ad;
  for (int k = 0; k < 2; ++k) {
    MOPO_COUNT_LAUNCH; sac_cols_kernel<<<32, 256, 0, st>>>(s->piW[2], tmp + k * blk, H, 2 * ad, ad, k * ad, 0);
    MOPO_COUNT_LAUNCH; sac_cols_kernel<<<1, 64, 0, st>>>(s->pib[2], tmp + k * blk + (size_t)H * ad, 1, 2 * ad, ad, k * ad, 0);
  }
  CUDA_TRY(cudaGetLastError());
  const float* W[4] = {s->piW[0], s->piW[1], tmp, tmp + blk};
  const float* b[4] = {s->pib[0], s->pib[1], tmp + (size_t)H * ad, tmp + blk + (size_t)H * ad};
  return mopo_policy_set_params(p, W, b, 1, stream);
}

static int sac_reserve(mopo_sac* s, int n) {
  const int od = s->od, ad = s->ad, H = s->H, in_q = od + ad;
  const size_t N = (size_t)n;
  struct { DevBuf* b; size_t n; } want[] = {
      {&s->Xpi, 2 * N * od}, {&s->H0p, 2 * N * H}, {&s->H1p, 2 * N * H}, {&s->HD, 2 * N * 2 * ad}, {&s->eps, 2 * N * ad}, {&s->logp, 2 * N},
      {&s->Xq, 2 * N * in_q}, {&s->Xt, N * in_q}, {&s->H0q, 4 * N * H}, {&s->H1q, 4 * N * H}, {&s->Qo, 4 * N}, {&s->T0, 2 * N * H},
      {&s->T1, 2 * N * H}, {&s->To, 2 * N}, {&s->dQo, 4 * N}, {&s->dZ1q, 4 * N * H}, {&s->dZ0q, 4 * N * H}, {&s->dXa, 2 * N * ad},
      {&s->dHD, N * 2 * ad}, {&s->dZ1p, N * H}, {&s->dZ0p, N * H}};
  for (auto& w : want) {
    void* before = w.b->p;
    TRY(w.b->reserve(w.n * sizeof(float)));
    if (before != w.b->p) s->drop_graph();          // a moved buffer invalidates the captured iteration
  }
  return MOPO_OK;
}

// One _do_training step on a batch already in device memory.
static int sac_update_impl(mopo_sac* s, const float* obs, const float* act, const float* rew, const uint8_t* term, const float* next_obs,
                           int n, const float* eps_in, uint64_t seed, float* losses_out, cudaStream_t st, bool ticked = false) {
  const int od = s->od, ad = s->ad, H = s->H, in_q = od + ad;
  TRY(sac_reserve(s, n));
  float *Xpi = s->Xpi.as<float>(), *H0p = s->H0p.as<float>(), *H1p = s->H1p.as<float>(), *HD = s->HD.as<float>();
  float *Xq = s->Xq.as<float>(), *Xt = s->Xt.as<float>(), *H0q = s->H0q.as<float>(), *H1q = s->H1q.as<float>(), *Qo = s->Qo.as<float>();
  float *T0 = s->T0.as<float>(), *T1 = s->T1.as<float>(), *To = s->To.as<float>(), *dQo = s->dQo.as<float>();
  float *dZ1q = s->dZ1q.as<float>(), *dZ0q = s->dZ0q.as<float>(), *dXa = s->dXa.as<float>(), *dHD = s->dHD.as<float>();
  float *dZ1p = s->dZ1p.as<float>(), *dZ0p = s->dZ0p.as<float>();
  if (!ticked) {
    MOPO_COUNT_LAUNCH;
    sac_tick_kernel<<<1, 1, 0, st>>>(s->t, s->d_step, s->lr_t, s->hp.policy_lr, s->hp.q_lr);
  }
  // policy on [s; s']
  if (obs != Xpi) {               // (the pools path samples straight into Xpi)
    CUDA_TRY(cudaMemcpyAsync(Xpi, obs, (size_t)n * od * sizeof(float), cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(Xpi + (size_t)n * od, next_obs, (size_t)n * od * sizeof(float), cudaMemcpyDeviceToDevice, st));
  }
  const int n2 = 2 * n;
  gemm<4>(GemmP{Xpi, 0, od, 0, s->piW[0], 0, H, 0, H0p, 0, H, n2, H, od, s->pib[0], nullptr, nullptr}, 1, st);
  gemm<4>(GemmP{H0p, 0, H, 0, s->piW[1], 0, H, 0, H1p, 0, H, n2, H, H, s->pib[1], nullptr, nullptr}, 1, st);
  gemm<2>(GemmP{H1p, 0, H, 0, s->piW[2], 0, 2 * ad, 0, HD, 0, 2 * ad, n2, 2 * ad, H, s->pib[2], nullptr, nullptr}, 1, st);
  MOPO_COUNT_LAUNCH;
  sac_pi_head_kernel<<<(n2 + 127) / 128, 128, 0, st>>>(HD, eps_in, seed, s->d_step, obs, act, next_obs, n, od, ad, s->eps.as<float>(),
                                                       s->logp.as<float>(), Xq, Xt);
  // twin critics on [(s, a); (s, pi(s))] and their targets on (s', pi(s')): nets batched over blockIdx.z, inputs shared
  const int64_t sH2 = (int64_t)n2 * H, sH1 = (int64_t)n * H;
  CUDA_TRY(cudaEventRecord(s->ev_fork2, st));
  CUDA_TRY(cudaStreamWaitEvent(s->side, s->ev_fork2, 0));
  gemm<4>(GemmP{Xt, 0, in_q, 0, s->tW[0], (int64_t)in_q * H, H, 0, T0, sH1, H, n, H, in_q, s->tb[0], nullptr, nullptr}, 2, s->side);
  gemm<4>(GemmP{T0, sH1, H, 0, s->tW[1], (int64_t)H * H, H, 0, T1, sH1, H, n, H, H, s->tb[1], nullptr, nullptr}, 2, s->side);
  gemm<2>(GemmP{T1, sH1, H, 0, s->tW[2], (int64_t)H, 1, 0, To, n, 1, n, 1, H, s->tb[2], nullptr, nullptr}, 2, s->side);
  CUDA_TRY(cudaEventRecord(s->ev_join2, s->side));
  gemm<4>(GemmP{Xq, 0, in_q, 0, s->qW[0], (int64_t)in_q * H, H, 0, H0q, sH2, H, n2, H, in_q, s->qb[0], nullptr, nullptr}, 2, st);
  gemm<4>(GemmP{H0q, sH2, H, 0, s->qW[1], (int64_t)H * H, H, 0, H1q, sH2, H, n2, H, H, s->qb[1], nullptr, nullptr}, 2, st);
  gemm<2>(GemmP{H1q, sH2, H, 0, s->qW[2], (int64_t)H, 1, 0, Qo, n2, 1, n2, 1, H, s->qb[2], nullptr, nullptr}, 2, st);
  CUDA_TRY(cudaStreamWaitEvent(st, s->ev_join2, 0));            // target critics: beside the main critics (see above)
  MOPO_COUNT_LAUNCH;
  sac_loss_kernel<<<1, 256, 0, st>>>(Qo, To, s->logp.as<float>(), rew, term, s->log_alpha, n, s->hp.reward_scale, s->hp.discount,
                                     s->hp.target_entropy, dQo, s->out, s->g[12]);
  if (losses_out) CUDA_TRY(cudaMemcpyAsync(losses_out, s->out, 4 * sizeof(float), cudaMemcpyDeviceToDevice, st));
  // critic backward over all 2n rows (data rows feed the critic gradients, pi rows the action gradient)
  gemm<5>(GemmP{dQo, n2, 1, 0, s->qW[2], (int64_t)H, 1, 1, dZ1q, sH2, H, n2, H, 1, nullptr, nullptr, H1q}, 2, st);
  gemm<5>(GemmP{dZ1q, sH2, H, 0, s->qW[1], (int64_t)H * H, H, 1, dZ0q, sH2, H, n2, H, H, nullptr, nullptr, H0q}, 2, st);
  // fork: critic parameter gradients (first n rows of every net) on the side stream
  CUDA_TRY(cudaEventRecord(s->ev_fork, st));
  CUDA_TRY(cudaStreamWaitEvent(s->side, s->ev_fork, 0));
  cudaStream_t sd = s->side;
  gemm<0>(GemmP{H1q, sH2, H, 1, dQo, n2, 1, 0, s->g[10], (int64_t)H, 1, H, 1, n, nullptr, nullptr, nullptr}, 2, sd);
  MOPO_COUNT_LAUNCH; colsum_kernel<<<dim3(1, 2), 256, 0, sd>>>(dQo, s->g[11], n, 1, (int64_t)n2);
  gemm<0>(GemmP{H0q, sH2, H, 1, dZ1q, sH2, H, 0, s->g[8], (int64_t)H * H, H, H, H, n, nullptr, nullptr, nullptr}, 2, sd);
  MOPO_COUNT_LAUNCH; colsum_kernel<<<dim3((H + 31) / 32, 2), 256, 0, sd>>>(dZ1q, s->g[9], n, H, sH2);
  gemm<0>(GemmP{Xq, 0, in_q, 1, dZ0q, sH2, H, 0, s->g[6], (int64_t)in_q * H, H, in_q, H, n, nullptr, nullptr, nullptr}, 2, sd);
  MOPO_COUNT_LAUNCH; colsum_kernel<<<dim3((H + 31) / 32, 2), 256, 0, sd>>>(dZ0q, s->g[7], n, H, sH2);
  CUDA_TRY(cudaEventRecord(s->ev_join, sd));
  // main: d(-min q)/d a of the pi rows -> policy head gradient -> policy backward
  gemm<0>(GemmP{dZ0q + (size_t)n * H, sH2, H, 0, s->qW[0] + (size_t)od * H, (int64_t)in_q * H, H, 1, dXa, (int64_t)n * ad, ad, n, ad, H,
                nullptr, nullptr, nullptr}, 2, st);
  MOPO_COUNT_LAUNCH;
  sac_pi_head_grad_kernel<<<(n + 127) / 128, 128, 0, st>>>(HD, s->eps.as<float>(), dXa, s->log_alpha, n, ad, dHD);
  // actor backward: the chain dHD -> dZ1p -> dZ0p stays on the main stream, each stage's weight / bias gradients go to side2
  cudaStream_t s2 = s->side2;
  CUDA_TRY(cudaEventRecord(s->ev_pi[0], st));
  gemm<5>(GemmP{dHD, 0, 2 * ad, 0, s->piW[2], 0, 2 * ad, 1, dZ1p, 0, H, n, H, 2 * ad, nullptr, nullptr, H1p}, 1, st);
  CUDA_TRY(cudaEventRecord(s->ev_pi[1], st));
  gemm<5>(GemmP{dZ1p, 0, H, 0, s->piW[1], 0, H, 1, dZ0p, 0, H, n, H, H, nullptr, nullptr, H0p}, 1, st);
  CUDA_TRY(cudaEventRecord(s->ev_pi[2], st));
  CUDA_TRY(cudaStreamWaitEvent(s2, s->ev_pi[0], 0));
  gemm<0>(GemmP{H1p, 0, H, 1, dHD, 0, 2 * ad, 0, s->g[4], 0, 2 * ad, H, 2 * ad, n, nullptr, nullptr, nullptr}, 1, s2);
  MOPO_COUNT_LAUNCH; colsum_kernel<<<dim3((2 * ad + 31) / 32, 1), 256, 0, s2>>>(dHD, s->g[5], n, 2 * ad, 0);
  CUDA_TRY(cudaStreamWaitEvent(s2, s->ev_pi[1], 0));
  gemm<0>(GemmP{H0p, 0, H, 1, dZ1p, 0, H, 0, s->g[2], 0, H, H, H, n, nullptr, nullptr, nullptr}, 1, s2);
  MOPO_COUNT_LAUNCH; colsum_kernel<<<dim3((H + 31) / 32, 1), 256, 0, s2>>>(dZ1p, s->g[3], n, H, 0);
  CUDA_TRY(cudaStreamWaitEvent(s2, s->ev_pi[2], 0));
  gemm<0>(GemmP{Xpi, 0, od, 1, dZ0p, 0, H, 0, s->g[0], 0, H, od, H, n, nullptr, nullptr, nullptr}, 1, s2);
  MOPO_COUNT_LAUNCH; colsum_kernel<<<dim3((H + 31) / 32, 1), 256, 0, s2>>>(dZ0p, s->g[1], n, H, 0);
  CUDA_TRY(cudaEventRecord(s->ev_pi[3], s2));
  CUDA_TRY(cudaStreamWaitEvent(st, s->ev_pi[3], 0));
  CUDA_TRY(cudaStreamWaitEvent(st, s->ev_join, 0));
  // the four Adam optimisers in one launch, then the Polyak update (every target_update_interval-th update, mopo.py:839-841)
  SacAdam ab;
  size_t biggest = 0;
  for (int i = 0; i < SAC_TENSORS; ++i) {
    ab.p[i] = s->param(i); ab.g[i] = s->g[i]; ab.m[i] = s->m[i]; ab.v[i] = s->v[i]; ab.n[i] = (int64_t)s->psize(i);
    ab.which_lr[i] = (i < 6 || i == 12) ? 0 : 1;
    if (s->psize(i) > biggest) biggest = s->psize(i);
  }
  ab.lr_t = s->lr_t;
  int gx = (int)((biggest + 1023) / 1024); if (gx > 148) gx = 148;
  MOPO_COUNT_LAUNCH;
  sac_adam_kernel<<<dim3(gx, SAC_TENSORS), 256, 0, st>>>(ab);
  const int interval = s->hp.target_update_interval > 0 ? s->hp.target_update_interval : 1;
  if (s->updates % interval == 0) {
    SacPolyak pk;
    for (int l = 0; l < 3; ++l) {
      pk.targ[2 * l] = s->tW[l]; pk.main[2 * l] = s->qW[l]; pk.n[2 * l] = (int64_t)s->psize(6 + 2 * l);
      pk.targ[2 * l + 1] = s->tb[l]; pk.main[2 * l + 1] = s->qb[l]; pk.n[2 * l + 1] = (int64_t)s->psize(7 + 2 * l);
    }
    pk.tau = s->hp.tau;
    MOPO_COUNT_LAUNCH;
    sac_polyak_kernel<<<dim3(gx, 6), 256, 0, st>>>(pk);
  }
  s->updates += 1;
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

MOPO_API int mopo_sac_update(mopo_sac* s, const mopo_sac_batch* b, void* stream) {
  REQUIRE(s && b, MOPO_EINVAL, "mopo_sac_update: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_update: parameters not set");
  REQUIRE(b->obs && b->act && b->rew && b->term && b->next_obs, MOPO_EINVAL, "mopo_sac_update: null batch field");
  REQUIRE(b->n > 0 && b->n <= (1 << 20), MOPO_EINVAL, "mopo_sac_update: n=%lld", (long long)b->n);
  CUDA_TRY(cudaSetDevice(s->device));
  return sac_update_impl(s, b->obs, b->act, b->rew, b->term, b->next_obs, (int)b->n, b->eps, b->seed, b->losses_out, (cudaStream_t)stream);
}

// n_updates x { _training_batch (mopo.py:801-821): n_env rows of the env pool then n_model rows of the model pool, drawn
// on the device (Philox stream 2, offset = update index); one update }.  losses_out: optional device float[n_updates][4].
MOPO_API int mopo_sac_update_from_pools(mopo_sac* s, const mopo_pool* env_pool, const mopo_pool* model_pool, int32_t n_env,
                                          int32_t n_model, uint64_t seed, int32_t n_updates, float* losses_out, void* stream) {
  REQUIRE(s && env_pool, MOPO_EINVAL, "mopo_sac_update_from_pools: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_update_from_pools: parameters not set");
  REQUIRE(n_env >= 0 && n_model >= 0 && n_env + n_model > 0 && n_updates >= 0, MOPO_EINVAL, "mopo_sac_update_from_pools: n_env=%d n_model=%d", n_env, n_model);
  REQUIRE(n_model == 0 || model_pool, MOPO_EINVAL, "mopo_sac_update_from_pools: model rows requested without a model pool");
  REQUIRE(env_pool->obs_dim == s->od && env_pool->act_dim == s->ad && (!model_pool || (model_pool->obs_dim == s->od && model_pool->act_dim == s->ad)),
          MOPO_EINVAL, "mopo_sac_update_from_pools: pool dims differ from the SAC networks'");
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int n = n_env + n_model, od = s->od, ad = s->ad;
  TRY(sac_reserve(s, n));
  TRY(s->ba.reserve((size_t)n * ad * sizeof(float)));
  TRY(s->br.reserve((size_t)n * sizeof(float))); TRY(s->bt.reserve((size_t)n));
  TRY(s->bidx.reserve((size_t)n * sizeof(int64_t)));
  // observations / next observations are sampled straight into the actor's stacked input [s; s']
  float *bo = s->Xpi.as<float>(), *ba = s->ba.as<float>(), *br = s->br.as<float>(), *bno = s->Xpi.as<float>() + (size_t)n * od;
  uint8_t* bt = s->bt.as<uint8_t>();
  int64_t* bidx = s->bidx.as<int64_t>();
  REQUIRE((n_env == 0 || env_pool->size > 0) && (n_model == 0 || model_pool->size > 0), MOPO_ESTATE, "mopo_sac_update_from_pools: pool is empty");
  const mopo_pool* pools[2] = {env_pool, n_model ? model_pool : env_pool};
  SacSample smp;
  memset(&smp, 0, sizeof(smp));                     // compared bytewise as the graph's key
  for (int w = 0; w < 2; ++w) {
    smp.obs[w] = pools[w]->obs; smp.act[w] = pools[w]->actions; smp.nobs[w] = pools[w]->next_obs; smp.rew[w] = pools[w]->rewards;
    smp.term[w] = pools[w]->terminals; smp.size[w] = pools[w]->size;
  }
  smp.n_env = n_env; smp.n = n; smp.od = od; smp.ad = ad;
  const int interval = s->hp.target_update_interval > 0 ? s->hp.target_update_interval : 1;
  // one iteration: tick (update index, lr_t) -> draw env rows, draw model rows (offsets 2u, 2u + 1 read on the device) -> update
  auto iteration = [&](float* lo) -> int {
    MOPO_COUNT_LAUNCH;
    sac_tick_kernel<<<1, 1, 0, st>>>(s->t, s->d_step, s->lr_t, s->hp.policy_lr, s->hp.q_lr);
    MOPO_COUNT_LAUNCH;
    sac_sample_kernel<<<(n * (2 * od + ad + 2) + 255) / 256, 256, 0, st>>>(smp, seed, s->d_step, bo, ba, bno, br, bt, bidx);
    return sac_update_impl(s, bo, ba, br, bt, bno, n, nullptr, seed, lo, st, true);
  };
  int k = 0;
  // ~45 small launches per update: capture the iteration once per call and replay it (needs a capturable stream, no
  // per-update host-side variation: Polyak every update, no loss read-back)
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (st) CUDA_TRY(cudaStreamIsCapturing(st, &cap));
  const bool can_graph = st && cap == cudaStreamCaptureStatusNone && !losses_out && interval == 1 && n_updates > 0;
  const bool key_match = s->graph && s->g_seed == seed && memcmp(&s->g_smp, &smp, sizeof(smp)) == 0;
  const bool recurring = s->have_last && s->last_seed == seed && memcmp(&s->last_smp, &smp, sizeof(smp)) == 0;
  s->last_smp = smp; s->last_seed = seed; s->have_last = true;
  if (can_graph && (key_match || n_updates >= 8 || recurring)) {
    TRY(sac_reserve(s, n));
    TRY(s->ba.reserve((size_t)n * ad * sizeof(float)));
    bool fresh = false;
    if (!(s->graph && s->g_seed == seed && memcmp(&s->g_smp, &smp, sizeof(smp)) == 0)) {
      s->drop_graph();
      cudaGraph_t g = nullptr;
      const long long before = (long long)mopo_launch_count();
      CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      int rc = iteration(nullptr);
      cudaError_t ce = cudaStreamEndCapture(st, &g);
      s->updates -= 1;                                // the capture enqueued nothing
      if (rc != MOPO_OK) { if (g) cudaGraphDestroy(g); return rc; }
      CUDA_TRY(ce);
      s->g_nodes = (int)((long long)mopo_launch_count() - before);
      cudaError_t ci = cudaGraphInstantiate(&s->graph, g, 0);
      cudaGraphDestroy(g);
      if (ci != cudaSuccess) s->graph = nullptr;
      CUDA_TRY(ci);
      s->g_smp = smp; s->g_seed = seed;
      fresh = true;                                   // the launches counted during the capture stand for the first replay
    }
    for (; k < n_updates; ++k) {
      CUDA_TRY(cudaGraphLaunch(s->graph, st));
      if (!(fresh && k == 0)) for (int i = 0; i < s->g_nodes; ++i) mopo_count_launch();
      s->updates += 1;
    }
    return MOPO_OK;
  }
  for (; k < n_updates; ++k) TRY(iteration(losses_out ? losses_out + 4 * k : nullptr));
  return MOPO_OK;
}
