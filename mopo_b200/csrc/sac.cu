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

// ---- the update as ONE cooperative kernel --------------------------------------------------------------------------------------
// Stages separated by grid barriers; inside a stage every CTA takes work units (32 x 32 GEMM tiles with a 4-way k split inside the
// CTA, bias-gradient column sums, 256-row blocks of the per-row maps) round robin over all the stage's ops, so the weight gradients
// run beside the next link of the data-gradient chain.
struct SacSample {
  const float *obs[2], *act[2], *nobs[2], *rew[2]; const uint8_t* term[2];
  int64_t size[2]; int n_env, n, od, ad;
};
constexpr int SAC_TENSORS = 13;   // pi 3 W + 3 b, q 3 W + 3 b (twins stacked), log_alpha
// the update's program: ops in stage order, a grid barrier after every stage (the host writes it, sac_program below)
enum { OP_GEMM = 0, OP_COLSUM, OP_OUTER, OP_SAMPLE, OP_PI_HEAD, OP_LOSS, OP_PI_GRAD };
struct FOp { GemmP g; int kind, epi, stage, units, tn, tmn, flags; };
constexpr int SAC_MAX_OPS = 40;
struct SacFused {
  int n, od, ad, adp, H, sample, polyak;        // adp: lanes per policy row (power of two >= ad)
  SacSample smp; uint64_t seed;
  int64_t* t; uint64_t* d_step; float* lr_t; float lr_pi, lr_q;
  float reward_scale, discount, target_entropy, tau;
  const float* eps_in;
  const float* act; const float* rew; const uint8_t* term;          // batch fields (the pools path: ba / br / bt, written by stage 0)
  float* ba; float* br; uint8_t* bt; int64_t* bidx;
  float *piW[3], *pib[3], *qW[3], *qb[3], *tW[3], *tb[3], *log_alpha;
  float *g[SAC_TENSORS], *m[SAC_TENSORS], *v[SAC_TENSORS]; int64_t psize[SAC_TENSORS];
  float *par[SAC_TENSORS], *targ[SAC_TENSORS]; int64_t cpref[SAC_TENSORS + 1];   // variables, their Polyak targets (or null), 4-element chunk prefix
  float *Xpi, *H0p, *H1p, *HD, *eps, *logp, *Xq, *Xt, *H0q, *H1q, *Qo, *T0, *T1, *To, *dQo, *dZ1q, *dZ0q, *dXa, *dHD, *dZ1p, *dZ0p;
  float* out;
  unsigned long long* trace;                   // null, or where CTA 0 records %globaltimer at its start and after every barrier
  unsigned long long* bar;                     // grid barrier counter, device-only state (gemm_fp32.cuh grid_bar)
  int n_ops; FOp ops[SAC_MAX_OPS];
};

// _training_batch (mopo.py:801-821): rows [0, n_env) from the env pool, [n_env, n) from the model pool, indices as
// mopo_pool_sample draws them (Philox stream 2, counter = (row, offset), offset = 2 (update index) + pool: every update and
// each pool has its own counter plane, so consecutive minibatches are independent draws) -- all five fields.
__device__ __forceinline__ void sac_sample(const SacFused& a, uint64_t step) {
  const SacSample& sp = a.smp;
  const int W = 2 * sp.od + sp.ad + 2;
  const uint64_t base = 2 * (step - 1);
  float* bo = a.Xpi; float* bno = a.Xpi + (size_t)sp.n * sp.od;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < sp.n * W; i += gridDim.x * blockDim.x) {
    const int j = i / W, c = i % W;
    const int w = j >= sp.n_env;
    Philox4 r = philox_pool_draw((uint32_t)(w ? j - sp.n_env : j), base + (uint64_t)w, a.seed);
    const int64_t idx = (int64_t)__umul64hi(((uint64_t)r.x << 32) | r.y, (uint64_t)sp.size[w]);
    if (c < sp.od) bo[(int64_t)j * sp.od + c] = sp.obs[w][idx * sp.od + c];
    else if (c < 2 * sp.od) bno[(int64_t)j * sp.od + (c - sp.od)] = sp.nobs[w][idx * sp.od + (c - sp.od)];
    else if (c < 2 * sp.od + sp.ad) a.ba[(int64_t)j * sp.ad + (c - 2 * sp.od)] = sp.act[w][idx * sp.ad + (c - 2 * sp.od)];
    else if (c == W - 2) { a.br[j] = sp.rew[w][idx]; a.bidx[j] = idx; }
    else a.bt[j] = sp.term[w][idx];
  }
}

// adp lanes (a power of two >= act_dim) per policy row r of [s; s'] (2n rows), lane d owns action dimension d: head
// [mu | log_std] -> u = mu + eps std, a = tanh(u), logp (gaussian_likelihood + squashing correction, mopo.py:281-296); writes the
// critic inputs:   r <  n: Xq[r] = (s_r, a_data_r),  Xq[n + r] = (s_r, a)      r >= n: Xt[r - n] = (s'_r, a)
__device__ __forceinline__ void sac_pi_head(const SacFused& a, uint64_t step64, int unit) {
  const int n = a.n, od = a.od, ad = a.ad, adp = a.adp;
  const int r = unit * (FUSED_THREADS / adp) + (int)threadIdx.x / adp, d = (int)threadIdx.x % adp;
  const bool on = r < 2 * n && d < ad;
  const int in_q = od + ad;
  float lp = 0.f;
  if (r < 2 * n) {
    const float* s = a.Xpi + (int64_t)r * od;                        // [s; s'] stacked
    float* xo = r < n ? a.Xq + (int64_t)(n + r) * in_q : a.Xt + (int64_t)(r - n) * in_q;
    float* xd = a.Xq + (int64_t)r * in_q;
    for (int k = d; k < od; k += adp) { const float sv = s[k]; xo[k] = sv; if (r < n) xd[k] = sv; }
    if (on) {
      float z;
      if (a.eps_in) z = a.eps_in[(int64_t)r * ad + d];
      else {
        float gg[4];
        philox_normal4((uint32_t)r, (uint32_t)step64, (uint32_t)(d >> 2), a.seed, gg, MOPO_STREAM_SAC);
        z = gg[d & 3];
      }
      if (r < n) xd[od + d] = a.act[(int64_t)r * ad + d];
      const float mu = __ldcg(a.HD + (int64_t)r * 2 * ad + d);
      const float ls = fminf(fmaxf(__ldcg(a.HD + (int64_t)r * 2 * ad + ad + d), LOG_STD_MIN), LOG_STD_MAX);
      const float sd = expf(ls);
      const float u = mu + z * sd;
      const float q = z * sd / (sd + PI_EPS);        // (u - mu) / (std + EPS) with u - mu = eps std exactly
      lp = -0.5f * (q * q + 2.0f * ls + 1.8378770664093453f) - 2.0f * (0.6931471805599453f - u - softplus_f(-2.0f * u));
      xo[od + d] = tanhf(u);
      a.eps[(int64_t)r * ad + d] = z;
    }
  }
  for (int o = adp >> 1; o > 0; o >>= 1) lp += __shfl_xor_sync(0xffffffffu, lp, o);   // over the row's lanes, fixed order
  if (r < 2 * n && d == 0) a.logp[r] = lp;
}

// One CTA: q targets, dQ for the stacked critic rows, the four losses and d alpha_loss / d log_alpha.
//   Qo [2, 2n]: twin e, rows [0,n) = q_e(s, a), [n,2n) = q_e(s, pi(s));  To [2, n] = target twins on (s', pi(s'))
//   out[0..3] = q1_loss, q2_loss, policy_loss, alpha_loss;  out[4] = d alpha_loss / d log_alpha
__device__ __forceinline__ void sac_loss(const SacFused& a, float* sh) {
  const int n = a.n;
  const float *Qo = a.Qo, *To = a.To, *logp = a.logp;
  float* dQo = a.dQo;
  const float la = *a.log_alpha, alpha = expf(la), inv_n = 1.0f / (float)n;
  float s1 = 0.f, s2 = 0.f, sp = 0.f, sa = 0.f;
  for (int b = threadIdx.x; b < n; b += blockDim.x) {
    const float next_value = fminf(To[b], To[n + b]) - alpha * logp[n + b];
    const float y = a.reward_scale * a.rew[b] + a.discount * (a.term[b] ? 0.0f : 1.0f) * next_value;
    const float q1 = Qo[b], q2 = Qo[2 * n + b], q1p = Qo[n + b], q2p = Qo[3 * n + b];
    dQo[b] = (q1 - y) * inv_n;
    dQo[2 * n + b] = (q2 - y) * inv_n;
    const bool first = q1p <= q2p;                 // tf.minimum sends a tie's gradient to its first argument
    dQo[n + b] = first ? -inv_n : 0.0f;
    dQo[3 * n + b] = first ? 0.0f : -inv_n;
    s1 += (q1 - y) * (q1 - y);
    s2 += (q2 - y) * (q2 - y);
    sp += alpha * logp[b] - fminf(q1p, q2p);
    sa += logp[b] + a.target_entropy;
  }
  s1 = block_sum(s1, sh); s2 = block_sum(s2, sh); sp = block_sum(sp, sh); sa = block_sum(sa, sh);
  if (threadIdx.x == 0) {
    a.out[0] = 0.5f * s1 * inv_n; a.out[1] = 0.5f * s2 * inv_n; a.out[2] = sp * inv_n; a.out[3] = -la * sa * inv_n; a.out[4] = -sa * inv_n;
    *a.g[12] = -sa * inv_n;
  }
}

// d policy_loss / d [mu | log_std_raw], one thread per (row b of s, action dimension d); dXa [2, n, ad] = d(-min q)/d a per twin.
__device__ __forceinline__ void sac_pi_head_grad(const SacFused& a, int i) {
  const int n = a.n, ad = a.ad;
  if (i >= n * ad) return;
  const int b = i / ad, d = i - b * ad;
  const float c = expf(*a.log_alpha) / (float)n;
  const float mu = a.HD[(int64_t)b * 2 * ad + d], ls_raw = a.HD[(int64_t)b * 2 * ad + ad + d];
  const float ls = fminf(fmaxf(ls_raw, LOG_STD_MIN), LOG_STD_MAX);
  const float sd = expf(ls), z = a.eps[i];
  const float u = mu + z * sd, av = tanhf(u);
  const float den = sd + PI_EPS;
  const float dev = z * sd;                          // u - mu
  const float r2 = dev / (den * den);
  const float d_a = __ldcg(a.dXa + i) + __ldcg(a.dXa + (int64_t)n * ad + i);
  const float g_u = c * (-r2 + 2.0f * av) + d_a * (1.0f - av * av);
  const float d_mu = g_u + c * r2;
  float d_ls = g_u * (z * sd) + c * (dev * dev * sd / (den * den * den) - 1.0f);
  if (ls_raw < LOG_STD_MIN || ls_raw > LOG_STD_MAX) d_ls = 0.0f;      // tf.clip_by_value passes no gradient outside
  a.dHD[(int64_t)b * 2 * ad + d] = d_mu;
  a.dHD[(int64_t)b * 2 * ad + ad + d] = d_ls;
}

// C[e, m, n] = x[e, m] w[e, n] (Zin[e, m, n] > 0): the data gradient through a one-unit head (K = 1) as an elementwise map;
// GemmP fields: A = x (stride sA per net), B = w (stride sB), C / Zin [E, M, N] contiguous rows (ldc = N), four columns per thread
__device__ __forceinline__ void outer_unit(const GemmP& p, int u) {
  const int nq = p.N >> 2;                                   // host guarantees N % 4 == 0 and 16-byte alignment
  const int64_t i = (int64_t)u * FUSED_THREADS + threadIdx.x;
  const int64_t rows = i / nq; const int c = (int)(i - rows * nq) * 4;
  const int e = (int)(rows / p.M); const int m = (int)(rows - (int64_t)e * p.M);
  if (e >= p.K) return;                                      // (K carries the net count here)
  const float x = __ldcg(p.A + (int64_t)e * p.sA + m);
  const float4 w = __ldcg(reinterpret_cast<const float4*>(p.B + (int64_t)e * p.sB + c));
  const int64_t o = (int64_t)e * p.sC + (int64_t)m * p.ldc + c;
  const float4 z = __ldcg(reinterpret_cast<const float4*>(p.Zin + o));
  *reinterpret_cast<float4*>(p.C + o) = make_float4(z.x > 0.f ? x * w.x : 0.f, z.y > 0.f ? x * w.y : 0.f, z.z > 0.f ? x * w.z : 0.f, z.w > 0.f ? x * w.w : 0.f);
}

// units [off, off + count) of a stage go to the CTAs (one per SM) round robin; f(u) runs unit u of this op on the calling CTA
template <class F>
__device__ __forceinline__ void stage_units(int& off, int count, F f) {
  const int G = gridDim.x;
  for (int u = ((int)blockIdx.x - off % G + G) % G; u < count; u += G) f(u);
  off += count;
}

__global__ void __launch_bounds__(FUSED_THREADS, 1) sac_fused_kernel(const __grid_constant__ SacFused a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FusedSmem& sm = *reinterpret_cast<FusedSmem*>(smem_raw);
  __shared__ float s_lr[2];
  const int64_t k_adam = *a.t + 1;                 // tf.train.AdamOptimizer's step being applied
  const uint64_t step = *a.d_step + 1;             // Philox step id of this update's draws
  if (threadIdx.x == 0) {                          // lr_t of both optimiser families (sqrt(1 - b2^k) / (1 - b1^k))
    const double c = sqrt(1.0 - pow((double)ADAM_B2, (double)k_adam)) / (1.0 - pow((double)ADAM_B1, (double)k_adam));
    s_lr[0] = (float)((double)a.lr_pi * c); s_lr[1] = (float)((double)a.lr_q * c);
  }
  unsigned long long bar_target = grid_bar_begin(a.bar);
  int tr_i = 0;
  auto TRC = [&]() { if (a.trace && blockIdx.x == 0 && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); a.trace[tr_i] = t; } ++tr_i; };
  TRC();
  int i = 0;
  while (i < a.n_ops) {
    const int stage = a.ops[i].stage;
    int off = 0;
    for (; i < a.n_ops && a.ops[i].stage == stage; ++i) {
      const FOp& op = a.ops[i];
      switch (op.kind) {
        case OP_GEMM: stage_units(off, op.units, [&](int u) { tile_unit(sm, op.g, op.epi, op.flags, op.tn, op.tmn, u); }); break;
        case OP_COLSUM: stage_units(off, op.units, [&](int u) { colsum_unit(sm, op.g.A, op.g.C, op.g.M, op.g.N, op.g.sA, u); }); break;
        case OP_OUTER: stage_units(off, op.units, [&](int u) { outer_unit(op.g, u); }); break;
        case OP_SAMPLE: sac_sample(a, step); break;
        case OP_PI_HEAD: stage_units(off, op.units, [&](int u) { sac_pi_head(a, step, u); }); break;
        case OP_LOSS: stage_units(off, 1, [&](int) { sac_loss(a, &sm.Bs[0][0][0]); }); break;
        case OP_PI_GRAD: stage_units(off, op.units, [&](int u) { sac_pi_head_grad(a, u * FUSED_THREADS + (int)threadIdx.x); }); break;
        default: break;
      }
    }
    grid_bar(a.bar, bar_target); TRC();
    if (stage == 0 && blockIdx.x == 0 && threadIdx.x == 0) { *a.t = k_adam; *a.d_step = step; a.lr_t[0] = s_lr[0]; a.lr_t[1] = s_lr[1]; }
  }
  // last: the four Adam optimisers, and target <- (1 - tau) target + tau main on the freshly written critic variables; all 13
  // tensors as one list of 4-element chunks, so that a thread's loads of one chunk are the only round trip
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gn = (int64_t)gridDim.x * blockDim.x;
  for (int64_t c = gtid; c < a.cpref[SAC_TENSORS]; c += gn) {
    int t = 0;
    while (c >= a.cpref[t + 1]) ++t;
    const int64_t j0 = (c - a.cpref[t]) * 4;
    const int cnt = (int)min((int64_t)4, a.psize[t] - j0);
    const float lr_t = s_lr[(t < 6 || t == 12) ? 0 : 1];
    float* p = a.par[t] + j0; float* targ = a.polyak && a.targ[t] ? a.targ[t] + j0 : nullptr;
    const float* g = a.g[t] + j0; float* m = a.m[t] + j0; float* v = a.v[t] + j0;
    float gi[4], mi[4], vi[4], pi[4], ti[4] = {0.f, 0.f, 0.f, 0.f};
    if (cnt == 4) {
      const float4 g4 = *reinterpret_cast<const float4*>(g), m4 = *reinterpret_cast<const float4*>(m), v4 = *reinterpret_cast<const float4*>(v),
                   p4 = *reinterpret_cast<const float4*>(p);
      gi[0] = g4.x; gi[1] = g4.y; gi[2] = g4.z; gi[3] = g4.w; mi[0] = m4.x; mi[1] = m4.y; mi[2] = m4.z; mi[3] = m4.w;
      vi[0] = v4.x; vi[1] = v4.y; vi[2] = v4.z; vi[3] = v4.w; pi[0] = p4.x; pi[1] = p4.y; pi[2] = p4.z; pi[3] = p4.w;
      if (targ) { const float4 t4 = *reinterpret_cast<const float4*>(targ); ti[0] = t4.x; ti[1] = t4.y; ti[2] = t4.z; ti[3] = t4.w; }
    } else {
      for (int j = 0; j < cnt; ++j) { gi[j] = g[j]; mi[j] = m[j]; vi[j] = v[j]; pi[j] = p[j]; if (targ) ti[j] = targ[j]; }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mi[j] = ADAM_B1 * mi[j] + (1.0f - ADAM_B1) * gi[j];
      vi[j] = ADAM_B2 * vi[j] + (1.0f - ADAM_B2) * gi[j] * gi[j];
      pi[j] = pi[j] - lr_t * mi[j] / (sqrtf(vi[j]) + ADAM_EPS);
      ti[j] = (1.0f - a.tau) * ti[j] + a.tau * pi[j];
    }
    if (cnt == 4) {
      *reinterpret_cast<float4*>(m) = make_float4(mi[0], mi[1], mi[2], mi[3]); *reinterpret_cast<float4*>(v) = make_float4(vi[0], vi[1], vi[2], vi[3]);
      *reinterpret_cast<float4*>(p) = make_float4(pi[0], pi[1], pi[2], pi[3]);
      if (targ) *reinterpret_cast<float4*>(targ) = make_float4(ti[0], ti[1], ti[2], ti[3]);
    } else {
      for (int j = 0; j < cnt; ++j) { m[j] = mi[j]; v[j] = vi[j]; p[j] = pi[j]; if (targ) targ[j] = ti[j]; }
    }
  }
  TRC();
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
  DevBuf ba, br, bt, bidx;                   // batch drawn from the pools
  DevBuf Xpi, H0p, H1p, HD, eps, logp, Xq, Xt, H0q, H1q, Qo, T0, T1, To, dQo, dZ1q, dZ0q, dXa, dHD, dZ1p, dZ0p;
  uint64_t* d_step = nullptr;                         // device Philox step id of the policy noise (advanced by the update kernel)
  int grid = 0;                                       // CTAs of sac_fused_kernel: one per SM (cooperative launch)
  unsigned long long* bar = nullptr;                  // its grid barrier counter (device-only state)
  unsigned long long* trace = nullptr;                // probe build: device buffer for CTA 0's globaltimer after every stage
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
  int per_sm = 0, sms = 0;
  CUDA_TRY(cudaFuncSetAttribute(sac_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FusedSmem)));
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sac_fused_kernel, FUSED_THREADS, sizeof(FusedSmem)));
  CUDA_TRY(cudaMalloc(&s->bar, sizeof(unsigned long long))); CUDA_TRY(cudaMemset(s->bar, 0, sizeof(unsigned long long)));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  REQUIRE(per_sm > 0 && sms > 0, MOPO_ECUDA, "mopo_sac_create: the update kernel does not fit on an SM");
  s->grid = sms;
  *out = s;
  return MOPO_OK;
}

MOPO_API int mopo_sac_destroy(mopo_sac* s) {
  if (!s) return MOPO_OK;
  cudaSetDevice(s->device);
  for (int i = 0; i < 3; ++i) { cudaFree(s->piW[i]); cudaFree(s->pib[i]); cudaFree(s->qW[i]); cudaFree(s->qb[i]); cudaFree(s->tW[i]); cudaFree(s->tb[i]); }
  for (int i = 0; i < SAC_TENSORS; ++i) { cudaFree(s->g[i]); cudaFree(s->m[i]); cudaFree(s->v[i]); }
  cudaFree(s->log_alpha); cudaFree(s->t); cudaFree(s->d_step); cudaFree(s->lr_t); cudaFree(s->out); cudaFree(s->bar);
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

MOPO_API int mopo_sac_opt_sizes(mopo_sac* s, int64_t* sizes) {
  REQUIRE(s && sizes, MOPO_EINVAL, "mopo_sac_opt_sizes: null argument");
  static_assert(SAC_TENSORS == MOPO_SAC_TENSORS, "header and kernel disagree on the tensor count");
  for (int i = 0; i < SAC_TENSORS; ++i) sizes[i] = (int64_t)s->psize(i);
  return MOPO_OK;
}

MOPO_API int mopo_sac_get_opt_state(mopo_sac* s, float* const* m, float* const* v, int64_t* t, uint64_t* draw, void* stream) {
  REQUIRE(s && m && v && t && draw, MOPO_EINVAL, "mopo_sac_get_opt_state: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_get_opt_state: parameters not set");
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  for (int i = 0; i < SAC_TENSORS; ++i) {
    REQUIRE(m[i] && v[i], MOPO_EINVAL, "mopo_sac_get_opt_state: tensor %d null", i);
    TRY(sac_copy(m[i], s->m[i], s->psize(i), cudaMemcpyDeviceToHost, st));
    TRY(sac_copy(v[i], s->v[i], s->psize(i), cudaMemcpyDeviceToHost, st));
  }
  CUDA_TRY(cudaMemcpyAsync(t, s->t, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(draw, s->d_step, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return MOPO_OK;
}

MOPO_API int mopo_sac_set_opt_state(mopo_sac* s, const float* const* m, const float* const* v, int64_t t, uint64_t draw, void* stream) {
  REQUIRE(s && m && v, MOPO_EINVAL, "mopo_sac_set_opt_state: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_set_opt_state: parameters not set");
  REQUIRE(t >= 0, MOPO_EINVAL, "mopo_sac_set_opt_state: t=%lld", (long long)t);
  CUDA_TRY(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  for (int i = 0; i < SAC_TENSORS; ++i) {
    REQUIRE(m[i] && v[i], MOPO_EINVAL, "mopo_sac_set_opt_state: tensor %d null", i);
    TRY(sac_copy(s->m[i], m[i], s->psize(i), cudaMemcpyHostToDevice, st));
    TRY(sac_copy(s->v[i], v[i], s->psize(i), cudaMemcpyHostToDevice, st));
  }
  CUDA_TRY(cudaMemcpyAsync(s->t, &t, sizeof(int64_t), cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaMemcpyAsync(s->d_step, &draw, sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  s->updates = t;                                  // (the Polyak interval counts updates, mopo.py:839-841)
  return MOPO_OK;
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
    TRY(w.b->reserve(w.n * sizeof(float)));
  }
  return MOPO_OK;
}

// One _do_training step: ONE cooperative launch.  smp != nullptr: the batch is drawn from the pools inside the kernel
// (into Xpi / ba / br / bt); else obs / next_obs are already in Xpi or copied there, act / rew / term are read in place.
static int sac_update_impl(mopo_sac* s, const SacSample* smp, const float* obs, const float* act, const float* rew, const uint8_t* term,
                           const float* next_obs, int n, const float* eps_in, uint64_t seed, float* losses_out, cudaStream_t st) {
  const int od = s->od;
  TRY(sac_reserve(s, n));
  SacFused a;
  memset(&a, 0, sizeof(a));
  a.n = n; a.od = od; a.ad = s->ad; a.H = s->H;
  a.adp = 1; while (a.adp < s->ad) a.adp *= 2;
  a.seed = seed; a.t = s->t; a.d_step = s->d_step; a.lr_t = s->lr_t; a.lr_pi = s->hp.policy_lr; a.lr_q = s->hp.q_lr;
  a.reward_scale = s->hp.reward_scale; a.discount = s->hp.discount; a.target_entropy = s->hp.target_entropy; a.tau = s->hp.tau;
  a.eps_in = eps_in;
  float* Xpi = s->Xpi.as<float>();
  if (smp) {
    a.sample = 1; a.smp = *smp;
    a.ba = s->ba.as<float>(); a.br = s->br.as<float>(); a.bt = s->bt.as<uint8_t>(); a.bidx = s->bidx.as<int64_t>();
    a.act = a.ba; a.rew = a.br; a.term = a.bt;
  } else {
    CUDA_TRY(cudaMemcpyAsync(Xpi, obs, (size_t)n * od * sizeof(float), cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(Xpi + (size_t)n * od, next_obs, (size_t)n * od * sizeof(float), cudaMemcpyDeviceToDevice, st));
    a.act = act; a.rew = rew; a.term = term;
  }
  const int interval = s->hp.target_update_interval > 0 ? s->hp.target_update_interval : 1;
  a.polyak = s->updates % interval == 0;             // every target_update_interval-th update (mopo.py:839-841)
  for (int l = 0; l < 3; ++l) { a.piW[l] = s->piW[l]; a.pib[l] = s->pib[l]; a.qW[l] = s->qW[l]; a.qb[l] = s->qb[l]; a.tW[l] = s->tW[l]; a.tb[l] = s->tb[l]; }
  a.log_alpha = s->log_alpha;
  a.cpref[0] = 0;
  for (int i = 0; i < SAC_TENSORS; ++i) {
    a.g[i] = s->g[i]; a.m[i] = s->m[i]; a.v[i] = s->v[i]; a.psize[i] = (int64_t)s->psize(i);
    a.par[i] = s->param(i);
    a.targ[i] = (i >= 6 && i < 12) ? ((i & 1) ? s->tb[(i - 6) / 2] : s->tW[(i - 6) / 2]) : nullptr;
    a.cpref[i + 1] = a.cpref[i] + (a.psize[i] + 3) / 4;
  }
  a.Xpi = Xpi; a.H0p = s->H0p.as<float>(); a.H1p = s->H1p.as<float>(); a.HD = s->HD.as<float>(); a.eps = s->eps.as<float>();
  a.logp = s->logp.as<float>(); a.Xq = s->Xq.as<float>(); a.Xt = s->Xt.as<float>(); a.H0q = s->H0q.as<float>(); a.H1q = s->H1q.as<float>();
  a.Qo = s->Qo.as<float>(); a.T0 = s->T0.as<float>(); a.T1 = s->T1.as<float>(); a.To = s->To.as<float>(); a.dQo = s->dQo.as<float>();
  a.dZ1q = s->dZ1q.as<float>(); a.dZ0q = s->dZ0q.as<float>(); a.dXa = s->dXa.as<float>(); a.dHD = s->dHD.as<float>();
  a.dZ1p = s->dZ1p.as<float>(); a.dZ0p = s->dZ0p.as<float>();
  a.out = s->out;
  // the program: stages of independent ops (the weight / bias gradients sit beside the next link of the data-gradient chain)
  {
    const int ad = s->ad, H = s->H, in_q = od + ad, n2 = 2 * n;
    const int64_t sH2 = (int64_t)n2 * H, sH1 = (int64_t)n * H;
    int stage = 0, k = 0;
    bool full = false;
    auto push = [&](FOp op) { if (k < SAC_MAX_OPS) a.ops[k++] = op; else full = true; };
    auto gemm_op = [&](const GemmP& g, int E, int epi) {
      FOp op; memset(&op, 0, sizeof(op));
      op.g = g; op.kind = OP_GEMM; op.epi = epi; op.stage = stage;
      op.tn = (g.N + FT - 1) / FT; op.tmn = op.tn * ((g.M + FT - 1) / FT); op.units = op.tmn * E; op.flags = gemm_flags(g);
      push(op);
    };
    auto colsum_op = [&](const float* dZ, float* db, int rows, int N, int64_t stride, int E) {
      FOp op; memset(&op, 0, sizeof(op));
      op.g.A = dZ; op.g.C = db; op.g.M = rows; op.g.N = N; op.g.sA = stride; op.kind = OP_COLSUM; op.stage = stage; op.units = ((N + 31) / 32) * E;
      push(op);
    };
    auto plain_op = [&](int kind, int units) { FOp op; memset(&op, 0, sizeof(op)); op.kind = kind; op.stage = stage; op.units = units; push(op); };
    if (smp) { plain_op(OP_SAMPLE, 0); ++stage; }
    // actor on [s; s']
    gemm_op(GemmP{a.Xpi, 0, od, 0, a.piW[0], 0, H, 0, a.H0p, 0, H, n2, H, od, a.pib[0], nullptr, nullptr}, 1, 4); ++stage;
    gemm_op(GemmP{a.H0p, 0, H, 0, a.piW[1], 0, H, 0, a.H1p, 0, H, n2, H, H, a.pib[1], nullptr, nullptr}, 1, 4); ++stage;
    gemm_op(GemmP{a.H1p, 0, H, 0, a.piW[2], 0, 2 * ad, 0, a.HD, 0, 2 * ad, n2, 2 * ad, H, a.pib[2], nullptr, nullptr}, 1, 2); ++stage;
    plain_op(OP_PI_HEAD, (n2 + FUSED_THREADS / a.adp - 1) / (FUSED_THREADS / a.adp)); ++stage;                 // squash, log-likelihood, critic inputs
    // twin critics on [(s, a); (s, pi(s))] and the target twins on (s', pi(s')), side by side
    gemm_op(GemmP{a.Xq, 0, in_q, 0, a.qW[0], (int64_t)in_q * H, H, 0, a.H0q, sH2, H, n2, H, in_q, a.qb[0], nullptr, nullptr}, 2, 4);
    gemm_op(GemmP{a.Xt, 0, in_q, 0, a.tW[0], (int64_t)in_q * H, H, 0, a.T0, sH1, H, n, H, in_q, a.tb[0], nullptr, nullptr}, 2, 4); ++stage;
    gemm_op(GemmP{a.H0q, sH2, H, 0, a.qW[1], (int64_t)H * H, H, 0, a.H1q, sH2, H, n2, H, H, a.qb[1], nullptr, nullptr}, 2, 4);
    gemm_op(GemmP{a.T0, sH1, H, 0, a.tW[1], (int64_t)H * H, H, 0, a.T1, sH1, H, n, H, H, a.tb[1], nullptr, nullptr}, 2, 4); ++stage;
    gemm_op(GemmP{a.H1q, sH2, H, 0, a.qW[2], (int64_t)H, 1, 0, a.Qo, n2, 1, n2, 1, H, a.qb[2], nullptr, nullptr}, 2, 2);
    gemm_op(GemmP{a.T1, sH1, H, 0, a.tW[2], (int64_t)H, 1, 0, a.To, n, 1, n, 1, H, a.tb[2], nullptr, nullptr}, 2, 2); ++stage;
    plain_op(OP_LOSS, 1); ++stage;                                   // targets, losses, dQ
    // critic backward over all 2n rows (data rows feed the critic gradients, pi rows the action gradient)
    if (H % 4 == 0) {                                                // K = 1: an elementwise outer product
      FOp op; memset(&op, 0, sizeof(op));
      op.g = GemmP{a.dQo, n2, 1, 0, a.qW[2], (int64_t)H, 1, 1, a.dZ1q, sH2, H, n2, H, 2, nullptr, nullptr, a.H1q};
      op.kind = OP_OUTER; op.stage = stage; op.units = (int)(((int64_t)2 * n2 * (H / 4) + FUSED_THREADS - 1) / FUSED_THREADS);
      push(op);
    } else {
      gemm_op(GemmP{a.dQo, n2, 1, 0, a.qW[2], (int64_t)H, 1, 1, a.dZ1q, sH2, H, n2, H, 1, nullptr, nullptr, a.H1q}, 2, 5);
    }
    gemm_op(GemmP{a.H1q, sH2, H, 1, a.dQo, n2, 1, 0, a.g[10], (int64_t)H, 1, H, 1, n, nullptr, nullptr, nullptr}, 2, 0);
    colsum_op(a.dQo, a.g[11], n, 1, (int64_t)n2, 2); ++stage;
    gemm_op(GemmP{a.dZ1q, sH2, H, 0, a.qW[1], (int64_t)H * H, H, 1, a.dZ0q, sH2, H, n2, H, H, nullptr, nullptr, a.H0q}, 2, 5);
    gemm_op(GemmP{a.H0q, sH2, H, 1, a.dZ1q, sH2, H, 0, a.g[8], (int64_t)H * H, H, H, H, n, nullptr, nullptr, nullptr}, 2, 0);
    colsum_op(a.dZ1q, a.g[9], n, H, sH2, 2); ++stage;
    gemm_op(GemmP{a.dZ0q + (size_t)n * H, sH2, H, 0, a.qW[0] + (size_t)od * H, (int64_t)in_q * H, H, 1, a.dXa, (int64_t)n * ad, ad, n, ad, H,
                  nullptr, nullptr, nullptr}, 2, 0);
    gemm_op(GemmP{a.Xq, 0, in_q, 1, a.dZ0q, sH2, H, 0, a.g[6], (int64_t)in_q * H, H, in_q, H, n, nullptr, nullptr, nullptr}, 2, 0);
    colsum_op(a.dZ0q, a.g[7], n, H, sH2, 2); ++stage;
    plain_op(OP_PI_GRAD, (n * ad + FUSED_THREADS - 1) / FUSED_THREADS); ++stage;                  // policy head gradient
    // actor backward
    gemm_op(GemmP{a.dHD, 0, 2 * ad, 0, a.piW[2], 0, 2 * ad, 1, a.dZ1p, 0, H, n, H, 2 * ad, nullptr, nullptr, a.H1p}, 1, 5);
    gemm_op(GemmP{a.H1p, 0, H, 1, a.dHD, 0, 2 * ad, 0, a.g[4], 0, 2 * ad, H, 2 * ad, n, nullptr, nullptr, nullptr}, 1, 0);
    colsum_op(a.dHD, a.g[5], n, 2 * ad, 0, 1); ++stage;
    gemm_op(GemmP{a.dZ1p, 0, H, 0, a.piW[1], 0, H, 1, a.dZ0p, 0, H, n, H, H, nullptr, nullptr, a.H0p}, 1, 5);
    gemm_op(GemmP{a.H0p, 0, H, 1, a.dZ1p, 0, H, 0, a.g[2], 0, H, H, H, n, nullptr, nullptr, nullptr}, 1, 0);
    colsum_op(a.dZ1p, a.g[3], n, H, 0, 1); ++stage;
    gemm_op(GemmP{a.Xpi, 0, od, 1, a.dZ0p, 0, H, 0, a.g[0], 0, H, od, H, n, nullptr, nullptr, nullptr}, 1, 0);
    colsum_op(a.dZ0p, a.g[1], n, H, 0, 1); ++stage;
    REQUIRE(!full, MOPO_EINVAL, "sac update: program table too small");
    a.n_ops = k;
    a.bar = s->bar;
  }
  a.trace = s->trace;
  void* args[] = {&a};
  MOPO_COUNT_LAUNCH;
  CUDA_TRY(cudaLaunchCooperativeKernel((const void*)sac_fused_kernel, dim3(s->grid), dim3(FUSED_THREADS), args, sizeof(FusedSmem), st));
  if (losses_out) CUDA_TRY(cudaMemcpyAsync(losses_out, s->out, 4 * sizeof(float), cudaMemcpyDeviceToDevice, st));
  s->updates += 1;
  return MOPO_OK;
}

MOPO_API int mopo_sac_update(mopo_sac* s, const mopo_sac_batch* b, void* stream) {
  REQUIRE(s && b, MOPO_EINVAL, "mopo_sac_update: null argument");
  REQUIRE(s->params_set, MOPO_ESTATE, "mopo_sac_update: parameters not set");
  REQUIRE(b->obs && b->act && b->rew && b->term && b->next_obs, MOPO_EINVAL, "mopo_sac_update: null batch field");
  REQUIRE(b->n > 0 && b->n <= (1 << 20), MOPO_EINVAL, "mopo_sac_update: n=%lld", (long long)b->n);
  CUDA_TRY(cudaSetDevice(s->device));
  return sac_update_impl(s, nullptr, b->obs, b->act, b->rew, b->term, b->next_obs, (int)b->n, b->eps, b->seed, b->losses_out, (cudaStream_t)stream);
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
  TRY(s->ba.reserve((size_t)n * ad * sizeof(float)));
  TRY(s->br.reserve((size_t)n * sizeof(float))); TRY(s->bt.reserve((size_t)n));
  TRY(s->bidx.reserve((size_t)n * sizeof(int64_t)));
  REQUIRE((n_env == 0 || env_pool->size > 0) && (n_model == 0 || model_pool->size > 0), MOPO_ESTATE, "mopo_sac_update_from_pools: pool is empty");
  const mopo_pool* pools[2] = {env_pool, n_model ? model_pool : env_pool};
  SacSample smp;
  memset(&smp, 0, sizeof(smp));
  for (int w = 0; w < 2; ++w) {
    smp.obs[w] = pools[w]->obs; smp.act[w] = pools[w]->actions; smp.nobs[w] = pools[w]->next_obs; smp.rew[w] = pools[w]->rewards;
    smp.term[w] = pools[w]->terminals; smp.size[w] = pools[w]->size;
  }
  smp.n_env = n_env; smp.n = n; smp.od = od; smp.ad = ad;
  // observations / next observations are drawn straight into the actor's stacked input [s; s']; the update index and the Adam
  // step live on the device, so the n_updates launches queue back to back with nothing read back in between
  for (int k = 0; k < n_updates; ++k)
    TRY(sac_update_impl(s, &smp, nullptr, nullptr, nullptr, nullptr, nullptr, n, nullptr, seed, losses_out ? losses_out + 4 * k : nullptr, st));
  return MOPO_OK;
}

#ifdef MOPO_DEBUG_PROBES
// Stage timeline of the update kernel: every later update writes CTA 0's %globaltimer at kernel start, after each stage's
// barrier and at its end into buf (device, >= 64 entries); NULL switches it off.  tools/sac_probe.py --trace prints it.
MOPO_API int mopo_debug_sac_trace(mopo_sac* s, unsigned long long* buf) {
  REQUIRE(s, MOPO_EINVAL, "mopo_debug_sac_trace: null handle");
  s->trace = buf;
  return MOPO_OK;
}
#endif
