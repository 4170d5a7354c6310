// FakeEnv.step tail (everything after model.predict), pool ring-buffer kernels and the
// order-preserving compaction between rollout steps.
//
// Reference: mopo/models/fake_env.py:63-131 (mean += obs, std, sample, elite gather, log-prob,
// termination, penalty), mopo/static/<env>.py (termination_fn), flexible_replay_pool.py:45-83
// (ring index), mopo/algorithms/mopo.py:753-758 (obs = next_obs[nonterm_mask]).
#include "internal.cuh"
#include "philox.cuh"

namespace {

constexpr int TAIL_ROWS = 128;     // == the forward kernel's 128-row input block (the fused tail writes the next step's tiles)

__device__ __forceinline__ uint32_t pack_bf16x2_rn(float a, float b) {   // lo = a, hi = b (as tc::pack_bf16x2)
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}

__device__ __forceinline__ bool term_fn(int kind, const float* s /* s[0]=rew, s[1..]=next_obs */, int obs_dim) {
  // Comparisons in double: the reference compares (float64) samples with Python float literals.
  const float* o = s + 1;
  switch (kind) {
    case MOPO_TERM_HOPPER: {   // hopper.py:9-18; :12 is abs(bool) -> upper bound only
      bool fin = true, lt = true;
      for (int d = 0; d < obs_dim; ++d) fin = fin && isfinite(o[d]);
      for (int d = 1; d < obs_dim; ++d) lt = lt && ((double)o[d] < 100.0);
      bool nd = fin && lt && ((double)o[0] > 0.7) && (fabs((double)o[1]) < 0.2);
      return !nd;
    }
    case MOPO_TERM_WALKER2D: { // walker2d.py:9-17
      double hgt = o[0], ang = o[1];
      bool nd = (hgt > 0.8) && (hgt < 2.0) && (ang > -1.0) && (ang < 1.0);
      return !nd;
    }
    case MOPO_TERM_ANT: {      // ant.py:9-16
      bool fin = true;
      for (int d = 0; d < obs_dim; ++d) fin = fin && isfinite(o[d]);
      double x = o[0];
      return !(fin && (x >= 0.2) && (x <= 1.0));
    }
    case MOPO_TERM_HUMANOID: { // humanoid.py:10-14
      double z = o[0];
      return (z < 1.0) || (z > 2.0);
    }
    default: return false;     // halfcheetah.py:9-11
  }
}

__global__ void __launch_bounds__(TAIL_ROWS) step_tail_kernel(TailArgs a) {
  const int64_t n = a.d_n ? min((int64_t)*a.d_n, a.n_max) : a.n_max;
  const int64_t j = (int64_t)blockIdx.x * TAIL_ROWS + threadIdx.x;
  const bool valid = j < n;
  const int E = a.E, D1 = a.D1;
  bool alive = false;
  if (valid) {
    const uint32_t orow = a.origin ? (uint32_t)a.origin[j] : (uint32_t)(a.origin_base + j);
    const float* ob = a.obs + j * a.obs_dim;
    float smp[MOPO_MAX_D1];
    int m = 0;
    float pen = 0.f;
    if (!a.deterministic) {
      m = a.model_inds ? a.model_inds[j]
                       : a.elites[philox_elite(orow, (uint32_t)a.step_id, a.seed, (uint32_t)a.n_elites)];
      const float* mm = a.mean + ((int64_t)m * a.ldb + j) * D1;
      const float* ll = a.lv + ((int64_t)m * a.ldb + j) * D1;
      float z[4];
      for (int d = 0; d < D1; ++d) {
        float mu = mm[d] + (d ? ob[d - 1] : 0.0f);          // fake_env.py:66
        float sd = sqrtf(expf(ll[d]));                        // bnn.py:675 + fake_env.py:67
        float zz;
        if (a.noise) zz = a.noise[j * D1 + d];
        else { if ((d & 3) == 0) philox_normal4(orow, (uint32_t)a.step_id, (uint32_t)(d >> 2), a.seed, z); zz = z[d & 3]; }
        smp[d] = fmaf(zz, sd, mu);                            // fake_env.py:72,79
        if (a.model_means) a.model_means[j * D1 + d] = mu;
        if (a.model_stds) a.model_stds[j * D1 + d] = sd;
      }
    } else {                                                  // fake_env.py:84-86: mean over ALL nets
      for (int d = 0; d < D1; ++d) {
        float sm = 0.f, ss = 0.f, o = d ? ob[d - 1] : 0.0f;
        for (int e = 0; e < E; ++e) {
          sm += a.mean[((int64_t)e * a.ldb + j) * D1 + d] + o;
          ss += sqrtf(expf(a.lv[((int64_t)e * a.ldb + j) * D1 + d]));
        }
        smp[d] = sm / (float)E;
        if (a.model_means) a.model_means[j * D1 + d] = smp[d];
        if (a.model_stds) a.model_stds[j * D1 + d] = ss / (float)E;
      }
      m = -1;
    }
    // penalty (fake_env.py:97-115)
    if (a.penalty_kind == MOPO_PEN_LEARNED_VAR) {
      for (int e = 0; e < E; ++e) {
        const float* ll = a.lv + ((int64_t)e * a.ldb + j) * D1;
        float s2 = 0.f;
        for (int d = 0; d < D1; ++d) { float sd = sqrtf(expf(ll[d])); s2 = fmaf(sd, sd, s2); }
        pen = fmaxf(pen, s2);
      }
      pen = sqrtf(pen);
    } else if (a.penalty_kind == MOPO_PEN_DISAGREEMENT) {
      float d2[MOPO_MAX_E];
      for (int e = 0; e < E; ++e) d2[e] = 0.f;
      for (int d = 1; d < D1; ++d) {
        float o = ob[d - 1], mbar = 0.f;
        for (int e = 0; e < E; ++e) mbar += a.mean[((int64_t)e * a.ldb + j) * D1 + d] + o;
        mbar /= (float)E;
        for (int e = 0; e < E; ++e) {
          float df = (a.mean[((int64_t)e * a.ldb + j) * D1 + d] + o) - mbar;
          d2[e] = fmaf(df, df, d2[e]);
        }
      }
      for (int e = 0; e < E; ++e) pen = fmaxf(pen, d2[e]);
      pen = sqrtf(pen);
    }
    // diagnostics (fake_env.py:20-35), only on request
    if (a.log_prob || a.dev) {
      float prob = 0.f;
      for (int e = 0; e < E; ++e) {
        float slog = 0.f, sq = 0.f;
        for (int d = 0; d < D1; ++d) {
          float mu = a.mean[((int64_t)e * a.ldb + j) * D1 + d] + (d ? ob[d - 1] : 0.0f);
          float l = a.lv[((int64_t)e * a.ldb + j) * D1 + d];
          float df = smp[d] - mu;
          slog += l; sq += df * df * expf(-l);
        }
        prob += expf(-0.5f * ((float)D1 * 1.8378770664093453f + slog + sq));
      }
      if (a.log_prob) a.log_prob[j] = logf(prob);
      if (a.dev) {
        float acc = 0.f;
        for (int d = 0; d < D1; ++d) {
          float s1 = 0.f, s2 = 0.f;
          for (int e = 0; e < E; ++e) { float v = a.mean[((int64_t)e * a.ldb + j) * D1 + d]; s1 += v; }
          s1 /= (float)E;
          for (int e = 0; e < E; ++e) { float v = a.mean[((int64_t)e * a.ldb + j) * D1 + d] - s1; s2 += v * v; }
          acc += sqrtf(s2 / (float)E);
        }
        a.dev[j] = acc / (float)D1;
      }
    }
    const bool term = term_fn(a.term_kind, smp, a.obs_dim);   // fake_env.py:91
    alive = !term;
    const float unpen = smp[0];
    const float rew = (a.penalty_kind == MOPO_PEN_NONE) ? unpen : unpen - a.coeff * pen;   // fake_env.py:115
    if (a.next_obs) for (int d = 0; d < a.obs_dim; ++d) a.next_obs[j * a.obs_dim + d] = smp[d + 1];
    if (a.rewards) a.rewards[j] = rew;
    if (a.terminals) a.terminals[j] = term ? 1 : 0;
    if (a.penalty) a.penalty[j] = pen;
    if (a.unpenalized) a.unpenalized[j] = unpen;
    if (a.model_inds_out) a.model_inds_out[j] = m;
    if (a.p_obs && (n - j) <= a.pool_max) {                   // flexible_replay_pool.py:57-83, last write wins
      const int64_t off = a.d_pool_off ? *a.d_pool_off : 0;
      const int64_t r = (a.pool_base + off + j) % a.pool_max;
      for (int d = 0; d < a.obs_dim; ++d) { a.p_obs[r * a.obs_dim + d] = ob[d]; a.p_nobs[r * a.obs_dim + d] = smp[d + 1]; }
      const float* ac = a.act + j * a.act_dim;
      for (int d = 0; d < a.act_dim; ++d) a.p_act[r * a.act_dim + d] = ac[d];
      a.p_rew[r] = rew;
      a.p_term[r] = term ? 1 : 0;
    }
  }
  if (a.blk_counts) {
    int c = __syncthreads_count(alive);
    if (threadIdx.x == 0) a.blk_counts[blockIdx.x] = c;
  }
}

// Fused-head step tail (stochastic step, learned-variance or no penalty; tensor-core path): the forward kernel kept
// only the chosen net's (mean, variance) per output dim [D1, ldb] and each net's variance sums [2, E, ldb]
// (fake_env.py:78-81,110).
// FOUR threads per row, in four warp-uniform groups (thread = group * 128 + row): one thread per row ran ~2900
// dependent instructions at 10 warps per SM (ncu: 28 % issue utilisation, latency bound).  Group q draws the Philox
// blocks g = q, q+4, .. (4 output dims each: Gaussian sample, fake_env.py:66-79), group 1 also picks the next step's
// elite, group 3 reduces the penalty (max_e ||std_e||_2), group 0 finishes the row (termination_fn, r - lambda*u,
// pool scalars), and group q builds K-panel q of the next step's input tile.  The row-major obs / act / next_obs /
// pool traffic goes through shared memory so that every global access of the block is a contiguous span.
// D1C: D1 known at compile time (0 = a.D1).
// When a.xt_next is set (static-row rollouts) the tail is also the NEXT step's front end (fake_env.py:46 concat,
// models/utils.py:96 scaler, bf16 K-panel input tiles, bnn.py:342-344 elite pick): this block's 128 rows are exactly
// one input block of the forward kernel, so the next step needs no separate tiling pass.
constexpr int TQ = 4;
// sqrt.approx (MUFU.RSQ based, ~1 ulp): the bf16 tensor-core path's variances carry ~3 significant digits
__device__ __forceinline__ float sqrt_fast(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
// contiguous span copy, 16 bytes per thread when both ends are 16-byte aligned (n floats, NT threads)
// __noinline__: ten call sites x (vector path + scalar path) inlined made the fused tail 100 KB of code that every CTA runs
// once -- a fifth of its stall samples were instruction fetch (profiles/r01_step_tail_final_ncu_stalls.txt)
template <int NT>
__device__ __noinline__ void span_copy(float* __restrict__ dst, const float* __restrict__ src, int n, int tid) {
  if ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0) {
    const int n4 = n >> 2;
    for (int i = tid; i < n4; i += NT) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
    for (int i = (n4 << 2) + tid; i < n; i += NT) dst[i] = src[i];
  } else {
    for (int i = tid; i < n; i += NT) dst[i] = src[i];
  }
}
template <int D1C>
__global__ void __launch_bounds__(TAIL_ROWS * TQ, 3) step_tail_fused_kernel(TailArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int od = a.obs_dim, ad = a.act_dim, D1 = D1C ? D1C : a.D1, E = a.E;
  constexpr int GMAX = D1C ? ((D1C + 3) / 4 + TQ - 1) / TQ : (MOPO_MAX_D1 / 4 + TQ - 1) / TQ;   // Philox blocks per group
  float* sObs = sm;                       // [TAIL_ROWS, od]
  float* sNobs = sObs + TAIL_ROWS * od;   // [TAIL_ROWS, od]
  float* sAct = sNobs + TAIL_ROWS * od;   // [TAIL_ROWS, ad]
  float* sActN = sAct + TAIL_ROWS * ad;   // [TAIL_ROWS, ad] next step's actions   (only with xt_next)
  float* sMu = sActN + TAIL_ROWS * ad;    // [in]
  float* sSg = sMu + od + ad;             // [in]
  float* sUnpen = sSg + od + ad;          // [TAIL_ROWS] sampled reward
  float* sPen = sUnpen + TAIL_ROWS;       // [TAIL_ROWS] penalty
  const int64_t n = a.d_n ? min((int64_t)*a.d_n, a.n_max) : a.n_max;
  const int64_t j0 = (int64_t)blockIdx.x * TAIL_ROWS;
  const int rows = (int)max((int64_t)0, min((int64_t)TAIL_ROWS, n - j0));
  const int tid = threadIdx.x, r = tid % TAIL_ROWS, q = tid / TAIL_ROWS;
  constexpr int NT = TAIL_ROWS * TQ;
  const int64_t j = j0 + r;
  const bool ok = r < rows;
  const int nG = (D1 + 3) / 4;
  // --- part 1: everything that does not depend on the forward kernel's outputs (under a programmatic dependent launch
  // this runs while the forward kernel's last CTA pairs are still busy): input spans -> smem, Philox draws, elite pick
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (a.ready.flags) {                   // obs / act may still be arriving over PCIe (mopo_rollout_host): one look at this block's
    int c = 0;                           // chunk flag -- not there yet: read the inputs behind the forward, which has waited for them
    const int tile = (int)((j0 + TAIL_ROWS - 1) / 256);
    while (c + 1 < a.ready.n && tile >= a.ready.end_tile[c]) ++c;
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(a.ready.flags + c) : "memory");
    if (!__syncthreads_and(v == a.ready.value)) asm volatile("griddepcontrol.wait;" ::: "memory");
    else __threadfence();
  }
  uint32_t orow = 0;
  float z[GMAX][4];
  int32_t pick = 0;
  {
    if (!a.obs_behind) span_copy<NT>(sObs, a.obs + j0 * od, rows * od, tid);
    span_copy<NT>(sAct, a.act + j0 * ad, rows * ad, tid);
    if (a.xt_next) {
      span_copy<NT>(sActN, a.act_next + j0 * ad, rows * ad, tid);
      for (int i = tid; i < od + ad; i += NT) { sMu[i] = a.mu[i]; sSg[i] = 1.0f / a.sigma[i]; }   // as prep_tiles_kernel
    }
  }
  if (ok) {
    orow = a.origin ? (uint32_t)a.origin[j] : (uint32_t)(a.origin_base + j);
#pragma unroll
    for (int i = 0; i < GMAX; ++i) {                               // 4 output dims per Philox block
      const int g = q + TQ * i;
      if (g < nG) {
        if (a.noise) {
#pragma unroll
          for (int u = 0; u < 4; ++u) z[i][u] = 4 * g + u < D1 ? a.noise[j * D1 + 4 * g + u] : 0.f;
        } else philox_normal4(orow, (uint32_t)a.step_id, (uint32_t)g, a.seed, z[i]);
      }
    }
    if (q == 1 && a.sel_next)                                      // next step's elite pick (bnn.py:342-344)
      pick = a.model_inds_next ? a.model_inds_next[j]
                               : a.elites[philox_elite(orow, (uint32_t)a.step_id_next, a.seed, (uint32_t)a.n_elites)];
  }
  // --- part 2: the forward kernel's outputs
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // obs written by the PREVIOUS step's tail: that grid is only known to be complete behind this wait (the forward between the
  // two tails waited for it; this grid may have started before it ended)
  if (a.obs_behind) span_copy<NT>(sObs, a.obs + j0 * od, rows * od, tid);
  float2 mv[GMAX][4];
  float p2a[MOPO_MAX_E], p2b[MOPO_MAX_E];
  if (ok) {
#pragma unroll
    for (int i = 0; i < GMAX; ++i) {
      const int g = q + TQ * i;
#pragma unroll
      for (int u = 0; u < 4; ++u) if (g < nG && 4 * g + u < D1) mv[i][u] = a.selmv[(int64_t)(4 * g + u) * a.ldb + j];
    }
    if (q == 3 && a.penalty_kind == MOPO_PEN_LEARNED_VAR) {
#pragma unroll
      for (int e = 0; e < MOPO_MAX_E; ++e)
        if (e < E) { p2a[e] = a.pen2[(int64_t)e * a.ldb + j]; p2b[e] = a.pen2[((int64_t)E + e) * a.ldb + j]; }
    }
  }
  __syncthreads();
  if (ok) {
    const float* ob = sObs + r * od;
    float* nob = sNobs + r * od;
#pragma unroll
    for (int i = 0; i < GMAX; ++i) {
      const int g = q + TQ * i;
      if (g < nG) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int d = 4 * g + u;
          if (d < D1) {
            const float mu = mv[i][u].x + (d ? ob[d - 1] : 0.0f);    // fake_env.py:66
            const float sd = sqrt_fast(mv[i][u].y);                  // fake_env.py:67
            const float smp = fmaf(z[i][u], sd, mu);                 // fake_env.py:72,79
            if (d == 0) sUnpen[r] = smp; else nob[d - 1] = smp;      // fake_env.py:90
            if (a.model_means) a.model_means[j * D1 + d] = mu;
            if (a.model_stds) a.model_stds[j * D1 + d] = sd;
          }
        }
      }
    }
    if (q == 1 && a.sel_next) a.sel_next[j] = pick;
    if (q == 3) {
      float pen = 0.f;
      if (a.penalty_kind == MOPO_PEN_LEARNED_VAR) {                // max_e ||std_e||_2  (fake_env.py:110)
#pragma unroll
        for (int e = 0; e < MOPO_MAX_E; ++e) if (e < E) pen = fmaxf(pen, p2a[e] + p2b[e]);
        pen = sqrt_fast(pen);
      }
      sPen[r] = pen;
    }
  }
  __syncthreads();
  bool alive = false;
  if (ok && q == 0) {
    const float* nob = sNobs + r * od;
    const float unpen = sUnpen[r], pen = sPen[r];
    const bool term = term_fn(a.term_kind, nob - 1, od);           // fake_env.py:91 (term_fn skips the reward slot)
    alive = !term;
    const float rew = (a.penalty_kind == MOPO_PEN_NONE) ? unpen : unpen - a.coeff * pen;   // fake_env.py:115
    if (a.rewards) a.rewards[j] = rew;
    if (a.terminals) a.terminals[j] = term ? 1 : 0;
    if (a.penalty) a.penalty[j] = pen;
    if (a.unpenalized) a.unpenalized[j] = unpen;
    if (a.model_inds_out) a.model_inds_out[j] = a.sel[j];
    if (a.p_obs && (n - j) <= a.pool_max) {                        // flexible_replay_pool.py:57-83, last write wins
      const int64_t off = a.d_pool_off ? *a.d_pool_off : 0;
      const int64_t pr = (a.pool_base + off + j) % a.pool_max;
      a.p_rew[pr] = rew;
      a.p_term[pr] = term ? 1 : 0;
    }
  }
  if (a.xt_next) {
    // next step's input block (same layout as prep_tiles_kernel): K-panel kc, row r -> 8 bf16; rows >= n are zero
    const int in = od + ad, kcs = a.K0 / 8;
    const float* nob = sNobs + r * od;
    const float* an = sActN + r * ad;
    for (int kc = q; kc < kcs; kc += TQ) {
      uint32_t p[4] = {0, 0, 0, 0};
      if (ok) {
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          float v[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int k = kc * 8 + 2 * w + u;
            if (k < in) v[u] = ((k < od ? nob[k] : an[k - od]) - sMu[k]) * sSg[k];   // models/utils.py:96
            else v[u] = k < in + 2 ? 1.f : 0.f;
          }
          p[w] = pack_bf16x2_rn(v[0], v[1]);
        }
      }
      a.xt_next[((int64_t)blockIdx.x * kcs + kc) * TAIL_ROWS + r] = make_uint4(p[0], p[1], p[2], p[3]);
      // the forward kernel reads whole 256-row tiles: an odd block count leaves one all-zero block to write
      if (blockIdx.x == gridDim.x - 1 && (gridDim.x & 1))
        a.xt_next[((int64_t)gridDim.x * kcs + kc) * TAIL_ROWS + r] = make_uint4(0, 0, 0, 0);
    }
  }
  if (a.next_obs) {
    span_copy<NT>(a.next_obs + j0 * od, sNobs, rows * od, tid);
  }
  if (a.p_obs) {
    const int64_t off = a.d_pool_off ? *a.d_pool_off : 0;
    const int64_t pr0 = (a.pool_base + off + j0) % a.pool_max;
    if (n - j0 <= a.pool_max && pr0 + rows <= a.pool_max) {        // common case: no wrap, nothing skipped -> flat copies
      span_copy<NT>(a.p_obs + pr0 * od, sObs, rows * od, tid);
      span_copy<NT>(a.p_nobs + pr0 * od, sNobs, rows * od, tid);
      span_copy<NT>(a.p_act + pr0 * ad, sAct, rows * ad, tid);
    } else {
      for (int i = tid; i < rows * od; i += NT) {
        const int rr = i / od, c = i - rr * od;
        if ((n - (j0 + rr)) <= a.pool_max) {
          const int64_t pr = (a.pool_base + off + j0 + rr) % a.pool_max;
          a.p_obs[pr * od + c] = sObs[i];
          a.p_nobs[pr * od + c] = sNobs[i];
        }
      }
      for (int i = tid; i < rows * ad; i += NT) {
        const int rr = i / ad, c = i - rr * ad;
        if ((n - (j0 + rr)) <= a.pool_max) a.p_act[((a.pool_base + off + j0 + rr) % a.pool_max) * ad + c] = sAct[i];
      }
    }
  }
  if (a.blk_counts) {
    int c = __syncthreads_count(alive);
    if (tid == 0) a.blk_counts[blockIdx.x] = c;
  }
}

__global__ void prep_kernel(const float* __restrict__ obs, const float* __restrict__ act,
                            const float* __restrict__ mu, const float* __restrict__ sigma, float* __restrict__ xn,
                            int64_t n_max, const int* __restrict__ d_n, int obs_dim, int act_dim) {
  const int64_t n = d_n ? min((int64_t)*d_n, n_max) : n_max;
  const int in = obs_dim + act_dim;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n * in; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t j = i / in; int c = (int)(i % in);
    float v = c < obs_dim ? obs[j * obs_dim + c] : act[j * act_dim + (c - obs_dim)];
    xn[i] = (v - mu[c]) / sigma[c];                           // mopo/models/utils.py:96
  }
}

// mopo.py:753-758: obs = next_obs[~term], order preserved.  Same row->block map as the tail.
// With `ta.xt` set it is also the next step's front end (as prep_tiles_kernel: concat with the next step's actions,
// scaler, bf16 K-panel tiles, elite picks), so terminating envs need no separate tiling pass after the first step either.
__global__ void __launch_bounds__(TAIL_ROWS)
compact_kernel(const float* __restrict__ next_obs, const uint8_t* __restrict__ term,
               const int32_t* __restrict__ origin_in, int64_t origin_base, int64_t n_max, const int* __restrict__ d_n,
               const int* __restrict__ blk_counts, int obs_dim, float* __restrict__ obs_out,
               int32_t* __restrict__ origin_out, int* __restrict__ d_n_next, int64_t* __restrict__ d_pool_off, CompactTiles ta) {
  const int64_t n = min((int64_t)*d_n, n_max);
  __shared__ int s_part[TAIL_ROWS / 32];
  __shared__ int s_base;
  __shared__ float s_mu[MOPO_MAX_D1 + 64], s_rs[MOPO_MAX_D1 + 64];
  if (ta.xt) for (int i = threadIdx.x; i < obs_dim + ta.act_dim; i += TAIL_ROWS) { s_mu[i] = ta.mu[i]; s_rs[i] = 1.0f / ta.sigma[i]; }
  int part = 0;
  for (int b = threadIdx.x; b < (int)blockIdx.x; b += TAIL_ROWS) part += blk_counts[b];
  for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (lane == 0) s_part[warp] = part;
  __syncthreads();
  if (threadIdx.x == 0) { int s = 0; for (int w = 0; w < TAIL_ROWS / 32; ++w) s += s_part[w]; s_base = s; }
  __syncthreads();
  const int base = s_base;
  const int64_t j = (int64_t)blockIdx.x * TAIL_ROWS + threadIdx.x;
  const bool alive = (j < n) && (term[j] == 0);
  const unsigned bal = __ballot_sync(0xffffffffu, alive);
  __syncthreads();
  if (lane == 0) s_part[warp] = __popc(bal);
  __syncthreads();
  int woff = 0;
  for (int w = 0; w < warp; ++w) woff += s_part[w];
  if (alive) {
    const int64_t dst = base + woff + __popc(bal & ((1u << lane) - 1u));
    for (int d = 0; d < obs_dim; ++d) obs_out[dst * obs_dim + d] = next_obs[j * obs_dim + d];
    const int32_t org = origin_in ? origin_in[j] : (int32_t)(origin_base + j);
    origin_out[dst] = org;
    if (ta.xt) {
      const int od = obs_dim, ad = ta.act_dim, in = od + ad, kcs = ta.K0 / 8;
      const float* an = ta.act_next + dst * ad;
      for (int kc = 0; kc < kcs; ++kc) {
        uint32_t p[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          float v[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int k = kc * 8 + 2 * w + u;
            if (k < in) v[u] = ((k < od ? next_obs[j * od + k] : an[k - od]) - s_mu[k]) * s_rs[k];   // models/utils.py:96
            else v[u] = k < in + 2 ? 1.f : 0.f;
          }
          p[w] = pack_bf16x2_rn(v[0], v[1]);
        }
        ta.xt[((dst / TAIL_ROWS) * kcs + kc) * TAIL_ROWS + dst % TAIL_ROWS] = make_uint4(p[0], p[1], p[2], p[3]);
      }
    }
    if (ta.sel)
      ta.sel[dst] = ta.model_inds ? ta.model_inds[dst]
                                  : ta.elites[philox_elite((uint32_t)org, (uint32_t)ta.step_id, ta.seed, (uint32_t)ta.n_elites)];
  }
  if (blockIdx.x == gridDim.x - 1) {
    int tot = base;
    for (int w = 0; w < TAIL_ROWS / 32; ++w) tot += s_part[w];
    if (threadIdx.x == 0) {
      *d_n_next = tot;
      if (d_pool_off) *d_pool_off += n;
    }
    if (ta.xt) {                                        // rows tot .. next multiple of 256 of the last tile: zeros
      const int kcs = ta.K0 / 8;
      const int64_t end = ((int64_t)tot + 255) / 256 * 256;
      for (int64_t rdst = tot + threadIdx.x; rdst < end; rdst += TAIL_ROWS)
        for (int kc = 0; kc < kcs; ++kc)
          ta.xt[((rdst / TAIL_ROWS) * kcs + kc) * TAIL_ROWS + rdst % TAIL_ROWS] = make_uint4(0, 0, 0, 0);
    }
  }
}

__global__ void term_only_kernel(int kind, const float* __restrict__ next_obs, int64_t n, int obs_dim, uint8_t* __restrict__ out) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x)
    out[j] = term_fn(kind, next_obs + j * obs_dim - 1, obs_dim) ? 1 : 0;     // term_fn skips the reward slot
}

__global__ void exp_kernel(const float* __restrict__ lv, float* __restrict__ var, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    var[i] = expf(lv[i]);
}

}  // namespace

int tail_rows_per_block() { return TAIL_ROWS; }

int launch_prep(const float* obs, const float* act, const float* mu, const float* sigma, float* xn,
                int64_t n_max, const int* d_n, int obs_dim, int act_dim, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  int64_t tot = n_max * (obs_dim + act_dim);
  int64_t g64 = (tot + 255) / 256; int grid = (int)(g64 > 148 * 16 ? 148 * 16 : g64);
  MOPO_COUNT_LAUNCH, prep_kernel<<<grid, 256, 0, st>>>(obs, act, mu, sigma, xn, n_max, d_n, obs_dim, act_dim);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int launch_tail(const TailArgs& a, cudaStream_t st) {
  if (a.n_max == 0) return MOPO_OK;
  int grid = (int)((a.n_max + TAIL_ROWS - 1) / TAIL_ROWS);
  if (a.sel) {
    const size_t smem = sizeof(float) * (TAIL_ROWS * (2 * a.obs_dim + 2 * a.act_dim + 2) + 2 * (a.obs_dim + a.act_dim));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(TAIL_ROWS * TQ); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = a.pdl ? 1 : 0;      // rollout chain: may be scheduled while the forward kernel drains
    if (smem > 48 * 1024) {                           // wide shapes (humanoid: obs 45, act 17 -> 65 KB): opt in
      CUDA_TRY(cudaFuncSetAttribute(step_tail_fused_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));   // per device: set on every such launch
      REQUIRE(smem <= 96 * 1024, MOPO_EUNSUPPORTED, "step tail: obs_dim %d / act_dim %d need %zu B of shared memory", a.obs_dim, a.act_dim, smem);
    }
    MOPO_COUNT_LAUNCH;
    if (a.D1 == 18) CUDA_TRY(cudaLaunchKernelEx(&cfg, step_tail_fused_kernel<18>, a));        // obs 17 (+1)
    else if (a.D1 == 12) CUDA_TRY(cudaLaunchKernelEx(&cfg, step_tail_fused_kernel<12>, a));   // obs 11 (+1)
    else CUDA_TRY(cudaLaunchKernelEx(&cfg, step_tail_fused_kernel<0>, a));
  } else {
    MOPO_COUNT_LAUNCH, step_tail_kernel<<<grid, TAIL_ROWS, 0, st>>>(a);
  }
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int launch_compact(const float* next_obs, const uint8_t* term, const int32_t* origin_in, int64_t origin_base,
                   int64_t n_max, const int* d_n, const int* blk_counts, int obs_dim,
                   float* obs_out, int32_t* origin_out, int* d_n_next, int64_t* d_pool_off, const CompactTiles& ta, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  int grid = (int)((n_max + TAIL_ROWS - 1) / TAIL_ROWS);
  MOPO_COUNT_LAUNCH, compact_kernel<<<grid, TAIL_ROWS, 0, st>>>(next_obs, term, origin_in, origin_base, n_max, d_n, blk_counts, obs_dim,
                                             obs_out, origin_out, d_n_next, d_pool_off, ta);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int launch_term_only(int kind, const float* next_obs, int64_t n, int obs_dim, uint8_t* out, cudaStream_t st) {
  if (n == 0) return MOPO_OK;
  int64_t g64 = (n + 255) / 256; int grid = (int)(g64 > 148 * 16 ? 148 * 16 : g64);
  MOPO_COUNT_LAUNCH, term_only_kernel<<<grid, 256, 0, st>>>(kind, next_obs, n, obs_dim, out);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int launch_exp(const float* lv, float* var, int64_t n, cudaStream_t st) {
  if (n == 0) return MOPO_OK;
  int64_t g64 = (n + 255) / 256; int grid = (int)(g64 > 148 * 16 ? 148 * 16 : g64);
  MOPO_COUNT_LAUNCH, exp_kernel<<<grid, 256, 0, st>>>(lv, var, n);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}
