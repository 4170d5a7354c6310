// Ensemble training on the device: BNN.train's inner loop (reference mopo/models/bnn.py:421-427) -- the Adam step on
//   train_loss = sum_e [ mean((mean_e - t)^2 exp(-logvar_e)) + mean(logvar_e) ]                (bnn.py:242, 693-696)
//              + sum_l weight_decay_l * sum(W_l^2) / 2                                          (bnn.py:243, fc.py:155-156)
//              + 0.01 sum(max_logvar) - 0.01 sum(min_logvar)                                    (bnn.py:244)
// over all W, b, max_logvar, min_logvar (bnn.py:212-224, 249), plus mse_loss (bnn.py:248) for the holdout bookkeeping
// and the per-net snapshot / restore of _save_state / _set_state (bnn.py:287-307).
//
// The master weights are the fp32 images the fp32 forward reads (mopo_ens::W / b, reference layout); training updates
// them in place and mopo_ens_train_end rebuilds the bf16 tcgen05 images from them.  Batches are small (E x 256 rows x
// 200 wide), so the step is launch-bound: plain fp32 FFMA tiled GEMMs (forward NN, data gradient NT, weight gradient
// TN) with the bias / swish / swish' work in their epilogues, a deterministic loss + logvar-bound gradient reduction
// (no float atomics: a training run is reproducible bit for bit), and one Adam launch per tensor.
#include "gemm_fp32.cuh"

namespace {

// x[e,b,:] = (inputs[r,:] - mu) / sigma, tg[e,b,:] = targets[r,:] with r = idx ? idx[e*ld + col0 + b] : b; col0 lives on the
// device (tick_kernel) so that one captured graph serves every batch of an epoch
__global__ void gather_scale_kernel(const float* __restrict__ inputs, const float* __restrict__ targets,
                                    const int32_t* __restrict__ idx, int64_t ld_idx, const int64_t* __restrict__ d_col0,
                                    int n, int E, int in, int D1, const float* __restrict__ mu,
                                    const float* __restrict__ sigma, float* __restrict__ x, float* __restrict__ tg) {
  const int W = in + D1;
  const int64_t col0 = d_col0 ? *d_col0 : 0;
  const int64_t tot = (int64_t)E * n * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % W);
    int64_t eb = i / W;
    int b = (int)(eb % n), e = (int)(eb / n);
    int64_t r = idx ? (int64_t)idx[(int64_t)e * ld_idx + col0 + b] : (int64_t)b;
    if (c < in) x[eb * in + c] = (inputs[r * in + c] - mu[c]) / sigma[c];
    else tg[eb * D1 + (c - in)] = targets[r * D1 + (c - in)];
  }
}

// Head backward (bnn.py:669-670, 693-696): out [E,n,2 D1] = [mean | raw] -> dOut (same layout), plus per-element
// contributions to the loss and to d max_logvar / d min_logvar in part [3][E,n,D1] (reduced by head_reduce_kernel).
__global__ void head_grad_kernel(const float* __restrict__ out, const float* __restrict__ tg,
                                 const float* __restrict__ max_lv, const float* __restrict__ min_lv, int E, int n, int D1,
                                 float* __restrict__ dOut, float* __restrict__ part, int deterministic) {
  const int64_t tot = (int64_t)E * n * D1;
  const float c = 1.0f / ((float)n * (float)D1);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    int d = (int)(i % D1);
    int64_t eb = i / D1;
    float mean = out[eb * 2 * D1 + d], raw = out[eb * 2 * D1 + D1 + d], t = tg[i];
    if (deterministic) {                               // bnn.py:246-247: the loss is the plain MSE; the log-variance half gets no gradient
      const float diff = mean - t;
      dOut[eb * 2 * D1 + d] = 2.0f * diff * c;
      dOut[eb * 2 * D1 + D1 + d] = 0.0f;
      part[i] = diff * diff * c;
      part[tot + i] = 0.0f;
      part[2 * tot + i] = 0.0f;
      continue;
    }
    float mx = max_lv[d], mn = min_lv[d];
    float l1 = mx - softplus_f(mx - raw);
    float lv = mn + softplus_f(l1 - mn);
    float inv = expf(-lv), diff = mean - t;
    float d_lv = (1.0f - diff * diff * inv) * c;
    float s_min = sigmoid_f(l1 - mn), s_max = sigmoid_f(mx - raw);
    dOut[eb * 2 * D1 + d] = 2.0f * diff * inv * c;
    dOut[eb * 2 * D1 + D1 + d] = d_lv * s_min * s_max;
    part[i] = (diff * diff * inv + lv) * c;
    part[tot + i] = d_lv * s_min * (1.0f - s_max);
    part[2 * tot + i] = d_lv * (1.0f - s_min);
  }
}

// blocks [0, D1): g_lv[d] = 0.01 + sum_{e,b} part1, g_lv[D1 + d] = -0.01 + sum part2;  blocks [D1, D1 + E): loss_e
__global__ void __launch_bounds__(256) head_reduce_kernel(const float* __restrict__ part, int E, int n, int D1,
                                                          float* __restrict__ g_lv, float* __restrict__ loss_e, float bound_grad) {
  __shared__ float sh[8];
  const int64_t tot = (int64_t)E * n * D1;
  if ((int)blockIdx.x < D1) {
    const int d = blockIdx.x;
    float a = 0.f, b = 0.f;
    for (int64_t r = threadIdx.x; r < (int64_t)E * n; r += blockDim.x) {
      a += part[tot + r * D1 + d];
      b += part[2 * tot + r * D1 + d];
    }
    a = block_sum(a, sh);
    b = block_sum(b, sh);
    if (threadIdx.x == 0) { g_lv[d] = a + bound_grad; g_lv[D1 + d] = b - bound_grad; }
  } else {
    const int e = blockIdx.x - D1;
    float a = 0.f;
    for (int64_t r = threadIdx.x; r < (int64_t)n * D1; r += blockDim.x) a += part[(int64_t)e * n * D1 + r];
    a = block_sum(a, sh);
    if (threadIdx.x == 0) loss_e[e] = a;
  }
}

// mse_e = mean_{b,d} (mean - t)^2   (bnn.py:698)
__global__ void __launch_bounds__(256) mse_kernel(const float* __restrict__ out, const float* __restrict__ tg, int n, int D1,
                                                  float* __restrict__ mse) {
  __shared__ float sh[8];
  const int e = blockIdx.x;
  float a = 0.f;
  for (int64_t r = threadIdx.x; r < (int64_t)n * D1; r += blockDim.x) {
    int64_t b = r / D1; int d = (int)(r % D1);
    float diff = out[((int64_t)e * n + b) * 2 * D1 + d] - tg[(int64_t)e * n * D1 + r];
    a += diff * diff;
  }
  a = block_sum(a, sh);
  if (threadIdx.x == 0) mse[e] = a / ((float)n * (float)D1);
}

// sum of squares of a tensor slice (columns [c0, c1) of rows of length ncols) -> out[slot]; one block, deterministic
__global__ void __launch_bounds__(1024) sumsq_kernel(const float* __restrict__ w, int64_t rows, int ncols, int c0, int c1,
                                                     float scale, float* __restrict__ out) {
  __shared__ float sh[32];
  const int wc = c1 - c0;
  float a = 0.f;
  for (int64_t i = threadIdx.x; i < rows * wc; i += blockDim.x) {
    float v = w[(i / wc) * ncols + c0 + (i % wc)];
    a = fmaf(v, v, a);
  }
  a = block_sum(a, sh);
  if (threadIdx.x == 0) *out = a * scale;
}

// total = sum(loss_e) + sum(decay terms) + 0.01 (sum max_lv - sum min_lv)
__global__ void loss_total_kernel(const float* __restrict__ loss_e, int E, const float* __restrict__ decay_terms, int nd,
                                  const float* __restrict__ max_lv, const float* __restrict__ min_lv, int D1,
                                  float* __restrict__ out, float bound_coeff) {
  float t = 0.f;
  for (int e = 0; e < E; ++e) t += loss_e[e];
  for (int i = 0; i < nd; ++i) t += decay_terms[i];
  float a = 0.f, b = 0.f;
  for (int d = 0; d < D1; ++d) { a += max_lv[d]; b += min_lv[d]; }
  *out = t + bound_coeff * a - bound_coeff * b;
}

// Per-batch bookkeeping on the device: ctr = {next batch's first column, Adam step t, this batch's first column};
// lr_t = lr sqrt(1 - b2^t) / (1 - b1^t) for the step being applied.
__global__ void tick_kernel(int64_t* __restrict__ ctr, float* __restrict__ lr_t, float lr, int bs) {
  const int64_t t = ++ctr[1];
  ctr[2] = ctr[0];
  ctr[0] += bs;
  *lr_t = (float)((double)lr * sqrt(1.0 - pow((double)ADAM_B2, (double)t)) / (1.0 - pow((double)ADAM_B1, (double)t)));
}

// tf.train.AdamOptimizer._apply_dense for every trainable tensor in ONE launch (blockIdx.y = tensor); the weight-decay
// gradient decay * w is added here (columns < split use decay_a, the rest decay_b: the fused [mean | logvar] head).
constexpr int ADAM_MAX_TENSORS = 2 * MOPO_MAX_LAYERS + 2;
struct AdamBatch {
  float* p[ADAM_MAX_TENSORS]; const float* g[ADAM_MAX_TENSORS]; float* m[ADAM_MAX_TENSORS]; float* v[ADAM_MAX_TENSORS];
  int64_t n[ADAM_MAX_TENSORS]; int ncols[ADAM_MAX_TENSORS], split[ADAM_MAX_TENSORS];
  float decay_a[ADAM_MAX_TENSORS], decay_b[ADAM_MAX_TENSORS];
  const float* lr_t;
};
__global__ void __launch_bounds__(256) adam_kernel(AdamBatch a) {
  const int t = blockIdx.y;
  const float lr_t = *a.lr_t;
  float* __restrict__ p = a.p[t]; const float* __restrict__ g = a.g[t];
  float* __restrict__ m = a.m[t]; float* __restrict__ v = a.v[t];
  const int64_t n = a.n[t];
  const int ncols = a.ncols[t], split = a.split[t];
  const float da = a.decay_a[t], db = a.decay_b[t];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float w = p[i];
    float dec = ((int)(i % ncols) < split) ? da : db;
    float gi = fmaf(dec, w, g[i]);
    float mi = ADAM_B1 * m[i] + (1.0f - ADAM_B1) * gi;
    float vi = ADAM_B2 * v[i] + (1.0f - ADAM_B2) * gi * gi;
    m[i] = mi; v[i] = vi;
    p[i] = w - lr_t * mi / (sqrtf(vi) + ADAM_EPS);
  }
}

// dst[e] = src[e] for one net of an [E, per_net] tensor
__global__ void copy_net_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t per_net, int e) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per_net; i += (int64_t)gridDim.x * blockDim.x)
    dst[(int64_t)e * per_net + i] = src[(int64_t)e * per_net + i];
}

// src[r, c0 + c] -> dst[r, c]  (split the fused head back into the reference's mean / var heads)
__global__ void take_cols_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t rows, int n_src, int n_dst, int c0) {
  const int64_t tot = rows * n_dst;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = src[(i / n_dst) * n_src + c0 + (i % n_dst)];
}

}  // namespace

struct mopo_train {
  float lr = 1e-3f;
  float decay[MOPO_MAX_LAYERS + 1] = {};   // hidden 0..L-1, mean head [L], var head [L+1]
  int64_t t = 0;
  float *gW[MOPO_MAX_LAYERS] = {}, *gb[MOPO_MAX_LAYERS] = {};
  float *mW[MOPO_MAX_LAYERS] = {}, *vW[MOPO_MAX_LAYERS] = {}, *mb[MOPO_MAX_LAYERS] = {}, *vb[MOPO_MAX_LAYERS] = {};
  float *sW[MOPO_MAX_LAYERS] = {}, *sb[MOPO_MAX_LAYERS] = {};   // per-net snapshots (bnn.py:287-290)
  float *g_lv = nullptr, *m_lv = nullptr, *v_lv = nullptr;   // [2 D1]: max | min
  bool snap[MOPO_MAX_E] = {};
  bool deterministic = false;  // bnn.py:246-247: MSE + decays instead of the Gaussian NLL + log-variance bounds
  DevBuf x, tg, Z, Hh, out, dA, dB, part, scal;
  cudaStream_t side = nullptr;   // weight / bias gradients run here, beside the data-gradient chain
  cudaEvent_t ev_fork[MOPO_MAX_LAYERS] = {}, ev_join = nullptr;
  int64_t* ctr = nullptr;      // device {next col0, Adam t, current col0}
  float* lr_t = nullptr;       // device
  // the captured batch step (full batches on a non-default stream): valid for exactly these arguments
  cudaGraphExec_t graph = nullptr;
  const void *g_in = nullptr, *g_tg = nullptr, *g_idx = nullptr; int64_t g_ld = 0; int g_bs = 0; int g_nodes = 0;
};

static size_t wsize(const mopo_ens* h, int l) { return (size_t)h->E * (l == 0 ? h->in : h->H) * (l == h->L ? h->HO : h->H); }
static size_t bsize(const mopo_ens* h, int l) { return (size_t)h->E * (l == h->L ? h->HO : h->H); }

static void train_free(mopo_train* tr) {
  if (!tr) return;
  for (int l = 0; l < MOPO_MAX_LAYERS; ++l) {
    float* ps[] = {tr->gW[l], tr->gb[l], tr->mW[l], tr->vW[l], tr->mb[l], tr->vb[l], tr->sW[l], tr->sb[l]};
    for (float* p : ps) if (p) cudaFree(p);
  }
  float* qs[] = {tr->g_lv, tr->m_lv, tr->v_lv};
  for (float* p : qs) if (p) cudaFree(p);
  DevBuf* bufs[] = {&tr->x, &tr->tg, &tr->Z, &tr->Hh, &tr->out, &tr->dA, &tr->dB, &tr->part, &tr->scal};
  for (DevBuf* b : bufs) b->release();
  if (tr->ctr) cudaFree(tr->ctr);
  if (tr->lr_t) cudaFree(tr->lr_t);
  if (tr->graph) cudaGraphExecDestroy(tr->graph);
  for (cudaEvent_t e : tr->ev_fork) if (e) cudaEventDestroy(e);
  if (tr->ev_join) cudaEventDestroy(tr->ev_join);
  if (tr->side) cudaStreamDestroy(tr->side);
  delete tr;
}

void mopo_train_release(mopo_ens* h) { train_free(h->tr); h->tr = nullptr; }

MOPO_API int mopo_ens_train_begin(mopo_ens* h, float lr, const float* weight_decays) {
  REQUIRE(h && weight_decays, MOPO_EINVAL, "mopo_ens_train_begin: null argument");
  REQUIRE(h->params_set, MOPO_ESTATE, "mopo_ens_train_begin: parameters not set");
  REQUIRE(lr > 0.f, MOPO_EINVAL, "mopo_ens_train_begin: lr=%g", (double)lr);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  // The optimiser lives as long as the model: the reference builds ONE tf.train.AdamOptimizer in finalize() (bnn.py:180, 252),
  // so its moments and step count carry over from one BNN.train() call to the next (periodic model training,
  // model_train_freq); only mopo_ens_train_reset / mopo_ens_destroy drop them.  A new run just clears the snapshot store.
  if (h->tr) {
    bool same = h->tr->lr == lr;
    for (int l = 0; l < h->L + 2; ++l) same = same && h->tr->decay[l] == weight_decays[l];
    h->tr->lr = lr;
    for (int l = 0; l < h->L + 2; ++l) h->tr->decay[l] = weight_decays[l];
    for (int e = 0; e < MOPO_MAX_E; ++e) h->tr->snap[e] = false;
    if (!same && h->tr->graph) { cudaGraphExecDestroy(h->tr->graph); h->tr->graph = nullptr; }   // lr / decays are baked into the captured nodes
    return MOPO_OK;
  }
  mopo_train* tr = new mopo_train();
  h->tr = tr;
  tr->lr = lr;
  for (int l = 0; l < h->L + 2; ++l) tr->decay[l] = weight_decays[l];
  for (int l = 0; l <= h->L; ++l) {
    float** ws[] = {&tr->gW[l], &tr->mW[l], &tr->vW[l], &tr->sW[l]};
    float** bs[] = {&tr->gb[l], &tr->mb[l], &tr->vb[l], &tr->sb[l]};
    for (float** p : ws) { CUDA_TRY(cudaMalloc(p, wsize(h, l) * sizeof(float))); CUDA_TRY(cudaMemset(*p, 0, wsize(h, l) * sizeof(float))); }
    for (float** p : bs) { CUDA_TRY(cudaMalloc(p, bsize(h, l) * sizeof(float))); CUDA_TRY(cudaMemset(*p, 0, bsize(h, l) * sizeof(float))); }
  }
  float** qs[] = {&tr->g_lv, &tr->m_lv, &tr->v_lv};
  for (float** p : qs) { CUDA_TRY(cudaMalloc(p, 2 * h->D1 * sizeof(float))); CUDA_TRY(cudaMemset(*p, 0, 2 * h->D1 * sizeof(float))); }
  CUDA_TRY(cudaMalloc(&tr->ctr, 4 * sizeof(int64_t)));
  CUDA_TRY(cudaMemset(tr->ctr, 0, 4 * sizeof(int64_t)));
  CUDA_TRY(cudaMalloc(&tr->lr_t, sizeof(float)));
  CUDA_TRY(cudaStreamCreateWithFlags(&tr->side, cudaStreamNonBlocking));
  for (int l = 0; l <= h->L; ++l) CUDA_TRY(cudaEventCreateWithFlags(&tr->ev_fork[l], cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&tr->ev_join, cudaEventDisableTiming));
  return MOPO_OK;
}

// Scratch for n rows per net.  Growing a buffer moves it, which invalidates the captured graph.
static int train_reserve(mopo_ens* h, int n) {
  mopo_train* tr = h->tr;
  const int E = h->E, H = h->H, L = h->L, HO = h->HO, D1 = h->D1, in = h->in;
  const size_t rows = (size_t)E * n;
  const int wide = H > HO ? H : HO;
  DevBuf* bufs[] = {&tr->x, &tr->tg, &tr->Z, &tr->Hh, &tr->out, &tr->dA, &tr->dB, &tr->part, &tr->scal};
  const size_t want[] = {rows * in, rows * D1, rows * H * L, rows * H * L, rows * HO, rows * wide, rows * wide * L, 3 * rows * D1,
                         (size_t)(2 * E + L + 8)};
  bool moved = false;
  for (int i = 0; i < 9; ++i) {
    void* before = bufs[i]->p;
    TRY(bufs[i]->reserve(want[i] * sizeof(float)));
    moved |= before != bufs[i]->p;
  }
  if (moved && tr->graph) { cudaGraphExecDestroy(tr->graph); tr->graph = nullptr; }
  return MOPO_OK;
}

// Forward of n rows per net into tr->out ([E,n,2 D1] = [mean | raw log-variance]); keeps Z / H of every hidden layer.
static int train_forward(mopo_ens* h, const float* inputs, const float* targets, const int32_t* idx, int64_t ld_idx,
                         const int64_t* d_col0, int n, cudaStream_t st) {
  mopo_train* tr = h->tr;
  const int E = h->E, H = h->H, L = h->L, HO = h->HO, D1 = h->D1, in = h->in;
  const size_t rows = (size_t)E * n;
  TRY(train_reserve(h, n));
  MOPO_COUNT_LAUNCH;
  gather_scale_kernel<<<grid1d((int64_t)rows * (in + D1)), 256, 0, st>>>(inputs, targets, idx, ld_idx, d_col0, n, E, in, D1, h->mu,
                                                                       h->sigma, tr->x.as<float>(), tr->tg.as<float>());
  const int64_t sAct = (int64_t)n * H;
  const float* A = tr->x.as<float>();
  int64_t sA = (int64_t)n * in;
  int K = in;
  for (int l = 0; l < L; ++l) {
    float* Hl = tr->Hh.as<float>() + (size_t)l * rows * H;
    float* Zl = tr->Z.as<float>() + (size_t)l * rows * H;
    GemmP p{A, sA, K, 0, h->W[l], (int64_t)K * H, H, 0, Hl, sAct, H, n, H, K, h->b[l], Zl, nullptr};
    gemm<1>(p, E, st);
    A = Hl; sA = sAct; K = H;
  }
  GemmP p{A, sA, K, 0, h->W[L], (int64_t)H * HO, HO, 0, tr->out.as<float>(), (int64_t)n * HO, HO, n, HO, H, h->b[L], nullptr, nullptr};
  gemm<2>(p, E, st);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

// One Adam update on the batch inputs[idx[e, col0 : col0 + n]] (bnn.py:423-426); col0 and the Adam step count advance on
// the device (tick_kernel), so the same launch sequence -- or its captured graph -- serves every batch.
static int train_one_batch(mopo_ens* h, const float* inputs, const float* targets, const int32_t* idx, int64_t ld_idx,
                           int bs, int n, float* loss_out, cudaStream_t st) {
  mopo_train* tr = h->tr;
  const int E = h->E, H = h->H, L = h->L, HO = h->HO, D1 = h->D1, in = h->in;
  const size_t rows = (size_t)E * n;
  MOPO_COUNT_LAUNCH;
  tick_kernel<<<1, 1, 0, st>>>(tr->ctr, tr->lr_t, tr->lr, bs);
  TRY(train_forward(h, inputs, targets, idx, ld_idx, tr->ctr + 2, n, st));
  float* scal = tr->scal.as<float>();          // [0,E): loss_e, [E, E+L+2): decay terms
  float* dOut = tr->dA.as<float>();
  MOPO_COUNT_LAUNCH;
  head_grad_kernel<<<grid1d((int64_t)rows * D1), 256, 0, st>>>(tr->out.as<float>(), tr->tg.as<float>(), h->max_lv, h->min_lv, E, n,
                                                             D1, dOut, tr->part.as<float>(), tr->deterministic ? 1 : 0);
  MOPO_COUNT_LAUNCH;
  head_reduce_kernel<<<D1 + E, 256, 0, st>>>(tr->part.as<float>(), E, n, D1, tr->g_lv, scal, tr->deterministic ? 0.0f : 0.01f);
  if (loss_out) {
    for (int l = 0; l < L; ++l) {
      MOPO_COUNT_LAUNCH;
      sumsq_kernel<<<1, 1024, 0, st>>>(h->W[l], (int64_t)(wsize(h, l) / H), H, 0, H, 0.5f * tr->decay[l], scal + E + l);
    }
    MOPO_COUNT_LAUNCH;
    sumsq_kernel<<<1, 1024, 0, st>>>(h->W[L], (int64_t)E * H, HO, 0, D1, 0.5f * tr->decay[L], scal + E + L);
    MOPO_COUNT_LAUNCH;
    sumsq_kernel<<<1, 1024, 0, st>>>(h->W[L], (int64_t)E * H, HO, D1, HO, 0.5f * tr->decay[L + 1], scal + E + L + 1);
    MOPO_COUNT_LAUNCH;
    loss_total_kernel<<<1, 1, 0, st>>>(scal, E, scal + E, L + 2, h->max_lv, h->min_lv, D1, loss_out, tr->deterministic ? 0.0f : 0.01f);
  }
  // backward.  dZ_l feeds three independent consumers: gW_l = In_l^T dZ_l, gb_l = colsum(dZ_l) (side stream) and the
  // chain dZ_{l-1} = (dZ_l W_l^T) .* swish'(Z_{l-1}) (main stream); every dZ_l has its own buffer so the branches overlap.
  const int wide = H > HO ? H : HO;
  float* dZ[MOPO_MAX_LAYERS];
  dZ[L] = dOut;
  for (int l = 0; l < L; ++l) dZ[l] = tr->dB.as<float>() + (size_t)l * rows * wide;
  for (int l = L; l >= 0; --l) {
    const int N = l == L ? HO : H, K = l == 0 ? in : H;
    const float* In = l == 0 ? tr->x.as<float>() : tr->Hh.as<float>() + (size_t)(l - 1) * rows * H;
    CUDA_TRY(cudaEventRecord(tr->ev_fork[l], st));
    CUDA_TRY(cudaStreamWaitEvent(tr->side, tr->ev_fork[l], 0));
    // gW[l] (K x N) = In^T (K x n) * dZ (n x N)
    GemmP pw{In, (int64_t)n * K, K, 1, dZ[l], (int64_t)n * N, N, 0, tr->gW[l], (int64_t)K * N, N, K, N, n, nullptr, nullptr, nullptr};
    gemm<0>(pw, E, tr->side);
    MOPO_COUNT_LAUNCH;
    colsum_kernel<<<dim3((N + 31) / 32, E), 256, 0, tr->side>>>(dZ[l], tr->gb[l], n, N, (int64_t)n * N);
    if (l > 0) {
      // dZ_{l-1} (n x K) = (dZ_l (n x N) * W_l^T (N x K)) .* swish'(Z_{l-1})
      const float* Zp = tr->Z.as<float>() + (size_t)(l - 1) * rows * H;
      GemmP pd{dZ[l], (int64_t)n * N, N, 0, h->W[l], (int64_t)K * N, N, 1, dZ[l - 1], (int64_t)n * K, K, n, K, N, nullptr, nullptr, Zp};
      gemm<3>(pd, E, st);
    }
  }
  CUDA_TRY(cudaEventRecord(tr->ev_join, tr->side));
  CUDA_TRY(cudaStreamWaitEvent(st, tr->ev_join, 0));
  AdamBatch ab;
  int nt = 0;
  size_t biggest = 0;
  auto add = [&](float* p, const float* g, float* m, float* v, size_t cnt, int ncols, int split, float da, float db) {
    ab.p[nt] = p; ab.g[nt] = g; ab.m[nt] = m; ab.v[nt] = v; ab.n[nt] = (int64_t)cnt; ab.ncols[nt] = ncols; ab.split[nt] = split;
    ab.decay_a[nt] = da; ab.decay_b[nt] = db; ++nt;
    if (cnt > biggest) biggest = cnt;
  };
  for (int l = 0; l <= L; ++l) {
    const int N = l == L ? HO : H;
    add(h->W[l], tr->gW[l], tr->mW[l], tr->vW[l], wsize(h, l), N, l == L ? D1 : N, tr->decay[l], l == L ? tr->decay[L + 1] : tr->decay[l]);
    add(h->b[l], tr->gb[l], tr->mb[l], tr->vb[l], bsize(h, l), N, N, 0.f, 0.f);
  }
  add(h->max_lv, tr->g_lv, tr->m_lv, tr->v_lv, D1, D1, D1, 0.f, 0.f);
  add(h->min_lv, tr->g_lv + D1, tr->m_lv + D1, tr->v_lv + D1, D1, D1, D1, 0.f, 0.f);
  ab.lr_t = tr->lr_t;
  int gx = (int)((biggest + 1023) / 1024); if (gx < 1) gx = 1; if (gx > 148) gx = 148;
  MOPO_COUNT_LAUNCH;
  adam_kernel<<<dim3(gx, nt), 256, 0, st>>>(ab);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

MOPO_API int mopo_ens_train_batches(mopo_ens* h, const float* inputs, const float* targets, int64_t n_rows, const int32_t* idx,
                                      int64_t ld_idx, int64_t n_cols, int32_t batch_size, float* loss_out, void* stream) {
  REQUIRE(h && inputs && targets && idx, MOPO_EINVAL, "mopo_ens_train_batches: null argument");
  REQUIRE(h->tr, MOPO_ESTATE, "mopo_ens_train_batches: call mopo_ens_train_begin first");
  REQUIRE(n_rows > 0 && n_cols >= 0 && ld_idx >= n_cols && batch_size > 0, MOPO_EINVAL,
          "mopo_ens_train_batches: n_rows=%lld n_cols=%lld ld_idx=%lld batch_size=%d", (long long)n_rows, (long long)n_cols,
          (long long)ld_idx, batch_size);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaMemsetAsync(h->tr->ctr, 0, sizeof(int64_t), st));      // first column of this call's first batch
  mopo_train* tr = h->tr;
  const int64_t n_full = n_cols / batch_size;
  int64_t k = 0;
  // Launch-bound loop (35 small kernels per batch): capture one full batch once, replay it for the others.
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (st) CUDA_TRY(cudaStreamIsCapturing(st, &cap));
  if (st && cap == cudaStreamCaptureStatusNone && !loss_out && n_full >= 4) {
    TRY(train_reserve(h, batch_size));
    bool fresh = false;
    if (!(tr->graph && tr->g_in == inputs && tr->g_tg == targets && tr->g_idx == idx && tr->g_ld == ld_idx && tr->g_bs == batch_size)) {
      if (tr->graph) { cudaGraphExecDestroy(tr->graph); tr->graph = nullptr; }
      cudaGraph_t g = nullptr;
      const long long before = (long long)mopo_launch_count();
      CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      int rc = train_one_batch(h, inputs, targets, idx, ld_idx, batch_size, batch_size, nullptr, st);
      cudaError_t ce = cudaStreamEndCapture(st, &g);
      if (rc != MOPO_OK) { if (g) cudaGraphDestroy(g); return rc; }
      CUDA_TRY(ce);
      tr->g_nodes = (int)((long long)mopo_launch_count() - before);
      cudaError_t ci = cudaGraphInstantiate(&tr->graph, g, 0);
      cudaGraphDestroy(g);
      CUDA_TRY(ci);
      tr->g_in = inputs; tr->g_tg = targets; tr->g_idx = idx; tr->g_ld = ld_idx; tr->g_bs = batch_size;
      fresh = true;                     // the launches counted during the capture stand for the first replay
    }
    for (; k < n_full; ++k) {
      CUDA_TRY(cudaGraphLaunch(tr->graph, st));
      if (!(fresh && k == 0)) for (int i = 0; i < tr->g_nodes; ++i) mopo_count_launch();
    }
    tr->t += n_full;
  }
  for (int64_t c0 = k * batch_size; c0 < n_cols; c0 += batch_size, ++k) {
    const int n = (int)(n_cols - c0 < batch_size ? n_cols - c0 : batch_size);
    TRY(train_one_batch(h, inputs, targets, idx, ld_idx, batch_size, n, loss_out ? loss_out + k : nullptr, st));
    tr->t += 1;
  }
  return MOPO_OK;
}

MOPO_API int mopo_ens_train_mse(mopo_ens* h, const float* inputs, const float* targets, const int32_t* idx, int64_t ld_idx,
                                  int64_t n, float* mse_out, void* stream) {
  REQUIRE(h && inputs && targets && mse_out, MOPO_EINVAL, "mopo_ens_train_mse: null argument");
  REQUIRE(h->tr, MOPO_ESTATE, "mopo_ens_train_mse: call mopo_ens_train_begin first");
  REQUIRE(n > 0 && n < (1 << 24) && (!idx || ld_idx >= n), MOPO_EINVAL, "mopo_ens_train_mse: n=%lld ld_idx=%lld", (long long)n, (long long)ld_idx);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  TRY(train_forward(h, inputs, targets, idx, ld_idx, nullptr, (int)n, st));
  MOPO_COUNT_LAUNCH;
  mse_kernel<<<h->E, 256, 0, st>>>(h->tr->out.as<float>(), h->tr->tg.as<float>(), (int)n, h->D1, mse_out);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

MOPO_API int mopo_ens_train_snapshot(mopo_ens* h, int32_t net, void* stream) {
  REQUIRE(h && h->tr, MOPO_ESTATE, "mopo_ens_train_snapshot: call mopo_ens_train_begin first");
  REQUIRE(net >= 0 && net < h->E, MOPO_EINVAL, "mopo_ens_train_snapshot: net %d outside [0,%d)", net, h->E);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  for (int l = 0; l <= h->L; ++l) {
    const size_t pw = wsize(h, l) / h->E, pb = bsize(h, l) / h->E;
    CUDA_TRY(cudaMemcpyAsync(h->tr->sW[l] + net * pw, h->W[l] + net * pw, pw * sizeof(float), cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(h->tr->sb[l] + net * pb, h->b[l] + net * pb, pb * sizeof(float), cudaMemcpyDeviceToDevice, st));
  }
  h->tr->snap[net] = true;
  return MOPO_OK;
}

MOPO_API int mopo_ens_train_restore(mopo_ens* h, void* stream) {
  REQUIRE(h && h->tr, MOPO_ESTATE, "mopo_ens_train_restore: call mopo_ens_train_begin first");
  for (int e = 0; e < h->E; ++e)     // bnn.py:298: _set_state indexes self._state[net] for every net (KeyError otherwise)
    REQUIRE(h->tr->snap[e], MOPO_ESTATE, "mopo_ens_train_restore: net %d has no saved state", e);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  for (int l = 0; l <= h->L; ++l) {
    CUDA_TRY(cudaMemcpyAsync(h->W[l], h->tr->sW[l], wsize(h, l) * sizeof(float), cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(h->b[l], h->tr->sb[l], bsize(h, l) * sizeof(float), cudaMemcpyDeviceToDevice, st));
  }
  return MOPO_OK;
}

MOPO_API int mopo_ens_train_end(mopo_ens* h, void* stream) {
  REQUIRE(h, MOPO_EINVAL, "mopo_ens_train_end: null argument");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  if (tc_supported(h)) TRY(tc_build_images(h, st));
  if (tc3_supported(h)) TRY(tc3_build_images(h, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return MOPO_OK;             // the optimiser state stays on the handle (see mopo_ens_train_begin)
}

MOPO_API int mopo_ens_train_reset(mopo_ens* h) {
  REQUIRE(h, MOPO_EINVAL, "mopo_ens_train_reset: null argument");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  mopo_train_release(h);
  return MOPO_OK;
}

MOPO_API int mopo_ens_train_set_deterministic(mopo_ens* h, int32_t deterministic) {
  REQUIRE(h && h->tr, MOPO_ESTATE, "mopo_ens_train_set_deterministic: call mopo_ens_train_begin first");
  if (h->tr->deterministic != (deterministic != 0) && h->tr->graph) { cudaGraphExecDestroy(h->tr->graph); h->tr->graph = nullptr; }
  h->tr->deterministic = deterministic != 0;
  return MOPO_OK;
}

MOPO_API int mopo_ens_get_params(mopo_ens* h, float* const* weights, float* const* biases, float* var_w, float* var_b,
                                   float* max_logvar, float* min_logvar, void* stream) {
  REQUIRE(h && weights && biases && max_logvar && min_logvar, MOPO_EINVAL, "mopo_ens_get_params: null argument");
  REQUIRE(h->params_set, MOPO_ESTATE, "mopo_ens_get_params: parameters not set");
  REQUIRE(!h->cfg.separate_mean_var || (var_w && var_b), MOPO_EINVAL, "mopo_ens_get_params: separate_mean_var needs var_w/var_b");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  const int E = h->E, H = h->H, L = h->L, D1 = h->D1, HO = h->HO;
  for (int l = 0; l < L; ++l) {
    CUDA_TRY(cudaMemcpyAsync(weights[l], h->W[l], wsize(h, l) * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(biases[l], h->b[l], bsize(h, l) * sizeof(float), cudaMemcpyDeviceToHost, st));
  }
  if (!h->cfg.separate_mean_var) {
    CUDA_TRY(cudaMemcpyAsync(weights[L], h->W[L], wsize(h, L) * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(biases[L], h->b[L], bsize(h, L) * sizeof(float), cudaMemcpyDeviceToHost, st));
  } else {
    TRY(h->misc.reserve(sizeof(float) * ((size_t)E * H * D1 + (size_t)E * D1)));
    float* tw = h->misc.as<float>();
    float* tb = tw + (size_t)E * H * D1;
    float* dstw[2] = {weights[L], var_w};
    float* dstb[2] = {biases[L], var_b};
    for (int s = 0; s < 2; ++s) {
      MOPO_COUNT_LAUNCH;
      take_cols_kernel<<<256, 256, 0, st>>>(h->W[L], tw, (int64_t)E * H, HO, D1, s * D1);
      MOPO_COUNT_LAUNCH;
      take_cols_kernel<<<8, 256, 0, st>>>(h->b[L], tb, (int64_t)E, HO, D1, s * D1);
      CUDA_TRY(cudaGetLastError());
      CUDA_TRY(cudaMemcpyAsync(dstw[s], tw, (size_t)E * H * D1 * sizeof(float), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaMemcpyAsync(dstb[s], tb, (size_t)E * D1 * sizeof(float), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaStreamSynchronize(st));
    }
  }
  CUDA_TRY(cudaMemcpyAsync(max_logvar, h->max_lv, D1 * sizeof(float), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(min_logvar, h->min_lv, D1 * sizeof(float), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return MOPO_OK;
}
