// C ABI of libmopo_b200.so (see include/mopo_b200.h for the reference interfaces replaced).
#include "internal.cuh"
#include "philox.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>

static thread_local char g_err[512] = "";

void mopo_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}


#include <atomic>
static std::atomic<long long> g_launches{0};
void mopo_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
MOPO_API int64_t mopo_launch_count(void) { return g_launches.load(); }

MOPO_API const char* mopo_last_error(void) { return g_err; }
MOPO_API int mopo_version(void) { return 200; }
#ifndef MOPO_BUILD_HASH
#define MOPO_BUILD_HASH "unhashed"
#endif
MOPO_API const char* mopo_build_hash(void) { return MOPO_BUILD_HASH; }

// ------------------------------------------------------------------------------------------
// ensemble
// ------------------------------------------------------------------------------------------
MOPO_API int mopo_ens_create(const mopo_ens_cfg* cfg, mopo_ens** out) {
  REQUIRE(cfg && out, MOPO_EINVAL, "mopo_ens_create: null argument");
  REQUIRE(cfg->obs_dim > 0 && cfg->act_dim > 0 && cfg->rew_dim >= 0 && cfg->hidden_dim > 0, MOPO_EINVAL,
          "mopo_ens_create: bad dims obs=%d act=%d rew=%d hidden=%d", cfg->obs_dim, cfg->act_dim, cfg->rew_dim, cfg->hidden_dim);
  REQUIRE(cfg->n_hidden >= 0 && cfg->n_hidden < MOPO_MAX_LAYERS - 1, MOPO_EINVAL, "mopo_ens_create: n_hidden=%d unsupported", cfg->n_hidden);
  // n_hidden == 0 is construct_model(model_type='linear') (constructor.py:25-27): one FC whose output finalize() doubles (bnn.py:185)
  REQUIRE(cfg->n_hidden >= 1 || !cfg->separate_mean_var, MOPO_EINVAL, "mopo_ens_create: a linear model has no hidden activation for a separate variance head");
  REQUIRE(cfg->num_networks >= 1 && cfg->num_networks <= MOPO_MAX_E, MOPO_EINVAL, "mopo_ens_create: num_networks=%d (max %d)", cfg->num_networks, MOPO_MAX_E);
  REQUIRE(cfg->num_elites >= 1 && cfg->num_elites <= cfg->num_networks, MOPO_EINVAL, "mopo_ens_create: num_elites=%d", cfg->num_elites);
  REQUIRE(cfg->obs_dim + cfg->rew_dim <= MOPO_MAX_D1, MOPO_EINVAL, "mopo_ens_create: obs_dim+rew_dim > %d", MOPO_MAX_D1);
  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  REQUIRE(cfg->device >= 0 && cfg->device < ndev, MOPO_EINVAL, "mopo_ens_create: device %d of %d", cfg->device, ndev);
  CUDA_TRY(cudaSetDevice(cfg->device));
  mopo_ens* h = new mopo_ens();
  h->cfg = *cfg;
  h->E = cfg->num_networks; h->in = cfg->obs_dim + cfg->act_dim; h->D1 = cfg->obs_dim + cfg->rew_dim;
  h->H = cfg->hidden_dim; h->L = cfg->n_hidden; h->HO = 2 * h->D1;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, cfg->device));
  h->sm_count = prop.multiProcessorCount;
  for (int l = 0; l <= h->L; ++l) {
    int K = l == 0 ? h->in : h->H, N = l == h->L ? h->HO : h->H;
    CUDA_TRY(cudaMalloc(&h->W[l], sizeof(float) * h->E * K * N));
    CUDA_TRY(cudaMalloc(&h->b[l], sizeof(float) * h->E * N));
  }
  CUDA_TRY(cudaMalloc(&h->mu, sizeof(float) * h->in));
  CUDA_TRY(cudaMalloc(&h->sigma, sizeof(float) * h->in));
  CUDA_TRY(cudaMalloc(&h->max_lv, sizeof(float) * h->D1));
  CUDA_TRY(cudaMalloc(&h->min_lv, sizeof(float) * h->D1));
  CUDA_TRY(cudaMalloc(&h->elites, sizeof(int32_t) * h->E));
  *out = h;
  return MOPO_OK;
}

MOPO_API int mopo_ens_destroy(mopo_ens* h) {
  if (!h) return MOPO_OK;
  cudaSetDevice(h->cfg.device);
  for (int l = 0; l < MOPO_MAX_LAYERS; ++l) { if (h->W[l]) cudaFree(h->W[l]); if (h->b[l]) cudaFree(h->b[l]); }
  cudaFree(h->mu); cudaFree(h->sigma); cudaFree(h->max_lv); cudaFree(h->min_lv); cudaFree(h->elites);
  if (h->tc_img) cudaFree(h->tc_img);
  if (h->tc3_img) cudaFree(h->tc3_img);
  mopo_train_release(h);
  for (cudaEvent_t e : h->act_ev) cudaEventDestroy(e);
  if (h->ev_start) cudaEventDestroy(h->ev_start);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  h->st_obs.release(); h->st_act.release(); h->ready_flags.release();
  if (h->flag_src) cudaFreeHost(h->flag_src);
  for (cudaEvent_t e : h->ev) cudaEventDestroy(e);
  DevBuf* bufs[] = {&h->sel, &h->selmv, &h->pen2, &h->xn, &h->xtiles, &h->actA, &h->actB, &h->mean, &h->lv, &h->blk_counts, &h->misc,
                    &h->ro_obs[0], &h->ro_obs[1], &h->ro_origin[0], &h->ro_origin[1], &h->ro_counts, &h->ro_act};
  for (DevBuf* b : bufs) b->release();
  delete h;
  return MOPO_OK;
}

// dst[e, k, col0 + n] = src[e, k, n]   (fuse the mean head and the separate logvar head side by side)
__global__ void place_cols_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int n_src, int n_dst, int col0) {
  int64_t tot = (int64_t)rows * n_src;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / n_src; int c = (int)(i % n_src);
    dst[r * n_dst + col0 + c] = src[i];
  }
}

static int upload(float* dst, const float* src, size_t n, int on_device, cudaStream_t st) {
  CUDA_TRY(cudaMemcpyAsync(dst, src, n * sizeof(float), on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
  return MOPO_OK;
}

MOPO_API int mopo_ens_set_params(mopo_ens* h, const float* const* weights, const float* const* biases,
                                   const float* var_w, const float* var_b, const float* mu, const float* sigma,
                                   const float* max_logvar, const float* min_logvar, int on_device, void* stream) {
  REQUIRE(h && weights && biases && mu && sigma && max_logvar && min_logvar, MOPO_EINVAL, "mopo_ens_set_params: null argument");
  REQUIRE(!h->cfg.separate_mean_var || (var_w && var_b), MOPO_EINVAL, "mopo_ens_set_params: separate_mean_var needs var_w/var_b");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const int E = h->E, H = h->H, D1 = h->D1;
  for (int l = 0; l < h->L; ++l) {
    REQUIRE(weights[l] && biases[l], MOPO_EINVAL, "mopo_ens_set_params: layer %d null", l);
    int K = l == 0 ? h->in : H;
    TRY(upload(h->W[l], weights[l], (size_t)E * K * H, on_device, st));
    TRY(upload(h->b[l], biases[l], (size_t)E * H, on_device, st));
  }
  REQUIRE(weights[h->L] && biases[h->L], MOPO_EINVAL, "mopo_ens_set_params: head null");
  if (!h->cfg.separate_mean_var) {                       // bnn.py:185: last layer already 2*D1 wide
    TRY(upload(h->W[h->L], weights[h->L], (size_t)E * (h->L == 0 ? h->in : H) * h->HO, on_device, st));
    TRY(upload(h->b[h->L], biases[h->L], (size_t)E * h->HO, on_device, st));
  } else {                                               // bnn.py:657-667: both heads read the last hidden activation
    TRY(h->misc.reserve(sizeof(float) * ((size_t)E * H * D1 + (size_t)E * D1)));
    float* tw = h->misc.as<float>();
    float* tb = tw + (size_t)E * H * D1;
    const float* srcw[2] = {weights[h->L], var_w};
    const float* srcb[2] = {biases[h->L], var_b};
    for (int s = 0; s < 2; ++s) {
      TRY(upload(tw, srcw[s], (size_t)E * H * D1, on_device, st));
      TRY(upload(tb, srcb[s], (size_t)E * D1, on_device, st));
      MOPO_COUNT_LAUNCH, place_cols_kernel<<<256, 256, 0, st>>>(tw, h->W[h->L], E * H, D1, h->HO, s * D1);
      MOPO_COUNT_LAUNCH, place_cols_kernel<<<8, 256, 0, st>>>(tb, h->b[h->L], E, D1, h->HO, s * D1);
      CUDA_TRY(cudaGetLastError());
      if (!on_device) CUDA_TRY(cudaStreamSynchronize(st));   // host staging may be reused by the caller
    }
  }
  TRY(upload(h->mu, mu, h->in, on_device, st));
  TRY(upload(h->sigma, sigma, h->in, on_device, st));
  TRY(upload(h->max_lv, max_logvar, D1, on_device, st));
  TRY(upload(h->min_lv, min_logvar, D1, on_device, st));
  if (tc_supported(h)) TRY(tc_build_images(h, st));
  if (tc3_supported(h)) TRY(tc3_build_images(h, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  h->params_set = true;
  return MOPO_OK;
}

// Device addresses of the fp32 master images (reference layout), for in-place collectives (NCCL broadcast of the ensemble after
// model training: mopo_b200/dist.py) -- followed by mopo_ens_refresh_images, which rebuilds the tensor-core images from them.
MOPO_API int mopo_ens_param_image(mopo_ens* h, int32_t which, int32_t layer, void** ptr, int64_t* count) {
  REQUIRE(h && ptr && count, MOPO_EINVAL, "mopo_ens_param_image: null argument");
  const int E = h->E;
  switch (which) {
    case MOPO_IMG_W:
      REQUIRE(layer >= 0 && layer <= h->L, MOPO_EINVAL, "mopo_ens_param_image: layer %d", layer);
      *ptr = h->W[layer]; *count = (int64_t)E * (layer == 0 ? h->in : h->H) * (layer == h->L ? h->HO : h->H); return MOPO_OK;
    case MOPO_IMG_B:
      REQUIRE(layer >= 0 && layer <= h->L, MOPO_EINVAL, "mopo_ens_param_image: layer %d", layer);
      *ptr = h->b[layer]; *count = (int64_t)E * (layer == h->L ? h->HO : h->H); return MOPO_OK;
    case MOPO_IMG_MU: *ptr = h->mu; *count = h->in; return MOPO_OK;
    case MOPO_IMG_SIGMA: *ptr = h->sigma; *count = h->in; return MOPO_OK;
    case MOPO_IMG_MAX_LOGVAR: *ptr = h->max_lv; *count = h->D1; return MOPO_OK;
    case MOPO_IMG_MIN_LOGVAR: *ptr = h->min_lv; *count = h->D1; return MOPO_OK;
  }
  mopo_set_error("mopo_ens_param_image: which=%d", which);
  return MOPO_EINVAL;
}

MOPO_API int mopo_ens_refresh_images(mopo_ens* h, void* stream) {
  REQUIRE(h, MOPO_EINVAL, "mopo_ens_refresh_images: null handle");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  if (tc_supported(h)) TRY(tc_build_images(h, st));
  if (tc3_supported(h)) TRY(tc3_build_images(h, st));
  h->params_set = true;
  return MOPO_OK;
}

MOPO_API int mopo_ens_set_elites(mopo_ens* h, const int32_t* elites, int32_t n) {
  REQUIRE(h && elites, MOPO_EINVAL, "mopo_ens_set_elites: null argument");
  REQUIRE(n >= 1 && n <= h->E, MOPO_EINVAL, "mopo_ens_set_elites: n=%d outside [1,%d]", n, h->E);
  for (int i = 0; i < n; ++i) REQUIRE(elites[i] >= 0 && elites[i] < h->E, MOPO_EINVAL, "mopo_ens_set_elites: index %d out of range", elites[i]);
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  CUDA_TRY(cudaMemcpy(h->elites, elites, sizeof(int32_t) * n, cudaMemcpyHostToDevice));
  h->n_elites = n;
  h->elites_set = true;
  return MOPO_OK;
}

// Which kernel a precision request runs on.  MOPO_PREC_FP32: the fp32-accurate tcgen05 kernel (split bf16, ens_tc3.cu)
// where the shape allows, else the FFMA kernel; MOPO_PREC_FP32_REF: always FFMA (the checker); MOPO_PREC_BF16: the
// net-stationary bf16 kernel (ens_tc.cu), or the streamed-weights kernel in bf16 mode for wider shapes (ant, humanoid).
enum { PATH_FFMA = 0, PATH_TC = 1, PATH_TC3 = 2, PATH_TC1 = 3, PATH_NONE = -1 };
static int ens_path(const mopo_ens* h, int precision) {
  static const bool no_tc3 = getenv("MOPO_NO_TC3") != nullptr;      // developer switch: MOPO_PREC_FP32 on the FFMA kernel
  if (precision == MOPO_PREC_FP32_REF) return PATH_FFMA;
  if (precision == MOPO_PREC_FP32) return (!no_tc3 && tc3_supported(h)) ? PATH_TC3 : PATH_FFMA;
  if (precision == MOPO_PREC_BF16) return tc_supported(h) ? PATH_TC : tc3_supported(h) ? PATH_TC1 : PATH_NONE;
  return PATH_NONE;
}
#define REQUIRE_PATH(path, precision)                                                                                   \
  do {                                                                                                                  \
    REQUIRE((precision) == MOPO_PREC_FP32 || (precision) == MOPO_PREC_BF16 || (precision) == MOPO_PREC_FP32_REF, MOPO_EINVAL, \
            "unknown precision %d", (precision));                                                                       \
    REQUIRE((path) != PATH_NONE, MOPO_EUNSUPPORTED,                                                                     \
            "the tcgen05 paths support hidden_dim=200, n_hidden=4, obs+act<=63, 2*(obs+rew)<=96 only (use MOPO_PREC_FP32)"); \
  } while (0)

static int forward(mopo_ens* h, int precision, const float* xn, int64_t n_max, const int* d_n, float* mean, float* lv,
                   int64_t ldb, cudaStream_t st) {
  const int path = ens_path(h, precision);
  REQUIRE_PATH(path, precision);
  if (path == PATH_FFMA) return fp32_forward(h, xn, n_max, d_n, mean, lv, ldb, st);
  if (path == PATH_TC) return tc_forward(h, xn, n_max, d_n, mean, lv, ldb, st);
  Tc3In x = {xn, nullptr, h->in, 0, false};
  TcOut o = {};
  o.mean = mean; o.lv = lv; o.ldb = ldb;
  return tc3_launch(h, path == PATH_TC3 ? 3 : 1, x, n_max, d_n, o, st);
}

MOPO_API int mopo_ens_forward(mopo_ens* h, const float* xn, int64_t B, float* mean, float* logvar, int precision, void* stream) {
  REQUIRE(h && (B == 0 || (xn && mean && logvar)), MOPO_EINVAL, "mopo_ens_forward: null argument");
  REQUIRE(B >= 0 && B < (1ll << 31), MOPO_EINVAL, "mopo_ens_forward: B=%lld", (long long)B);
  REQUIRE(h->params_set, MOPO_ESTATE, "mopo_ens_forward: parameters not set");
  if (B == 0) return MOPO_OK;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  return forward(h, precision, xn, B, nullptr, mean, logvar, B, (cudaStream_t)stream);
}

// Per-launch device timing of the ensemble-forward kernel as it runs inside steps / rollouts: CUDA events recorded
// on the launching stream right around each launch.  enable != 0 starts (and clears) collection; the query
// synchronises the events, returns the summed milliseconds and the number of launches, and clears.
MOPO_API int mopo_kernel_timing(mopo_ens* h, int enable) {
  REQUIRE(h, MOPO_EINVAL, "mopo_kernel_timing: null handle");
  h->timing = enable == 1;            // 1: the forward kernel (tc_launch)
  h->timing_tail = enable == 2;       // 2: the step tail inside mopo_rollout (measurement of the rest of the step)
  h->ev_used = 0;
  return MOPO_OK;
}
MOPO_API int mopo_kernel_time(mopo_ens* h, double* ms_total, int64_t* launches) {
  REQUIRE(h && ms_total && launches, MOPO_EINVAL, "mopo_kernel_time: null argument");
  double tot = 0;
  for (size_t i = 0; i + 1 < h->ev_used; i += 2) {
    CUDA_TRY(cudaEventSynchronize(h->ev[i + 1]));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
    tot += ms;
  }
  *ms_total = tot; *launches = (int64_t)(h->ev_used / 2);
  h->ev_used = 0;
  return MOPO_OK;
}

#ifdef MOPO_DEBUG_PROBES
__global__ void fill_mod_kernel(int32_t* p, int64_t n, int m) {   // m > 0: i % m ; m <= 0: the constant -m
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = m > 0 ? (int32_t)(i % m) : -m;
}

MOPO_API int mopo_debug_tc_trace(mopo_ens* h, const float* xn, int64_t B, float* mean, float* logvar,
                                 unsigned long long* trace, void* stream) {
  REQUIRE(h && xn && trace && B > 0, MOPO_EINVAL, "mopo_debug_tc_trace: bad argument");
  REQUIRE(h->params_set && tc_supported(h), MOPO_ESTATE, "mopo_debug_tc_trace: tensor-core path unavailable");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaMemsetAsync(trace, 0, sizeof(unsigned long long) * 5 * 2048, st));
  h->tc_trace = trace;
  int rc;
  if (mean && logvar) {
    rc = mopo_ens_forward(h, xn, B, mean, logvar, MOPO_PREC_BF16, stream);
  } else {                                       // fused-head mode: row i "chooses" net i % 5
    rc = h->sel.reserve(sizeof(int32_t) * B);
    if (!rc) rc = h->selmv.reserve(sizeof(float2) * B * h->D1);
    if (!rc) rc = h->pen2.reserve(sizeof(float) * 2 * h->E * B);
    if (!rc) { const char* md = getenv("MOPO_TRACE_SELMOD"); fill_mod_kernel<<<148, 256, 0, st>>>(h->sel.as<int32_t>(), B, md ? atoi(md) : 5); rc = tc_tiles_from_xn(h, xn, B, nullptr, st); }
    if (!rc) {
      TcOut o = {};
      o.sel = h->sel.as<int32_t>(); o.selmv = h->selmv.as<float2>(); o.pen2 = h->pen2.as<float>(); o.ldb = B;
      rc = tc_launch(h, B, nullptr, o, st);
    }
  }
  h->tc_trace = nullptr;
  return rc;
}

// the same for the streamed-weights kernel (ens_tc3.cu): precision MOPO_PREC_FP32 (three-term) or MOPO_PREC_BF16 on a shape
// the resident kernel rejects; roles 0 MMA issuer, 1/2 epilogue warps 0 / 4, 3 weight streamer, 4 head warp 12
MOPO_API int mopo_debug_tc3_trace(mopo_ens* h, const float* xn, int64_t B, float* mean, float* logvar, int terms,
                                  unsigned long long* trace, void* stream) {
  REQUIRE(h && xn && trace && B > 0 && (terms == 1 || terms == 3), MOPO_EINVAL, "mopo_debug_tc3_trace: bad argument");
  REQUIRE(h->params_set && tc3_supported(h), MOPO_ESTATE, "mopo_debug_tc3_trace: streamed tensor-core path unavailable");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaMemsetAsync(trace, 0, sizeof(unsigned long long) * 5 * 2048, st));
  h->tc_trace = trace;
  Tc3In x = {xn, nullptr, h->in, 0, false};
  TcOut o = {};
  int rc = MOPO_OK;
  if (mean && logvar) { o.mean = mean; o.lv = logvar; o.ldb = B; }
  else {
    rc = h->sel.reserve(sizeof(int32_t) * B);
    if (!rc) rc = h->selmv.reserve(sizeof(float2) * B * h->D1);
    if (!rc) rc = h->pen2.reserve(sizeof(float) * 2 * h->E * B);
    if (!rc) { fill_mod_kernel<<<148, 256, 0, st>>>(h->sel.as<int32_t>(), B, 5); o.sel = h->sel.as<int32_t>(); o.selmv = h->selmv.as<float2>(); o.pen2 = h->pen2.as<float>(); o.ldb = B; }
  }
  if (!rc) rc = tc3_launch(h, terms, x, B, nullptr, o, st);
  h->tc_trace = nullptr;
  return rc;
}
#endif  // MOPO_DEBUG_PROBES

MOPO_API int mopo_ens_predict(mopo_ens* h, const float* x, int64_t B, float* mean, float* var, int precision, void* stream) {
  REQUIRE(h && (B == 0 || (x && mean && var)), MOPO_EINVAL, "mopo_ens_predict: null argument");
  REQUIRE(B >= 0 && B < (1ll << 31), MOPO_EINVAL, "mopo_ens_predict: B=%lld", (long long)B);
  REQUIRE(h->params_set, MOPO_ESTATE, "mopo_ens_predict: parameters not set");
  if (B == 0) return MOPO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  TRY(h->xn.reserve(sizeof(float) * B * h->in));
  // prep with act_dim folded: x is already concat(obs, act)
  TRY(launch_prep(x, x, h->mu, h->sigma, h->xn.as<float>(), B, nullptr, h->in, 0, st));
  TRY(h->lv.reserve(sizeof(float) * h->E * B * h->D1));
  TRY(forward(h, precision, h->xn.as<float>(), B, nullptr, mean, h->lv.as<float>(), B, st));
  TRY(launch_exp(h->lv.as<float>(), var, (int64_t)h->E * B * h->D1, st));
  return MOPO_OK;
}

// BNN.predict with 3-D inputs (bnn.py:541-545 -> fc.py:100-101): net e sees its own rows x[e].  x [E,B,in] device fp32 ->
// mean / var [E,B,D1].  The scaler runs over all E*B rows, the contraction reads them with a per-net stride.
MOPO_API int mopo_ens_predict3d(mopo_ens* h, const float* x, int64_t B, float* mean, float* var, int precision, void* stream) {
  REQUIRE(h && (B == 0 || (x && mean && var)), MOPO_EINVAL, "mopo_ens_predict3d: null argument");
  REQUIRE(B >= 0 && (int64_t)h->E * B < (1ll << 31), MOPO_EINVAL, "mopo_ens_predict3d: B=%lld", (long long)B);
  REQUIRE(h->params_set, MOPO_ESTATE, "mopo_ens_predict3d: parameters not set");
  if (B == 0) return MOPO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  int path = ens_path(h, precision);
  if (path == PATH_TC) path = tc3_supported(h) ? PATH_TC1 : PATH_NONE;     // the net-stationary kernel tiles ONE input for all nets
  REQUIRE_PATH(path, precision);
  const int64_t rows = (int64_t)h->E * B;
  TRY(h->xn.reserve(sizeof(float) * rows * h->in));
  TRY(launch_prep(x, x, h->mu, h->sigma, h->xn.as<float>(), rows, nullptr, h->in, 0, st));
  TRY(h->lv.reserve(sizeof(float) * rows * h->D1));
  if (path == PATH_FFMA) TRY(fp32_forward(h, h->xn.as<float>(), B, nullptr, mean, h->lv.as<float>(), B, st, B * (int64_t)h->in));
  else {
    Tc3In in = {h->xn.as<float>(), nullptr, h->in, 0, false, B};
    TcOut o = {};
    o.mean = mean; o.lv = h->lv.as<float>(); o.ldb = B;
    TRY(tc3_launch(h, path == PATH_TC3 ? 3 : 1, in, B, nullptr, o, st));
  }
  return launch_exp(h->lv.as<float>(), var, rows * h->D1, st);
}

// One ensemble pass of a FakeEnv.step: scaler + forward.  On the tensor-core path a stochastic step with the
// learned-variance (or no) penalty and no mixture diagnostics takes the fused head: only the chosen net's mean /
// variance row and the per-net variance sums leave the SM.  Fills the TailArgs inputs accordingly.
static int step_forward(mopo_ens* h, int precision, bool want_diag, int penalty_kind, int deterministic,
                        const float* obs, const float* act, int64_t B, const int* d_n, const int32_t* model_inds,
                        uint64_t seed, uint64_t step_id, const int32_t* origin, int64_t origin_base,
                        TailArgs* t, cudaStream_t st, bool tiles_ready = false, bool pdl = false, const InputReady* ready = nullptr) {
  const int E = h->E, D1 = h->D1;
  const int path = ens_path(h, precision);
  REQUIRE_PATH(path, precision);
  if (path != PATH_FFMA) {
    const bool fused = !deterministic && !want_diag && penalty_kind != MOPO_PEN_DISAGREEMENT;
    TcPrep q = {};
    q.obs = obs; q.act = act; q.n_max = B; q.d_n = d_n; q.model_inds = model_inds;
    q.seed = seed; q.step_id = step_id; q.origin = origin; q.origin_base = origin_base;
    TcOut o = {};
    if (fused) {
      TRY(h->sel.reserve(sizeof(int32_t) * B));
      TRY(h->selmv.reserve(sizeof(float2) * B * D1));
      TRY(h->pen2.reserve(sizeof(float) * 2 * E * B));
      q.sel = h->sel.as<int32_t>();
      o.sel = q.sel; o.selmv = h->selmv.as<float2>(); o.pen2 = h->pen2.as<float>(); o.ldb = B;
      t->sel = o.sel; t->selmv = o.selmv; t->pen2 = o.pen2;
    } else {
      TRY(h->mean.reserve(sizeof(float) * E * B * D1));
      TRY(h->lv.reserve(sizeof(float) * E * B * D1));
      o.mean = h->mean.as<float>(); o.lv = h->lv.as<float>(); o.ldb = B;
      t->mean = o.mean; t->lv = o.lv;
    }
    t->ldb = B;
    o.pdl = pdl && tiles_ready;                                // ... behind which this launch is a programmatic dependent
    if (path == PATH_TC) {
      if (!tiles_ready) TRY(tc_tiles_from_step(h, q, st));     // else: written by the previous step's fused tail
      return tc_launch(h, B, d_n, o, st);
    }
    // streamed-weights kernel: reads obs / act itself (scaler and hi/lo split inside); only the elite picks come first
    if (fused && !tiles_ready) TRY(tc3_pick(h, q.sel, model_inds, B, d_n, seed, step_id, origin, origin_base, st));
    Tc3In x = {obs, act, h->cfg.obs_dim, h->cfg.act_dim, true};
    if (ready) x.ready = *ready;
    return tc3_launch(h, path == PATH_TC3 ? 3 : 1, x, B, d_n, o, st);
  }
  TRY(h->xn.reserve(sizeof(float) * B * h->in));
  TRY(h->mean.reserve(sizeof(float) * E * B * D1));
  TRY(h->lv.reserve(sizeof(float) * E * B * D1));
  TRY(launch_prep(obs, act, h->mu, h->sigma, h->xn.as<float>(), B, d_n, h->cfg.obs_dim, h->cfg.act_dim, st));
  TRY(fp32_forward(h, h->xn.as<float>(), B, d_n, h->mean.as<float>(), h->lv.as<float>(), B, st));
  t->mean = h->mean.as<float>(); t->lv = h->lv.as<float>(); t->ldb = B;
  return MOPO_OK;
}

static int check_step_common(const mopo_ens* h, int term_kind, int penalty_kind, int deterministic, const char* who) {
  REQUIRE(h->params_set, MOPO_ESTATE, "%s: parameters not set", who);
  REQUIRE(term_kind >= MOPO_TERM_NEVER && term_kind <= MOPO_TERM_HUMANOID, MOPO_EINVAL, "%s: term_kind=%d", who, term_kind);
  REQUIRE(penalty_kind >= MOPO_PEN_NONE && penalty_kind <= MOPO_PEN_DISAGREEMENT, MOPO_EINVAL, "%s: penalty_kind=%d", who, penalty_kind);
  REQUIRE(term_kind == MOPO_TERM_NEVER || term_kind == MOPO_TERM_HUMANOID || h->cfg.obs_dim >= 2, MOPO_EINVAL, "%s: termination needs obs_dim >= 2", who);
  (void)deterministic;
  return MOPO_OK;
}

MOPO_API int mopo_env_step(mopo_ens* h, const mopo_step_args* a, void* stream) {
  REQUIRE(h && a, MOPO_EINVAL, "mopo_env_step: null argument");
  REQUIRE(a->B >= 0 && a->B < (1ll << 31), MOPO_EINVAL, "mopo_env_step: B=%lld", (long long)a->B);
  TRY(check_step_common(h, a->term_kind, a->penalty_kind, a->deterministic, "mopo_env_step"));
  REQUIRE(a->deterministic || a->model_inds || h->elites_set, MOPO_ESTATE, "mopo_env_step: elites not set (BNN._model_inds exists only after train, bnn.py:337-340)");
  if (a->B == 0) return MOPO_OK;
  REQUIRE(a->obs && a->act && a->next_obs && a->rewards && a->terminals, MOPO_EINVAL, "mopo_env_step: null obs/act/outputs");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const int64_t B = a->B;
  TailArgs t;
  memset(&t, 0, sizeof(t));
  TRY(step_forward(h, a->precision, a->log_prob || a->dev, a->penalty_kind, a->deterministic, a->obs, a->act, B, nullptr,
                   a->model_inds, a->seed, a->step_id, a->origin, 0, &t, st));
  t.obs = a->obs; t.act = a->act; t.n_max = B; t.d_n = nullptr;
  t.E = h->E; t.D1 = h->D1; t.obs_dim = h->cfg.obs_dim; t.act_dim = h->cfg.act_dim;
  t.term_kind = a->term_kind; t.penalty_kind = a->penalty_kind; t.coeff = a->penalty_coeff; t.deterministic = a->deterministic;
  t.noise = a->noise; t.model_inds = a->model_inds; t.elites = h->elites; t.n_elites = h->n_elites;
  t.seed = a->seed; t.step_id = a->step_id; t.origin = a->origin; t.origin_base = 0;
  t.next_obs = a->next_obs; t.rewards = a->rewards; t.terminals = a->terminals;
  t.penalty = a->penalty; t.unpenalized = a->unpenalized; t.model_means = a->model_means; t.model_stds = a->model_stds;
  t.log_prob = a->log_prob; t.dev = a->dev; t.model_inds_out = a->model_inds_out;
  TRY(launch_tail(t, st));
  return MOPO_OK;
}

MOPO_API int mopo_termination_fn(int32_t term_kind, const float* next_obs, int64_t B, int32_t obs_dim, uint8_t* out, void* stream) {
  REQUIRE(term_kind >= MOPO_TERM_NEVER && term_kind <= MOPO_TERM_HUMANOID, MOPO_EINVAL, "mopo_termination_fn: term_kind=%d", term_kind);
  REQUIRE(B >= 0 && obs_dim >= 1 && (B == 0 || (next_obs && out)), MOPO_EINVAL, "mopo_termination_fn: bad argument");
  REQUIRE(term_kind == MOPO_TERM_NEVER || term_kind == MOPO_TERM_HUMANOID || obs_dim >= 2, MOPO_EINVAL, "mopo_termination_fn: needs obs_dim >= 2");
  return launch_term_only(term_kind, next_obs, B, obs_dim, out, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------
// pool
// ------------------------------------------------------------------------------------------
MOPO_API int mopo_pool_create(int32_t obs_dim, int32_t act_dim, int64_t max_size, int32_t device, mopo_pool** out) {
  REQUIRE(out && obs_dim > 0 && act_dim > 0 && max_size > 0, MOPO_EINVAL, "mopo_pool_create: bad argument");
  CUDA_TRY(cudaSetDevice(device));
  mopo_pool* p = new mopo_pool();
  p->obs_dim = obs_dim; p->act_dim = act_dim; p->device = device; p->max_size = max_size; p->size = 0; p->pointer = 0;
  CUDA_TRY(cudaMalloc(&p->actions, sizeof(float) * max_size * act_dim));
  CUDA_TRY(cudaMalloc(&p->rewards, sizeof(float) * max_size));
  CUDA_TRY(cudaMalloc(&p->terminals, max_size));
  CUDA_TRY(cudaMalloc(&p->obs, sizeof(float) * max_size * obs_dim));
  CUDA_TRY(cudaMalloc(&p->next_obs, sizeof(float) * max_size * obs_dim));
  // flexible_replay_pool.py:39-43: fields are zero-initialised
  CUDA_TRY(cudaMemset(p->actions, 0, sizeof(float) * max_size * act_dim));
  CUDA_TRY(cudaMemset(p->rewards, 0, sizeof(float) * max_size));
  CUDA_TRY(cudaMemset(p->terminals, 0, max_size));
  CUDA_TRY(cudaMemset(p->obs, 0, sizeof(float) * max_size * obs_dim));
  CUDA_TRY(cudaMemset(p->next_obs, 0, sizeof(float) * max_size * obs_dim));
  *out = p;
  return MOPO_OK;
}

MOPO_API int mopo_pool_destroy(mopo_pool* p) {
  if (!p) return MOPO_OK;
  cudaSetDevice(p->device);
  cudaFree(p->actions); cudaFree(p->rewards); cudaFree(p->terminals); cudaFree(p->obs); cudaFree(p->next_obs);
  delete p;
  return MOPO_OK;
}
MOPO_API int64_t mopo_pool_size(const mopo_pool* p) { return p ? p->size : -1; }
MOPO_API int64_t mopo_pool_pointer(const mopo_pool* p) { return p ? p->pointer : -1; }
MOPO_API int64_t mopo_pool_max_size(const mopo_pool* p) { return p ? p->max_size : -1; }
MOPO_API void* mopo_pool_field(const mopo_pool* p, int32_t f) {
  if (!p) return nullptr;
  switch (f) {
    case MOPO_F_ACTIONS: return p->actions;
    case MOPO_F_REWARDS: return p->rewards;
    case MOPO_F_TERMINALS: return p->terminals;
    case MOPO_F_OBS: return p->obs;
    case MOPO_F_NEXT_OBS: return p->next_obs;
  }
  return nullptr;
}

// rows [skip, n) of src -> ring rows (base + j) % max ; flattened so stores are coalesced per field
template <class T>
__global__ void ring_scatter_kernel(const T* __restrict__ src, T* __restrict__ dst, int64_t n, int64_t skip, int width,
                                    int64_t base, int64_t max_size) {
  int64_t tot = (n - skip) * width;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t j = skip + i / width; int c = (int)(i % width);
    dst[((base + j) % max_size) * width + c] = src[j * width + c];
  }
}
template <class T>
__global__ void gather_rows_kernel(const T* __restrict__ src, const int64_t* __restrict__ idx, T* __restrict__ dst, int64_t n, int width) {
  int64_t tot = n * width;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t j = i / width; int c = (int)(i % width);
    dst[i] = src[idx[j] * width + c];
  }
}
__global__ void philox_index_kernel(int64_t* idx, int64_t n, uint64_t seed, uint64_t offset, int64_t size) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    // counter = (row of this draw, draw id): two draws with different `offset` share no counter, so their index sets are
    // independent (np.random.randint per batch, flexible_replay_pool.py:85-87)
    Philox4 r = philox_pool_draw((uint32_t)i, offset, seed);
    uint64_t u = ((uint64_t)r.x << 32) | r.y;
    idx[i] = (int64_t)__umul64hi(u, (uint64_t)size);     // floor(u * size / 2^64): np.random.randint(0, size)
  }
}

static int grid_for(int64_t tot) { int64_t g = (tot + 255) / 256; if (g < 1) g = 1; if (g > 148 * 16) g = 148 * 16; return (int)g; }

MOPO_API int mopo_pool_add(mopo_pool* p, const float* obs, const float* act, const float* next_obs,
                             const float* rew, const uint8_t* term, int64_t n, void* stream) {
  REQUIRE(p && n >= 0, MOPO_EINVAL, "mopo_pool_add: bad argument");
  if (n == 0) return MOPO_OK;
  REQUIRE(obs && act && next_obs && rew && term, MOPO_EINVAL, "mopo_pool_add: null field");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(p->device));
  const int64_t skip = n > p->max_size ? n - p->max_size : 0;   // duplicates: the last write wins
  MOPO_COUNT_LAUNCH, ring_scatter_kernel<<<grid_for((n - skip) * p->act_dim), 256, 0, st>>>(act, p->actions, n, skip, p->act_dim, p->pointer, p->max_size);
  MOPO_COUNT_LAUNCH, ring_scatter_kernel<<<grid_for(n - skip), 256, 0, st>>>(rew, p->rewards, n, skip, 1, p->pointer, p->max_size);
  MOPO_COUNT_LAUNCH, ring_scatter_kernel<<<grid_for(n - skip), 256, 0, st>>>(term, p->terminals, n, skip, 1, p->pointer, p->max_size);
  MOPO_COUNT_LAUNCH, ring_scatter_kernel<<<grid_for((n - skip) * p->obs_dim), 256, 0, st>>>(obs, p->obs, n, skip, p->obs_dim, p->pointer, p->max_size);
  MOPO_COUNT_LAUNCH, ring_scatter_kernel<<<grid_for((n - skip) * p->obs_dim), 256, 0, st>>>(next_obs, p->next_obs, n, skip, p->obs_dim, p->pointer, p->max_size);
  CUDA_TRY(cudaGetLastError());
  p->pointer = (p->pointer + n) % p->max_size;                 // flexible_replay_pool.py:45-48
  p->size = p->size + n < p->max_size ? p->size + n : p->max_size;
  return MOPO_OK;
}

// Rows [pointer, pointer + n) were written in place (e.g. by an NCCL all-gather straight into the field arrays): advance
// the ring exactly as add_samples would (flexible_replay_pool.py:45-48).  The caller keeps pointer + n <= max_size.
MOPO_API int mopo_pool_advance(mopo_pool* p, int64_t n) {
  REQUIRE(p && n >= 0 && p->pointer + n <= p->max_size, MOPO_EINVAL, "mopo_pool_advance: n=%lld from pointer %lld exceeds max_size %lld",
          (long long)n, p ? (long long)p->pointer : 0ll, p ? (long long)p->max_size : 0ll);
  p->pointer = (p->pointer + n) % p->max_size;
  p->size = p->size + n < p->max_size ? p->size + n : p->max_size;
  return MOPO_OK;
}

static int pool_gather_impl(const mopo_pool* p, const int64_t* idx, int64_t n, float* obs, float* act, float* next_obs,
                            float* rew, uint8_t* term, cudaStream_t st) {
  if (act) MOPO_COUNT_LAUNCH, gather_rows_kernel<<<grid_for(n * p->act_dim), 256, 0, st>>>(p->actions, idx, act, n, p->act_dim);
  if (rew) MOPO_COUNT_LAUNCH, gather_rows_kernel<<<grid_for(n), 256, 0, st>>>(p->rewards, idx, rew, n, 1);
  if (term) MOPO_COUNT_LAUNCH, gather_rows_kernel<<<grid_for(n), 256, 0, st>>>(p->terminals, idx, term, n, 1);
  if (obs) MOPO_COUNT_LAUNCH, gather_rows_kernel<<<grid_for(n * p->obs_dim), 256, 0, st>>>(p->obs, idx, obs, n, p->obs_dim);
  if (next_obs) MOPO_COUNT_LAUNCH, gather_rows_kernel<<<grid_for(n * p->obs_dim), 256, 0, st>>>(p->next_obs, idx, next_obs, n, p->obs_dim);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

MOPO_API int mopo_pool_gather(const mopo_pool* p, const int64_t* idx, int64_t n, float* obs, float* act,
                                float* next_obs, float* rew, uint8_t* term, void* stream) {
  REQUIRE(p && n >= 0 && (n == 0 || idx), MOPO_EINVAL, "mopo_pool_gather: bad argument");
  if (n == 0) return MOPO_OK;
  CUDA_TRY(cudaSetDevice(p->device));
  return pool_gather_impl(p, idx, n, obs, act, next_obs, rew, term, (cudaStream_t)stream);
}

MOPO_API int mopo_pool_sample(const mopo_pool* p, int64_t n, uint64_t seed, uint64_t offset, int64_t* idx_out,
                                float* obs, float* act, float* next_obs, float* rew, uint8_t* term, void* stream) {
  REQUIRE(p && n >= 0 && (n == 0 || idx_out), MOPO_EINVAL, "mopo_pool_sample: bad argument");
  REQUIRE(p->size > 0 || n == 0, MOPO_ESTATE, "mopo_pool_sample: pool is empty");
  if (n == 0) return MOPO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(p->device));
  MOPO_COUNT_LAUNCH, philox_index_kernel<<<grid_for(n), 256, 0, st>>>(idx_out, n, seed, offset, p->size);
  CUDA_TRY(cudaGetLastError());
  return pool_gather_impl(p, idx_out, n, obs, act, next_obs, rew, term, st);
}

// ------------------------------------------------------------------------------------------
// rollout
// ------------------------------------------------------------------------------------------
__global__ void rollout_init_kernel(int* counts, int64_t* pool_off, int n0, int h) {
  if (threadIdx.x == 0) { counts[0] = n0; for (int i = 1; i <= h; ++i) counts[i] = 0; *pool_off = 0; }
}

MOPO_API int mopo_rollout(mopo_ens* h, mopo_pool* pool, const mopo_rollout_args* a, void* stream) {
  REQUIRE(h && pool && a, MOPO_EINVAL, "mopo_rollout: null argument");
  REQUIRE(a->B >= 0 && a->B < (1ll << 31) && a->rollout_length >= 0 && a->rollout_length <= 1024, MOPO_EINVAL, "mopo_rollout: B=%lld h=%d", (long long)a->B, a->rollout_length);
  REQUIRE(pool->obs_dim == h->cfg.obs_dim && pool->act_dim == h->cfg.act_dim && pool->device == h->cfg.device, MOPO_EINVAL, "mopo_rollout: pool/ensemble mismatch");
  TRY(check_step_common(h, a->term_kind, a->penalty_kind, a->deterministic, "mopo_rollout"));
  REQUIRE(a->deterministic || a->model_inds || h->elites_set, MOPO_ESTATE, "mopo_rollout: elites not set");
  const int H = a->rollout_length;
  const int64_t B = a->B;
  if (a->steps_added) for (int i = 0; i < H; ++i) a->steps_added[i] = 0;
  if (B == 0 || H == 0) return MOPO_OK;
  REQUIRE(a->start_obs && (a->actions || a->policy) && a->steps_added, MOPO_EINVAL, "mopo_rollout: null start_obs/actions/steps_added");
  for (int i = 0; i < H; ++i) REQUIRE(a->policy || a->actions[i], MOPO_EINVAL, "mopo_rollout: actions[%d] null (and no policy)", i);
  if (a->policy) {
    REQUIRE(a->policy->obs_dim == h->cfg.obs_dim && a->policy->act_dim == h->cfg.act_dim && a->policy->device == h->cfg.device, MOPO_EINVAL, "mopo_rollout: policy/ensemble mismatch");
    REQUIRE(a->policy->params_set || a->rollout_random, MOPO_ESTATE, "mopo_rollout: policy parameters not set");
  }
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const int od = h->cfg.obs_dim, ad = h->cfg.act_dim;
  const int nblk = (int)((B + tail_rows_per_block() - 1) / tail_rows_per_block());
  TRY(h->blk_counts.reserve(sizeof(int) * nblk));
  // ro_obs[0/1]: dense obs of the current step / next_obs of the current step ; ro_origin likewise
  for (int k = 0; k < 2; ++k) { TRY(h->ro_obs[k].reserve(sizeof(float) * B * od)); TRY(h->ro_origin[k].reserve(sizeof(int32_t) * B)); }
  TRY(h->misc.reserve(sizeof(float) * B * od + B));                       // dense next_obs + terminals of the step
  TRY(h->ro_counts.reserve(sizeof(int) * (H + 1) + sizeof(int64_t) * 2));
  int64_t* d_pool_off = h->ro_counts.as<int64_t>();                          // [0] = rows appended so far
  int* d_counts = reinterpret_cast<int*>(d_pool_off + 2);
  float* nobs_dense = h->misc.as<float>();
  uint8_t* term_dense = reinterpret_cast<uint8_t*>(nobs_dense + B * od);
  // halfcheetah-style termination_fn (all False, halfcheetah.py:9-11): nothing ever leaves the batch, so the row count,
  // the row order and the ring offsets of every step are known on the host: no device counters, no compaction pass,
  // and the step tail writes next_obs straight into the next step's observation buffer.
  const bool static_rows = a->term_kind == MOPO_TERM_NEVER;
  if (!static_rows) MOPO_COUNT_LAUNCH, rollout_init_kernel<<<1, 32, 0, st>>>(d_counts, d_pool_off, (int)B, H);
  const float* obs_cur = a->start_obs;
  const int32_t* origin_cur = nullptr;
  // static rows on the fused tensor-core path: the tail of step i also tiles step i+1's inputs and picks its elites
  // (with a device policy, act[i+1] = policy(next_obs[i]) does not exist yet when the tail of step i runs: no chaining)
  const int path = ens_path(h, a->precision);
  REQUIRE_PATH(path, a->precision);
  const bool fused_ok = path != PATH_FFMA && !a->deterministic && a->penalty_kind != MOPO_PEN_DISAGREEMENT;
  const bool streamed = path == PATH_TC3 || path == PATH_TC1;   // the forward reads obs / act itself: "tiles" = the elite picks
  const bool chain = static_rows && fused_ok && !a->policy;
  if (a->policy) TRY(h->ro_act.reserve(sizeof(float) * B * ad));
  // terminating envs on the fused tensor-core path with supplied actions: the compaction tiles the next step's inputs
  const bool compact_tiles = !static_rows && fused_ok && !a->policy;
  const bool policy_tiles = a->policy && !a->rollout_random && path == PATH_TC && fused_ok && policy_tc_supported(a->policy);
  // programmatic dependent launches between the chained kernels (not while per-kernel timing events sit between them)
  static const bool pdl_env = getenv("MOPO_NO_PDL") == nullptr;
  const bool pdl = chain && pdl_env && !h->timing && !h->timing_tail;
  // the same between policy -> forward -> tail -> policy when the policy kernel is the tiling pass of a static-row rollout
  const bool pdl_pol = policy_tiles && static_rows && pdl_env && !h->timing && !h->timing_tail;
  for (int i = 0; i < H; ++i) {
    const int* d_n = static_rows ? nullptr : d_counts + i;
    TailArgs t;
    memset(&t, 0, sizeof(t));
    const uint64_t step_id = (a->rollout_id << 16) ^ (uint64_t)i;
    // actions[i] is first read by step i's tiling pass -- or, when steps are chained, by the tail of step i-1, so the
    // chained rollout waits for actions[i+1] ahead of step i's forward (which leaves forward -> tail adjacent in the
    // stream for the programmatic dependent launch; the upload has had a whole step to arrive by then)
    if (a->action_events) {
      if (a->action_events[i] && (i == 0 || !chain)) CUDA_TRY(cudaStreamWaitEvent(st, (cudaEvent_t)a->action_events[i], 0));
      if (chain && i + 1 < H && a->action_events[i + 1]) CUDA_TRY(cudaStreamWaitEvent(st, (cudaEvent_t)a->action_events[i + 1], 0));
    }
    const float* act_i = (a->actions && a->actions[i]) ? a->actions[i] : nullptr;
    bool tiles_done = false;
    if (!act_i) {                                      // act = policy(obs) on the device (mopo.py:734 / :736-738)
      PolicyRun pr = {};
      pr.obs = obs_cur; pr.n_max = B; pr.d_n = d_n;
      pr.deterministic = a->policy_deterministic; pr.random_uniform = a->rollout_random; pr.precision = a->precision;
      pr.seed = a->seed; pr.step_id = step_id; pr.origin = origin_cur; pr.origin_base = a->origin_base;
      pr.actions = h->ro_act.as<float>();
      if (policy_tiles) {                              // the tcgen05 policy kernel is also this step's tiling pass
        int K0;
        TRY(tc_reserve_xtiles(h, B, &K0));
        TRY(h->sel.reserve(sizeof(int32_t) * B));
        pr.xt = h->xtiles.as<uint4>(); pr.ens_mu = h->mu; pr.ens_sigma = h->sigma; pr.ens_K0 = K0;
        pr.sel = h->sel.as<int32_t>(); pr.model_inds = a->model_inds ? a->model_inds[i] : nullptr;
        pr.elites = h->elites; pr.n_elites = h->n_elites; pr.ens_step_id = step_id;
        pr.pdl = pdl_pol && i > 0;                     // behind the previous step's tail
      }
      TRY(policy_run(a->policy, pr, st));
      act_i = pr.actions;
      tiles_done = policy_tiles;
    }
    TRY(step_forward(h, a->precision, false, a->penalty_kind, a->deterministic, obs_cur, act_i, B, d_n,
                     a->model_inds ? a->model_inds[i] : nullptr, a->seed, step_id, origin_cur, a->origin_base, &t, st,
                     (chain && i > 0) || tiles_done || (compact_tiles && i > 0), (pdl && !a->action_events) || (pdl_pol && tiles_done),
                     (h->host_ready_on && streamed && (int)h->host_ready.size() == H) ? &h->host_ready[i] : nullptr));
    if (i == 0 && h->after_first_forward) {                 // (mopo_rollout_host: the kernel is polling the first chunk's flag by now)
      int (*hook)(void*) = h->after_first_forward;
      h->after_first_forward = nullptr;
      TRY(hook(h->after_first_forward_ctx));
    }
    t.pdl = (chain && pdl) || (pdl_pol && tiles_done);
    t.obs_behind = t.pdl && i > 0;
    if (h->host_ready_on && streamed && (int)h->host_ready.size() == H) t.ready = h->host_ready[i];
    if (chain && i + 1 < H) {
      if (!streamed) {
        t.xt_next = h->xtiles.as<uint4>(); t.act_next = a->actions[i + 1]; t.mu = h->mu; t.sigma = h->sigma;
        t.K0 = (h->in + 2 + 15) / 16 * 16;
      }
      t.sel_next = h->sel.as<int32_t>(); t.model_inds_next = a->model_inds ? a->model_inds[i + 1] : nullptr;
      t.step_id_next = (a->rollout_id << 16) ^ (uint64_t)(i + 1);
    }
    t.obs = obs_cur; t.act = act_i; t.n_max = B; t.d_n = d_n;
    t.E = h->E; t.D1 = h->D1; t.obs_dim = od; t.act_dim = ad;
    t.term_kind = a->term_kind; t.penalty_kind = a->penalty_kind; t.coeff = a->penalty_coeff; t.deterministic = a->deterministic;
    t.noise = a->noise ? a->noise[i] : nullptr; t.model_inds = a->model_inds ? a->model_inds[i] : nullptr;
    t.elites = h->elites; t.n_elites = h->n_elites;
    t.seed = a->seed; t.step_id = step_id; t.origin = origin_cur; t.origin_base = a->origin_base;
    float* obs_next = h->ro_obs[i & 1].as<float>();
    t.next_obs = static_rows ? obs_next : nobs_dense; t.rewards = nullptr; t.terminals = static_rows ? nullptr : term_dense;
    t.p_act = pool->actions; t.p_rew = pool->rewards; t.p_obs = pool->obs; t.p_nobs = pool->next_obs; t.p_term = pool->terminals;
    t.pool_max = pool->max_size;
    if (static_rows) { t.pool_base = (pool->pointer + (int64_t)i * B) % pool->max_size; t.d_pool_off = nullptr; t.blk_counts = nullptr; }
    else { t.pool_base = pool->pointer; t.d_pool_off = d_pool_off; t.blk_counts = h->blk_counts.as<int>(); }
    cudaEvent_t tv0 = nullptr, tv1 = nullptr;
    if (h->timing_tail) {
      if (h->ev_used + 2 > h->ev.size()) for (int k = 0; k < 2; ++k) { cudaEvent_t e; CUDA_TRY(cudaEventCreate(&e)); h->ev.push_back(e); }
      tv0 = h->ev[h->ev_used]; tv1 = h->ev[h->ev_used + 1]; h->ev_used += 2;
      CUDA_TRY(cudaEventRecord(tv0, st));
    }
    TRY(launch_tail(t, st));
    if (tv1) CUDA_TRY(cudaEventRecord(tv1, st));
    if (static_rows) { obs_cur = obs_next; continue; }
    int32_t* origin_next = h->ro_origin[i & 1].as<int32_t>();
    CompactTiles ct = {};
    if (compact_tiles && i + 1 < H) {                  // the compaction is also step i+1's tiling pass / elite pick
      TRY(h->sel.reserve(sizeof(int32_t) * B));
      if (!streamed) {
        int K0;
        TRY(tc_reserve_xtiles(h, B, &K0));
        ct.xt = h->xtiles.as<uint4>(); ct.act_next = a->actions[i + 1]; ct.mu = h->mu; ct.sigma = h->sigma; ct.act_dim = ad; ct.K0 = K0;
      }
      ct.sel = h->sel.as<int32_t>(); ct.model_inds = a->model_inds ? a->model_inds[i + 1] : nullptr;
      ct.elites = h->elites; ct.n_elites = h->n_elites; ct.seed = a->seed; ct.step_id = (a->rollout_id << 16) ^ (uint64_t)(i + 1);
      if (!streamed && a->action_events && a->action_events[i + 1]) CUDA_TRY(cudaStreamWaitEvent(st, (cudaEvent_t)a->action_events[i + 1], 0));
    }
    TRY(launch_compact(nobs_dense, term_dense, origin_cur, a->origin_base, B, d_n, h->blk_counts.as<int>(), od,
                       obs_next, origin_next, d_counts + i + 1, d_pool_off, ct, st));
    obs_cur = obs_next; origin_cur = origin_next;
  }
  std::vector<int> counts(H + 1, (int)B);
  // static rows: every count is known on the host, so the call returns without synchronising (stream-ordered like every
  // other entry point); otherwise one D2H of the h counts + one synchronise, at the end
  if (!static_rows) {
    CUDA_TRY(cudaMemcpyAsync(counts.data(), d_counts, sizeof(int) * (H + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
  }
  int64_t total = 0;
  for (int i = 0; i < H; ++i) { a->steps_added[i] = counts[i]; total += counts[i]; }
  pool->pointer = (pool->pointer + total) % pool->max_size;      // == h successive add_samples calls
  pool->size = pool->size + total < pool->max_size ? pool->size + total : pool->max_size;
  return MOPO_OK;
}

// mopo_rollout with HOST inputs (the reference-facing call: MOPO hands NumPy arrays).  start_obs [B,obs] and every
// actions[i] [B,act] are host pointers (pinned memory makes the copies asynchronous).  They are staged into device buffers
// owned by the handle: the start states and step 0's actions on `stream`, the later steps' actions on an internal copy
// stream behind one event each -- the rollout waits for event i only ahead of the first kernel that reads actions[i], so
// those uploads overlap the earlier steps.  noise / model_inds (optional) stay device pointers.  Everything else as mopo_rollout.
MOPO_API int mopo_rollout_host(mopo_ens* h, mopo_pool* pool, const mopo_rollout_args* a, void* stream) {
  REQUIRE(h && pool && a, MOPO_EINVAL, "mopo_rollout_host: null argument");
  REQUIRE(a->B >= 0 && a->B < (1ll << 31) && a->rollout_length >= 0 && a->rollout_length <= 1024, MOPO_EINVAL, "mopo_rollout_host: B=%lld h=%d", (long long)a->B, a->rollout_length);
  const int H = a->rollout_length;
  const int64_t B = a->B;
  if (B == 0 || H == 0) return mopo_rollout(h, pool, a, stream);
  REQUIRE(a->start_obs && (a->actions || a->policy), MOPO_EINVAL, "mopo_rollout_host: null start_obs/actions");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  const int od = h->cfg.obs_dim, ad = h->cfg.act_dim;
  if (!h->copy_stream) {
    CUDA_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&h->ev_start, cudaEventDisableTiming));
  }
  while ((int)h->act_ev.size() < H) {
    cudaEvent_t e;
    CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->act_ev.push_back(e);
  }
  const bool host_actions = a->actions != nullptr && !a->policy;
  // growing a staging buffer frees the old one: wait for its readers first
  if (h->st_obs.cap < sizeof(float) * B * od || (host_actions && h->st_act.cap < sizeof(float) * B * ad * H)) CUDA_TRY(cudaDeviceSynchronize());
  TRY(h->st_obs.reserve(sizeof(float) * B * od));
  if (host_actions) TRY(h->st_act.reserve(sizeof(float) * B * ad * H));
  std::vector<const float*> d_act(H, nullptr);
  std::vector<void*> evs(H, nullptr);
  const int path = ens_path(h, a->precision);
  // Streamed-weights kernel: its forward reads obs / act itself, so the rollout need not wait for the uploads at all -- the
  // inputs go up in chunks on the copy stream, each followed by a stream-ordered 4-byte copy into its flag; step 0's forward
  // starts at once, runs its units tile-major (every pair strides through the tiles in upload order) and polls the flag of a
  // tile's chunk before it touches the rows (first chunks small: the kernel is launched behind the first 0.3 MB, the other chunks
  // are enqueued while it already runs); the later steps' actions likewise, in fewer chunks.  The step tail, which may start
  // under the forward (programmatic launch), looks at the flag once and otherwise reads its inputs behind the forward.
  // (MOPO_NO_INPUT_FLAGS: developer A/B switch back to whole-array uploads behind events.)
  // (not under stream capture: a graph would hold the kernel and the copies it polls for as unordered branches)
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  CUDA_TRY(cudaStreamIsCapturing(st, &cap));
  const bool flagged = host_actions && (path == PATH_TC3 || path == PATH_TC1) && cap == cudaStreamCaptureStatusNone &&
                       getenv("MOPO_NO_INPUT_FLAGS") == nullptr;
  h->host_ready_on = false;
  if (flagged) {
    if (h->ready_flags.cap < sizeof(uint32_t) * 8 * H) {
      CUDA_TRY(cudaDeviceSynchronize());
      TRY(h->ready_flags.reserve(sizeof(uint32_t) * 8 * H));
      CUDA_TRY(cudaMemset(h->ready_flags.p, 0, h->ready_flags.cap));
    }
    if (!h->flag_src) {                                              // pinned constants 0x01010101 b: the sources of the flag writes
      CUDA_TRY(cudaHostAlloc(reinterpret_cast<void**>(&h->flag_src), 256 * sizeof(uint32_t), cudaHostAllocDefault));
      for (uint32_t bb = 0; bb < 256; ++bb) h->flag_src[bb] = 0x01010101u * bb;
    }
    h->ready_epoch = h->ready_epoch % 255 + 1;                       // 1 .. 255: never the byte the flags hold from the call before
    const int byte = (int)h->ready_epoch;
    const int nT = (int)((B + 255) / 256);
    h->host_ready.assign(H, InputReady());
    // the copy stream starts behind the previous rollout on `st`, which may still read the staging buffers
    CUDA_TRY(cudaEventRecord(h->ev_start, st));
    CUDA_TRY(cudaStreamWaitEvent(h->copy_stream, h->ev_start, 0));
    uint32_t* flags = h->ready_flags.as<uint32_t>();
    for (int i = 0; i < H; ++i) {
      REQUIRE(a->actions[i], MOPO_EINVAL, "mopo_rollout_host: actions[%d] null", i);
      d_act[i] = h->st_act.as<float>() + (size_t)i * B * ad;
      InputReady& r = h->host_ready[i];
      r.flags = flags + 8 * i; r.value = 0x01010101u * (uint32_t)byte;
      const int div0[5] = {16, 8, 4, 2, 1}, div1[4] = {8, 4, 2, 1}, div2[2] = {2, 1};   // later steps have had time: fewer chunks
      const int* dv = i == 0 ? div0 : i == 1 ? div1 : div2;
      const int n_frac = i == 0 ? 5 : i == 1 ? 4 : 2;
      r.n = 0;
      for (int c = 0; c < n_frac; ++c) {
        const int end = (nT + dv[c] - 1) / dv[c];
        if (r.n == 0 || end > r.end_tile[r.n - 1]) r.end_tile[r.n++] = end;
      }
    }
    // chunk c of step i: [obs rows,] action rows, then the flag
    struct Upload {
      mopo_ens* h; const mopo_rollout_args* a; int64_t B; int od, ad, H, byte; const float* const* d_act; const uint32_t* flag_src;
      int chunk(int i, int c) const {
        const InputReady& r = h->host_ready[i];
        const int64_t row0 = c ? std::min<int64_t>(B, (int64_t)r.end_tile[c - 1] * 256) : 0, row1 = std::min<int64_t>(B, (int64_t)r.end_tile[c] * 256);
        const int64_t rows = row1 - row0;
        cudaStream_t cs = h->copy_stream;
        if (i == 0) CUDA_TRY(cudaMemcpyAsync(h->st_obs.as<float>() + row0 * od, a->start_obs + row0 * od, sizeof(float) * rows * od, cudaMemcpyHostToDevice, cs));
        CUDA_TRY(cudaMemcpyAsync(const_cast<float*>(d_act[i]) + row0 * ad, a->actions[i] + row0 * ad, sizeof(float) * rows * ad, cudaMemcpyHostToDevice, cs));
        CUDA_TRY(cudaMemcpyAsync(const_cast<uint32_t*>(r.flags) + c, flag_src + byte, sizeof(uint32_t), cudaMemcpyHostToDevice, cs));   // (a copy-engine write)
        return MOPO_OK;
      }
      int rest() const {
        for (int i = 0; i < H; ++i)
          for (int c = i == 0 ? 1 : 0; c < h->host_ready[i].n; ++c) TRY(chunk(i, c));
        return MOPO_OK;
      }
      static int hook(void* self) { return static_cast<const Upload*>(self)->rest(); }
    } up = {h, a, B, od, ad, H, byte, d_act.data(), h->flag_src};
    TRY(up.chunk(0, 0));                               // the first 1/16 of step 0's rows; everything else once its forward is launched
    h->after_first_forward = &Upload::hook; h->after_first_forward_ctx = &up;
    h->host_ready_on = true;
    mopo_rollout_args b = *a;
    b.start_obs = h->st_obs.as<float>();
    b.actions = d_act.data(); b.action_events = nullptr;
    int rc = mopo_rollout(h, pool, &b, stream);
    h->host_ready_on = false;
    if (h->after_first_forward) {                      // the rollout failed before its first forward: the caller's buffers must still not be left half read
      h->after_first_forward = nullptr;
      const int rc2 = up.rest();
      if (rc == MOPO_OK) rc = rc2;
    }
    return rc;
  }
  CUDA_TRY(cudaMemcpyAsync(h->st_obs.p, a->start_obs, sizeof(float) * B * od, cudaMemcpyHostToDevice, st));
  if (host_actions) {
    for (int i = 0; i < H; ++i) {
      REQUIRE(a->actions[i], MOPO_EINVAL, "mopo_rollout_host: actions[%d] null", i);
      float* dst = h->st_act.as<float>() + (size_t)i * B * ad;
      d_act[i] = dst;
      if (i == 0) {
        CUDA_TRY(cudaMemcpyAsync(dst, a->actions[0], sizeof(float) * B * ad, cudaMemcpyHostToDevice, st));
        // The copy stream starts behind step 0's inputs: they are on the critical path and get the whole link (uploads on
        // two streams would share it), and behind the previous rollout on `st`, which may still read the staging buffers.
        CUDA_TRY(cudaEventRecord(h->ev_start, st));
        CUDA_TRY(cudaStreamWaitEvent(h->copy_stream, h->ev_start, 0));
      } else {
        CUDA_TRY(cudaMemcpyAsync(dst, a->actions[i], sizeof(float) * B * ad, cudaMemcpyHostToDevice, h->copy_stream));
        CUDA_TRY(cudaEventRecord(h->act_ev[i], h->copy_stream));
        evs[i] = h->act_ev[i];
      }
    }
  }
  mopo_rollout_args b = *a;
  b.start_obs = h->st_obs.as<float>();
  if (host_actions) { b.actions = d_act.data(); b.action_events = evs.data(); }
  return mopo_rollout(h, pool, &b, stream);
}
