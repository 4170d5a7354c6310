// The SAC actor inside the model rollout: MOPO.get_action_meta (reference mopo/algorithms/mopo.py:468-485) over the
// graph of mlp_gaussian_policy + apply_squashing_func (mopo.py:281-312):
//   h = relu(relu(obs @ W0 + b0) @ W1 + b1);  mu = h @ Wmu + bmu;  log_std = clip(h @ Wls + bls, -20, 2)
//   pi = mu + N(0,1) * exp(log_std);  action = tanh(pi)      (deterministic: tanh(mu))
// so that act = policy(obs) at mopo.py:734 happens on the device between two model steps.
// This file: the C ABI (mopo_policy_*), the fp32 FFMA path (parity, rtol 1e-3) and the shared head math; the tcgen05
// path is policy_tc.cu.
#include "internal.cuh"
#include "philox.cuh"

namespace {

constexpr int PR = 16;          // rows per block of the fp32 layer kernel

// y[b, n] = act(x[b, :] @ W[:, n] + bias[n]) for PR rows per block, one thread per output column (W reads coalesce,
// x rows broadcast from shared memory).  Simple on purpose: this is the exact-fp32 parity path.
template <bool RELU>
__global__ void __launch_bounds__(256) policy_layer_fp32(const float* __restrict__ x, int ldx, const float* __restrict__ W,
                                                         const float* __restrict__ bias, float* __restrict__ y, int ldy,
                                                         int64_t n_max, const int* __restrict__ d_n, int K, int N) {
  extern __shared__ float sx[];                                   // [PR, K]
  const int64_t n = d_n ? min((int64_t)*d_n, n_max) : n_max;
  const int64_t r0 = (int64_t)blockIdx.x * PR;
  if (r0 >= n) return;
  const int rows = (int)min((int64_t)PR, n - r0);
  for (int i = threadIdx.x; i < PR * K; i += blockDim.x) {
    const int r = i / K, k = i - r * K;
    sx[i] = r < rows ? x[(r0 + r) * ldx + k] : 0.f;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < N; c += blockDim.x) {
    float acc[PR];
#pragma unroll
    for (int r = 0; r < PR; ++r) acc[r] = 0.f;
    for (int k = 0; k < K; ++k) {
      const float w = W[(int64_t)k * N + c];
#pragma unroll
      for (int r = 0; r < PR; ++r) acc[r] = fmaf(sx[r * K + k], w, acc[r]);
    }
    const float bb = bias[c];
    for (int r = 0; r < rows; ++r) {
      const float v = acc[r] + bb;
      y[(r0 + r) * ldy + c] = RELU ? fmaxf(v, 0.f) : v;
    }
  }
}

// Heads + sampling + squashing of one row (mopo.py:298-303, 287-289); h = last hidden activation [H].
struct HeadArgs {
  const float* h; int ldh;
  const float *Wmu, *bmu, *Wls, *bls;
  int H, act_dim;
  int64_t n_max; const int* d_n;
  int deterministic, random_uniform;
  const float* noise; uint64_t seed, step_id; const int32_t* origin; int64_t origin_base;
  float *actions, *mu_out, *std_out;
};

__device__ __forceinline__ float policy_noise(const HeadArgs& a, int64_t j, uint32_t orow, int d, float (&z)[4]) {
  if (a.noise) return a.noise[j * a.act_dim + d];
  if ((d & 3) == 0) philox_normal4(orow, (uint32_t)a.step_id, (uint32_t)(d >> 2), a.seed, z, MOPO_STREAM_POLICY);
  return z[d & 3];
}

__global__ void __launch_bounds__(128) policy_head_fp32(HeadArgs a) {
  const int64_t n = a.d_n ? min((int64_t)*a.d_n, a.n_max) : a.n_max;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t orow = a.origin ? (uint32_t)a.origin[j] : (uint32_t)(a.origin_base + j);
  float z[4];
  if (a.random_uniform) {                                          // mopo.py:738: np.random.uniform(-1, 1)
    for (int d = 0; d < a.act_dim; d += 4) {
      Philox4 r = philox4x32_10(orow, (uint32_t)a.step_id, (uint32_t)(d >> 2), MOPO_STREAM_POLICY, (uint32_t)a.seed, (uint32_t)(a.seed >> 32));
      const uint32_t w[4] = {r.x, r.y, r.z, r.w};
      for (int u = 0; u < 4 && d + u < a.act_dim; ++u) a.actions[j * a.act_dim + d + u] = 2.0f * mopo_u01(w[u]) - 1.0f;
    }
    return;
  }
  const float* hr = a.h + j * a.ldh;
  for (int d = 0; d < a.act_dim; ++d) {
    float mu = a.bmu[d], ls = a.bls[d];
    for (int k = 0; k < a.H; ++k) {
      const float hv = hr[k];
      mu = fmaf(hv, a.Wmu[(int64_t)k * a.act_dim + d], mu);
      ls = fmaf(hv, a.Wls[(int64_t)k * a.act_dim + d], ls);
    }
    ls = fminf(fmaxf(ls, -20.0f), 2.0f);                           // mopo.py:299 (LOG_STD_MIN / MAX)
    const float sd = expf(ls);                                      // mopo.py:300
    const float zz = a.deterministic ? 0.f : policy_noise(a, j, orow, d, z);
    a.actions[j * a.act_dim + d] = tanhf(fmaf(zz, sd, mu));         // mopo.py:301, 288-289
    if (a.mu_out) a.mu_out[j * a.act_dim + d] = mu;
    if (a.std_out) a.std_out[j * a.act_dim + d] = sd;
  }
}

}  // namespace

static int policy_fp32(mopo_policy* p, const PolicyRun& r, cudaStream_t st) {
  const int H = p->H;
  const int64_t B = r.n_max;
  HeadArgs ha = {};
  ha.H = H; ha.act_dim = p->act_dim; ha.n_max = B; ha.d_n = r.d_n;
  ha.deterministic = r.deterministic; ha.random_uniform = r.random_uniform;
  ha.noise = r.noise; ha.seed = r.seed; ha.step_id = r.step_id; ha.origin = r.origin; ha.origin_base = r.origin_base;
  ha.actions = r.actions; ha.mu_out = r.mu_out; ha.std_out = r.std_out;
  ha.Wmu = p->W[2]; ha.bmu = p->b[2]; ha.Wls = p->W[3]; ha.bls = p->b[3];
  if (!r.random_uniform) {
    TRY(p->h0.reserve(sizeof(float) * B * H));
    TRY(p->h1.reserve(sizeof(float) * B * H));
    const int grid = (int)((B + PR - 1) / PR);
    MOPO_COUNT_LAUNCH;
    policy_layer_fp32<true><<<grid, 256, sizeof(float) * PR * p->obs_dim, st>>>(r.obs, p->obs_dim, p->W[0], p->b[0], p->h0.as<float>(), H,
                                                                               B, r.d_n, p->obs_dim, H);
    MOPO_COUNT_LAUNCH;
    policy_layer_fp32<true><<<grid, 256, sizeof(float) * PR * H, st>>>(p->h0.as<float>(), H, p->W[1], p->b[1], p->h1.as<float>(), H,
                                                                       B, r.d_n, H, H);
    ha.h = p->h1.as<float>(); ha.ldh = H;
  }
  MOPO_COUNT_LAUNCH;
  policy_head_fp32<<<(int)((B + 127) / 128), 128, 0, st>>>(ha);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int policy_run(mopo_policy* p, const PolicyRun& r, cudaStream_t st) {
  if (r.n_max == 0) return MOPO_OK;
  if (r.precision == MOPO_PREC_BF16 && !r.random_uniform) {
    REQUIRE(policy_tc_supported(p), MOPO_EUNSUPPORTED, "policy: the tcgen05 path supports hidden_dim=256, obs_dim+2<=64, 2*act_dim<=32");
    return policy_tc_launch(p, r, st);
  }
  REQUIRE(r.precision == MOPO_PREC_FP32 || r.precision == MOPO_PREC_BF16 || r.precision == MOPO_PREC_FP32_REF, MOPO_EINVAL, "policy: unknown precision %d", r.precision);
  return policy_fp32(p, r, st);
}

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
MOPO_API int mopo_policy_create(int32_t obs_dim, int32_t act_dim, int32_t hidden_dim, int32_t device, mopo_policy** out) {
  REQUIRE(out && obs_dim > 0 && act_dim > 0 && hidden_dim > 0, MOPO_EINVAL, "mopo_policy_create: bad dims obs=%d act=%d hidden=%d", obs_dim, act_dim, hidden_dim);
  REQUIRE(act_dim <= 64, MOPO_EINVAL, "mopo_policy_create: act_dim=%d (max 64)", act_dim);
  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  REQUIRE(device >= 0 && device < ndev, MOPO_EINVAL, "mopo_policy_create: device %d of %d", device, ndev);
  CUDA_TRY(cudaSetDevice(device));
  mopo_policy* p = new mopo_policy();
  p->obs_dim = obs_dim; p->act_dim = act_dim; p->H = hidden_dim; p->device = device;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  p->sm_count = prop.multiProcessorCount;
  const int K[4] = {obs_dim, hidden_dim, hidden_dim, hidden_dim}, N[4] = {hidden_dim, hidden_dim, act_dim, act_dim};
  for (int l = 0; l < 4; ++l) {
    CUDA_TRY(cudaMalloc(&p->W[l], sizeof(float) * K[l] * N[l]));
    CUDA_TRY(cudaMalloc(&p->b[l], sizeof(float) * N[l]));
  }
  *out = p;
  return MOPO_OK;
}

MOPO_API int mopo_policy_destroy(mopo_policy* p) {
  if (!p) return MOPO_OK;
  cudaSetDevice(p->device);
  for (int l = 0; l < 4; ++l) { if (p->W[l]) cudaFree(p->W[l]); if (p->b[l]) cudaFree(p->b[l]); }
  if (p->tc_img) cudaFree(p->tc_img);
  p->h0.release(); p->h1.release();
  delete p;
  return MOPO_OK;
}

MOPO_API int mopo_policy_set_params(mopo_policy* p, const float* const* weights, const float* const* biases, int on_device, void* stream) {
  REQUIRE(p && weights && biases, MOPO_EINVAL, "mopo_policy_set_params: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(p->device));
  const int K[4] = {p->obs_dim, p->H, p->H, p->H}, N[4] = {p->H, p->H, p->act_dim, p->act_dim};
  const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  for (int l = 0; l < 4; ++l) {
    REQUIRE(weights[l] && biases[l], MOPO_EINVAL, "mopo_policy_set_params: layer %d null", l);
    CUDA_TRY(cudaMemcpyAsync(p->W[l], weights[l], sizeof(float) * K[l] * N[l], kind, st));
    CUDA_TRY(cudaMemcpyAsync(p->b[l], biases[l], sizeof(float) * N[l], kind, st));
  }
  if (policy_tc_supported(p)) TRY(policy_tc_build_image(p, st));
  if (!on_device) CUDA_TRY(cudaStreamSynchronize(st));     // the host arrays may be freed by the caller on return
  p->params_set = true;
  return MOPO_OK;
}

MOPO_API int mopo_policy_act(mopo_policy* p, const mopo_policy_args* a, void* stream) {
  REQUIRE(p && a, MOPO_EINVAL, "mopo_policy_act: null argument");
  REQUIRE(a->B >= 0 && a->B < (1ll << 31), MOPO_EINVAL, "mopo_policy_act: B=%lld", (long long)a->B);
  REQUIRE(p->params_set || a->random_uniform, MOPO_ESTATE, "mopo_policy_act: parameters not set");
  if (a->B == 0) return MOPO_OK;
  REQUIRE(a->obs && a->actions, MOPO_EINVAL, "mopo_policy_act: null obs/actions");
  CUDA_TRY(cudaSetDevice(p->device));
  PolicyRun r = {};
  r.obs = a->obs; r.n_max = a->B; r.d_n = nullptr;
  r.deterministic = a->deterministic; r.random_uniform = a->random_uniform; r.precision = a->precision;
  r.noise = a->noise; r.seed = a->seed; r.step_id = a->step_id; r.origin = a->origin; r.origin_base = a->origin_base;
  r.actions = a->actions; r.mu_out = a->mu_out; r.std_out = a->std_out;
  return policy_run(p, r, (cudaStream_t)stream);
}
