// fp32 FFMA ensemble forward: the rtol-1e-3 parity path (MOPO_PREC_FP32).
//
// Computes BNN._compile_outputs (reference mopo/models/bnn.py:631-675) layer by layer:
//   hidden l : out[e,b,:] = swish(in[e,b,:] @ W_l[e] + b_l[e])          (fc.py:96-106, swish fc.py:21)
//   head     : [mean | raw] = h4[e,b,:] @ Whead[e] + bhead[e]
//              logvar = max_lv - softplus(max_lv - raw); logvar = min_lv + softplus(logvar - min_lv)
// Layer 0 reads the scaled input shared by all nets (the 2-D einsum path, fc.py:98-99).
// One launch per layer, activations [E,chunk,H] ping-pong through HBM: simple and exact;
// the tensor-core path in ens_tc.cu is the fast one.
#include "gemm_fp32.cuh"   // load4, softplus_f, TM / TN
#include <mutex>

namespace {

constexpr int BM = 64, BN = 64;

__device__ __forceinline__ float swish_f(float x) { return x / (1.0f + expf(-x)); }

// MODE 0: swish -> C[e, m, n] (ldc = N);  MODE 1: head -> mean/lv [e, m, D1]
// 64 x 64 x 32 tiles, 256 threads x (4 x 4) outputs; 16-byte global loads along each operand's contiguous dimension where
// the pitch allows, the next k-tile's loads in flight under the current tile's FMAs, 16-byte shared-memory operand reads.
// Every output is one fmaf chain over k = 0 .. K-1 in order: the result does not depend on the tiling.
template <int MODE>
__global__ void __launch_bounds__(256)
ens_layer_fp32(const float* __restrict__ A, int64_t strideA, int lda,
               const float* __restrict__ W, const float* __restrict__ bias,
               float* __restrict__ C, int64_t strideC,
               float* __restrict__ mean, float* __restrict__ lv, int64_t ld_out,
               const float* __restrict__ max_lv, const float* __restrict__ min_lv,
               int M_max, const int* __restrict__ d_n, int N, int K, int D1) {
  constexpr int BK = 32, NQ = 2;
  const int M = d_n ? min(*d_n, M_max) : M_max;
  const int e = blockIdx.z;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  if (m0 >= M) return;
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const float* Ae = A + (int64_t)e * strideA;
  const float* We = W + (int64_t)e * K * N;
  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;     // 16 x 16 threads, each TM x TN
  const bool vA = (lda % 4 == 0) && (strideA % 4 == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
  const bool vB = (N % 4 == 0) && ((reinterpret_cast<uintptr_t>(W) & 15) == 0);
  // A tile: row m = f / 8, k4 = 4 (f % 8);  W tile: row k = f / 16, n4 = 4 (f % 16);  f = tid + 256 q
  auto fetchA = [&](int k0, int q) { const int f = tid + 256 * q; return load4(Ae, m0 + f / 8, lda, k0 + 4 * (f % 8), M, K, vA); };
  auto fetchB = [&](int k0, int q) { const int f = tid + 256 * q; return load4(We, k0 + f / 16, N, n0 + 4 * (f % 16), K, N, vB); };
  float acc[TM][TN] = {};
  float4 ra[NQ], rb[NQ];
#pragma unroll
  for (int q = 0; q < NQ; ++q) { ra[q] = fetchA(0, q); rb[q] = fetchB(0, q); }
  for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      const int f = tid + 256 * q;
      const int r = f / 8, c = 4 * (f % 8);
      As[c][r] = ra[q].x; As[c + 1][r] = ra[q].y; As[c + 2][r] = ra[q].z; As[c + 3][r] = ra[q].w;
      *reinterpret_cast<float4*>(&Bs[f / 16][4 * (f % 16)]) = rb[q];
    }
    __syncthreads();
    if (k0 + BK < K) {
#pragma unroll
      for (int q = 0; q < NQ; ++q) { ra[q] = fetchA(k0 + BK, q); rb[q] = fetchB(k0 + BK, q); }
    }
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * TM]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * TN]);
      const float a[TM] = {a4.x, a4.y, a4.z, a4.w}, b[TN] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int m = m0 + ty * TM + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int n = n0 + tx * TN + j;
      if (n >= N) continue;
      float v = acc[i][j] + bias[(int64_t)e * N + n];
      if (MODE == 0) {
        C[(int64_t)e * strideC + (int64_t)m * N + n] = swish_f(v);
      } else {
        if (n < D1) {
          mean[((int64_t)e * ld_out + m) * D1 + n] = v;
        } else {
          int d = n - D1;
          float mx = max_lv[d], mn = min_lv[d];
          float l = mx - softplus_f(mx - v);
          l = mn + softplus_f(l - mn);
          lv[((int64_t)e * ld_out + m) * D1 + d] = l;
        }
      }
    }
  }
}

}  // namespace

int fp32_forward(mopo_ens* h, const float* xn, int64_t n_max, const int* d_n, float* mean, float* lv,
                 int64_t ldb, cudaStream_t st, int64_t x_net_stride) {
  if (n_max == 0) return MOPO_OK;
  const int E = h->E, H = h->H;
  TRY(h->actA.reserve((size_t)E * n_max * H * sizeof(float)));
  TRY(h->actB.reserve((size_t)E * n_max * H * sizeof(float)));
  float* bufs[2] = {h->actA.as<float>(), h->actB.as<float>()};
  const int64_t strideAct = n_max * (int64_t)H;
  dim3 blk(256);
  const float* A = xn;
  int64_t strideA = x_net_stride;        // 0: every net reads the same rows (fc.py:98-99); else per-net inputs (fc.py:100-101)
  int lda = h->in, K = h->in;
  for (int l = 0; l < h->L; ++l) {
    dim3 grid((H + BN - 1) / BN, (unsigned)((n_max + BM - 1) / BM), E);
    float* out = bufs[l & 1];
    MOPO_COUNT_LAUNCH, ens_layer_fp32<0><<<grid, blk, 0, st>>>(A, strideA, lda, h->W[l], h->b[l], out, strideAct, nullptr, nullptr, 0,
                                            nullptr, nullptr, (int)n_max, d_n, H, K, 0);
    A = out; strideA = strideAct; lda = H; K = H;
  }
  dim3 grid((h->HO + BN - 1) / BN, (unsigned)((n_max + BM - 1) / BM), E);
  MOPO_COUNT_LAUNCH, ens_layer_fp32<1><<<grid, blk, 0, st>>>(A, strideA, lda, h->W[h->L], h->b[h->L], nullptr, 0, mean, lv, ldb,
                                          h->max_lv, h->min_lv, (int)n_max, d_n, h->HO, K, h->D1);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

// ---- mopo_ens_predict_noalloc: BNN.predict(factored=True) on this file's FFMA kernels with NO allocation ----------------------
// (No counterpart in the reference: added for callers that cannot call cudaMalloc --
// code under CUDA graph capture, tight serving loops, harnesses that forbid it.)  fp32_forward above keeps its scaled inputs and
// the [E, B, H] activation ping-pong in DevBufs that grow with cudaMalloc; this variant carves the same buffers out of ONE
// file-scope device array of MOPO_NOALLOC_SCRATCH_BYTES, which every GPU that loads the library holds until exit.  All calls, on
// every handle and stream, share the array and it keeps the last call's contents: calls that overlap on the GPU race (serialise
// them on one stream or with events).  Nothing in it has to start at zero -- every word a call reads (xn, the activations, the
// log-variances) is written earlier in the same call by its own kernels -- so nothing is cleared.
namespace {

constexpr size_t NOALLOC_FLOATS = MOPO_NOALLOC_SCRATCH_BYTES / sizeof(float);
static __device__ __align__(256) float g_noalloc_scratch[NOALLOC_FLOATS];

constexpr int NOALLOC_MAX_DEV = 64;
float* g_noalloc_ptr[NOALLOC_MAX_DEV] = {};     // device address of the array per device (the symbol itself is a host-side stand-in)
std::mutex g_noalloc_mu;

// once per device, with that device current
int noalloc_scratch(int device, float** out) {
  if (device < 0 || device >= NOALLOC_MAX_DEV) { mopo_set_error("mopo_ens_predict_noalloc: device %d out of range", device); return MOPO_EINVAL; }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    float* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    g_noalloc_ptr[device] = p;
  }
  *out = g_noalloc_ptr[device];
  return MOPO_OK;
}

inline size_t round64(size_t n) { return (n + 63) & ~(size_t)63; }      // 256-byte aligned sub-buffers

}  // namespace

MOPO_API int mopo_ens_predict_noalloc(mopo_ens* h, const float* x, int64_t B, float* mean, float* var, void* stream) {
  REQUIRE(h && (B == 0 || (x && mean && var)), MOPO_EINVAL, "mopo_ens_predict_noalloc: null argument");
  REQUIRE(B >= 0 && B < (1ll << 31), MOPO_EINVAL, "mopo_ens_predict_noalloc: B=%lld", (long long)B);
  REQUIRE(h->params_set, MOPO_ESTATE, "mopo_ens_predict_noalloc: parameters not set");
  if (B == 0) return MOPO_OK;
  const int E = h->E, H = h->H, D1 = h->D1;
  // scratch: xn [B, in] | act A [E, B, H] | act B [E, B, H] | logvar [E, B, D1]
  const size_t n_x = round64((size_t)B * h->in), n_act = round64((size_t)E * B * H), n_lv = round64((size_t)E * B * D1);
  const size_t need = n_x + (h->L > 0 ? (h->L > 1 ? 2 : 1) : 0) * n_act + n_lv;
  REQUIRE(need <= NOALLOC_FLOATS, MOPO_EINVAL,
          "mopo_ens_predict_noalloc: B=%lld rows need %zu bytes of scratch, the static array holds %zu (MOPO_NOALLOC_SCRATCH_BYTES); "
          "split the batch or use mopo_ens_predict", (long long)B, need * sizeof(float), (size_t)MOPO_NOALLOC_SCRATCH_BYTES);
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  float* base = nullptr;
  TRY(noalloc_scratch(h->cfg.device, &base));
  float* xn = base;
  float* bufs[2] = {base + n_x, base + n_x + n_act};
  float* lv = base + n_x + (h->L > 0 ? (h->L > 1 ? 2 : 1) : 0) * n_act;
  TRY(launch_prep(x, x, h->mu, h->sigma, xn, B, nullptr, h->in, 0, st));          // x is already concat(obs, act)
  const int64_t strideAct = B * (int64_t)H;
  dim3 blk(256);
  const float* A = xn;
  int64_t strideA = 0;
  int lda = h->in, K = h->in;
  for (int l = 0; l < h->L; ++l) {
    dim3 grid((H + BN - 1) / BN, (unsigned)((B + BM - 1) / BM), E);
    float* out = bufs[l & 1];
    MOPO_COUNT_LAUNCH, ens_layer_fp32<0><<<grid, blk, 0, st>>>(A, strideA, lda, h->W[l], h->b[l], out, strideAct, nullptr, nullptr, 0,
                                                            nullptr, nullptr, (int)B, nullptr, H, K, 0);
    A = out; strideA = strideAct; lda = H; K = H;
  }
  dim3 grid((h->HO + BN - 1) / BN, (unsigned)((B + BM - 1) / BM), E);
  MOPO_COUNT_LAUNCH, ens_layer_fp32<1><<<grid, blk, 0, st>>>(A, strideA, lda, h->W[h->L], h->b[h->L], nullptr, 0, mean, lv, B,
                                                          h->max_lv, h->min_lv, (int)B, nullptr, h->HO, K, D1);
  CUDA_TRY(cudaGetLastError());
  return launch_exp(lv, var, (int64_t)E * B * D1, st);
}
