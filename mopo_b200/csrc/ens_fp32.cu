// fp32 FFMA ensemble forward: the rtol-1e-3 parity path (MOPO_PREC_FP32).
//
// Computes BNN._compile_outputs (reference mopo/models/bnn.py:631-675) layer by layer:
//   hidden l : out[e,b,:] = swish(in[e,b,:] @ W_l[e] + b_l[e])          (fc.py:96-106, swish fc.py:21)
//   head     : [mean | raw] = h4[e,b,:] @ Whead[e] + bhead[e]
//              logvar = max_lv - softplus(max_lv - raw); logvar = min_lv + softplus(logvar - min_lv)
// Layer 0 reads the scaled input shared by all nets (the 2-D einsum path, fc.py:98-99).
// One launch per layer, activations [E,chunk,H] ping-pong through HBM: simple and exact;
// the tensor-core path in ens_tc.cu is the fast one.
#include "internal.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4;

__device__ __forceinline__ float swish_f(float x) { return x / (1.0f + expf(-x)); }
__device__ __forceinline__ float softplus_f(float x) { return fmaxf(x, 0.0f) + log1pf(expf(-fabsf(x))); }

// MODE 0: swish -> C[e, m, n] (ldc = N);  MODE 1: head -> mean/lv [e, m, D1]
template <int MODE>
__global__ void __launch_bounds__(256)
ens_layer_fp32(const float* __restrict__ A, int64_t strideA, int lda,
               const float* __restrict__ W, const float* __restrict__ bias,
               float* __restrict__ C, int64_t strideC,
               float* __restrict__ mean, float* __restrict__ lv, int64_t ld_out,
               const float* __restrict__ max_lv, const float* __restrict__ min_lv,
               int M_max, const int* __restrict__ d_n, int N, int K, int D1) {
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
  float acc[TM][TN] = {};
  for (int k0 = 0; k0 < K; k0 += BK) {
    // A tile: BM x BK (row-major, K contiguous) -> As[k][m]
    for (int i = tid; i < BM * BK; i += 256) {
      int m = i / BK, k = i % BK;
      float v = 0.f;
      if (m0 + m < M && k0 + k < K) v = Ae[(int64_t)(m0 + m) * lda + k0 + k];
      As[k][m] = v;
    }
    for (int i = tid; i < BK * BN; i += 256) {
      int k = i / BN, n = i % BN;
      float v = 0.f;
      if (k0 + k < K && n0 + n < N) v = We[(int64_t)(k0 + k) * N + n0 + n];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * TM]);     // one 16-byte shared load per operand
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
                 int64_t ldb, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  const int E = h->E, H = h->H;
  TRY(h->actA.reserve((size_t)E * n_max * H * sizeof(float)));
  TRY(h->actB.reserve((size_t)E * n_max * H * sizeof(float)));
  float* bufs[2] = {h->actA.as<float>(), h->actB.as<float>()};
  const int64_t strideAct = n_max * (int64_t)H;
  dim3 blk(256);
  const float* A = xn;
  int64_t strideA = 0;
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
