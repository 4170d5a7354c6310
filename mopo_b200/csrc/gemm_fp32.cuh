// fp32 FFMA batched GEMM + small reduction helpers shared by the training paths (train.cu: BNN.train, sac.cu: the SAC update).
#pragma once
#include "internal.cuh"
#include <math.h>

namespace {

constexpr int TM = 4, TN = 4;
constexpr float ADAM_B1 = 0.9f, ADAM_B2 = 0.999f, ADAM_EPS = 1e-8f;

__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float softplus_f(float x) { return fmaxf(x, 0.0f) + log1pf(expf(-fabsf(x))); }

// C[e] (M x N) = opA(A[e]) (M x K) * opB(B[e]) (K x N), fp32 FFMA, 64 x 64 x 32 or 32 x 32 x 32 tiles, 4 x 4 per thread.
//   tA: A stored [K, M] (lda = M's row length);  tB: B stored [N, K].
//   EPI 0: C = acc                                     (weight gradient)
//   EPI 1: z = acc + bias[e,n]; Z = z; C = z sigmoid(z) (hidden layer forward; Z kept for the backward pass)
//   EPI 2: C = acc + bias[e,n]                          (head forward)
//   EPI 3: C = acc * swish'(Z[e,m,n])                   (data gradient through the previous layer's activation)
//   EPI 4: C = max(acc + bias[e,n], 0)                  (ReLU layer forward)
//   EPI 5: C = acc * (Zin[e,m,n] > 0)                   (data gradient through a ReLU whose OUTPUT is Zin)
struct GemmP {
  const float* A; int64_t sA; int lda; int tA;
  const float* B; int64_t sB; int ldb; int tB;
  float* C; int64_t sC; int ldc;
  int M, N, K;
  const float* bias; float* Zout; const float* Zin;
};

// Four elements of a tile along the operand's contiguous dimension: one 16-byte load when the row pitch allows it.
__device__ __forceinline__ float4 load4(const float* __restrict__ base, int64_t row, int ld, int c, int row_lim, int c_lim, bool vec) {
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row >= row_lim) return v;
  const float* q = base + row * ld + c;
  if (vec && c + 3 < c_lim) return __ldg(reinterpret_cast<const float4*>(q));
  if (c < c_lim) v.x = __ldg(q);
  if (c + 1 < c_lim) v.y = __ldg(q + 1);
  if (c + 2 < c_lim) v.z = __ldg(q + 2);
  if (c + 3 < c_lim) v.w = __ldg(q + 3);
  return v;
}

// Tile TB x TB x BK with (TB / 4)^2 threads, 4 x 4 outputs per thread: 64 x 64 x 32 (256 threads) when the grid fills the GPU,
// 32 x 32 x 32 (64 threads, 4 x the CTAs) for the small grids of the latency-bound training steps.
template <int EPI, int TB>
__global__ void __launch_bounds__((TB / 4) * (TB / 4)) train_gemm_kernel(GemmP p) {
  constexpr int BK = 32;                           // k depth of a tile (64 was slower for both tile sizes: measured)
  constexpr int NT = (TB / 4) * (TB / 4);          // threads
  constexpr int TX = TB / 4;                       // threads along n (and along m)
  constexpr int NQ = TB * BK / 4 / NT;             // 16-byte loads per operand, tile and thread
  const int e = blockIdx.z;
  const int m0 = blockIdx.y * TB, n0 = blockIdx.x * TB;
  __shared__ __align__(16) float As[BK][TB + 4];
  __shared__ __align__(16) float Bs[BK][TB + 4];
  const float* Ae = p.A + (int64_t)e * p.sA;
  const float* Be = p.B + (int64_t)e * p.sB;
  const int tid = threadIdx.x, tx = tid % TX, ty = tid / TX;
  const bool vA = (p.lda % 4 == 0) && (p.sA % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0);
  const bool vB = (p.ldb % 4 == 0) && (p.sB % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.B) & 15) == 0);
  // load q of this thread covers 4 elements along the operand's contiguous dimension:
  //   k contiguous (A as stored [M, K], B stored [N, K]): (row, k4) = (f / (BK/4), 4 (f % (BK/4)))
  //   k is the row (A stored [K, M], B as stored [K, N]):  (k, c4)  = (f / (TB/4), 4 (f % (TB/4)))      f = tid + NT q
  auto fetch = [&](const float* base, int ld, bool t_is_kmajor, bool vec, int k0, int q, int o0, int o_lim) {
    const int f = tid + NT * q;
    if (t_is_kmajor) return load4(base, k0 + f / (TB / 4), ld, o0 + 4 * (f % (TB / 4)), p.K, o_lim, vec);
    return load4(base, o0 + f / (BK / 4), ld, k0 + 4 * (f % (BK / 4)), o_lim, p.K, vec);
  };
  auto stash = [&](float (*S)[TB + 4], bool t_is_kmajor, int q, const float4& v) {
    const int f = tid + NT * q;
    if (t_is_kmajor) *reinterpret_cast<float4*>(&S[f / (TB / 4)][4 * (f % (TB / 4))]) = v;
    else {
      const int r = f / (BK / 4), c = 4 * (f % (BK / 4));
      S[c][r] = v.x; S[c + 1][r] = v.y; S[c + 2][r] = v.z; S[c + 3][r] = v.w;
    }
  };
  const bool ka = p.tA != 0, kb = p.tB == 0;       // operand stored with k as the row index
  float acc[TM][TN] = {};
  float4 ra[NQ], rb[NQ];
#pragma unroll
  for (int q = 0; q < NQ; ++q) { ra[q] = fetch(Ae, p.lda, ka, vA, 0, q, m0, p.M); rb[q] = fetch(Be, p.ldb, kb, vB, 0, q, n0, p.N); }
  for (int k0 = 0; k0 < p.K; k0 += BK) {
#pragma unroll
    for (int q = 0; q < NQ; ++q) { stash(As, ka, q, ra[q]); stash(Bs, kb, q, rb[q]); }
    __syncthreads();
    if (k0 + BK < p.K) {                                                  // next tile in flight under this tile's FMAs
#pragma unroll
      for (int q = 0; q < NQ; ++q) { ra[q] = fetch(Ae, p.lda, ka, vA, k0 + BK, q, m0, p.M); rb[q] = fetch(Be, p.ldb, kb, vB, k0 + BK, q, n0, p.N); }
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
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int n = n0 + tx * TN + j;
      if (n >= p.N) continue;
      const int64_t o = (int64_t)e * p.sC + (int64_t)m * p.ldc + n;
      float v = acc[i][j];
      if (EPI == 1) {
        v += p.bias[(int64_t)e * p.N + n];
        p.Zout[o] = v;
        v = v * sigmoid_f(v);
      } else if (EPI == 2) {
        v += p.bias[(int64_t)e * p.N + n];
      } else if (EPI == 3) {
        float z = p.Zin[o], s = sigmoid_f(z);
        v *= s + z * s * (1.0f - s);
      } else if (EPI == 4) {
        v = fmaxf(v + p.bias[(int64_t)e * p.N + n], 0.0f);
      } else if (EPI == 5) {
        v = p.Zin[o] > 0.0f ? v : 0.0f;
      }
      p.C[o] = v;
    }
  }
}

template <int EPI>
int gemm(const GemmP& p, int E, cudaStream_t st) {
  MOPO_COUNT_LAUNCH;
  const int big = ((p.N + 63) / 64) * ((p.M + 63) / 64) * E;
  if (big >= 2 * 148) {
    dim3 grid((p.N + 63) / 64, (p.M + 63) / 64, E);
    train_gemm_kernel<EPI, 64><<<grid, 256, 0, st>>>(p);
  } else {
    dim3 grid((p.N + 31) / 32, (p.M + 31) / 32, E);
    train_gemm_kernel<EPI, 32><<<grid, 64, 0, st>>>(p);
  }
  return MOPO_OK;
}

// db[e, c] = sum_{b<n} dZ[e*stride + b*N + c]: 32 columns x 8 row groups per block, fixed summation order (deterministic)
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ dZ, float* __restrict__ db, int n, int N, int64_t stride) {
  __shared__ float sh[8][33];
  const int e = blockIdx.y, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (c < N) {
    const float* p = dZ + (int64_t)e * stride + c;
    for (int b = w; b < n; b += 8) s += p[(int64_t)b * N];
  }
  sh[w][lane] = s;
  __syncthreads();
  if (w == 0 && c < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sh[i][lane];
    db[(int64_t)e * N + c] = t;
  }
}

__device__ float block_sum(float v, float* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x < 32) {
    r = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  }
  __syncthreads();
  return r;   // valid in thread 0
}

int grid1d(int64_t tot) { int64_t g = (tot + 255) / 256; return (int)(g < 1 ? 1 : g > 148 * 8 ? 148 * 8 : g); }


}  // namespace
