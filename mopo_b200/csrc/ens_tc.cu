// tcgen05 ensemble forward (MOPO_PREC_BF16): BNN._compile_outputs (reference mopo/models/bnn.py:631-675,
// fc.py:96-106) for all 7 nets on the 5th-gen tensor cores.
//
// Design ("net-stationary pairs, two row tiles in flight"):
//   * A CTA PAIR (cluster of 2, tcgen05 cta_group::2, M = 256 rows per MMA) holds ONE net's bf16 weights
//     (4 hidden layers + fused [mean|logvar] head) resident in its shared memory; each CTA keeps the N-half of
//     every layer that cta_group::2 expects from it.  (net, 256-row tile) units are range-partitioned over the
//     74 pairs; a pair reloads weights only when its range crosses into the next net.
//   * TWO row tiles X and Y are in flight per pair, one layer apart in phase: while the 8 epilogue warps turn
//     tile X's accumulator into the next layer's A operand (swish), the tensor pipe computes tile Y's layer,
//     and vice versa.  The epilogue is MUFU-bound (one tanh.approx per activation, 16/clk/SM, measured), the MMA
//     of a layer is ~15 % shorter, so in steady state the special-function unit never waits.
//   * Tensor memory (512 columns) cannot hold two TS-form tiles (2 x (208 acc + 104 packed A) = 624), so the
//     tiles differ: X chains its activations through TENSOR MEMORY (TS-form MMA, A = packed bf16x2 columns),
//     Y chains them through SHARED MEMORY (SS-form MMA, K-panel bf16).  Column map (per CTA, 128 lanes):
//         [  0,104) X.A   packed activations of X (K = 208)
//         [104,216) X.D1  accumulator, output features   0..111
//         [216,312) X.D2  accumulator, output features 112..199 + 8 padding columns  [304,312)
//         [304,400) Y.D2  8 padding columns [304,312) + output features 112..199
//         [400,512) Y.D1  accumulator, output features   0..111
//     The N = 208 hidden width (200 features + 8 zero-weight padding rows) is issued as N = 112 + N = 96; the
//     padding columns of X.D2 and Y.D2 are the SAME eight columns: both tiles write zeros there and nobody reads
//     them, which is what makes 208 + 104 + 208 - 8 = 512 fit.  The weight image therefore orders a layer's
//     output rows as [f0..111 | pad | f112..199 | pad]: Y's second half starts at the first pad, X's 8 rows later.
//   * bias is folded into the GEMM through two constant-1 K columns (bias split hi + lo bf16), the 1/2 of
//     swish(x) = xh + xh * tanh(xh), xh = x/2, is folded into the weights.
//
// Warp roles (448 threads): warps 0-3 hidden-layer epilogue of output features 0..111, warps 4-7 of features
// 112..199, warp 8 of the leader CTA issues every MMA of the pair (elect.sync, warp-uniform control flow ->
// descriptors stay in uniform registers), warp 9 loads weights and input tiles with bulk async copies (the inputs
// arrive pre-tiled in bf16 K-panels from prep_tiles_kernel), warps 10-13 own the head stage (soft clamp, variance,
// penalty partials, elite gather, global stores): the 8 hidden-epilogue warps, whose MUFU time is the kernel's
// critical resource, go straight from a tile's last hidden layer to the next tile's first.
#include "internal.cuh"
#include "tc_common.cuh"
#include "tc_head.cuh"
#include "philox.cuh"

namespace {

using namespace tc;

constexpr int H = 200;
constexpr int KH = 208;                     // hidden K: 200 features + 2 bias-one columns + 6 zeros
constexpr int KPANELS = KH / 8;             // 26
constexpr int NH1 = 112, NH2 = 96;          // the two MMAs of a hidden layer
constexpr int WROWS = 112;                  // weight rows per CTA per hidden layer: 56 (half 1) + 56 (half 2 window)
constexpr int W_PANEL = WROWS * 16;         // 1792 B
constexpr int W_LAYER = KPANELS * W_PANEL;  // 46592 B
constexpr int A_PANEL = 128 * 16;           // 2048 B: one K-panel of a 128-row A tile
constexpr int AY_BYTES = KPANELS * A_PANEL; // 53248 B
constexpr int NTHREADS = 448;                // 8 hidden-epilogue warps, MMA issuer, loader, 4 head warps
// tensor-memory column map
constexpr int C_XA = 0, C_XD1 = 104, C_XD2 = 216, C_YD2 = 304, C_YD1 = 400;

struct TcShape {
  int in, D1, HO, K0, NHEAD;     // K0 = roundup(in + 2, 16), NHEAD = roundup(2*D1, 16)
  __host__ __device__ int w0_bytes() const { return (K0 / 8) * W_PANEL; }
  __host__ __device__ int head_bytes() const { return KPANELS * (NHEAD / 2) * 16; }
  __host__ __device__ int img_bytes() const { return w0_bytes() + 3 * W_LAYER + head_bytes(); }
  __host__ __device__ int x_bytes() const { return (K0 / 8) * A_PANEL; }      // one 128-row input block
};

__host__ __device__ inline TcShape make_shape(int in, int D1) {
  TcShape s;
  s.in = in; s.D1 = D1; s.HO = 2 * D1;
  s.K0 = (in + 2 + 15) / 16 * 16;
  s.NHEAD = (2 * D1 + 15) / 16 * 16;
  return s;
}

// ---------------------------------------------------------------------------------------------
// weight image: [net e][cta rank r] -> W0 | W1 | W2 | W3 | head, each in K-panel layout
//   element (row j, k) of a block at byte (k/8) * (rows*16) + j*16 + (k%8)*2
// hidden-layer block of CTA r, local row j:
//   j <  56 : output feature 56 r + j
//   j >= 56 : virtual row g = 112 + 48 r + (j - 56) of [f0..111 | pad 112..119 | f112..199 at 120..207 | pad 208..215]
// ---------------------------------------------------------------------------------------------
struct ImgSrc {
  const float* W[5];
  const float* b[5];
};

__device__ __forceinline__ float bias_part(float v, int part) {
  __nv_bfloat16 hi = __float2bfloat16_rn(v);
  return part == 0 ? __bfloat162float(hi) : v - __bfloat162float(hi);
}

__global__ void build_image_kernel(ImgSrc src, __nv_bfloat16* __restrict__ img, TcShape s, int E) {
  const int per = s.img_bytes() / 2;
  const int64_t total = (int64_t)E * 2 * per;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i / (2 * per));
    const int r = (int)((i / per) % 2);
    int o = (int)(i % per);                       // element offset inside this CTA's image
    int layer, rows;
    const int w0 = s.w0_bytes() / 2, wl = W_LAYER / 2;
    if (o < w0) { layer = 0; rows = WROWS; }
    else if (o < w0 + 3 * wl) { o -= w0; layer = 1 + o / wl; o %= wl; rows = WROWS; }
    else { o -= w0 + 3 * wl; layer = 4; rows = s.NHEAD / 2; }
    const int kc = o / (rows * 8), j = (o / 8) % rows, kk = o % 8;
    const int k = kc * 8 + kk;
    const int Kin = layer == 0 ? s.in : H;
    const int Nout = layer == 4 ? s.HO : H;
    int n;                                        // output feature of this row, or -1 for a padding row
    if (layer == 4) {
      // head rows of CTA r = epilogue warpgroup r's block: [mean(d) x P | raw logvar(d) x P], d = r*dh + i, P = NHEAD/4
      const int P = s.NHEAD / 4, dh = (s.D1 + 1) / 2;
      const int part = j / P, ii = j % P, d = r * dh + ii;
      n = d < (r ? s.D1 : dh) ? part * s.D1 + d : -1;
    }
    else if (j < 56) n = 56 * r + j;
    else {
      const int g = 112 + 48 * r + (j - 56);
      n = (g >= 120 && g < 208) ? g - 8 : -1;
    }
    const float scale = layer == 4 ? 1.0f : 0.5f; // hidden layers produce xh = x/2 for the tanh form of swish
    float v = 0.f;
    if (n >= 0 && n < Nout) {
      if (k < Kin) v = scale * src.W[layer][((int64_t)e * Kin + k) * Nout + n];
      else if (k < Kin + 2) v = bias_part(scale * src.b[layer][(int64_t)e * Nout + n], k - Kin);
    }
    img[i] = __float2bfloat16_rn(v);
  }
}

// Head column order (see build_image_kernel): warpgroup w of the epilogue owns output dims d = w*dh + i, i < P,
// dh = ceil(D1/2), P = NHEAD/4: its accumulator block is columns [w*NHEAD/2, (w+1)*NHEAD/2) = [mean x P | raw x P].
// ---------------------------------------------------------------------------------------------
// input tiles: fp32 rows [n, in] -> bf16 K-panel blocks of 128 rows: xt[block][kc][r][8], with the two
// constant-1 bias columns at k = in, in+1; rows >= n are zero.  One thread per (row, kc) 16-byte slot.
// ---------------------------------------------------------------------------------------------
__global__ void panelize_kernel(const float* __restrict__ xn, uint4* __restrict__ xt, int n_max, const int* __restrict__ d_n,
                                int in, int K0) {
  const int n = d_n ? min(*d_n, n_max) : n_max;
  const int kcs = K0 / 8;
  const int64_t slots = (int64_t)((n + 255) / 256) * 256 * kcs;     // whole 256-row tiles (both CTAs' blocks)
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < slots; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i % 128);
    const int kc = (int)((i / 128) % kcs);
    const int64_t blk = i / (128 * kcs);
    const int64_t row = blk * 128 + r;
    uint32_t p[4] = {0, 0, 0, 0};
    if (row < n) {
      const float* src = xn + row * in;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k0 = kc * 8 + 2 * j, k1 = k0 + 1;
        const float v0 = k0 < in ? src[k0] : (k0 < in + 2 ? 1.f : 0.f);
        const float v1 = k1 < in ? src[k1] : (k1 < in + 2 ? 1.f : 0.f);
        p[j] = pack_bf16x2(v0, v1);
      }
    }
    xt[i] = make_uint4(p[0], p[1], p[2], p[3]);     // i == (blk * kcs + kc) * 128 + r
  }
}

// FakeEnv.step front end for the tensor-core path: concat(obs, act) (fake_env.py:46), scaler (models/utils.py:96),
// bf16 K-panel tiling, and the per-row elite pick (bnn.py:342-344: supplied indices, or Philox over the elite list)
// in one pass.  sel == nullptr skips the pick (deterministic / full-output steps).
struct PrepArgs {
  const float *obs, *act, *mu, *sigma;
  uint4* xt;
  int n_max; const int* d_n;
  int obs_dim, act_dim, K0;
  int32_t* sel; const int32_t* model_inds; const int32_t* elites; int n_elites;
  uint64_t seed, step_id; const int32_t* origin; int64_t origin_base;
};

// one CTA per 128-row block; rows go through shared memory so that global reads and the 16-byte tile writes coalesce
__global__ void __launch_bounds__(128) prep_tiles_kernel(PrepArgs a) {
  extern __shared__ float sIn[];                    // [128, in] scaled inputs
  const int n = a.d_n ? min(*a.d_n, a.n_max) : a.n_max;
  const int od = a.obs_dim, ad = a.act_dim, in = od + ad, kcs = a.K0 / 8;
  const int n_blocks = ((n + 255) / 256) * 2;
  const int tid = threadIdx.x;
  float* sMu = sIn + 128 * in;            // [in]
  float* sSg = sMu + in;                  // [in]
  for (int i = tid; i < in; i += 128) { sMu[i] = a.mu[i]; sSg[i] = 1.0f / a.sigma[i]; }   // reciprocal once; the step tail tiles with the same arithmetic
  for (int blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
    const int64_t j0 = (int64_t)blk * 128;
    const int rows = (int)max((int64_t)0, min((int64_t)128, n - j0));
    __syncthreads();
    // raw rows -> smem (flat contiguous spans, loads batched in flight); the scaler is applied on the way out
    float* sObs = sIn;                      // [128, od]
    float* sAct = sIn + 128 * od;           // [128, ad]
    {
      const float* go = a.obs + j0 * od;
      const float* ga = a.act + j0 * ad;
#pragma unroll 8
      for (int i = tid; i < rows * od; i += 128) sObs[i] = go[i];
#pragma unroll 4
      for (int i = tid; i < rows * ad; i += 128) sAct[i] = ga[i];
    }
    __syncthreads();
    const bool ok = tid < rows;
    for (int kc = 0; kc < kcs; ++kc) {
      uint32_t p[4] = {0, 0, 0, 0};
      if (ok) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int k0 = kc * 8 + 2 * q, k1 = k0 + 1;
          float v[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int k = u ? k1 : k0;
            if (k < in) {
              const float x = k < od ? sObs[tid * od + k] : sAct[tid * ad + (k - od)];
              v[u] = (x - sMu[k]) * sSg[k];                                          // models/utils.py:96
            } else v[u] = k < in + 2 ? 1.f : 0.f;
          }
          p[q] = pack_bf16x2(v[0], v[1]);
        }
      }
      a.xt[((int64_t)blk * kcs + kc) * 128 + tid] = make_uint4(p[0], p[1], p[2], p[3]);
    }
    if (ok && a.sel) {
      const int64_t row = j0 + tid;
      const uint32_t orow = a.origin ? (uint32_t)a.origin[row] : (uint32_t)(a.origin_base + row);
      a.sel[row] = a.model_inds ? a.model_inds[row]
                                : a.elites[philox_elite(orow, (uint32_t)a.step_id, a.seed, (uint32_t)a.n_elites)];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// forward kernel
// ---------------------------------------------------------------------------------------------
struct TcArgs {
  const uint8_t* img;      // weight images [E][2][img_bytes]
  const uint8_t* xt;       // input tiles (panelize_kernel)
  const float* max_lv;     // [D1]
  const float* min_lv;     // [D1]
  // FULL output (sel == nullptr): everything BNN.predict returns
  float* mean;             // [E, ldb, D1]
  float* lv;               // [E, ldb, D1]
  int64_t ldb;
  // FUSED output (sel != nullptr): what FakeEnv.step keeps of the [E,B,D1] arrays (fake_env.py:78-81,110)
  const int32_t* sel;      // [n] net chosen for each row (the elite pick)
  float2* selmv;           // [D1, ldb] (mean, variance) of the chosen net, dimension-major so that stores coalesce
  float* pen2;             // [2, E, ldb] sum_d var_d over each warpgroup's half of the output dims
  int n_max;
  const int* d_n;
  int E;
  TcShape s;
  unsigned long long* trace;   // debug timeline (mopo_debug_tc_trace) or nullptr
};

struct Bars {
  uint64_t w_full;         // weights of the current net are in smem (tx)
  uint64_t x_ld[2];        // this CTA's input slot t has landed (tx)
  uint64_t x_full[2];      // leader: both CTAs' input slot t are in smem                  (count 2)
  uint64_t x_free[2];      // each CTA: the L0 MMAs that read slot t have retired           (commit)
  uint64_t d_full[2];      // each CTA: the accumulator of tile t's current stage is ready  (commit)
  uint64_t a_ready[2];     // leader: tile t's hidden epilogue stage is done (A operand written / D consumed) (count 16)
  uint64_t h_full[2];      // each CTA: tile t's head accumulator is ready                  (commit)
  uint64_t h_done[2];      // leader: the head warps have read tile t's head accumulator    (count 8)
  uint32_t tmem_base;
  uint32_t pad;
};


// swish of two accumulator columns (xh = x/2 already) -> packed bf16x2
__device__ __forceinline__ uint32_t swish2(uint32_t a, uint32_t b) {
  const float x = __uint_as_float(a), y = __uint_as_float(b);
  return pack_bf16x2(fmaf(x, tanh_fast(x), x), fmaf(y, tanh_fast(y), y));
}

template <int N> __device__ __forceinline__ void tmem_ld_n(uint32_t taddr, uint32_t (&v)[32]) {
  if constexpr (N == 32) tmem_ld32(taddr, v);
  else if constexpr (N == 16) tmem_ld16(taddr, reinterpret_cast<uint32_t (&)[16]>(v));
  else tmem_ld8(taddr, reinterpret_cast<uint32_t (&)[8]>(v));
}

// N accumulator columns (output features feat..feat+N-1 of this thread's row) -> swish -> next layer's A operand:
// tile X: packed columns C_XA + feat/2 of tensor memory; tile Y: K-panels feat/8.. of the smem A tile (this row's 16 B)
template <bool IS_X, int N>
__device__ __forceinline__ void swish_store(const uint32_t (&v)[32], int feat, uint32_t tlane, uint8_t* ay_row) {
  uint32_t p[N / 2];
#pragma unroll
  for (int j = 0; j < N / 2; ++j) p[j] = swish2(v[2 * j], v[2 * j + 1]);
  if constexpr (IS_X) {
    if constexpr (N == 32) tmem_st16(tlane + C_XA + feat / 2, reinterpret_cast<const uint32_t (&)[16]>(p));
    else if constexpr (N == 16) tmem_st8(tlane + C_XA + feat / 2, reinterpret_cast<const uint32_t (&)[8]>(p));
    else tmem_st4(tlane + C_XA + feat / 2, reinterpret_cast<const uint32_t (&)[4]>(p));
  } else {
#pragma unroll
    for (int q = 0; q < N / 8; ++q)
      st_shared_v4(ay_row + (feat / 8 + q) * A_PANEL, p[4 * q], p[4 * q + 1], p[4 * q + 2], p[4 * q + 3]);
  }
}

// one hidden-layer epilogue of one warp: its 32 rows x (112 or 88) accumulator columns, software pipelined
// (the tcgen05.ld of the next chunk is in flight while the current chunk goes through the MUFU)
template <bool IS_X, int WG>
__device__ __forceinline__ void epi_hidden(uint32_t tlane, uint8_t* ay_row) {
  constexpr int src = WG == 0 ? (IS_X ? C_XD1 : C_YD1) : (IS_X ? C_XD2 : C_YD2 + 8);
  constexpr int f0 = WG == 0 ? 0 : 112;
  uint32_t va[32], vb[32];
  tmem_ld_n<32>(tlane + src, va);
  tmem_ld_wait();
  tmem_ld_n<32>(tlane + src + 32, vb);
  swish_store<IS_X, 32>(va, f0, tlane, ay_row);
  tmem_ld_wait();
  if constexpr (WG == 0) {
    tmem_ld_n<32>(tlane + src + 64, va);
    swish_store<IS_X, 32>(vb, f0 + 32, tlane, ay_row);
    tmem_ld_wait();
    tmem_ld_n<16>(tlane + src + 96, vb);
    swish_store<IS_X, 32>(va, f0 + 64, tlane, ay_row);
    tmem_ld_wait();
    swish_store<IS_X, 16>(vb, f0 + 96, tlane, ay_row);
  } else {
    tmem_ld_n<16>(tlane + src + 64, va);
    swish_store<IS_X, 32>(vb, f0 + 32, tlane, ay_row);
    tmem_ld_wait();
    tmem_ld_n<8>(tlane + src + 80, vb);
    swish_store<IS_X, 16>(va, f0 + 64, tlane, ay_row);
    tmem_ld_wait();
    swish_store<IS_X, 8>(vb, f0 + 80, tlane, ay_row);
  }
  if constexpr (IS_X) tmem_st_wait();
  else fence_proxy_async_smem();
}


constexpr int TRACE_CAP = 2048;   // events per role
// timeline probe: role r (0 issuer, 1 epilogue WG0 warp 0, 2 epilogue WG1 warp 4, 3 loader, 4 head warp 12) of pair 0 / leader CTA
#define TR(role, tag)                                                                                   \
  do {                                                                                                  \
    if (TRACE && a.trace && pair == 0 && rank == 0 && lane == 0 && tr_n < TRACE_CAP)                    \
      a.trace[(role) * TRACE_CAP + tr_n++] = ((unsigned long long)(tag) << 48) | ((unsigned long long)clock64() & 0xFFFFFFFFFFFFull); \
  } while (0)

template <bool TRACE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1) ens_tc_kernel(TcArgs a) {
  int tr_n = 0;
  (void)tr_n;
  extern __shared__ __align__(128) uint8_t smem[];
  const TcShape s = a.s;
  uint8_t* sW0 = smem;
  uint8_t* sW = sW0 + s.w0_bytes();                 // 3 hidden layers
  uint8_t* sWh = sW + 3 * W_LAYER;
  uint8_t* sAY = sWh + s.head_bytes();              // tile Y's activations (K-panel bf16)
  uint8_t* sX = sAY + AY_BYTES;                     // input slots of tiles X, Y
  float* sLv = reinterpret_cast<float*>(sX + 2 * s.x_bytes());     // max_lv[D1], min_lv[D1]
  Bars* bars = reinterpret_cast<Bars*>(reinterpret_cast<uint8_t*>(sLv) + 2 * MOPO_MAX_D1 * sizeof(float));

  const int tid = threadIdx.x, lane = tid % 32;
  const int warp = __shfl_sync(0xffffffffu, tid / 32, 0);    // warp-uniform for the compiler (uniform registers)
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x / 2, n_pairs = gridDim.x / 2;
  const int n = __shfl_sync(0xffffffffu, a.d_n ? min(*a.d_n, a.n_max) : a.n_max, 0);
  const int nT = (n + 255) / 256;                            // 256-row tiles
  const int64_t U = (int64_t)a.E * nT;                       // (net, tile) units, net-major
  const int u0 = (int)(U * pair / n_pairs), u1 = (int)(U * (pair + 1) / n_pairs);

#ifdef MOPO_TC_WATCHDOG
  if (tid == 0 && blockIdx.x == 0) printf("watchdog: Bars at smem %u (w_full +0, x_ld +8, x_full +24, x_free +40, d_full +56, a_ready +72, h_full +88, h_done +104)\n", smem_u32(bars));
#endif
  if (tid == 0) {
    mbar_init(&bars->w_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->x_ld[i], 1); mbar_init(&bars->x_full[i], 2); mbar_init(&bars->x_free[i], 1);
      mbar_init(&bars->d_full[i], 1); mbar_init(&bars->a_ready[i], 16);
      mbar_init(&bars->h_full[i], 1); mbar_init(&bars->h_done[i], 8);
    }
    fence_barrier_init();
  }
  for (int i = tid; i < 2 * (s.NHEAD / 4); i += NTHREADS) {          // head constants per warpgroup block (see epi_head)
    const int P = s.NHEAD / 4, dh = (s.D1 + 1) / 2;
    const int w = i / P, d = w * dh + i % P;
    const bool ok = d < (w ? s.D1 : dh);
    sLv[i] = ok ? expf(a.min_lv[d]) : 0.0f;
    sLv[MOPO_MAX_D1 + i] = ok ? fminf(expf(-a.max_lv[d]), 1e30f) : 1e30f;
  }
  if (warp == 8) { tmem_alloc<2>(&bars->tmem_base, 512); tmem_relinquish<2>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = __shfl_sync(0xffffffffu, bars->tmem_base, 0);
  if (warp < 4) {
    // constant K columns 200..207 of both tiles' A operands: [1, 1, 0, 0, 0, 0, 0, 0] (the bias rows); written once
    const uint32_t ones[4] = {0x3F803F80u, 0u, 0u, 0u};
    tmem_st4(tbase + ((uint32_t)(warp * 32) << 16) + C_XA + 100, ones);
    tmem_st_wait();
    st_shared_v4(sAY + 25 * A_PANEL + tid * 16, 0x3F803F80u, 0u, 0u, 0u);
    fence_proxy_async_smem();
    tc_fence_before();
  }
  __syncthreads();

  // PDL: the step tail that consumes this launch's outputs may be scheduled onto SMs as their CTAs exit (its input
  // copies and Philox draws do not depend on this kernel); everything this kernel reads from or writes to global memory
  // that the PREVIOUS grid (the previous step's tail: input tiles, elite picks, selmv / pen2 readers) touches comes after
  // pdl_wait() in the loader and head warps, so barrier setup, TMEM allocation and the first weight load overlap it.
  pdl_launch_dependents();
  bool pdl_waited = false;

  // phase bookkeeping that survives net switches
  // (use counts packed 16 bits per tile slot: a per-slot array indexed by t would live in local memory)
  uint32_t n_stage = 0;              // issuer: a_ready phases consumed / hidden epilogue: d_full phases / head warps: h_full phases
  uint32_t n_x = 0;                  // issuer / loader: input-slot uses
  uint32_t n_tile = 0;               // issuer: tiles whose head MMA has been issued
#define CNT(v, t) (((v) >> (16 * (t))) & 0xFFFFu)
#define INC(v, t) ((v) += 1u << (16 * (t)))
  uint32_t seg = 0;

  for (int u = u0; u < u1; ++seg) {
    const int e = u / nT;
    const int t_lo = u - e * nT;
    const int u_end = min(u1, (e + 1) * nT);
    const int cnt = u_end - u;                      // tiles of net e handled by this pair
    const int n_super = (cnt + 1) / 2;
    u = u_end;

    // ---- weights of net e -> smem (all MMAs of the previous net have retired: the epilogue saw their commits) ----
    cluster_sync_all();
    if (warp == 9 && lane == 0) {
      const uint8_t* src = a.img + ((size_t)e * 2 + rank) * s.img_bytes();
      mbar_expect_tx(&bars->w_full, (uint32_t)s.img_bytes());
      uint32_t off = 0;
      const uint32_t sizes[5] = {(uint32_t)s.w0_bytes(), W_LAYER, W_LAYER, W_LAYER, (uint32_t)s.head_bytes()};
      for (int i = 0; i < 5; ++i) { bulk_g2s(smem + off, src + off, sizes[i], &bars->w_full); off += sizes[i]; }
    }
    if (warp >= 9 && !pdl_waited) { pdl_wait(); pdl_waited = true; }   // weights are in flight; inputs / outputs come next
    mbar_wait(&bars->w_full, seg & 1);
    cluster_sync_all();                              // both CTAs' weights are in place before any MMA reads them

    if (warp == 9) {
      // ===== input loader: one bulk copy per 128-row block (bf16 K-panels from panelize_kernel) =====
      if (lane == 0) {
        for (int k = 0; k < n_super; ++k) {
          for (int t = 0; t < 2; ++t) {
            if (2 * k + t >= cnt) continue;
            if (CNT(n_x, t) > 0) mbar_wait(&bars->x_free[t], (CNT(n_x, t) - 1) & 1);
            TR(3, 70 + t);
            const int64_t blk = 2 * (int64_t)(t_lo + 2 * k + t) + rank;
            mbar_expect_tx(&bars->x_ld[t], (uint32_t)s.x_bytes());
            bulk_g2s(sX + t * s.x_bytes(), a.xt + blk * s.x_bytes(), (uint32_t)s.x_bytes(), &bars->x_ld[t]);
            mbar_wait(&bars->x_ld[t], CNT(n_x, t) & 1);
            mbar_arrive_remote(&bars->x_full[t], 0);
            TR(3, 72 + t);
            INC(n_x, t);
          }
        }
      }
      __syncwarp();
    } else if (warp == 8) {
      // ===== MMA issuer (leader CTA): all lanes run the loop, one elected lane issues =====
      if (rank == 0) {
        const uint32_t id1 = idesc_bf16(256, NH1), id2 = idesc_bf16(256, NH2), idh = idesc_bf16(256, s.NHEAD);
        const uint32_t w0a = smem_u32(sW0), wa = smem_u32(sW), wha = smem_u32(sWh), aya = smem_u32(sAY), xa = smem_u32(sX);
        const uint32_t hpanel = (uint32_t)(s.NHEAD / 2) * 16;
        // Issue order per super-tile k: L1..L3 of X, Y alternating, then head(X_k), L0(X_k+1), head(Y_k), L0(Y_k+1): a slot's
        // next tile starts its first layer as soon as the head warps have drained that slot's head accumulator, without
        // queueing behind the other slot's last hidden epilogue.  (Per slot the order is unchanged: L0..L3, head, L0' ...)
        for (int k = 0; k < n_super; ++k) {
          for (int step = (k == 0 ? 0 : 2); step < 12; ++step) {
            {
              const int t = step < 8 ? (step & 1) : (step >= 10);
              const int j = step < 2 ? 0 : step < 8 ? (step >> 1) : (step == 8 || step == 10) ? 4 : 0;
              const int kk = (step == 9 || step == 11) ? k + 1 : k;
              if (2 * kk + t >= cnt) continue;
              if (j == 0) {
                // L0 overwrites the columns the previous tile's head accumulator lives in: the head warps must have read it
                mbar_wait_cluster(&bars->x_full[t], CNT(n_x, t) & 1, 100 * k + step); INC(n_x, t);
                if (CNT(n_tile, t) > 0) mbar_wait_cluster(&bars->h_done[t], (CNT(n_tile, t) - 1) & 1, 100 * k + step);
              } else {
                mbar_wait_cluster(&bars->a_ready[t], CNT(n_stage, t) & 1, 100 * k + step); INC(n_stage, t);
              }
              tc_fence_after();
              TR(0, 10 * j + t);
              const uint32_t d1 = tbase + (t == 0 ? C_XD1 : C_YD1), d2 = tbase + (t == 0 ? C_XD2 : C_YD2);
              const uint32_t h2off = 56 * 16 + (t == 0 ? 8 * 16 : 0);       // half-2 window: X starts 8 rows later
              if (elect_one()) {
                if (j == 0) {
                  for (int ks = 0; ks < s.K0 / 16; ++ks) {
                    const uint64_t ad = smem_desc(xa + t * s.x_bytes() + ks * 2 * A_PANEL, A_PANEL, 128);
                    const uint32_t b = w0a + ks * 2 * W_PANEL;
                    mma_ss<2>(d1, ad, smem_desc(b, W_PANEL, 128), id1, ks > 0);
                    mma_ss<2>(d2, ad, smem_desc(b + h2off, W_PANEL, 128), id2, ks > 0);
                  }
                  mma_commit<2>(&bars->x_free[t]);
                } else if (j < 4) {
                  const uint32_t wl = wa + (j - 1) * W_LAYER;
                  if (t == 0) {
#pragma unroll
                    for (int ks = 0; ks < KPANELS / 2; ++ks) {
                      const uint32_t b = wl + ks * 2 * W_PANEL;
                      mma_ts<2>(d1, tbase + C_XA + 8 * ks, smem_desc(b, W_PANEL, 128), id1, ks > 0);
                      mma_ts<2>(d2, tbase + C_XA + 8 * ks, smem_desc(b + h2off, W_PANEL, 128), id2, ks > 0);
                    }
                  } else {
#pragma unroll
                    for (int ks = 0; ks < KPANELS / 2; ++ks) {
                      const uint32_t b = wl + ks * 2 * W_PANEL;
                      const uint64_t ad = smem_desc(aya + ks * 2 * A_PANEL, A_PANEL, 128);
                      mma_ss<2>(d1, ad, smem_desc(b, W_PANEL, 128), id1, ks > 0);
                      mma_ss<2>(d2, ad, smem_desc(b + h2off, W_PANEL, 128), id2, ks > 0);
                    }
                  }
                } else {
                  if (t == 0) {
#pragma unroll
                    for (int ks = 0; ks < KPANELS / 2; ++ks)
                      mma_ts<2>(d1, tbase + C_XA + 8 * ks, smem_desc(wha + ks * 2 * hpanel, hpanel, 128), idh, ks > 0);
                  } else {
#pragma unroll
                    for (int ks = 0; ks < KPANELS / 2; ++ks)
                      mma_ss<2>(d1, smem_desc(aya + ks * 2 * A_PANEL, A_PANEL, 128), smem_desc(wha + ks * 2 * hpanel, hpanel, 128), idh, ks > 0);
                  }
                }
                mma_commit<2>(j < 4 ? &bars->d_full[t] : &bars->h_full[t]);
              }
              __syncwarp();
              if (j == 4) INC(n_tile, t);
            }
          }
        }
      }
    } else if (warp < 8) {
      // ===== hidden-layer epilogue warps: 0-3 -> features 0..111, 4-7 -> features 112..199 =====
      const int wg = warp >> 2, q = warp & 3;
      const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
      const int rr = q * 32 + lane;                     // row inside this CTA's 128-row block
      uint8_t* ay_row = sAY + rr * 16;
      for (int k = 0; k < n_super; ++k) {
        for (int j = 0; j < 4; ++j) {
          for (int t = 0; t < 2; ++t) {
            if (2 * k + t >= cnt) continue;
            if (q == 0) TR(1 + wg, 100 + 10 * j + t);
            mbar_wait(&bars->d_full[t], CNT(n_stage, t) & 1, 100 * k + 10 * j + t);
            INC(n_stage, t);
            tc_fence_after();
            if (q == 0) TR(1 + wg, 200 + 10 * j + t);
            if (t == 0) { if (wg == 0) epi_hidden<true, 0>(tlane, ay_row); else epi_hidden<true, 1>(tlane, ay_row); }
            else        { if (wg == 0) epi_hidden<false, 0>(tlane, ay_row); else epi_hidden<false, 1>(tlane, ay_row); }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_remote(&bars->a_ready[t], 0);
            if (q == 0) TR(1 + wg, 300 + 10 * j + t);
          }
        }
      }
    } else if (warp >= 10) {
      // ===== head warps (10-13, tensor-memory quadrant warp % 4): soft clamp, variance, penalty partials, elite gather.
      // Off the hidden epilogue's critical path: they own a tile's head accumulator from h_full to h_done. =====
      const int q = warp & 3;
      const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
      const int rr = q * 32 + lane;
      const HeadShape hs = {s.D1, s.NHEAD};
      const HeadOut ho = {1.0f, a.mean, a.lv, a.ldb, a.sel, a.selmv, a.pen2, a.E};
      for (int k = 0; k < n_super; ++k) {
        for (int t = 0; t < 2; ++t) {
          if (2 * k + t >= cnt) continue;
          const int64_t row = 256 * (int64_t)(t_lo + 2 * k + t) + (int64_t)rank * 128 + rr;
          const bool valid = row < n;
          bool mine = false;                           // fused head: is net e the one this row keeps?  (in flight during the wait)
          if (a.sel && valid) mine = a.sel[row] == e;
          if (q == 0) TR(4, 140 + t);
          mbar_wait(&bars->h_full[t], CNT(n_stage, t) & 1, 100 * k + t);
          INC(n_stage, t);
          tc_fence_after();
          if (q == 0) TR(4, 240 + t);
          const uint32_t tblk = tlane + (t == 0 ? C_XD1 : C_YD1);
          bool arrived = false;
          if (a.sel) {
            if (s.D1 == 18) { head_fused<48, 18>(tblk, hs, sLv, ho, e, row, valid, mine, &bars->h_done[t], lane); arrived = true; }        // obs 17 (+1)
            else if (s.D1 == 12) { head_fused<32, 12>(tblk, hs, sLv, ho, e, row, valid, mine, &bars->h_done[t], lane); arrived = true; }   // obs 11 (+1)
            else if (s.NHEAD == 48) { head_fused<48, 0>(tblk, hs, sLv, ho, e, row, valid, mine, &bars->h_done[t], lane); arrived = true; }
            else if (s.NHEAD == 32) { head_fused<32, 0>(tblk, hs, sLv, ho, e, row, valid, mine, &bars->h_done[t], lane); arrived = true; }
            else { epi_head<true>(tblk, 0, hs, sLv, ho, e, row, valid, mine); epi_head<true>(tblk + s.NHEAD / 2, 1, hs, sLv, ho, e, row, valid, mine); }
          } else { epi_head<false>(tblk, 0, hs, sLv, ho, e, row, valid, false); epi_head<false>(tblk + s.NHEAD / 2, 1, hs, sLv, ho, e, row, valid, false); }
          if (!arrived) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_remote(&bars->h_done[t], 0);
          }
          if (q == 0) TR(4, 340 + t);
        }
      }
    }
  }
  if (warp >= 9 && !pdl_waited) pdl_wait();           // every CTA waits before it exits (CTAs without work too)
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                 // peer smem / TMEM must outlive the leader's MMAs
  if (warp == 8) tmem_dealloc<2>(tbase, 512);
}

size_t tc_smem_bytes(const TcShape& s) {
  return (size_t)s.img_bytes() + AY_BYTES + 2 * s.x_bytes() + 2 * MOPO_MAX_D1 * sizeof(float) + sizeof(Bars) + 128;
}

}  // namespace

int tc_supported(const mopo_ens* h) {
  if (h->H != H || h->L != 4) return 0;
  if (h->in + 2 > 64 || h->HO > 112) return 0;
  if (h->E > h->sm_count / 2) return 0;
  return tc_smem_bytes(make_shape(h->in, h->D1)) <= 227 * 1024;
}

int tc_build_images(mopo_ens* h, cudaStream_t st) {
  const TcShape s = make_shape(h->in, h->D1);
  const size_t bytes = (size_t)h->E * 2 * s.img_bytes();
  if (h->tc_img_bytes < bytes) {
    if (h->tc_img) cudaFree(h->tc_img);
    h->tc_img = nullptr;
    CUDA_TRY(cudaMalloc(&h->tc_img, bytes));
    h->tc_img_bytes = bytes;
  }
  ImgSrc src;
  for (int l = 0; l < 5; ++l) { src.W[l] = h->W[l]; src.b[l] = h->b[l]; }
  MOPO_COUNT_LAUNCH, build_image_kernel<<<296, 256, 0, st>>>(src, reinterpret_cast<__nv_bfloat16*>(h->tc_img), s, h->E);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

static int tc_reserve_tiles(mopo_ens* h, int64_t n_max, const TcShape& s, int* grid) {
  const int64_t tiles = (n_max + 255) / 256;
  TRY(h->xtiles.reserve((size_t)tiles * 2 * s.x_bytes()));
  const int64_t blocks = (tiles * 256 * (s.K0 / 8) + 255) / 256;
  *grid = (int)(blocks > 148 * 8 ? 148 * 8 : blocks);
  return MOPO_OK;
}

// for producers of the input tiles other than the kernels of this file (the policy kernel, the step tail)
int tc_reserve_xtiles(mopo_ens* h, int64_t n_max, int* K0) {
  const TcShape s = make_shape(h->in, h->D1);
  int grid;
  *K0 = s.K0;
  return tc_reserve_tiles(h, n_max, s, &grid);
}

int tc_tiles_from_xn(mopo_ens* h, const float* xn, int64_t n_max, const int* d_n, cudaStream_t st) {
  const TcShape s = make_shape(h->in, h->D1);
  int grid;
  TRY(tc_reserve_tiles(h, n_max, s, &grid));
  MOPO_COUNT_LAUNCH, panelize_kernel<<<grid, 256, 0, st>>>(xn, h->xtiles.as<uint4>(), (int)n_max, d_n, h->in, s.K0);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int tc_tiles_from_step(mopo_ens* h, const TcPrep& q, cudaStream_t st) {
  const TcShape s = make_shape(h->in, h->D1);
  int grid;
  TRY(tc_reserve_tiles(h, q.n_max, s, &grid));
  PrepArgs a;
  a.obs = q.obs; a.act = q.act; a.mu = h->mu; a.sigma = h->sigma; a.xt = h->xtiles.as<uint4>();
  a.n_max = (int)q.n_max; a.d_n = q.d_n; a.obs_dim = h->cfg.obs_dim; a.act_dim = h->cfg.act_dim; a.K0 = s.K0;
  a.sel = q.sel; a.model_inds = q.model_inds; a.elites = h->elites; a.n_elites = h->n_elites;
  a.seed = q.seed; a.step_id = q.step_id; a.origin = q.origin; a.origin_base = q.origin_base;
  const int64_t blocks = ((q.n_max + 255) / 256) * 2;
  (void)grid;
  MOPO_COUNT_LAUNCH, prep_tiles_kernel<<<(int)(blocks > 148 * 16 ? 148 * 16 : blocks), 128, sizeof(float) * (128 + 2) * h->in, st>>>(a);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int tc_launch(mopo_ens* h, int64_t n_max, const int* d_n, const TcOut& o, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  if (!h->tc_img) { mopo_set_error("tc_launch: weight images not built"); return MOPO_ESTATE; }
  TcArgs a;
  a.s = make_shape(h->in, h->D1);
  a.img = reinterpret_cast<const uint8_t*>(h->tc_img);
  a.xt = h->xtiles.as<uint8_t>();
  a.max_lv = h->max_lv; a.min_lv = h->min_lv;
  a.mean = o.mean; a.lv = o.lv; a.ldb = o.ldb;
  a.sel = o.sel; a.selmv = o.selmv; a.pen2 = o.pen2;
  a.n_max = (int)n_max; a.d_n = d_n; a.E = h->E;
  a.trace = reinterpret_cast<unsigned long long*>(h->tc_trace);
  const size_t smem = tc_smem_bytes(a.s);
  const int n_pairs = h->sm_count / 2;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  if (h->timing) {
    if (h->ev_used + 2 > h->ev.size()) {
      for (int i = 0; i < 2; ++i) { cudaEvent_t e; CUDA_TRY(cudaEventCreate(&e)); h->ev.push_back(e); }
    }
    ev0 = h->ev[h->ev_used]; ev1 = h->ev[h->ev_used + 1]; h->ev_used += 2;
    CUDA_TRY(cudaEventRecord(ev0, st));
  }
#ifdef MOPO_DEBUG_PROBES
  if (a.trace) {                                        // timeline variant: only in the probe build (mopo_debug_tc_trace)
    CUDA_TRY(cudaFuncSetAttribute(ens_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    MOPO_COUNT_LAUNCH, ens_tc_kernel<true><<<2 * n_pairs, NTHREADS, smem, st>>>(a);
  } else
#endif
  {
    CUDA_TRY(cudaFuncSetAttribute(ens_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    if (o.pdl) {
      // programmatic dependent launch behind the previous step's tail (static-row rollouts): see pdl_wait() in the kernel
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(2 * n_pairs); cfg.blockDim = dim3(NTHREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      MOPO_COUNT_LAUNCH;
      CUDA_TRY(cudaLaunchKernelEx(&cfg, ens_tc_kernel<false>, a));
    } else {
      MOPO_COUNT_LAUNCH, ens_tc_kernel<false><<<2 * n_pairs, NTHREADS, smem, st>>>(a);
    }
  }
  if (ev1) CUDA_TRY(cudaEventRecord(ev1, st));
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int tc_forward(mopo_ens* h, const float* xn, int64_t n_max, const int* d_n, float* mean, float* lv,
               int64_t ldb, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  TRY(tc_tiles_from_xn(h, xn, n_max, d_n, st));
  TcOut o = {};
  o.mean = mean; o.lv = lv; o.ldb = ldb;
  return tc_launch(h, n_max, d_n, o, st);
}
