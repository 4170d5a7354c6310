// tcgen05 ensemble forward (MOPO_PREC_BF16): BNN._compile_outputs (reference mopo/models/bnn.py:631-675,
// fc.py:96-106) for all 7 nets on the 5th-gen tensor cores.
//
// Design ("net-stationary"): a CTA PAIR (cluster of 2, tcgen05 cta_group::2, M = 256 rows per MMA) owns ONE
// net.  Its bf16 weights (4 hidden layers + fused [mean|logvar] head, 286 KB) stay resident in the pair's
// shared memory for the whole launch -- each CTA holds the N-half of every layer that cta_group::2 expects
// from it (143 KB) -- and row tiles of 256 scaled inputs stream through:
//
//   L0 : D0 = x0 (smem, K-panel bf16)        * W0^T     SS-form MMA
//   L1..L3 : D = A_l (TMEM, packed bf16x2)   * W_l^T    TS-form MMA, A never leaves tensor memory
//   head   : [mean | raw logvar] = A_4 * Wh^T           TS-form MMA
//
// Accumulators live in TMEM (two 208-column regions, ping-pong by layer).  The epilogue warps read a region
// with tcgen05.ld, apply bias (folded into the GEMM through two constant-1 K columns, bias split hi+lo),
// swish  h = xh + xh*tanh(xh), xh = x/2 (the 1/2 is folded into the weights), and write the next layer's
// A operand IN PLACE over the accumulator columns they have just consumed (tcgen05.st).  Each layer's MMA is
// issued as two N-halves (112 + 96 columns) with separate commits, so epilogue warpgroup 1 converts the
// first half while the tensor pipe is still producing the second, and the next layer's K-steps start as
// soon as the 16-column groups they need have been written (mbarriers grp[0..3]).
//
// Warp roles (320 threads): warps 0-3 epilogue of N-half 1 (+ head), warps 4-7 epilogue of N-half 2,
// warp 8 lane 0 of the leader CTA issues every MMA of the pair, warp 9 loads weights (cp.async.bulk) and
// the input tiles (fp32 -> bf16 K-panels, double buffered).
#include "internal.cuh"
#include "tc_common.cuh"

namespace {

using namespace tc;

constexpr int HP = 208;          // hidden width padded to a multiple of 16 (200 features + 2 bias-one columns)
constexpr int H = 200;
constexpr int NH1 = 112, NH2 = 96;                 // N-halves of a hidden layer (each a legal cta_group::2 N)
constexpr int ROWS1 = NH1 / 2, ROWS2 = NH2 / 2;    // B rows this CTA provides for each half (56, 48)
constexpr int WROWS = ROWS1 + ROWS2;               // 104 weight rows per CTA per hidden layer
constexpr int KPANELS = HP / 8;                    // 26 K-panels
constexpr int W_PANEL = WROWS * 16;                // 1664 B
constexpr int W_LAYER = KPANELS * W_PANEL;         // 43264 B
constexpr int X_PANEL = 128 * 16;                  // 2048 B
constexpr int NTHREADS = 320;
constexpr int REGION_COLS = 256;                   // TMEM columns between the two accumulator regions

struct TcShape {
  int in, D1, HO, K0, NHEAD;     // K0 = roundup(in + 2, 16), NHEAD = roundup(2*D1, 16)
  __host__ __device__ int w0_bytes() const { return (K0 / 8) * W_PANEL; }
  __host__ __device__ int head_bytes() const { return KPANELS * (NHEAD / 2) * 16; }
  __host__ __device__ int img_bytes() const { return w0_bytes() + 3 * W_LAYER + head_bytes(); }
  __host__ __device__ int x_bytes() const { return (K0 / 8) * X_PANEL; }
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
    int n;
    if (layer == 4) n = r * (s.NHEAD / 2) + j;
    else n = j < ROWS1 ? r * ROWS1 + j : NH1 + r * ROWS2 + (j - ROWS1);
    const float scale = layer == 4 ? 1.0f : 0.5f; // hidden layers produce xh = x/2 for the tanh form of swish
    float v = 0.f;
    if (n < Nout) {
      if (k < Kin) v = scale * src.W[layer][((int64_t)e * Kin + k) * Nout + n];
      else if (k < Kin + 2) v = bias_part(scale * src.b[layer][(int64_t)e * Nout + n], k - Kin);
    }
    img[i] = __float2bfloat16_rn(v);
  }
}

// ---------------------------------------------------------------------------------------------
// forward kernel
// ---------------------------------------------------------------------------------------------
struct TcArgs {
  const uint8_t* img;      // weight images
  const float* xn;         // [n, in] scaled inputs
  const float* max_lv;     // [D1]
  const float* min_lv;     // [D1]
  float* mean;             // [E, ldb, D1]
  float* lv;               // [E, ldb, D1]
  int64_t ldb;
  int n_max;
  const int* d_n;
  int E;
  TcShape s;
  unsigned long long* trace;   // debug timeline (mopo_debug_tc_trace) or nullptr
};

struct Bars {
  uint64_t w_full;
  uint64_t x_full[2];      // leader: both CTAs' input tiles are in smem          (count 2)
  uint64_t x_free[2];      // each CTA: L0 MMAs that read buffer s have retired    (count 1, commit)
  uint64_t grp[4];         // leader: 16-col groups of the next A operand written  (count 8 = 4 warps x 2 CTAs)
  uint64_t d_full[2];      // each CTA: N-half h of the current layer is in TMEM   (count 1, commit)
  uint64_t d_head;         // each CTA: head accumulator is in TMEM               (count 1, commit)
  uint32_t tmem_base;
  uint32_t pad;
};

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float softplus_f(float x) { return fmaxf(x, 0.0f) + log1pf(expf(-fabsf(x))); }

// 16 accumulator columns (xh, fp32) -> 8 packed bf16x2 of swish(x) = xh + xh * tanh(xh)
__device__ __forceinline__ void swish_pack16(const uint32_t (&v)[16], uint32_t (&p)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float a = __uint_as_float(v[2 * j]), b = __uint_as_float(v[2 * j + 1]);
    p[j] = pack_bf16x2(fmaf(a, tanh_fast(a), a), fmaf(b, tanh_fast(b), b));
  }
}

// TMEM column (inside a region) of K-chunk c of the in-place A operand
__device__ __forceinline__ int a_col(int c) { return c < 7 ? 8 * c : NH1 + 8 * (c - 7); }

constexpr int TRACE_CAP = 2048;   // events per role
// timeline probe: role r (0 issuer, 1 epilogue WG0, 2 epilogue WG1, 3 loader) of pair 0 / leader CTA appends (tag, clock)
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
  uint8_t* sX = sWh + s.head_bytes();               // 2 input buffers
  float* sLv = reinterpret_cast<float*>(sX + 2 * s.x_bytes());     // max_lv[D1], min_lv[D1]
  Bars* bars = reinterpret_cast<Bars*>(reinterpret_cast<uint8_t*>(sLv) + 2 * MOPO_MAX_D1 * sizeof(float));

  const int tid = threadIdx.x, lane = tid % 32;
  const int warp = __shfl_sync(0xffffffffu, tid / 32, 0);    // warp-uniform for the compiler (uniform registers)
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x / 2, n_pairs = gridDim.x / 2;
  const int n = __shfl_sync(0xffffffffu, a.d_n ? min(*a.d_n, a.n_max) : a.n_max, 0);
  const int n_tiles = (n + 255) / 256;

  // pair -> (net, slot among the pairs of that net)
  const int base = n_pairs / a.E, rem = n_pairs % a.E;
  int e, slot, cnt;
  if (pair < rem * (base + 1)) { e = pair / (base + 1); slot = pair % (base + 1); cnt = base + 1; }
  else { const int q = pair - rem * (base + 1); e = rem + q / base; slot = q % base; cnt = base; }
  const int my_tiles = slot < n_tiles ? (n_tiles - slot + cnt - 1) / cnt : 0;

  if (tid == 0) {
    mbar_init(&bars->w_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->x_full[i], 2); mbar_init(&bars->x_free[i], 1); mbar_init(&bars->d_full[i], 1); }
    for (int i = 0; i < 4; ++i) mbar_init(&bars->grp[i], 8);
    mbar_init(&bars->d_head, 1);
    fence_barrier_init();
  }
  for (int i = tid; i < s.D1; i += NTHREADS) { sLv[i] = a.max_lv[i]; sLv[MOPO_MAX_D1 + i] = a.min_lv[i]; }
  if (warp == 8) { tmem_alloc<2>(&bars->tmem_base, 512); tmem_relinquish<2>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 9 && lane == 0) {                      // resident weights: one bulk copy per layer block
    const uint8_t* src = a.img + ((size_t)e * 2 + rank) * s.img_bytes();
    mbar_expect_tx(&bars->w_full, (uint32_t)s.img_bytes());
    uint32_t off = 0;
    const uint32_t sizes[5] = {(uint32_t)s.w0_bytes(), W_LAYER, W_LAYER, W_LAYER, (uint32_t)s.head_bytes()};
    for (int i = 0; i < 5; ++i) { bulk_g2s(smem + off, src + off, sizes[i], &bars->w_full); off += sizes[i]; }
  }
  if (tid == 0) mbar_wait(&bars->w_full, 0);
  __syncthreads();
  cluster_sync_all();                                // both CTAs' weights and barriers are ready
  const uint32_t tbase = __shfl_sync(0xffffffffu, bars->tmem_base, 0);

  if (warp == 9) {
    // ===== input loader: fp32 [rows, in] -> bf16 K-panels, + two constant-1 columns for the bias rows =====
    for (int it = 0; it < my_tiles; ++it) {
      const int sbuf = it & 1;
      if (it >= 2) mbar_wait(&bars->x_free[sbuf], ((it >> 1) - 1) & 1);
      const int tile = slot + it * cnt;
      TR(3, 70);
      uint8_t* dst = sX + sbuf * s.x_bytes();
      for (int rr = lane; rr < 128; rr += 32) {
        const int row = tile * 256 + (int)rank * 128 + rr;
        const bool ok = row < n;
        const float* src = a.xn + (size_t)row * s.in;
        for (int kc = 0; kc < s.K0 / 8; ++kc) {
          uint32_t p[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int k0 = kc * 8 + 2 * j, k1 = k0 + 1;
            const float v0 = !ok ? 0.f : (k0 < s.in ? src[k0] : (k0 < s.in + 2 ? 1.f : 0.f));
            const float v1 = !ok ? 0.f : (k1 < s.in ? src[k1] : (k1 < s.in + 2 ? 1.f : 0.f));
            p[j] = pack_bf16x2(v0, v1);
          }
          *reinterpret_cast<uint4*>(dst + kc * X_PANEL + rr * 16) = make_uint4(p[0], p[1], p[2], p[3]);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&bars->x_full[sbuf], 0);
      TR(3, 71);
    }
  } else if (warp == 8) {
    // ===== MMA issuer: one thread of the leader CTA drives both SMs' tensor cores =====
    if (rank == 0) {                                  // all 32 lanes run the loop; elect_one() issues
      const uint32_t id1 = idesc_bf16(256, NH1), id2 = idesc_bf16(256, NH2), idh = idesc_bf16(256, s.NHEAD);
      uint32_t g = 0;                                 // global layer counter -> accumulator region
      uint32_t grp_par = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int sbuf = it & 1;
        TR(0, 1);
        mbar_wait_cluster(&bars->x_full[sbuf], (it >> 1) & 1);
        tc_fence_after();
        TR(0, 2);
        // ---- L0 (SS) ----
        {
          const uint32_t dcol = tbase + (g & 1) * REGION_COLS;
          const uint32_t xa = smem_u32(sX + sbuf * s.x_bytes());
          for (int half = 0; half < 2; ++half) {
            for (int ks = 0; ks < s.K0 / 16; ++ks) {
              const uint64_t ad = smem_desc(xa + ks * 2 * X_PANEL, X_PANEL, 128);
              const uint64_t bd = smem_desc(smem_u32(sW0) + ks * 2 * W_PANEL + (half ? ROWS1 * 16 : 0), W_PANEL, 128);
              if (elect_one()) mma_ss<2>(dcol + (half ? NH1 : 0), ad, bd, half ? id2 : id1, ks > 0);
            }
            if (elect_one()) mma_commit<2>(&bars->d_full[half]);
            TR(0, 30 + half);
          }
          if (elect_one()) mma_commit<2>(&bars->x_free[sbuf]);
          ++g;
        }
        // ---- L1..L3 and head (TS): A = in-place bf16 activations in the previous region ----
        for (int l = 1; l <= 4; ++l) {
          const uint32_t acol = tbase + ((g - 1) & 1) * REGION_COLS;
          const uint32_t dcol = tbase + (g & 1) * REGION_COLS;
          const bool head = l == 4;
          const uint8_t* wl = head ? sWh : sW + (l - 1) * W_LAYER;
          const int panel = head ? (s.NHEAD / 2) * 16 : W_PANEL;
          for (int half = 0; half < (head ? 1 : 2); ++half) {
            for (int ks = 0; ks < KPANELS / 2; ++ks) {
              if (half == 0 && (ks == 0 || ks == 4 || ks == 7 || ks == 10)) {
                const int gi = ks == 0 ? 0 : ks == 4 ? 1 : ks == 7 ? 2 : 3;
                mbar_wait_cluster(&bars->grp[gi], grp_par);
                tc_fence_after();
                TR(0, 100 + l * 10 + gi);
              }
              const uint64_t bd = smem_desc(smem_u32(wl) + ks * 2 * panel + ((!head && half) ? ROWS1 * 16 : 0), panel, 128);
              if (elect_one()) mma_ts<2>(dcol + (half ? NH1 : 0), acol + a_col(ks), bd, head ? idh : (half ? id2 : id1), ks > 0);
            }
            if (elect_one()) mma_commit<2>(head ? &bars->d_head : &bars->d_full[half]);
            TR(0, 30 + l * 2 + half);
          }
          grp_par ^= 1;
          ++g;
        }
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue warpgroups: warps 0-3 -> N-half 1 (+ head), warps 4-7 -> N-half 2 =====
    const int wg = warp / 4;                          // 0 or 1
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const int rr = (warp & 3) * 32 + lane;            // row inside this CTA's 128-row slice
    uint32_t g = 0, dpar = 0, hpar = 0;
    const int c_lo = wg == 0 ? 0 : 7, c_hi = wg == 0 ? 7 : 13;
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = slot + it * cnt;
      for (int l = 0; l < 4; ++l) {
        TR(1 + wg, 40 + l);
        mbar_wait(&bars->d_full[wg], dpar);
        dpar ^= 1;
        tc_fence_after();
        TR(1 + wg, 50 + l);
        const uint32_t reg = tbase + lane_off + (g & 1) * REGION_COLS;
        for (int c = c_lo; c < c_hi; ++c) {
          uint32_t v[16], p[8];
          tmem_ld16(reg + 16 * c, v);
          tmem_ld_wait();
          swish_pack16(v, p);
          if (c == 12) p[4] = 0x3F803F80u;            // columns 200/201 := 1.0 -> next layer's bias rows (hi + lo)
          tmem_st8(reg + a_col(c), p);
          const bool last = (c == 3 || c == 6 || c == 9 || c == 12);
          if (last) {
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&bars->grp[c == 3 ? 0 : c == 6 ? 1 : c == 9 ? 2 : 3], 0);
          }
        }
        ++g;
        TR(1 + wg, 60 + l);
      }
      if (wg == 0) {
        // ---- head epilogue: mean = cols [0,D1), logvar = clamp(cols [D1,2*D1))  (bnn.py:654-655 / 669-670) ----
        TR(1, 44);
        mbar_wait(&bars->d_head, hpar);
        hpar ^= 1;
        tc_fence_after();
        TR(1, 54);
        const uint32_t reg = tbase + lane_off + (g & 1) * REGION_COLS;
        const int row = tile * 256 + (int)rank * 128 + rr;
        float* mrow = a.mean + ((size_t)e * a.ldb + row) * s.D1;
        float* lrow = a.lv + ((size_t)e * a.ldb + row) * s.D1;
        for (int c = 0; c < s.NHEAD / 16; ++c) {
          uint32_t v[16];
          tmem_ld16(reg + 16 * c, v);
          tmem_ld_wait();
          if (row < n) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int col = 16 * c + j;
              const float x = __uint_as_float(v[j]);
              if (col < s.D1) mrow[col] = x;
              else if (col < s.HO) {
                const int d = col - s.D1;
                const float mx = sLv[d], mn = sLv[MOPO_MAX_D1 + d];
                float lg = mx - softplus_f(mx - x);
                lg = mn + softplus_f(lg - mn);
                lrow[d] = lg;
              }
            }
          }
        }
        tc_fence_before();
        TR(1, 64);
      }
      ++g;
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                 // peer smem / TMEM must outlive the leader's MMAs
  if (warp == 8) tmem_dealloc<2>(tbase, 512);
}

size_t tc_smem_bytes(const TcShape& s) {
  return (size_t)s.img_bytes() + 2 * s.x_bytes() + 2 * MOPO_MAX_D1 * sizeof(float) + sizeof(Bars) + 128;
}

}  // namespace

int tc_supported(const mopo_ens* h) {
  if (h->H != H || h->L != 4) return 0;
  if (h->in + 2 > 64 || h->HO > 96) return 0;
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

int tc_forward(mopo_ens* h, const float* xn, int64_t n_max, const int* d_n, float* mean, float* lv,
               int64_t ldb, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  if (!h->tc_img) { mopo_set_error("tc_forward: weight images not built"); return MOPO_ESTATE; }
  TcArgs a;
  a.img = reinterpret_cast<const uint8_t*>(h->tc_img);
  a.xn = xn; a.max_lv = h->max_lv; a.min_lv = h->min_lv; a.mean = mean; a.lv = lv; a.ldb = ldb;
  a.n_max = (int)n_max; a.d_n = d_n; a.E = h->E; a.s = make_shape(h->in, h->D1);
  a.trace = reinterpret_cast<unsigned long long*>(h->tc_trace);
  const size_t smem = tc_smem_bytes(a.s);
  const int n_pairs = h->sm_count / 2;
  if (a.trace) {
    CUDA_TRY(cudaFuncSetAttribute(ens_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    MOPO_COUNT_LAUNCH, ens_tc_kernel<true><<<2 * n_pairs, NTHREADS, smem, st>>>(a);
  } else {
    CUDA_TRY(cudaFuncSetAttribute(ens_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    MOPO_COUNT_LAUNCH, ens_tc_kernel<false><<<2 * n_pairs, NTHREADS, smem, st>>>(a);
  }
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}
