// tcgen05 policy forward (MOPO_PREC_BF16): the SAC actor of mopo.py:292-312 on the 5th-gen tensor cores, one CTA per SM.
//   * the bf16 weights of all three GEMM stages (hidden 0, hidden 1, [mu | log_std] head) stay resident in shared
//     memory (172 KB for obs 17 / act 6), bias folded in through two constant-1 K columns (hi + lo bf16 split);
//   * a CTA walks 128-row tiles: the 4 epilogue warps tile their observation rows into a bf16 K-panel block (SS-form MMA,
//     M = 128, N = 256), ReLU the accumulator into packed bf16 columns of TENSOR MEMORY (the A operand of the next,
//     TS-form, MMA: the 128 x 256 activations never touch shared or global memory), and finish the row from the head
//     accumulator: clip, exp, Philox (or supplied) noise, tanh -> actions.
// Tensor-memory columns: [0,136) packed activations (K = 272 = 256 features + 2 bias ones + 14 zeros),
// [160,192) head accumulator, [256,512) hidden accumulator.
// Warp roles (192 threads): 0-3 epilogue (TMEM quadrant = warp), 4 MMA issuer (elect.sync), 5 weight loader.
#include "internal.cuh"
#include "tc_common.cuh"
#include "philox.cuh"

namespace {

using namespace tc;

constexpr int PH = 256;                    // hidden width
constexpr int PKH = PH + 16;               // hidden K incl. the bias columns
constexpr int PKP = PKH / 8;               // 34 K-panels
constexpr int P_THREADS = 192;
constexpr int C_A = 0, C_HD = 160, C_D = 256;

struct PShape {
  int obs, act, K0, NHD;
  __host__ __device__ int w0_bytes() const { return (K0 / 8) * PH * 16; }
  __host__ __device__ int w1_bytes() const { return PKP * PH * 16; }
  __host__ __device__ int wh_bytes() const { return PKP * NHD * 16; }
  __host__ __device__ int img_bytes() const { return w0_bytes() + w1_bytes() + wh_bytes(); }
  __host__ __device__ int x_bytes() const { return (K0 / 8) * 128 * 16; }
};
__host__ __device__ inline PShape make_pshape(int obs, int act) {
  PShape s;
  s.obs = obs; s.act = act;
  s.K0 = (obs + 2 + 15) / 16 * 16;
  s.NHD = (2 * act + 15) / 16 * 16;
  return s;
}

struct PImgSrc { const float* W[4]; const float* b[4]; };

__device__ __forceinline__ float pbias_part(float v, int part) {
  const float hi = __bfloat162float(__float2bfloat16_rn(v));
  return part == 0 ? hi : v - hi;
}

// image: W0 | W1 | Wh, each K-panel: element (row n, k) at (k/8) * rows * 16 + n * 16 + (k%8) * 2 bytes
__global__ void policy_image_kernel(PImgSrc src, __nv_bfloat16* __restrict__ img, PShape s) {
  const int total = s.img_bytes() / 2;
  const int e0 = s.w0_bytes() / 2, e1 = e0 + s.w1_bytes() / 2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int o = i, layer, rows, Kin;
    if (o < e0) { layer = 0; rows = PH; Kin = s.obs; }
    else if (o < e1) { o -= e0; layer = 1; rows = PH; Kin = PH; }
    else { o -= e1; layer = 2; rows = s.NHD; Kin = PH; }
    const int kc = o / (rows * 8), n = (o / 8) % rows, k = kc * 8 + o % 8;
    float v = 0.f;
    if (layer < 2) {
      if (k < Kin) v = src.W[layer][(int64_t)k * PH + n];
      else if (k < Kin + 2) v = pbias_part(src.b[layer][n], k - Kin);
    } else if (n < 2 * s.act) {                       // head rows interleaved: 2d = mu_d, 2d+1 = log_std_d
      const int l = 2 + (n & 1), c = n >> 1;
      if (k < Kin) v = src.W[l][(int64_t)k * s.act + c];
      else if (k < Kin + 2) v = pbias_part(src.b[l][c], k - Kin);
    }
    img[i] = __float2bfloat16_rn(v);
  }
}

struct PArgs {
  const uint8_t* img;
  PolicyRun r;
  PShape s;
};

struct PBars {
  uint64_t w_full;      // weights landed (tx)
  uint64_t d_full;      // accumulator of the current stage is complete (commit)
  uint64_t a_ready;     // the 4 epilogue warps have written the next A operand / consumed D (count 4)
  uint32_t tmem_base, pad;
};

__device__ __forceinline__ uint32_t relu2(uint32_t a, uint32_t b) {
  return pack_bf16x2(fmaxf(__uint_as_float(a), 0.f), fmaxf(__uint_as_float(b), 0.f));
}

// ReLU epilogue of one warp: its 32 rows x 256 accumulator columns -> packed bf16 A columns [0,128)
__device__ __forceinline__ void relu_stage(uint32_t tlane) {
  uint32_t va[32], vb[32], p[16];
  tmem_ld32(tlane + C_D, va);
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    tmem_ld_wait();
    tmem_ld32(tlane + C_D + 32 * (c + 1), vb);
#pragma unroll
    for (int j = 0; j < 16; ++j) p[j] = relu2(va[2 * j], va[2 * j + 1]);
    tmem_st16(tlane + C_A + 16 * c, p);
    tmem_ld_wait();
    if (c + 2 < 8) tmem_ld32(tlane + C_D + 32 * (c + 2), va);
#pragma unroll
    for (int j = 0; j < 16; ++j) p[j] = relu2(vb[2 * j], vb[2 * j + 1]);
    tmem_st16(tlane + C_A + 16 * (c + 1), p);
  }
  tmem_st_wait();
}

__global__ void __launch_bounds__(P_THREADS, 1) policy_tc_kernel(PArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const PShape s = a.s;
  uint8_t* sW0 = smem;
  uint8_t* sW1 = sW0 + s.w0_bytes();
  uint8_t* sWh = sW1 + s.w1_bytes();
  uint8_t* sX = sWh + s.wh_bytes();
  PBars* bars = reinterpret_cast<PBars*>(sX + s.x_bytes());
  float* sAct = reinterpret_cast<float*>(bars + 1);       // [128, 16] this tile's actions (for the ensemble tiles)
  float* sEmu = sAct + 128 * 16;                           // [80] ensemble scaler: mu, 1 / sigma
  float* sErs = sEmu + 80;
  const int tid = threadIdx.x, lane = tid % 32;
  const int warp = __shfl_sync(0xffffffffu, tid / 32, 0);
  const PolicyRun& r = a.r;

  // PDL (rollout chain): the forward kernel behind this one may be scheduled as CTAs exit; everything the previous grid
  // (the previous step's tail) wrote -- obs, the row count -- is read after pdl_wait(), so barrier setup, TMEM allocation
  // and the weight load overlap that tail.
  pdl_launch_dependents();
  if (tid == 0) {
    mbar_init(&bars->w_full, 1); mbar_init(&bars->d_full, 1); mbar_init(&bars->a_ready, 4);
    fence_barrier_init();
  }
  if (warp == 4) { tmem_alloc<1>(&bars->tmem_base, 512); tmem_relinquish<1>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = __shfl_sync(0xffffffffu, bars->tmem_base, 0);
  if (warp == 5 && lane == 0) {
    mbar_expect_tx(&bars->w_full, (uint32_t)s.img_bytes());
    uint32_t off = 0;
    const uint32_t sizes[3] = {(uint32_t)s.w0_bytes(), (uint32_t)s.w1_bytes(), (uint32_t)s.wh_bytes()};
    for (int i = 0; i < 3; ++i) { bulk_g2s(smem + off, a.img + off, sizes[i], &bars->w_full); off += sizes[i]; }
  }
  pdl_wait();                                              // (the weight copies above are already in flight)
  const int64_t n = r.d_n ? min((int64_t)*r.d_n, r.n_max) : r.n_max;
  const int nT = (int)((n + 127) / 128);
  if (a.r.xt) for (int i = tid; i < s.obs + s.act; i += P_THREADS) { sEmu[i] = a.r.ens_mu[i]; sErs[i] = 1.0f / a.r.ens_sigma[i]; }
  if (warp < 4) {
    // constant K columns 256..271 of the A operand: [1, 1, 0, ...] (the bias rows); written once
    const uint32_t ones[8] = {0x3F803F80u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    tmem_st8(tbase + ((uint32_t)(warp * 32) << 16) + C_A + 128, ones);
    tmem_st_wait();
    tc_fence_before();
  }
  __syncthreads();

  uint32_t ph = 0;        // phases consumed (issuer: a_ready, epilogue: d_full)
  if (warp == 4) {
    // ===== MMA issuer =====
    const uint32_t idh = idesc_bf16(128, PH), idn = idesc_bf16(128, s.NHD);
    const uint32_t w0a = smem_u32(sW0), w1a = smem_u32(sW1), wha = smem_u32(sWh), xa = smem_u32(sX);
    mbar_wait(&bars->w_full, 0);
    for (int t = blockIdx.x; t < nT; t += gridDim.x) {
      for (int j = 0; j < 3; ++j) {
        mbar_wait(&bars->a_ready, ph & 1); ++ph;
        tc_fence_after();
        if (elect_one()) {
          if (j == 0) {
            for (int ks = 0; ks < s.K0 / 16; ++ks)
              mma_ss<1>(tbase + C_D, smem_desc(xa + ks * 2 * 2048, 2048, 128), smem_desc(w0a + ks * 2 * (PH * 16), PH * 16, 128), idh, ks > 0);
          } else if (j == 1) {
#pragma unroll
            for (int ks = 0; ks < PKP / 2; ++ks)
              mma_ts<1>(tbase + C_D, tbase + C_A + 8 * ks, smem_desc(w1a + ks * 2 * (PH * 16), PH * 16, 128), idh, ks > 0);
          } else {
            const uint32_t hp = (uint32_t)s.NHD * 16;
#pragma unroll
            for (int ks = 0; ks < PKP / 2; ++ks)
              mma_ts<1>(tbase + C_HD, tbase + C_A + 8 * ks, smem_desc(wha + ks * 2 * hp, hp, 128), idn, ks > 0);
          }
          mma_commit<1>(&bars->d_full);
        }
        __syncwarp();
      }
    }
  } else if (warp < 4) {
    // ===== epilogue warps: row = tile * 128 + warp * 32 + lane =====
    const uint32_t tlane = tbase + ((uint32_t)(warp * 32) << 16);
    const int rr = warp * 32 + lane;
    const int od = s.obs, ad = s.act, kcs = s.K0 / 8;
    for (int t = blockIdx.x; t < nT; t += gridDim.x) {
      const int64_t row = (int64_t)t * 128 + rr;
      const bool ok = row < n;
      // observation row -> bf16 K-panel block (k = obs, obs+1 are the constant-1 bias columns)
      for (int kc = 0; kc < kcs; ++kc) {
        uint32_t p[4] = {0, 0, 0, 0};
        if (ok) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int k0 = kc * 8 + 2 * q, k1 = k0 + 1;
            const float v0 = k0 < od ? r.obs[row * od + k0] : (k0 < od + 2 ? 1.f : 0.f);
            const float v1 = k1 < od ? r.obs[row * od + k1] : (k1 < od + 2 ? 1.f : 0.f);
            p[q] = pack_bf16x2(v0, v1);
          }
        }
        st_shared_v4(sX + kc * 2048 + rr * 16, p[0], p[1], p[2], p[3]);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->a_ready);
      for (int j = 0; j < 2; ++j) {                      // hidden 0, hidden 1: ReLU -> A
        mbar_wait(&bars->d_full, ph & 1); ++ph;
        tc_fence_after();
        relu_stage(tlane);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->a_ready);
      }
      mbar_wait(&bars->d_full, ph & 1); ++ph;            // head
      tc_fence_after();
      uint32_t hv[32];
      if (s.NHD == 16) tmem_ld16(tlane + C_HD, reinterpret_cast<uint32_t (&)[16]>(hv));
      else tmem_ld32(tlane + C_HD, hv);
      tmem_ld_wait();
      if (ok) {
        const uint32_t orow = r.origin ? (uint32_t)r.origin[row] : (uint32_t)(r.origin_base + row);
        float z[4];
#pragma unroll
        for (int d = 0; d < 16; ++d) {
          if (d < ad) {
            const float mu = __uint_as_float(hv[2 * d]);
            const float ls = fminf(fmaxf(__uint_as_float(hv[2 * d + 1]), -20.0f), 2.0f);  // mopo.py:299
            const float sd = __expf(ls);                                                   // mopo.py:300
            float zz = 0.f;
            if (!r.deterministic) {
              if (r.noise) zz = r.noise[row * ad + d];
              else { if ((d & 3) == 0) philox_normal4(orow, (uint32_t)r.step_id, (uint32_t)(d >> 2), r.seed, z, MOPO_STREAM_POLICY); zz = z[d & 3]; }
            }
            const float av = tanhf(fmaf(zz, sd, mu));                                      // mopo.py:301, 288-289
            r.actions[row * ad + d] = av;
            sAct[rr * 16 + d] = av;
            if (r.mu_out) r.mu_out[row * ad + d] = mu;
            if (r.std_out) r.std_out[row * ad + d] = sd;
          }
        }
      }
      if (r.xt) {
        // the ensemble step's front end for this row (as prep_tiles_kernel): fake_env.py:46 concat, models/utils.py:96
        // scaler, bf16 K-panel tile (block t, row rr), bnn.py:342-344 elite pick
        const int in = od + ad, ekcs = r.ens_K0 / 8;
        for (int kc = 0; kc < ekcs; ++kc) {
          uint32_t p[4] = {0, 0, 0, 0};
          if (ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float v[2];
#pragma unroll
              for (int u = 0; u < 2; ++u) {
                const int k = kc * 8 + 2 * q + u;
                if (k < in) v[u] = ((k < od ? r.obs[row * od + k] : sAct[rr * 16 + (k - od)]) - sEmu[k]) * sErs[k];
                else v[u] = k < in + 2 ? 1.f : 0.f;
              }
              p[q] = pack_bf16x2(v[0], v[1]);
            }
          }
          r.xt[((int64_t)t * ekcs + kc) * 128 + rr] = make_uint4(p[0], p[1], p[2], p[3]);
          if (t == nT - 1 && (nT & 1)) r.xt[((int64_t)nT * ekcs + kc) * 128 + rr] = make_uint4(0, 0, 0, 0);   // whole 256-row tiles
        }
        if (ok && r.sel) {
          const uint32_t orow = r.origin ? (uint32_t)r.origin[row] : (uint32_t)(r.origin_base + row);
          r.sel[row] = r.model_inds ? r.model_inds[row]
                                    : r.elites[philox_elite(orow, (uint32_t)r.ens_step_id, r.seed, (uint32_t)r.n_elites)];
        }
      }
      // (the next tile's input block + arrive hands the accumulators back to the issuer)
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<1>(tbase, 512);
}

size_t policy_smem_bytes(const PShape& s) { return (size_t)s.img_bytes() + s.x_bytes() + sizeof(PBars) + sizeof(float) * (128 * 16 + 160) + 128; }

}  // namespace

int policy_tc_supported(const mopo_policy* p) {
  if (p->H != PH) return 0;
  if (p->obs_dim + 2 > 64 || 2 * p->act_dim > 32 || p->act_dim > 16) return 0;
  return policy_smem_bytes(make_pshape(p->obs_dim, p->act_dim)) <= 227 * 1024;
}

int policy_tc_build_image(mopo_policy* p, cudaStream_t st) {
  const PShape s = make_pshape(p->obs_dim, p->act_dim);
  if (p->tc_img_bytes < (size_t)s.img_bytes()) {
    if (p->tc_img) cudaFree(p->tc_img);
    p->tc_img = nullptr;
    CUDA_TRY(cudaMalloc(&p->tc_img, s.img_bytes()));
    p->tc_img_bytes = s.img_bytes();
  }
  PImgSrc src;
  for (int l = 0; l < 4; ++l) { src.W[l] = p->W[l]; src.b[l] = p->b[l]; }
  MOPO_COUNT_LAUNCH;
  policy_image_kernel<<<148, 256, 0, st>>>(src, reinterpret_cast<__nv_bfloat16*>(p->tc_img), s);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int policy_tc_launch(mopo_policy* p, const PolicyRun& r, cudaStream_t st) {
  if (!p->tc_img) { mopo_set_error("policy_tc_launch: weight image not built"); return MOPO_ESTATE; }
  PArgs a;
  a.img = reinterpret_cast<const uint8_t*>(p->tc_img);
  a.r = r;
  a.s = make_pshape(p->obs_dim, p->act_dim);
  const size_t smem = policy_smem_bytes(a.s);
  const int64_t tiles = (r.n_max + 127) / 128;
  const int grid = (int)(tiles < p->sm_count ? tiles : p->sm_count);
  CUDA_TRY(cudaFuncSetAttribute(policy_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(P_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = r.pdl ? 1 : 0;
  MOPO_COUNT_LAUNCH;
  CUDA_TRY(cudaLaunchKernelEx(&cfg, policy_tc_kernel, a));
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}
