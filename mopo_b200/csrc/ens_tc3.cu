// tcgen05 ensemble forward at fp32 accuracy (MOPO_PREC_FP32): BNN._compile_outputs (reference mopo/models/bnn.py:631-675,
// fc.py:96-106) for all nets on the 5th-gen tensor cores, with every operand carried as an fp16 (hi, lo) pair.
//
// Arithmetic.  The reference contracts in fp32.  kind::f16 MMAs take 11-bit significands (fp16), so each fp32 operand
// x is split x = hi + lo, hi = f16(x), lo = f16(x - hi) (22 significant bits) and a layer is THREE MMAs per K step into
// the same fp32 accumulator in tensor memory:  Ah*Wh + Ah*Wl + Al*Wh  (the dropped Al*Wl term is 2^-22 relative).
// Weights and biases are stored x 64 (a power of two: exact) so that the lo parts of small weights stay normal fp16
// numbers; the epilogue multiplies the accumulator by 1/64.  Activations saturate at +-65504 (cvt.satfinite) instead
// of overflowing.  Measured against the float64 oracle the outputs agree to ~1e-6 absolute, as the FFMA checker path
// (ens_fp32.cu) does.  swish is x * rcp(1 + ex2(-x log2 e)) (two MUFU ops, ~3 ulp), not tanh.approx (2^-11).
// TERMS = 1 runs the same kernel as a plain bf16 forward (one MMA per K step, tanh-form swish) for shapes the
// net-stationary bf16 kernel (ens_tc.cu) cannot hold resident (ant / humanoid: mopo/static/ant.py, humanoid.py).
//
// Design ("one tile per CTA pair, activations never leave tensor memory, weights streamed"):
//   * A CTA PAIR (cluster of 2, cta_group::2, M = 256 rows per MMA) takes (net, 256-row tile) units, range-partitioned
//     over the 74 pairs.  hi + lo weights of one net are 2 x 2 x 157 KB: they do not fit two SMs, so a unit STREAMS its
//     net's five layers through a two-stage shared-memory ring (one layer per stage, 86.5 KB per CTA, bulk async copies
//     from the L2-resident image, refilled as soon as tcgen05.commit reports the layer's MMAs retired).
//   * Tensor memory holds the activations for the whole unit: two 208-column buffers ping-pong.  Layer j reads its A
//     operand (TS-form MMA) from one buffer and accumulates into the other; the epilogue then converts the accumulator
//     IN PLACE: the 16 fp32 columns of feature group g become [8 columns packed f16x2 hi | 8 columns packed lo] of the
//     same features, which is exactly the TS-form A operand of the next layer's K step g.  Column map (per CTA, 128 lanes):
//         [  0,208) buffer 0 : D of layers 0, 2   ->  A of layers 1, 3
//         [208,416) buffer 1 : D of layers 1, 3   ->  A of layer 2, head
//         [416,416+NHEAD)    : head accumulator [mean | raw logvar] (owned by the head warps from h_full to h_done)
//   * The N = 208 hidden width is issued as N = 112 + N = 96 with separate commits, and the 13 feature groups are dealt to
//     three epilogue warpgroups in five ROUNDS (groups 0-2 | 3-5 | 6 | 7-9 | 10-12); round r signals r_ready[r] and the MMA
//     issuer follows the epilogue through the next layer's K loop round by round: half 1's epilogue runs under half 2's
//     MMAs, the next layer's first K groups under half 2's epilogue.
//   * Layer 0's A operand comes from shared memory (SS form): the head warps read the raw fp32 obs / act rows, apply the
//     scaler (models/utils.py:96, true division as the reference), split and write the K-panel tile of the NEXT unit
//     while the current one computes -- no tiling pass, no HBM round trip.  Layer 0 of unit u+1 is issued ahead of unit
//     u's head MMA (it only needs buffer 0), on its own barrier set.
//   * Ring loads per unit: W1, W2, W3, [head | W0 of the next unit] -- an even count keeps every load a layer ahead.
//   * bias rides in the constant-1 K column (feature 200 of every hidden activation, k = in of the input tile).
//
// Warp roles (576 threads): warps 0-11 epilogue (three per SM sub-partition, each owns a 32-lane TMEM quadrant and one
// feature group per round), warp 12 of the leader CTA issues every MMA (elect.sync), warp 13 streams weights,
// warps 14-17 produce input tiles and run the head stage (shared with ens_tc.cu: tc_head.cuh).
// Measurements, what was tried and what limits it: DESIGN.md 4.1b.
#include "internal.cuh"
#include "tc_common.cuh"
#include "tc_head.cuh"
#include "philox.cuh"

namespace {

using namespace tc;

constexpr int H = 200;
constexpr int KH = 208;                 // hidden K: 200 features + the constant-1 bias column + 7 zeros
constexpr int NGRP = KH / 16;           // 13 feature groups (one K = 16 MMA step each)
constexpr int KP = KH / 8;              // 26 K-panels
constexpr int NH1 = 112, NH2 = 96;      // the two MMAs (N halves) of a hidden layer
constexpr int G1 = NH1 / 16;            // groups 0..6 are half-1 features
constexpr int WR = 104;                 // weight rows per CTA per hidden layer: 56 (half 1) + 48 (half 2)
constexpr int WP = WR * 16;             // 1664 B: one K-panel of a CTA's weight block
constexpr int WPART = KP * WP;          // 43264 B: the hi (or lo) part of one hidden layer
constexpr int STAGE = 2 * WPART;        // 86528 B: one ring stage
constexpr int XP = 128 * 16;            // 2048 B: one K-panel of a 128-row input block
constexpr int EPI_WG = 3;                // epilogue warpgroups: 3 warps per SM sub-partition keep the MUFU fed (2 left it ~35 % idle)
constexpr int EPI_WARPS = 4 * EPI_WG;    // warps 0 .. EPI_WARPS-1; then the MMA issuer, the weight streamer, 4 head warps
// Epilogue rounds: round r converts feature groups 3r .. 3r+2 (one per warpgroup; round 2 is group 6 alone, the last of half 1;
// rounds 3, 4 are half 2) and signals r_ready[r]: the next layer's K loop follows the epilogue round by round.
constexpr int NROUND = 5;
static_assert(EPI_WG == 3, "the round table below deals one feature group per warpgroup and round");
__host__ __device__ constexpr int round_g0(int r) { return r < 2 ? 3 * r : r == 2 ? 6 : 7 + 3 * (r - 3); }
__host__ __device__ constexpr int round_n(int r) { return r == 2 ? 1 : 3; }
constexpr int W_ISSUE = EPI_WARPS, W_LOAD = EPI_WARPS + 1, W_HEAD = EPI_WARPS + 2;
constexpr int NTHREADS = 32 * (EPI_WARPS + 6);
constexpr int C_BUF0 = 0, C_BUF1 = 208, C_HEAD = 416;
constexpr int MAX_IN = 64;
#ifndef MOPO_TC3_POLL_NS
#define MOPO_TC3_POLL_NS 0      // back-off between mbarrier polls of the issuer and the epilogue warps (see tc_common.cuh): 32 ns
                                // while the epilogue was MUFU-bound; with the shared reciprocals polling back to back is 0.7 % faster
#endif
constexpr float W_SCALE = 64.0f;         // TERMS == 3: weight / bias images are stored x 64 (see the header comment)

struct S3 {
  int in, D1, HO, K0, NHEAD;            // K0 = roundup(in + 1, 16), NHEAD = roundup(2*D1, 16)
  __host__ __device__ int w0_part() const { return (K0 / 8) * WP; }
  __host__ __device__ int head_part() const { return KP * (NHEAD / 2) * 16; }
  __host__ __device__ int blob_bytes(int j) const { return j == 0 ? 2 * w0_part() : j < 4 ? STAGE : 2 * head_part(); }
  __host__ __device__ int img_bytes() const { return 2 * w0_part() + 3 * STAGE + 2 * head_part(); }
  __host__ __device__ int x_part() const { return (K0 / 8) * XP; }
};

__host__ __device__ inline S3 make_s3(int in, int D1) {
  S3 s;
  s.in = in; s.D1 = D1; s.HO = 2 * D1;
  s.K0 = (in + 1 + 15) / 16 * 16;
  s.NHEAD = (2 * D1 + 15) / 16 * 16;
  return s;
}

// ---------------------------------------------------------------------------------------------
// weight image: [net e][cta rank r] -> blob 0 (layer 0) | blobs 1..3 (hidden) | blob 4 (head); a blob = [hi part | lo part],
// each part in K-panel layout: element (row j, k) at byte (k/8) * (rows*16) + j*16 + (k%8)*2.
// hidden-layer block of CTA r, local row j:  j < 56: output feature 56 r + j (half 1, N = 112);
//   j >= 56: feature 112 + 48 r + (j - 56) of half 2 (N = 96: features 112..199, then 8 zero rows).
// K index: k < Kin -> weight row k; k == Kin -> bias (the A operand carries a constant 1 there); else 0.
// ---------------------------------------------------------------------------------------------
struct ImgSrc3 {
  const float* W[5];
  const float* b[5];
};

// one image per mode: [0] fp16 (hi, lo) x 64 for TERMS == 3, [1] bf16 (hi, unused lo = 0) for TERMS == 1
__global__ void build_image3_kernel(ImgSrc3 src, uint16_t* __restrict__ img3, uint16_t* __restrict__ img1, S3 s, int E) {
  const int per = s.img_bytes() / 2;
  const int64_t total = (int64_t)E * 2 * per;
  const int w0e = s.w0_part() / 2, hle = WPART / 2, hde = s.head_part() / 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i / (2 * per));
    const int r = (int)((i / per) % 2);
    int o = (int)(i % per);
    int layer, pe, rows;
    if (o < 2 * w0e) { layer = 0; pe = w0e; rows = WR; }
    else if (o < 2 * w0e + 6 * hle) { o -= 2 * w0e; layer = 1 + o / (2 * hle); o %= 2 * hle; pe = hle; rows = WR; }
    else { o -= 2 * w0e + 6 * hle; layer = 4; pe = hde; rows = s.NHEAD / 2; }
    const int part = o / pe;
    o %= pe;
    const int kc = o / (rows * 8), j = (o / 8) % rows, kk = o % 8;
    const int k = kc * 8 + kk;
    const int Kin = layer == 0 ? s.in : H;
    const int Nout = layer == 4 ? s.HO : H;
    int n;
    if (layer == 4) {
      // head rows of CTA r = head block r: [mean(d) x P | raw logvar(d) x P], d = r*dh + i, P = NHEAD/4 (see tc_head.cuh)
      const int P = s.NHEAD / 4, dh = (s.D1 + 1) / 2;
      const int hp = j / P, ii = j % P, d = r * dh + ii;
      n = d < (r ? s.D1 : dh) ? hp * s.D1 + d : -1;
    } else if (j < 56) n = 56 * r + j;
    else {
      n = 112 + 48 * r + (j - 56);
      if (n >= H) n = -1;
    }
    float v = 0.f;
    if (n >= 0 && n < Nout) {
      if (k < Kin) v = src.W[layer][((int64_t)e * Kin + k) * Nout + n];
      else if (k == Kin) v = src.b[layer][(int64_t)e * Nout + n];
    }
    const float vs = v * W_SCALE;
    const __half hi = __float2half_rn(vs);
    img3[i] = __half_as_ushort(part == 0 ? hi : __float2half_rn(vs - __half2float(hi)));
    img1[i] = part == 0 ? __bfloat16_as_ushort(__float2bfloat16_rn(v)) : (uint16_t)0;
  }
}

// per-row elite pick (bnn.py:342-344): supplied indices, or Philox over the elite list (same draw as prep_tiles_kernel)
__global__ void pick_kernel(int32_t* __restrict__ sel, const int32_t* __restrict__ model_inds, const int32_t* __restrict__ elites,
                            int n_elites, int64_t n_max, const int* __restrict__ d_n, uint64_t seed, uint64_t step_id,
                            const int32_t* __restrict__ origin, int64_t origin_base) {
  const int64_t n = d_n ? min((int64_t)*d_n, n_max) : n_max;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t orow = origin ? (uint32_t)origin[j] : (uint32_t)(origin_base + j);
    sel[j] = model_inds ? model_inds[j] : elites[philox_elite(orow, (uint32_t)step_id, seed, (uint32_t)n_elites)];
  }
}

// ---------------------------------------------------------------------------------------------
// forward kernel
// ---------------------------------------------------------------------------------------------
struct T3Args {
  const uint8_t* img;      // weight images [E][2][img_bytes]
  const float *obs, *act;  // x = concat(obs [n, od], act [n, ad]) (fake_env.py:46); ad may be 0
  const float *mu, *sigma; // scaler (models/utils.py:96), or nullptr: x is already scaled
  int od, ad;
  int64_t net_stride;      // rows between consecutive nets' inputs (3-D predict, fc.py:100-101), 0: shared rows
  const float* max_lv;     // [D1]
  const float* min_lv;     // [D1]
  HeadOut out;
  int n_max;
  const int* d_n;
  int E;
  S3 s;
  unsigned long long* trace;   // debug timeline (mopo_debug_tc3_trace, probe build only) or nullptr
  InputReady ready;            // chunk flags of inputs still in flight over PCIe (units then run tile-major), or flags == nullptr
  int tile_major;
};

struct Bars3 {
  uint64_t w_full[2];      // this CTA's ring stage has landed (tx)
  uint64_t w_ready[2];     // leader: both CTAs' stage has landed                            (count 2)
  uint64_t w_free[2];      // each CTA: the MMAs that read the stage have retired            (commit)
  uint64_t x_full;         // leader: both CTAs' input block is in smem                      (count 8: head warps)
  uint64_t x_free;         // each CTA: the layer-0 MMAs that read the input block retired   (commit)
  uint64_t d_full[2];      // each CTA: half 1 / half 2 accumulator of hidden layer 1..3     (commit)
  uint64_t r_ready[NROUND];// leader: the feature groups of epilogue round r are written as the next A operand (count: ROUND_CNT)
  uint64_t d0_full[2];     // the same for layer 0, whose MMAs are issued a unit early (ahead of the previous unit's head):
  uint64_t r0_ready[NROUND]; // own barriers, so that its phases cannot lap or mix with layer 3's
  uint64_t h_full;         // each CTA: head accumulator ready                               (commit)
  uint64_t h_done;         // leader: the head warps have read it                            (count 8)
  uint32_t tmem_base;
  uint32_t pad;
};

__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// swish (fc.py:21: x * sigmoid(x)).  TERMS == 3: x * rcp(1 + ex2(-x log2 e)), two MUFU results, relative error ~3e-7.
// (Measured and rejected, profiles/r02_tc3_timeline_v2_newton.txt: taking every second reciprocal on the FMA pipe -- magic
// seed + 3 Newton steps -- to balance MUFU against issue slots made the epilogue 1.6x SLOWER: with two epilogue warps per
// SM sub-partition the dependent FMA chains are not hidden, as round 1 found for the bf16 kernel.)  TERMS == 1: tanh form.
template <int TERMS> __device__ __forceinline__ float swish_acc(float x) {
  if constexpr (TERMS == 1) { const float xh = 0.5f * x; return fmaf(xh, tanh_fast(xh), xh); }
  else return x * rcp_approx(1.0f + ex2_approx(-1.44269504f * x));
}

// operand formats: TERMS == 3 -> fp16 pairs, TERMS == 1 -> bf16
template <int TERMS> __device__ __forceinline__ uint32_t pack2(float a, float b) {
  if constexpr (TERMS == 3) return pack_f16x2(a, b); else return pack_bf16x2(a, b);
}
template <int TERMS> __device__ __forceinline__ uint32_t idesc_of(uint32_t M, uint32_t N) {
  return TERMS == 3 ? idesc_f16(M, N) : idesc_bf16(M, N);
}
// (hi, lo) split of two values: hi = f16x2(a, b), lo = f16x2(a - hi.a, b - hi.b)
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  hi = pack_f16x2(a, b);
  lo = pack_f16x2(a - f16lo_to_f32(hi), b - f16hi_to_f32(hi));
}

// One feature group (16 accumulator columns of this thread's row) at taddr -> swish -> split -> the SAME 16 columns as
// [hi x 8 | lo x 8].  ones: this is group 12, whose feature 200 is the constant 1 of the bias row.
// The epilogue is a LOOP over groups with this one body (two inlined copies, ~5 KB of code) shared by every warp, layer and
// half: the first version unrolled each (warpgroup, half) part on its own -- 40 KB of straight-line epilogue code that each
// warp ran once per layer -- and a quarter of all warp stall samples were instruction-fetch stalls on exactly those FMUL /
// MUFU / F2FP lines (profiles/r02_tc3_v3_ncu_stalls.txt; the L1.5 instruction cache is 32 KB).
template <int TERMS>
__device__ __forceinline__ void group_store(uint32_t taddr, const uint32_t (&v)[16], bool ones) {
  // Written pair by pair (swish of two features, then their split) so that the fp16 split of one pair sits between the
  // MUFU instructions of the following pairs in the instruction stream: with all 16 swishes first and all 8 splits after
  // them every warp had a ~60-instruction tail without a single MUFU, and the warps of a round run in phase.
  uint32_t o[16];
  constexpr float sc = TERMS == 3 ? 1.0f / W_SCALE : 1.0f;
#if !defined(MOPO_TC3_NO_SHARED_RCP)
  if constexpr (TERMS == 3) {
    // The sigmoids of FOUR features share ONE reciprocal (batch inversion): with d_i = 1 + 2^min(-x_i log2 e, 30),
    // r = rcp(d0 d1 d2 d3), 1/d0 = d1 (d2 d3 r), ... -- five MUFU results per four activations instead of eight (the MUFU is this
    // kernel's busiest pipe) for nine more FMULs.  The clamp keeps the product finite (<= 2^121); below x = -20.8 swish is
    // < 2e-8 in magnitude either way.  Relative error ~4e-7 (was 3e-7): the path's measured error is unchanged.
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float x[4], d[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        x[i] = sc * __uint_as_float(v[4 * q + i]);
        d[i] = 1.0f + ex2_approx(fminf(-1.44269504f * x[i], 30.0f));
      }
      const float p01 = d[0] * d[1], p23 = d[2] * d[3];
      const float r = rcp_approx(p01 * p23);
      const float r01 = r * p23, r23 = r * p01;
      float s0 = x[0] * (d[1] * r01);
      const float s1 = x[1] * (d[0] * r01), s2 = x[2] * (d[3] * r23), s3 = x[3] * (d[2] * r23);
      if (q == 2 && ones) s0 = 1.0f;                                   // feature 8 of group 12 = column 200: the bias row's constant
      split2(s0, s1, o[2 * q], o[8 + 2 * q]);
      split2(s2, s3, o[2 * q + 1], o[8 + 2 * q + 1]);
    }
  } else
#endif
  {
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      float s0 = swish_acc<TERMS>(sc * __uint_as_float(v[2 * p]));
      const float s1 = swish_acc<TERMS>(sc * __uint_as_float(v[2 * p + 1]));
      if (p == 4 && ones) s0 = 1.0f;
      if constexpr (TERMS == 3) split2(s0, s1, o[p], o[8 + p]);
      else { o[p] = pack_bf16x2(s0, s1); o[8 + p] = 0u; }
    }
  }
  tmem_st16(taddr, o);
}

// K groups [g0, g1) of one N half: D (+)= A[g] * B[g]^T with A from tensor memory (hi at column 16 g, lo at 16 g + 8)
template <int TERMS>
__device__ __forceinline__ void issue_ts(uint32_t d, uint32_t a_buf, uint32_t b_base, uint32_t b_lo, uint32_t bpanel,
                                         uint32_t idesc, int g0, int g1) {
  for (int g = g0; g < g1; ++g) {
    const uint32_t ah = a_buf + 16 * g;
    const uint32_t b = b_base + g * 2 * bpanel;
    mma_ts<2>(d, ah, smem_desc(b, bpanel, 128), idesc, g > 0);
    if constexpr (TERMS == 3) {
      mma_ts<2>(d, ah, smem_desc(b + b_lo, bpanel, 128), idesc, 1);
      mma_ts<2>(d, ah + 8, smem_desc(b, bpanel, 128), idesc, 1);
    }
  }
}

constexpr int TRACE_CAP = 2048;   // events per role
// timeline probe: role r (0 issuer, 1 epilogue warp 0, 2 epilogue warp 4, 3 weight streamer, 4 head warp 12) of pair 0 / leader CTA
#define TR(role, tag)                                                                                   \
  do {                                                                                                  \
    if (TRACE && a.trace && pair == 0 && rank == 0 && lane == 0 && tr_n < TRACE_CAP)                    \
      a.trace[(role) * TRACE_CAP + tr_n++] = ((unsigned long long)(tag) << 48) | ((unsigned long long)clock64() & 0xFFFFFFFFFFFFull); \
  } while (0)

template <int TERMS, bool TRACE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1) ens_tc3_kernel(T3Args a) {
  int tr_n = 0;
  (void)tr_n;
  extern __shared__ __align__(128) uint8_t smem[];
  const S3 s = a.s;
  uint8_t* sW = smem;                                   // two ring stages
  uint8_t* sX = sW + 2 * STAGE;                         // input block: hi panels | lo panels
  float* sLv = reinterpret_cast<float*>(sX + 2 * s.x_part());     // head constants (see tc_head.cuh)
  float* sMu = sLv + 2 * MOPO_MAX_D1;
  float* sSg = sMu + MAX_IN;
  Bars3* bars = reinterpret_cast<Bars3*>(sSg + MAX_IN);

  const int tid = threadIdx.x, lane = tid % 32;
  const int warp = __shfl_sync(0xffffffffu, tid / 32, 0);
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x / 2, n_pairs = gridDim.x / 2;
  const int n = __shfl_sync(0xffffffffu, a.d_n ? min(*a.d_n, a.n_max) : a.n_max, 0);
  const int nT = (n + 255) / 256;
  const int64_t U = (int64_t)a.E * nT;                  // (net, tile) units: net-major, or tile-major while the inputs arrive
  const bool tile_major = a.tile_major != 0;
  auto net_of = [&](int u) { return tile_major ? u % a.E : u / nT; };
  auto tile_of = [&](int u) { return tile_major ? u / a.E : u % nT; };
  // a pair's units: a contiguous block [U p / P, U (p + 1) / P), or -- while the inputs arrive tile by tile -- every P-th unit
  // from p on, so that all pairs move through the tiles in upload order
  const int ustep = tile_major ? n_pairs : 1;
  const int u0 = tile_major ? pair : (int)(U * pair / n_pairs), u1 = tile_major ? (int)U : (int)(U * (pair + 1) / n_pairs);

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->w_full[i], 1); mbar_init(&bars->w_ready[i], 2); mbar_init(&bars->w_free[i], 1);
      mbar_init(&bars->d_full[i], 1); mbar_init(&bars->d0_full[i], 1);
    }
    for (int r = 0; r < NROUND; ++r) { mbar_init(&bars->r_ready[r], 8 * round_n(r)); mbar_init(&bars->r0_ready[r], 8 * round_n(r)); }
    mbar_init(&bars->x_full, 8); mbar_init(&bars->x_free, 1);
    mbar_init(&bars->h_full, 1); mbar_init(&bars->h_done, 8);
    fence_barrier_init();
  }
  for (int i = tid; i < 2 * (s.NHEAD / 4); i += NTHREADS) {          // head constants per head block (tc_head.cuh)
    const int P = s.NHEAD / 4, dh = (s.D1 + 1) / 2;
    const int w = i / P, d = w * dh + i % P;
    const bool ok = d < (w ? s.D1 : dh);
    sLv[w * P + i % P] = ok ? expf(a.min_lv[d]) : 0.0f;
    sLv[MOPO_MAX_D1 + w * P + i % P] = ok ? fminf(expf(-a.max_lv[d]), 1e30f) : 1e30f;
  }
  for (int i = tid; i < MAX_IN; i += NTHREADS) {
    const bool sc = a.mu != nullptr && i < s.in;
    sMu[i] = sc ? a.mu[i] : 0.0f;
    sSg[i] = sc ? a.sigma[i] : 1.0f;
#ifndef MOPO_SCALER_DIVIDE
    sSg[i] = 1.0f / sSg[i];                            // the input warps multiply (one rounding away from models/utils.py:96's division)
#endif
  }
  if (warp == W_ISSUE) { tmem_alloc<2>(&bars->tmem_base, 512); tmem_relinquish<2>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = __shfl_sync(0xffffffffu, bars->tmem_base, 0);
  cluster_sync_all();                                   // both CTAs' barriers exist before any remote arrive / multicast commit

  // PDL: see ens_tc.cu.  Everything the previous grid (the previous step's tail) writes -- obs, the elite picks -- and
  // everything it reads that this grid writes (selmv / pen2) is touched by the head warps only, after pdl_wait().
  pdl_launch_dependents();

  if (warp == W_LOAD) {
    // ===== weight streamer.  Ring stage = load index & 1.  Loads: a prologue [W0 of the first unit], then per unit
    // W1, W2, W3 and ONE stage holding [head of this unit | W0 of the next unit] (an even number of loads per unit keeps
    // every load at least one layer ahead of its use; with five the second layer's weights could only be requested
    // once the previous unit's head had retired and arrived ~1500 cycles late, profiles/r02_tc3_timeline_first.txt) =====
    if (lane == 0) {
      uint32_t li = 0;
      const uint32_t hw_w0 = 2 * (uint32_t)s.head_part();            // offset of the W0 blob inside a [head | W0] stage
      const size_t off_w0 = 0, off_h = 2 * (size_t)s.w0_part() + 3 * (size_t)STAGE;
      auto img_of = [&](int u) { return a.img + ((size_t)net_of(u) * 2 + rank) * s.img_bytes(); };
      auto stage_wait = [&](int tag) {
        const uint32_t st = li & 1;
        if (li >= 2) mbar_wait_relaxed(&bars->w_free[st], ((li >> 1) - 1) & 1, 1000 + tag);
        TR(3, 70 + tag);
        return st;
      };
      auto stage_done = [&](uint32_t st, int tag) {
        mbar_wait_relaxed(&bars->w_full[st], (li >> 1) & 1, 1100 + tag);
        TR(3, 80 + tag);
        mbar_arrive_remote(&bars->w_ready[st], 0);
        ++li;
      };
      if (u0 < u1) {                                                  // prologue: W0 of the first unit
        const uint32_t st = stage_wait(0);
        mbar_expect_tx(&bars->w_full[st], 2 * (uint32_t)s.w0_part());
        bulk_g2s(sW + st * STAGE + hw_w0, img_of(u0) + off_w0, 2 * (uint32_t)s.w0_part(), &bars->w_full[st]);
        stage_done(st, 0);
      }
      for (int u = u0; u < u1; u += ustep) {
        const uint8_t* src = img_of(u);
        for (int j = 1; j < 4; ++j) {
          const uint32_t st = stage_wait(j);
          mbar_expect_tx(&bars->w_full[st], STAGE);
          const uint8_t* g = src + 2 * (size_t)s.w0_part() + (size_t)(j - 1) * STAGE;
          bulk_g2s(sW + st * STAGE, g, WPART, &bars->w_full[st]);
          bulk_g2s(sW + st * STAGE + WPART, g + WPART, WPART, &bars->w_full[st]);
          stage_done(st, j);
        }
        const uint32_t st = stage_wait(4);
        const bool more = u + ustep < u1;
        mbar_expect_tx(&bars->w_full[st], 2 * (uint32_t)s.head_part() + (more ? 2 * (uint32_t)s.w0_part() : 0u));
        bulk_g2s(sW + st * STAGE, src + off_h, 2 * (uint32_t)s.head_part(), &bars->w_full[st]);
        if (more) bulk_g2s(sW + st * STAGE + hw_w0, img_of(u + ustep) + off_w0, 2 * (uint32_t)s.w0_part(), &bars->w_full[st]);
        stage_done(st, 4);
      }
    }
    __syncwarp();
  } else if (warp == W_ISSUE) {
    // ===== MMA issuer (leader CTA): all lanes run the loop, one elected lane issues =====
    if (rank == 0) {
      if (TRACE && a.trace && lane == 0) {                       // wall clock beside the cycle counter: the SM clock under this load
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        if (pair == 0) { a.trace[5 * TRACE_CAP - 4] = (unsigned long long)clock64(); a.trace[5 * TRACE_CAP - 3] = ns; }
        if (pair < 256) a.trace[4 * TRACE_CAP + 1024 + 2 * pair] = ns;       // every pair: start / end of its issuer loop
      }
      const uint32_t id1 = idesc_of<TERMS>(256, NH1), id2 = idesc_of<TERMS>(256, NH2), idh = idesc_of<TERMS>(256, s.NHEAD);
      const uint32_t wa = smem_u32(sW), xa = smem_u32(sX);
      const uint32_t hpanel = (uint32_t)(s.NHEAD / 2) * 16;
      uint32_t li = 0, n_x = 0, n_r = 0, n_h = 0, n_u = 0;   // n_r / n_u: epilogue passes over hidden-layer / layer-0 accumulators consumed
      const uint32_t hw_w0 = 2 * (uint32_t)s.head_part();
      // layer 0 of a unit: A = the input block in shared memory (SS form), B = the W0 blob at wb, D -> buffer 0
      auto issue_l0 = [&](uint32_t wb) {
        mbar_wait_cluster_backoff<MOPO_TC3_POLL_NS>(&bars->x_full, n_x & 1, 1); ++n_x;
        tc_fence_after();
        TR(0, 1);
        if (elect_one()) {
          const uint32_t wlo = (uint32_t)s.w0_part(), xlo = (uint32_t)s.x_part();
          for (int hf = 0; hf < 2; ++hf) {
            const uint32_t d = tbase + C_BUF0 + (hf ? NH1 : 0);
            const uint32_t idesc = hf ? id2 : id1;
            const uint32_t boff = hf ? 56 * 16 : 0;
            for (int ks = 0; ks < s.K0 / 16; ++ks) {
              const uint32_t ax = xa + ks * 2 * XP, b = wb + boff + ks * 2 * WP;
              mma_ss<2>(d, smem_desc(ax, XP, 128), smem_desc(b, WP, 128), idesc, ks > 0);
              if constexpr (TERMS == 3) {
                mma_ss<2>(d, smem_desc(ax, XP, 128), smem_desc(b + wlo, WP, 128), idesc, 1);
                mma_ss<2>(d, smem_desc(ax + xlo, XP, 128), smem_desc(b, WP, 128), idesc, 1);
              }
            }
            mma_commit<2>(&bars->d0_full[hf]);
          }
          mma_commit<2>(&bars->x_free);
        }
        __syncwarp();
      };
      if (u0 < u1) {                                    // prologue: layer 0 of the first unit
        mbar_wait_cluster_backoff<MOPO_TC3_POLL_NS>(&bars->w_ready[0], 0, 0);
        TR(0, 0);
        issue_l0(wa + hw_w0);
        if (elect_one()) mma_commit<2>(&bars->w_free[0]);
        __syncwarp();
        ++li;
      }
      for (int u = u0; u < u1; u += ustep) {
        for (int j = 1; j < 5; ++j, ++li) {
          const uint32_t st = li & 1;
          const uint32_t wb = wa + st * STAGE;
          mbar_wait_cluster_backoff<MOPO_TC3_POLL_NS>(&bars->w_ready[st], (li >> 1) & 1, 10 * j);
          TR(0, 10 * j);
          // the A operand of layer j / the head = what the epilogue makes of layer j-1's accumulator, round by round
          uint64_t* rr = j == 1 ? bars->r0_ready : bars->r_ready;
          const uint32_t par = (j == 1 ? n_u : n_r) & 1;
          if (j < 4) {
            const uint32_t abuf = tbase + (((j - 1) & 1) ? C_BUF1 : C_BUF0);
            const uint32_t dbuf = tbase + ((j & 1) ? C_BUF1 : C_BUF0);
            for (int r = 0; r < NROUND; ++r) {
              mbar_wait_cluster_backoff<MOPO_TC3_POLL_NS>(&rr[r], par, 10 * j + r);
              tc_fence_after();
              TR(0, 50 + 10 * j + r);
              if (elect_one()) {
                issue_ts<TERMS>(dbuf, abuf, wb, WPART, WP, id1, round_g0(r), round_g0(r) + round_n(r));
                if (r == NROUND - 1) {
                  mma_commit<2>(&bars->d_full[0]);
                  issue_ts<TERMS>(dbuf + NH1, abuf, wb + 56 * 16, WPART, WP, id2, 0, NGRP);
                  mma_commit<2>(&bars->d_full[1]);
                  mma_commit<2>(&bars->w_free[st]);
                }
              }
              __syncwarp();
            }
            TR(0, 10 * j + 3);
          } else {
            // [head | W0'] stage.  Layer 0 of the NEXT unit goes to the tensor pipe first: it writes buffer 0 (free since
            // layer 3's MMAs were issued) while the epilogue still turns layer 3's accumulator (buffer 1) into the head's
            // A operand, and its own epilogue then runs under the head MMA.
            if (u + ustep < u1) issue_l0(wb + hw_w0);
            // the head MMA overwrites the previous unit's head accumulator: the head warps must have read it
            if (n_h > 0) mbar_wait_cluster_backoff<MOPO_TC3_POLL_NS>(&bars->h_done, (n_h - 1) & 1, 44);
            ++n_h;
            for (int r = 0; r < NROUND; ++r) {
              mbar_wait_cluster_backoff<MOPO_TC3_POLL_NS>(&rr[r], par, 40 + r);
              tc_fence_after();
              if (r == 0) TR(0, 42);
              if (elect_one()) {
                issue_ts<TERMS>(tbase + C_HEAD, tbase + C_BUF1, wb, (uint32_t)s.head_part(), hpanel, idh, round_g0(r), round_g0(r) + round_n(r));
                if (r == NROUND - 1) {
                  mma_commit<2>(&bars->h_full);
                  mma_commit<2>(&bars->w_free[st]);
                }
              }
              __syncwarp();
            }
          }
          if (j == 1) ++n_u; else ++n_r;
        }
      }
      if (TRACE && a.trace && lane == 0) {
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        if (pair == 0) { a.trace[5 * TRACE_CAP - 2] = (unsigned long long)clock64(); a.trace[5 * TRACE_CAP - 1] = ns; }
        if (pair < 256) a.trace[4 * TRACE_CAP + 1024 + 2 * pair + 1] = ns;
      }
    }
  } else if (warp < EPI_WARPS) {
    // ===== epilogue warps: accumulator half -> swish -> (hi, lo) A operand, in place =====
    const int wg = warp >> 2, q = warp & 3;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    uint32_t n_d = 0, n_u = 0;                          // hidden layers 1..3 / units consumed (every barrier flips once per layer)
    uint32_t va[16], vb[16];
    for (int u = u0; u < u1; u += ustep, ++n_u) {
      for (int j = 0; j < 4; ++j) {
        const uint32_t buf = tlane + ((j & 1) ? C_BUF1 : C_BUF0);
        uint64_t* dfull = j == 0 ? bars->d0_full : bars->d_full;
        uint64_t* rready = j == 0 ? bars->r0_ready : bars->r_ready;
        const uint32_t par = (j == 0 ? n_u : n_d) & 1;
        // round r: group round_g0(r) + wg (round 2 only has warpgroup 0's group 6).  Rounds alternate between the two
        // register sets so that the tcgen05.ld of this warp's next group is in flight while the current one is converted.
        bool pre = false;                                  // the current round's group is already on its way into registers
#pragma unroll 1
        for (int r = 0; r < NROUND; r += 2) {
#pragma unroll
          for (int ab = 0; ab < 2; ++ab) {
            const int rd = r + ab;
            if (rd >= NROUND) break;
            uint32_t (&cur)[16] = ab ? vb : va;
            uint32_t (&nxt)[16] = ab ? va : vb;
            const bool mine = wg < round_n(rd);
            if (rd == 0 || rd == 3) {                      // first round of a half: its accumulator must have been committed
              if (q == 0 && wg < 2) TR(1 + wg, 100 + 10 * j + (rd == 3));
              mbar_wait_backoff<MOPO_TC3_POLL_NS>(&dfull[rd == 3], par, 100 + 10 * j + rd);
              tc_fence_after();
              if (q == 0 && wg < 2) TR(1 + wg, 200 + 10 * j + (rd == 3));
            }
            if (mine) {
              const uint32_t t_cur = buf + 16 * (round_g0(rd) + wg);
              if (!pre) tmem_ld16(t_cur, cur);
              tmem_ld_wait();
              pre = rd + 1 < NROUND && rd + 1 != 3 && wg < round_n(rd + 1);      // same half, and this warpgroup has a group there
              if (pre) tmem_ld16(buf + 16 * (round_g0(rd + 1) + wg), nxt);
              group_store<TERMS>(t_cur, cur, round_g0(rd) + wg == NGRP - 1);
              tmem_st_wait();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_remote(&rready[rd], 0);
              if (q == 0 && wg < 2) TR(1 + wg, 400 + 10 * j + rd);
            } else pre = false;
            if ((rd == 2 || rd == 4) && q == 0 && wg < 2) TR(1 + wg, 300 + 10 * j + (rd == 4));
          }
        }
        if (j > 0) ++n_d;
      }
    }
  } else {
    // ===== head warps (10-13, TMEM quadrant warp % 4): input tiles of the next unit, then the head stage =====
    const int q = warp & 3;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    const int rr = q * 32 + lane;
    const HeadShape hs = {s.D1, s.NHEAD};
    constexpr bool PREC = TERMS == 3;
    const int kcs = s.K0 / 8;
    pdl_wait();
    // x block of unit u: this thread's row -> K-panels (hi | lo) of the SS-form A operand of layer 0
    int ready_tiles = 0;
    auto produce_x = [&](int u) {
      const int t = tile_of(u);
      if (a.ready.flags && t >= ready_tiles) {           // this tile's rows may still be on their way: wait for its chunk's flag
        int c = 0;
        while (c + 1 < a.ready.n && t >= a.ready.end_tile[c]) ++c;
        if (lane == 0) {
          uint32_t v, spins = 0;
          for (;;) {
            asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(a.ready.flags + c) : "memory");
            if (v == a.ready.value) break;
            __nanosleep(500);
            if (++spins > (1u << 24)) asm volatile("trap;");   // ~10 s: a lost upload fails the launch instead of computing on garbage
          }
          __threadfence();                               // the rows are read (ld.global.cg) after the flag
        }
        __syncwarp();
        ready_tiles = a.ready.end_tile[c];               // chunks arrive in order: nothing below this needs another look
      }
      const int64_t row = 256 * (int64_t)t + (int64_t)rank * 128 + rr;
      const bool ok = row < n;
      const int64_t src_row = row + (int64_t)net_of(u) * a.net_stride;
      const float* po = a.obs + src_row * a.od;
      const float* pa = a.act + src_row * a.ad;
      // four K-panels (32 input columns) per pass, all their loads issued before the first use: the first block of a launch
      // is produced behind pdl_wait() with nothing to hide its latency (one L2 round trip instead of one per panel)
      for (int kc0 = 0; kc0 < kcs; kc0 += 4) {
        float raw[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int k = kc0 * 8 + i;
          raw[i] = 0.0f;
          if (ok && k < s.in) raw[i] = __ldcg(k < a.od ? po + k : pa + (k - a.od));   // L2: the rows may have landed after a neighbour's line was cached
        }
#pragma unroll
        for (int pnl = 0; pnl < 4; ++pnl) {
          const int kc = kc0 + pnl;
          if (kc >= kcs) break;
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int k = kc * 8 + i;
            float x = 0.0f;
            if (ok) {
#ifdef MOPO_SCALER_DIVIDE
              if (k < s.in) x = (raw[pnl * 8 + i] - sMu[k]) / sSg[k];          // models/utils.py:96
#else
              if (k < s.in) x = (raw[pnl * 8 + i] - sMu[k]) * sSg[k];          // models/utils.py:96 (sSg = 1 / sigma)
#endif
              else if (k == s.in) x = 1.0f;                                    // the bias row's constant
            }
            v[i] = x;
          }
          uint32_t hi[4], lo[4];
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            if constexpr (TERMS == 3) split2(v[2 * p], v[2 * p + 1], hi[p], lo[p]);
            else { hi[p] = pack_bf16x2(v[2 * p], v[2 * p + 1]); lo[p] = 0u; }
          }
          st_shared_v4(sX + kc * XP + rr * 16, hi[0], hi[1], hi[2], hi[3]);
          st_shared_v4(sX + s.x_part() + kc * XP + rr * 16, lo[0], lo[1], lo[2], lo[3]);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(&bars->x_full, 0);
    };
    if (u0 < u1) produce_x(u0);
    uint32_t n_h = 0;
    for (int u = u0; u < u1; u += ustep, ++n_h) {
      const int e = net_of(u), t = tile_of(u);
      if (u + ustep < u1) {
        mbar_wait_relaxed(&bars->x_free, n_h & 1, 200);          // layer 0 of unit u has consumed the block
        if (q == 0) TR(4, 150);
        produce_x(u + ustep);
        if (q == 0) TR(4, 250);
      }
      const int64_t row = 256 * (int64_t)t + (int64_t)rank * 128 + rr;
      const bool valid = row < n;
      bool mine = false;
      if (a.out.sel && valid) mine = a.out.sel[row] == e;
      mbar_wait_relaxed(&bars->h_full, n_h & 1, 210);
      tc_fence_after();
      if (q == 0) TR(4, 240);
      const uint32_t tblk = tlane + C_HEAD;
      bool arrived = false;
      if (a.out.sel) {
        if (s.D1 == 18) { head_fused<48, 18, PREC>(tblk, hs, sLv, a.out, e, row, valid, mine, &bars->h_done, lane); arrived = true; }
        else if (s.D1 == 12) { head_fused<32, 12, PREC>(tblk, hs, sLv, a.out, e, row, valid, mine, &bars->h_done, lane); arrived = true; }
        else if (s.NHEAD == 48) { head_fused<48, 0, PREC>(tblk, hs, sLv, a.out, e, row, valid, mine, &bars->h_done, lane); arrived = true; }
        else if (s.NHEAD == 32) { head_fused<32, 0, PREC>(tblk, hs, sLv, a.out, e, row, valid, mine, &bars->h_done, lane); arrived = true; }
        else { epi_head<true, PREC>(tblk, 0, hs, sLv, a.out, e, row, valid, mine); epi_head<true, PREC>(tblk + s.NHEAD / 2, 1, hs, sLv, a.out, e, row, valid, mine); }
      } else { epi_head<false, PREC>(tblk, 0, hs, sLv, a.out, e, row, valid, false); epi_head<false, PREC>(tblk + s.NHEAD / 2, 1, hs, sLv, a.out, e, row, valid, false); }
      if (!arrived) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(&bars->h_done, 0);
      }
      if (q == 0) TR(4, 340);
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                   // peer smem / TMEM must outlive the leader's MMAs
  if (warp == W_ISSUE) tmem_dealloc<2>(tbase, 512);
}

size_t tc3_smem_bytes(const S3& s) {
  return (size_t)2 * STAGE + 2 * s.x_part() + (2 * MOPO_MAX_D1 + 2 * MAX_IN) * sizeof(float) + sizeof(Bars3) + 128;
}

}  // namespace

int tc3_supported(const mopo_ens* h) {
  if (h->H != H || h->L != 4) return 0;
  if (h->in + 1 > MAX_IN || h->HO > 96) return 0;
  const S3 s = make_s3(h->in, h->D1);
  return 2 * s.w0_part() <= STAGE && 2 * s.head_part() <= STAGE && tc3_smem_bytes(s) <= 227 * 1024;
}

int tc3_build_images(mopo_ens* h, cudaStream_t st) {
  const S3 s = make_s3(h->in, h->D1);
  const size_t one = (size_t)h->E * 2 * s.img_bytes();
  const size_t bytes = 2 * one;                         // [fp16 (hi, lo) image | bf16 image]
  if (h->tc3_img_bytes < bytes) {
    if (h->tc3_img) cudaFree(h->tc3_img);
    h->tc3_img = nullptr;
    CUDA_TRY(cudaMalloc(&h->tc3_img, bytes));
    h->tc3_img_bytes = bytes;
  }
  ImgSrc3 src;
  for (int l = 0; l < 5; ++l) { src.W[l] = h->W[l]; src.b[l] = h->b[l]; }
  MOPO_COUNT_LAUNCH, build_image3_kernel<<<592, 256, 0, st>>>(src, reinterpret_cast<uint16_t*>(h->tc3_img),
                                                              reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(h->tc3_img) + one), s, h->E);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int tc3_pick(mopo_ens* h, int32_t* sel, const int32_t* model_inds, int64_t n_max, const int* d_n, uint64_t seed,
             uint64_t step_id, const int32_t* origin, int64_t origin_base, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  const int64_t g64 = (n_max + 255) / 256;
  MOPO_COUNT_LAUNCH, pick_kernel<<<(int)(g64 > 148 * 8 ? 148 * 8 : g64), 256, 0, st>>>(sel, model_inds, h->elites, h->n_elites, n_max, d_n,
                                                                                  seed, step_id, origin, origin_base);
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

int tc3_launch(mopo_ens* h, int terms, const Tc3In& x, int64_t n_max, const int* d_n, const TcOut& o, cudaStream_t st) {
  if (n_max == 0) return MOPO_OK;
  if (!h->tc3_img) { mopo_set_error("tc3_launch: weight images not built"); return MOPO_ESTATE; }
  T3Args a;
  a.s = make_s3(h->in, h->D1);
  a.img = reinterpret_cast<const uint8_t*>(h->tc3_img) + (terms == 3 ? 0 : (size_t)h->E * 2 * a.s.img_bytes());
  a.obs = x.obs; a.act = x.act ? x.act : x.obs; a.od = x.od; a.ad = x.ad; a.net_stride = x.net_stride;
  a.mu = x.scale ? h->mu : nullptr; a.sigma = x.scale ? h->sigma : nullptr;
  a.max_lv = h->max_lv; a.min_lv = h->min_lv;
  a.out.acc_scale = terms == 3 ? 1.0f / W_SCALE : 1.0f;
  a.out.mean = o.mean; a.out.lv = o.lv; a.out.ldb = o.ldb;
  a.out.sel = o.sel; a.out.selmv = o.selmv; a.out.pen2 = o.pen2; a.out.E = h->E;
  a.n_max = (int)n_max; a.d_n = d_n; a.E = h->E;
  a.trace = reinterpret_cast<unsigned long long*>(h->tc_trace);
  a.ready = x.ready;
  a.tile_major = x.ready.flags != nullptr;
  const size_t smem = tc3_smem_bytes(a.s);
  const int n_pairs = h->sm_count / 2;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  if (h->timing) {
    if (h->ev_used + 2 > h->ev.size()) {
      for (int i = 0; i < 2; ++i) { cudaEvent_t e; CUDA_TRY(cudaEventCreate(&e)); h->ev.push_back(e); }
    }
    ev0 = h->ev[h->ev_used]; ev1 = h->ev[h->ev_used + 1]; h->ev_used += 2;
    CUDA_TRY(cudaEventRecord(ev0, st));
  }
  auto kern = terms == 3 ? ens_tc3_kernel<3, false> : ens_tc3_kernel<1, false>;
#ifdef MOPO_DEBUG_PROBES
  if (a.trace) kern = terms == 3 ? ens_tc3_kernel<3, true> : ens_tc3_kernel<1, true>;
#endif
  CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * n_pairs); cfg.blockDim = dim3(NTHREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = o.pdl ? 1 : 0;          // rollout chain: behind the previous step's tail (see pdl_wait())
  MOPO_COUNT_LAUNCH;
  CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, a));
  if (ev1) CUDA_TRY(cudaEventRecord(ev1, st));
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}
