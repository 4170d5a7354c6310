// Head stage shared by the tcgen05 ensemble kernels (ens_tc.cu: bf16, ens_tc3.cu: split-bf16 fp32-accurate):
// soft clamp + variance of BNN._compile_outputs (bnn.py:654-655 / 669-670, 675), penalty partials and elite gather
// (fake_env.py:78-81, 110), read from a tile's head accumulator in tensor memory.
#pragma once
#include "internal.cuh"
#include "tc_common.cuh"

namespace tc {

struct HeadShape { int D1, NHEAD; };     // NHEAD = roundup(2*D1, 16): accumulator columns of the fused [mean | raw logvar] head
struct HeadOut {
  float acc_scale;         // accumulator -> output units (the fp32-accurate kernel stores its weights x 64; else 1)
  // FULL output (sel == nullptr): everything BNN.predict returns
  float* mean;             // [E, ldb, D1]
  float* lv;               // [E, ldb, D1]
  int64_t ldb;
  // FUSED output (sel != nullptr): what FakeEnv.step keeps of the [E,B,D1] arrays (fake_env.py:78-81,110)
  const int32_t* sel;      // [n] net chosen for each row (the elite pick)
  float2* selmv;           // [D1, ldb] (mean, variance) of the chosen net, dimension-major so that stores coalesce
  float* pen2;             // [2, E, ldb] sum_d var_d over each warpgroup's half of the output dims
  int E;
};

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// softplus(x) = max(x,0) + log(1 + exp(-|x|)); ex2/lg2-based: absolute error < 1e-6 over the clamp's range
__device__ __forceinline__ float softplus_fast(float x) { return fmaxf(x, 0.0f) + __logf(1.0f + __expf(-fabsf(x))); }

// head epilogue.  Soft clamp + exp of bnn.py:654-655/669-670,675 in closed form:
//   lv1 = max - softplus(max - raw), lv = min + softplus(lv1 - min), var = exp(lv)
//   =>  var = e^min + e^lv1 = e^min + 1 / (e^-max + e^-raw)                  (exp(softplus(y)) = 1 + e^y)
// two MUFU ops per output dim instead of five.  sC[0][w*P+i] = e^min_d, sC[1][w*P+i] = e^-max_d for the dims of
// warpgroup w (padding dims: 0 and +inf, so they contribute var = 0).
// FUSED: pen2[w][e][row] = sum of var over this warpgroup's dims; rows whose chosen net is e (`mine`, prefetched by the
//        caller) also store (mean, var) to selmv[d][row].
// FULL : mean[e][row][d], lv[e][row][d] = log(var).
// PRECISE (the fp32-accurate kernel): expf / true division / logf instead of the fast-math forms.
template <bool FUSED, bool PRECISE = false>
__device__ __forceinline__ void epi_head(uint32_t tblk, int wg, const HeadShape& s, const float* sC, const HeadOut& a,
                                         int e, int64_t row, bool valid, bool mine) {
  const int P = s.NHEAD / 4, dh = (s.D1 + 1) / 2;
  const int d0 = wg * dh, nd = wg ? s.D1 - dh : dh;
  const float* cmin = sC + wg * P;
  const float* cnmax = sC + MOPO_MAX_D1 + wg * P;
  float pen = 0.f;
  for (int c = 0; c < P / 4; ++c) {
    uint32_t mv[4], rv[4];
    tmem_ld4(tblk + 4 * c, mv);
    tmem_ld4(tblk + P + 4 * c, rv);
    tmem_ld_wait();
    float var[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (PRECISE) var[i] = cmin[4 * c + i] + 1.0f / (cnmax[4 * c + i] + expf(-a.acc_scale * __uint_as_float(rv[i])));
      else var[i] = cmin[4 * c + i] + __fdividef(1.0f, cnmax[4 * c + i] + __expf(-a.acc_scale * __uint_as_float(rv[i])));
      pen += var[i];
    }
    if (FUSED) {
      if (mine) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (4 * c + i < nd) a.selmv[(int64_t)(d0 + 4 * c + i) * a.ldb + row] = make_float2(PRECISE ? a.acc_scale * __uint_as_float(mv[i]) : __uint_as_float(mv[i]), var[i]);
      }
    } else if (valid) {
      float* mrow = a.mean + ((size_t)e * a.ldb + row) * s.D1 + d0;
      float* lrow = a.lv + ((size_t)e * a.ldb + row) * s.D1 + d0;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (4 * c + i < nd) { mrow[4 * c + i] = PRECISE ? a.acc_scale * __uint_as_float(mv[i]) : __uint_as_float(mv[i]); lrow[4 * c + i] = PRECISE ? logf(var[i]) : __logf(var[i]); }
    }
  }
  if (FUSED && valid) a.pen2[((size_t)wg * a.E + e) * a.ldb + row] = pen;
}

// FMA-pipe exp2 and reciprocal for the head warps.  The MUFU is the forward kernel's saturated resource (one tanh per
// hidden activation): head arithmetic that queues behind it holds a tile's head accumulator columns for thousands of
// cycles, so the head uses the (idle) FMA / ALU pipes instead.
//   2^y, |y| <= 60: y = n + f, n = rint(y), 2^f by a degree-5 polynomial on [-0.5, 0.5] (rel. error 3e-6), exponent add.
__device__ __forceinline__ float exp2_fma(float y) {
  y = fminf(fmaxf(y, -60.0f), 60.0f);
  const float t = y + 12582912.0f;                       // 1.5 * 2^23: the integer part lands in the low mantissa bits
  const float f = y - (t - 12582912.0f);
  float p = 1.33335581e-3f;
  p = fmaf(p, f, 9.61812911e-3f);
  p = fmaf(p, f, 5.55041087e-2f);
  p = fmaf(p, f, 2.40226507e-1f);
  p = fmaf(p, f, 6.93147182e-1f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
//   1/d, d in [2^-60, 2^61]: magic-constant seed (12 % error) + 3 Newton steps (error^2 each: 1.4e-2, 2e-4, 4e-8)
__device__ __forceinline__ float rcp_fma(float d) {
  float x = __int_as_float(0x7EF311C7 - __float_as_int(d));
#pragma unroll
  for (int it = 0; it < 3; ++it) x = fmaf(x, fmaf(-d, x, 1.0f), x);
  return x;
}

// Fused head of one row with the column count known at compile time: the accumulator block is
// [w = 0: mean x P | raw x P][w = 1: mean x P | raw x P], P = NHEAD / 4 (see build_image_kernel).  All tensor-memory
// loads are issued before the one wait, the columns are handed back to the MMA warp (h_done) before any arithmetic or
// global store, and the ex2 / rcp chains of all dims overlap.
template <int NHEAD, int D1C, bool PRECISE = false>   // D1C: D1 known at compile time (0 = use s.D1): branch-free, fully interleaved
__device__ __forceinline__ void head_fused(uint32_t tblk, const HeadShape& s, const float* sC, const HeadOut& a,
                                           int e, int64_t row, bool valid, bool mine, uint64_t* done_bar, int lane) {
  constexpr int P = NHEAD / 4;
  uint32_t v[NHEAD];
#pragma unroll
  for (int c = 0; c < NHEAD / 16; ++c) tmem_ld16(tblk + 16 * c, reinterpret_cast<uint32_t (&)[16]>(v[16 * c]));
  tmem_ld_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive_remote(done_bar, 0);
  const int D1 = D1C ? D1C : s.D1;
  const int dh = (D1 + 1) / 2;
  float var[2][P], pen[2] = {0.f, 0.f};
  // padding dims have cmin = 0 and a huge (finite) cnmax: their variance is ~0, no predicate needed
#pragma unroll
  for (int w = 0; w < 2; ++w)
#pragma unroll
    for (int i = 0; i < P; ++i) {
      if constexpr (PRECISE) {
#ifdef MOPO_HEAD_LIBM
        var[w][i] = sC[w * P + i] + 1.0f / (sC[MOPO_MAX_D1 + w * P + i] + expf(-a.acc_scale * __uint_as_float(v[w * 2 * P + P + i])));
#else
        // MUFU forms (ex2.approx / rcp.approx: ~1e-6 relative for |raw logvar| <= 20) -- in the fp32-accurate kernel the MUFU has
        // room (37 % busy) and issue slots do not: libm's expf + an IEEE division were ~20 instructions per output
        float ex, rc;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"((-1.44269504f * a.acc_scale) * __uint_as_float(v[w * 2 * P + P + i])));
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(sC[MOPO_MAX_D1 + w * P + i] + ex));
        var[w][i] = sC[w * P + i] + rc;
#endif
      }
      else var[w][i] = sC[w * P + i] + rcp_fma(sC[MOPO_MAX_D1 + w * P + i] + exp2_fma(-1.44269504f * __uint_as_float(v[w * 2 * P + P + i])));
      pen[w] += var[w][i];
    }
  if (valid) {
    float* pp = a.pen2 + (size_t)e * a.ldb + row;
    pp[0] = pen[0];
    pp[(size_t)a.E * a.ldb] = pen[1];
  }
  if (mine) {
#pragma unroll
    for (int w = 0; w < 2; ++w) {
      float2* dst = a.selmv + (int64_t)(w * dh) * a.ldb + row;
      const int nd = w ? D1 - dh : dh;
#pragma unroll
      for (int i = 0; i < P; ++i) {
        if (i < nd) *dst = make_float2(PRECISE ? a.acc_scale * __uint_as_float(v[w * 2 * P + i]) : __uint_as_float(v[w * 2 * P + i]), var[w][i]);
        dst += a.ldb;
      }
    }
  }
}

}  // namespace tc
