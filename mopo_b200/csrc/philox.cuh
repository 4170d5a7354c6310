// Counter-based Philox4x32-10 (Salmon et al., SC'11) + Box-Muller; keyed by (seed), counter =
// (row id, step id, draw block, stream).  Replaces the reference's global MT19937 draws
// (np.random.normal at fake_env.py:72, np.random.choice at bnn.py:343); restated in NumPy in
// oracle/philox.py: elite picks are checked bit-exactly, normals to 1e-5 (approximate transcendentals).
#pragma once
#include <stdint.h>

#define MOPO_STREAM_ELITE 0u
#define MOPO_STREAM_NOISE 1u
#define MOPO_STREAM_POOL  2u
#define MOPO_STREAM_POLICY 3u   /* the policy's action noise (mopo.py:307) / uniform actions of rollout_random (mopo.py:738) */
#define MOPO_STREAM_SAC 4u      // policy noise inside the SAC update (rows 0..n-1: pi(s), n..2n-1: pi(s'))

struct Philox4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ uint32_t mopo_mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

__host__ __device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                          uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = mopo_mulhi32(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = mopo_mulhi32(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  return Philox4{c0, c1, c2, c3};
}

// uniform in (0,1): 24 high bits + half an ulp
__host__ __device__ __forceinline__ float mopo_u01(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

// one replay-pool index draw (np.random.randint(0, size), flexible_replay_pool.py:85-87): counter = (row i of the draw,
// the 64-bit id of the draw) -- draws with different ids (mopo_pool_sample's `offset`; the SAC update index and pool) are
// disjoint counter planes, i.e. independent.  index = floor(u64 * size / 2^64) with u64 = (x << 32) | y.
__host__ __device__ __forceinline__ Philox4 philox_pool_draw(uint32_t i, uint64_t draw, uint64_t seed) {
  return philox4x32_10(i, (uint32_t)draw, (uint32_t)(draw >> 32), MOPO_STREAM_POOL, (uint32_t)seed, (uint32_t)(seed >> 32));
}

#ifdef __CUDACC__
// 4 standard normals from one Philox block (two Box-Muller pairs).
__device__ __forceinline__ void philox_normal4(uint32_t row, uint32_t step, uint32_t blk, uint64_t seed, float out[4],
                                               uint32_t stream = MOPO_STREAM_NOISE) {
  Philox4 r = philox4x32_10(row, step, blk, stream, (uint32_t)seed, (uint32_t)(seed >> 32));
  float u0 = mopo_u01(r.x), u1 = mopo_u01(r.y), u2 = mopo_u01(r.z), u3 = mopo_u01(r.w);
  // fast intrinsics (lg2 / sin / cos MUFU forms): arguments stay in (0,1) and [0, 2*pi), where their absolute error
  // is ~1e-6 -- far below the sampling resolution of a 24-bit uniform
  float r0 = sqrtf(-2.0f * __logf(u0)), r1 = sqrtf(-2.0f * __logf(u2));
  float s0, c0, s1, c1;
  __sincosf(6.283185307179586f * u1, &s0, &c0);
  __sincosf(6.283185307179586f * u3, &s1, &c1);
  out[0] = r0 * c0; out[1] = r0 * s0; out[2] = r1 * c1; out[3] = r1 * s1;
}
// index into the elite list: floor(u * n) via mulhi (bias < n / 2^32)
__device__ __forceinline__ uint32_t philox_elite(uint32_t row, uint32_t step, uint64_t seed, uint32_t n) {
  Philox4 r = philox4x32_10(row, step, 0u, MOPO_STREAM_ELITE, (uint32_t)seed, (uint32_t)(seed >> 32));
  return __umulhi(r.x, n);
}
#endif
