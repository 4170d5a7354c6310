// Pipe-rate microbenchmarks used to size the epilogue (debug entry point, not on the product path):
// how many swish-relevant special-function results per clock an SM retires (tanh.approx.f32 vs the packed
// bf16x2 / f16x2 forms vs ex2), measured with all four sub-partitions saturated.
#include "internal.cuh"

namespace {

template <int KIND>
__device__ __forceinline__ uint32_t op(uint32_t x) {
  uint32_t y;
  if (KIND == 0) asm volatile("tanh.approx.f32 %0, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 1) asm volatile("tanh.approx.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 2) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 3) asm volatile("tanh.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 4) asm volatile("fma.rn.bf16x2 %0, %1, %1, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 5) asm volatile("fma.rn.f32 %0, %1, %1, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 6) asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 7) {         // the fp32-accurate epilogue's element: swish via ex2 + rcp (ens_tc3.cu swish_acc)
    float f = __uint_as_float(x) * 0.015625f, e, r;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.44269504f * f));
    asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    y = __float_as_uint(f * r);
  } else if (KIND == 8) asm volatile("cvt.rn.satfinite.f16x2.f32 %0, %1, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 10) asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x));
  else if (KIND == 11) asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
  else {                        // KIND 9: f16 -> f32 unpack + subtract (the lo part of the split)
    float f;
    asm volatile("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %1;\n\tcvt.f32.f16 %0, l;\n\t}" : "=f"(f) : "r"(x));
    y = __float_as_uint(__uint_as_float(x) - f);
  }
  return y;
}

template <int KIND>
__global__ void __launch_bounds__(512) mufu_bench_kernel(int iters, uint32_t seed, long long* cycles, uint32_t* sink) {
  uint32_t v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = seed + threadIdx.x * 8 + j;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = op<KIND>(v[j]);
  }
  __syncthreads();
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) s ^= v[j];
  if (s == 0x12345678u) sink[0] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

}  // namespace

// ops retired per clock per SM = 512 threads * 8 * iters / cycles   (x2 elements for the packed kinds)
extern "C" __attribute__((visibility("default")))
int mopo_debug_pipe_bench(int kind, int iters, long long* cycles_dev, int n_blocks, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* sink = reinterpret_cast<uint32_t*>(cycles_dev + n_blocks);
  switch (kind) {
    case 0: MOPO_COUNT_LAUNCH, mufu_bench_kernel<0><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 1: MOPO_COUNT_LAUNCH, mufu_bench_kernel<1><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 2: MOPO_COUNT_LAUNCH, mufu_bench_kernel<2><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 3: MOPO_COUNT_LAUNCH, mufu_bench_kernel<3><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 4: MOPO_COUNT_LAUNCH, mufu_bench_kernel<4><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 5: MOPO_COUNT_LAUNCH, mufu_bench_kernel<5><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 6: MOPO_COUNT_LAUNCH, mufu_bench_kernel<6><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 7: MOPO_COUNT_LAUNCH, mufu_bench_kernel<7><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 8: MOPO_COUNT_LAUNCH, mufu_bench_kernel<8><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 9: MOPO_COUNT_LAUNCH, mufu_bench_kernel<9><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 10: MOPO_COUNT_LAUNCH, mufu_bench_kernel<10><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    case 11: MOPO_COUNT_LAUNCH, mufu_bench_kernel<11><<<n_blocks, 512, 0, st>>>(iters, 1u, cycles_dev, sink); break;
    default: mopo_set_error("mopo_debug_pipe_bench: kind=%d", kind); return MOPO_EINVAL;
  }
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}

// ---------------------------------------------------------------------------------------------
// tcgen05.mma issue/execute rate probe: `ksteps` K=16 bf16 MMAs of shape (128*CG) x N, cycling round-robin
// over `nacc` independent accumulators, A from TMEM or shared memory (K-panel, no swizzle), B from shared
// memory.  Reports cycles from first issue to commit arrival (out[0]) and to last issue (out[1]).
// ---------------------------------------------------------------------------------------------
#include "tc_common.cuh"
namespace {
using namespace tc;

template <int CG, int N, int NACC, int A_TMEM>
__global__ void __launch_bounds__(128) mma_rate_kernel(int reps, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int rows = N / CG;                      // B rows held by this CTA
  uint8_t* sB = smem;                               // 26 panels x rows x 16 B
  uint8_t* sA = sB + 26 * 256 * 16;                 // 26 panels x 128 x 16 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sA + 26 * 128 * 16);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid / 32, 0);
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0;
  for (int i = tid; i < (26 * 256 * 16 + 26 * 128 * 16) / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (warp == 0) { tmem_alloc<CG>(tptr, 512); tmem_relinquish<CG>(); }
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = __shfl_sync(0xffffffffu, *tptr, 0);
  {                                                 // zero the A operand columns [408, 512)
    uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int c = 408; c < 512; c += 8) tmem_st8(tbase + ((uint32_t)(warp * 32) << 16) + c, z);
    tmem_st_wait();
    tc_fence_before();
  }
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  tc_fence_after();
  if (rank == 0 && warp == 0) {
    constexpr uint32_t idesc = idesc_bf16(128 * CG, N);
    const uint64_t bd0 = smem_desc(smem_u32(sB), rows * 16, 128), ad0 = smem_desc(smem_u32(sA), 128 * 16, 128);
    constexpr uint32_t bstep = (uint32_t)(2 * rows * 16) >> 4, astep = (uint32_t)(2 * 128 * 16) >> 4;
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 13; ++ks) {
#pragma unroll
          for (int acc = 0; acc < NACC; ++acc) {
            if (A_TMEM) mma_ts<CG>(tbase + acc * N, tbase + 408 + ks * 8, bd0 + (uint64_t)(ks * bstep), idesc, (r | ks) > 0);
            else mma_ss<CG>(tbase + acc * N, ad0 + (uint64_t)(ks * astep), bd0 + (uint64_t)(ks * bstep), idesc, (r | ks) > 0);
          }
        }
      }
      __syncwarp();
    }
    const long long t1 = clock64();
    if (elect_one()) mma_commit<CG>(bar, CG == 2 ? 3 : 1);
    __syncwarp();
    mbar_wait(bar, 0);
    const long long t2 = clock64();
    if (tid == 0) { out[0] = t2 - t0; out[1] = t1 - t0; }
  } else if (tid == 0) {
    mbar_wait(bar, 0);
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  if (warp == 0) tmem_dealloc<CG>(tbase, 512);
}

template <int N, int NACC, int A_TMEM>
int launch_rate(int reps, long long* out_dev, cudaStream_t st) {
  const size_t smem = 26 * 256 * 16 + 26 * 128 * 16 + 64;
  CUDA_TRY(cudaFuncSetAttribute(mma_rate_kernel<2, N, NACC, A_TMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  MOPO_COUNT_LAUNCH;
  CUDA_TRY(cudaLaunchKernelEx(&cfg, mma_rate_kernel<2, N, NACC, A_TMEM>, reps, out_dev));
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}
}  // namespace

// cta_group::2 only; (N, nacc) from a fixed menu; ksteps is rounded down to a multiple of 13 * nacc.
extern "C" __attribute__((visibility("default")))
int mopo_debug_mma_rate(int cta_group, int N, int nacc, int a_tmem, int ksteps, long long* out_dev, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const int reps = ksteps / (13 * (nacc > 0 ? nacc : 1));
#define RATE_CASE(n, k)                                                                    \
  if (cta_group == 2 && N == n && nacc == k)                                               \
    return a_tmem ? launch_rate<n, k, 1>(reps, out_dev, st) : launch_rate<n, k, 0>(reps, out_dev, st);
  RATE_CASE(208, 1) RATE_CASE(112, 1) RATE_CASE(96, 1) RATE_CASE(48, 1) RATE_CASE(256, 1)
  RATE_CASE(96, 4) RATE_CASE(48, 4) RATE_CASE(112, 2) RATE_CASE(208, 1)
#undef RATE_CASE
  mopo_set_error("mopo_debug_mma_rate: unsupported combination cg=%d N=%d nacc=%d", cta_group, N, nacc);
  return MOPO_EINVAL;
}

// ---------------------------------------------------------------------------------------------
// Interference probe: the MMA loop of mma_rate_kernel (cta_group::2, one accumulator, K = 16 steps back to back) while
// `n_warps` other warps of the same CTAs run an epilogue-like loop:
//   imode 0: idle; 1: ex2 + rcp chains in registers only (MUFU); 2: tcgen05.ld x16 + tcgen05.st x16 on TMEM columns the MMA
//   does not touch; 3: the fp32-accurate kernel's whole group body (ld, swish via ex2 + rcp, fp16 hi/lo split, st).
// out[0] = cycles of the MMA loop (to commit), out[1] = issue cycles, out[2] = group iterations the interferer warps of
// the leader CTA completed meanwhile, out[3] = the same warps' iterations per 1000 cycles when run WITHOUT the MMA loop.
// Answers whether the tensor pipe and the epilogue slow each other down (TMEM ports / MIO), which a timeline cannot.
// ---------------------------------------------------------------------------------------------
namespace {
using namespace tc;

template <int IMODE>
__device__ __forceinline__ void interferer_iter(uint32_t taddr, uint32_t (&v)[16]) {
  if (IMODE == 2 || IMODE == 3) { tmem_ld16(taddr, v); tmem_ld_wait(); }
  if (IMODE == 1 || IMODE == 3) {
    uint32_t o[16];
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      float s[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float x = 0.015625f * __uint_as_float(v[2 * p + u]);
        float e, r;
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.44269504f * x));
        asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
        s[u] = x * r;
      }
      const uint32_t hi = pack_f16x2(s[0], s[1]);
      o[p] = hi;
      o[8 + p] = pack_f16x2(s[0] - f16lo_to_f32(hi), s[1] - f16hi_to_f32(hi));
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = o[i] | 0x3c003c00u;      // keep the values finite and moving
  }
  if (IMODE == 2 || IMODE == 3) { tmem_st16(taddr, v); tmem_st_wait(); }
}

template <int N, int A_TMEM, int IMODE>
__global__ void __launch_bounds__(640) mma_interf_kernel(int reps, int with_mma, int icol, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int rows = N / 2;
  uint8_t* sB = smem;
  uint8_t* sA = sB + 26 * 256 * 16;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sA + 26 * 128 * 16);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 8);
  volatile int* stop = reinterpret_cast<volatile int*>(tptr + 1);
  int* iters = const_cast<int*>(stop) + 1;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid / 32, 0), lane = tid % 32;
  const uint32_t rank = cluster_ctarank();
  for (int i = tid; i < (26 * 256 * 16 + 26 * 128 * 16) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (warp == 0) { tmem_alloc<2>(tptr, 512); tmem_relinquish<2>(); }
  if (tid == 0) { mbar_init(bar, 1); for (int i = 4; i < 8; ++i) mbar_init(bar + i, (1 << 20) - 1); fence_barrier_init(); *stop = 0; *iters = 0; }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = __shfl_sync(0xffffffffu, *tptr, 0);
  if (warp < 4) {
    uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int c = 112; c < 512; c += 8) tmem_st8(tbase + ((uint32_t)(warp * 32) << 16) + c, z);
    tmem_st_wait();
    tc_fence_before();
  }
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  if (warp == 0) {
    if (rank == 0) {
      constexpr uint32_t idesc = idesc_bf16(256, N);
      const uint64_t bd0 = smem_desc(smem_u32(sB), rows * 16, 128), ad0 = smem_desc(smem_u32(sA), 128 * 16, 128);
      constexpr uint32_t bstep = (uint32_t)(2 * rows * 16) >> 4, astep = (uint32_t)(2 * 128 * 16) >> 4;
      const long long t0 = clock64();
      if (with_mma) {
        for (int r = 0; r < reps; ++r) {
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 13; ++ks) {
              if (A_TMEM) mma_ts<2>(tbase, tbase + 408 + ks * 8, bd0 + (uint64_t)(ks * bstep), idesc, (r | ks) > 0);
              else mma_ss<2>(tbase, ad0 + (uint64_t)(ks * astep), bd0 + (uint64_t)(ks * bstep), idesc, (r | ks) > 0);
            }
          }
          __syncwarp();
        }
      } else {
        while (clock64() - t0 < (long long)reps * 13 * (N / 2 + 3)) { }
      }
      const long long t1 = clock64();
      if (elect_one()) mma_commit<2>(bar, 3);
      __syncwarp();
      mbar_wait(bar, 0);
      const long long t2 = clock64();
      if (lane == 0) { out[0] = t2 - t0; out[1] = t1 - t0; }
    } else {
      mbar_wait(bar, 0);
    }
    if (lane == 0) *stop = 1;
  } else if (warp >= 4) {
    const uint32_t taddr = tbase + ((uint32_t)((warp & 3) * 32) << 16) + icol + 16 * ((warp - 4) >> 2);
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = 0x3c003c00u + tid + i;
    int n = 0;
    if (IMODE == 4) {
      // the fp32-accurate kernel's epilogue control flow: per "layer" a few rounds on consecutive groups, the next group's
      // tcgen05.ld in flight during the current group's conversion, and per round wait::st + fence + an mbarrier arrive
      uint32_t w[16];
      uint64_t* rbar = bar + 4 + (warp & 3);                 // nobody waits on it: arrivals only
      while (!*stop) {
        tmem_ld16(taddr, v);
#pragma unroll 1
        for (int r = 0; r < 4; r += 2) {
          tmem_ld_wait();
          tmem_ld16(taddr + 64 * ((r + 1) & 1), w);
          interferer_iter<1>(taddr, v);
          tmem_st16(taddr, v);
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(rbar, 0);
          tmem_ld_wait();
          if (r + 2 < 4) tmem_ld16(taddr, v);
          interferer_iter<1>(taddr, w);
          tmem_st16(taddr + 64, w);
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(rbar, 0);
          n += 2;
        }
      }
    } else {
      while (!*stop) { interferer_iter<IMODE>(taddr, v); ++n; }
    }
    uint32_t sink = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) sink ^= v[i];
    if (sink == 0x12345u) out[7] = sink;
    if (lane == 0 && rank == 0) atomicAdd(iters, n);
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0 && rank == 0) out[2] = *iters;
  cluster_sync_all();
  if (warp == 0) tmem_dealloc<2>(tbase, 512);
}

template <int N, int A_TMEM, int IMODE>
int launch_interf(int reps, int with_mma, int n_warps, int icol, long long* out_dev, cudaStream_t st) {
  const size_t smem = 26 * 256 * 16 + 26 * 128 * 16 + 128;
  CUDA_TRY(cudaFuncSetAttribute(mma_interf_kernel<N, A_TMEM, IMODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2); cfg.blockDim = dim3(128 + 32 * n_warps); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  MOPO_COUNT_LAUNCH;
  CUDA_TRY(cudaLaunchKernelEx(&cfg, mma_interf_kernel<N, A_TMEM, IMODE>, reps, with_mma, icol, out_dev));
  CUDA_TRY(cudaGetLastError());
  return MOPO_OK;
}
}  // namespace

// N in {112, 96}; imode 0..3; n_warps 0..16 interferer warps per CTA; with_mma 0 = the interferers alone for the same duration
extern "C" __attribute__((visibility("default")))
int mopo_debug_mma_interference(int N, int a_tmem, int ksteps, int imode, int n_warps, int with_mma, int icol, long long* out_dev, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const int reps = ksteps / 13;
  if (icol < N || icol + 16 * ((n_warps + 3) / 4) > 408 || (icol & 3)) { mopo_set_error("mopo_debug_mma_interference: icol=%d", icol); return MOPO_EINVAL; }
  if (n_warps < 0 || n_warps > 16) { mopo_set_error("mopo_debug_mma_interference: n_warps=%d", n_warps); return MOPO_EINVAL; }
#define INTERF_CASE(n, t, m) if (N == n && a_tmem == t && imode == m) return launch_interf<n, t, m>(reps, with_mma, n_warps, icol, out_dev, st);
  INTERF_CASE(112, 1, 0) INTERF_CASE(112, 1, 1) INTERF_CASE(112, 1, 2) INTERF_CASE(112, 1, 3)
  INTERF_CASE(96, 1, 0) INTERF_CASE(96, 1, 1) INTERF_CASE(96, 1, 2) INTERF_CASE(96, 1, 3)
  INTERF_CASE(112, 0, 0) INTERF_CASE(112, 0, 3) INTERF_CASE(112, 1, 4) INTERF_CASE(96, 1, 4)
#undef INTERF_CASE
  mopo_set_error("mopo_debug_mma_interference: unsupported combination N=%d a_tmem=%d imode=%d", N, a_tmem, imode);
  return MOPO_EINVAL;
}
