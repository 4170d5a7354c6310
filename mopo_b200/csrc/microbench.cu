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
  else asm volatile("fma.rn.f32 %0, %1, %1, %1;" : "=r"(y) : "r"(x));
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
