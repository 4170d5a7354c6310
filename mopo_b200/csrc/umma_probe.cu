// Known-answer probe for the tcgen05 conventions the ensemble kernel relies on (K-panel smem layout,
// descriptors, TS-form A operand packing, cta_group::2 operand split and commit multicast).
// D[M,N] = A[M,K] * B[N,K]^T with bf16 inputs and fp32 accumulation; tests/test_gpu_umma.py compares
// against NumPy.  Debug entry point (mopo_debug_umma_probe), not part of the product path.
#include "internal.cuh"
#include "tc_common.cuh"

namespace {

using namespace tc;

template <int CG, int A_TMEM>
__global__ void __launch_bounds__(128) umma_probe_kernel(const uint16_t* __restrict__ A, const uint16_t* __restrict__ B,
                                                        float* __restrict__ D, int N, int K) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int Kp = (K + 15) / 16 * 16;
  const int Nloc = N / CG;
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0;
  uint8_t* sA = smem;                                  // [Kp/8][128][16B]
  uint8_t* sB = sA + (Kp / 8) * 128 * 16;              // [Kp/8][Nloc][16B]
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + (Kp / 8) * Nloc * 16);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid / 32;
  const int tile = blockIdx.x / CG;
  const int row0 = tile * 128 * CG + rank * 128;

  if (warp == 0) { tmem_alloc<CG>(tptr, 512); tmem_relinquish<CG>(); }
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  for (int i = tid; i < (Kp / 8) * 128; i += 128) {    // A panels
    int kc = i / 128, r = i % 128;
    uint16_t v[8];
    for (int j = 0; j < 8; ++j) { int k = kc * 8 + j; v[j] = k < K ? A[(size_t)(row0 + r) * K + k] : 0; }
    *reinterpret_cast<uint4*>(sA + (size_t)kc * 128 * 16 + r * 16) = *reinterpret_cast<uint4*>(v);
  }
  for (int i = tid; i < (Kp / 8) * Nloc; i += 128) {   // B panels: this CTA's N-slice
    int kc = i / Nloc, r = i % Nloc;
    uint16_t v[8];
    for (int j = 0; j < 8; ++j) { int k = kc * 8 + j; v[j] = k < K ? B[(size_t)(rank * Nloc + r) * K + k] : 0; }
    *reinterpret_cast<uint4*>(sB + (size_t)kc * Nloc * 16 + r * 16) = *reinterpret_cast<uint4*>(v);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tptr;
  const uint32_t lane_addr = tbase + ((uint32_t)(warp * 32) << 16);
  if (A_TMEM) {                                        // row tid -> TMEM lane tid, columns 256 + k/2
    for (int kc = 0; kc < Kp / 16; ++kc) {
      uint32_t p[8];
      const uint4 lo = *reinterpret_cast<uint4*>(sA + (size_t)(2 * kc) * 128 * 16 + tid * 16);
      const uint4 hi = *reinterpret_cast<uint4*>(sA + (size_t)(2 * kc + 1) * 128 * 16 + tid * 16);
      p[0] = lo.x; p[1] = lo.y; p[2] = lo.z; p[3] = lo.w; p[4] = hi.x; p[5] = hi.y; p[6] = hi.z; p[7] = hi.w;
      tmem_st8(lane_addr + 256 + kc * 8, p);
    }
    tmem_st_wait();
    tc_fence_before();
  }
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  tc_fence_after();
  if (rank == 0 && tid == 0) {
    const uint32_t idesc = idesc_bf16(128 * CG, N);
    for (int ks = 0; ks < Kp / 16; ++ks) {
      const uint64_t bd = smem_desc(smem_u32(sB) + ks * 2 * Nloc * 16, Nloc * 16, 128);
      if (A_TMEM) mma_ts<CG>(tbase, tbase + 256 + ks * 8, bd, idesc, ks > 0);
      else mma_ss<CG>(tbase, smem_desc(smem_u32(sA) + ks * 2 * 128 * 16, 128 * 16, 128), bd, idesc, ks > 0);
    }
    mma_commit<CG>(bar, 3);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  for (int c = 0; c < N / 16; ++c) {
    uint32_t v[16];
    tmem_ld16(lane_addr + c * 16, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) D[(size_t)(row0 + tid) * N + c * 16 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  if (warp == 0) tmem_dealloc<CG>(tbase, 512);
}

template <int CG, int A_TMEM>
int launch_probe(const uint16_t* A, const uint16_t* B, float* D, int M, int N, int K, cudaStream_t st) {
  const int Kp = (K + 15) / 16 * 16;
  const size_t smem = (size_t)(Kp / 8) * 128 * 16 + (size_t)(Kp / 8) * (N / CG) * 16 + 64;
  CUDA_TRY(cudaFuncSetAttribute(umma_probe_kernel<CG, A_TMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(M / 128);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  mopo_count_launch();
  CUDA_TRY(cudaLaunchKernelEx(&cfg, umma_probe_kernel<CG, A_TMEM>, A, B, D, N, K));
  return MOPO_OK;
}

}  // namespace

extern "C" __attribute__((visibility("default")))
int mopo_debug_umma_probe(int cta_group, int a_in_tmem, const uint16_t* A, const uint16_t* B, float* D,
                          int M, int N, int K, void* stream) {
  if (!(cta_group == 1 || cta_group == 2) || M % (128 * cta_group) || N % (8 * cta_group) || N > 256 || K < 1 || K > 256) {
    mopo_set_error("mopo_debug_umma_probe: bad shape M=%d N=%d K=%d cg=%d", M, N, K, cta_group);
    return MOPO_EINVAL;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (cta_group == 1) return a_in_tmem ? launch_probe<1, 1>(A, B, D, M, N, K, st) : launch_probe<1, 0>(A, B, D, M, N, K, st);
  return a_in_tmem ? launch_probe<2, 1>(A, B, D, M, N, K, st) : launch_probe<2, 0>(A, B, D, M, N, K, st);
}
