/* mopo_b200 -- developer probes.  NOT part of the product ABI: these entry points exist only in libmopo_b200_dbg.so
 * (python -m mopo_b200.build --probes: the library's sources plus umma_probe.cu / microbench.cu, -DMOPO_DEBUG_PROBES).
 * tests/test_gpu_umma.py (known-answer tests of the tcgen05 conventions) and tools/ use them. */
#ifndef MOPO_B200_DEBUG_H
#define MOPO_B200_DEBUG_H
#include "mopo_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* tcgen05 known-answer probe: D[M,N] = A[M,K] * B[N,K]^T (bf16 bit patterns in, fp32 out; device pointers),
 * cta_group 1|2, A operand from shared memory (0) or tensor memory (1).  Used by tests/test_gpu_umma.py. */
int mopo_debug_umma_probe(int cta_group, int a_in_tmem, const uint16_t* A, const uint16_t* B, float* D,
                          int M, int N, int K, void* stream);

/* Timeline probe of the tcgen05 forward: runs mopo_ens_forward(bf16) once with per-role clock64 stamps of CTA
 * pair 0; trace = device uint64[5*2048] (role-major: 0 MMA issuer, 1/2 hidden-epilogue warpgroups, 3 input loader, 4 head warps),
 * entry = tag << 48 | clock.  tools/tc_timeline.py decodes it. */
int mopo_debug_tc_trace(mopo_ens* h, const float* xn, int64_t B, float* mean, float* logvar,
                        unsigned long long* trace, void* stream);
/* The same for the streamed-weights kernel (ens_tc3.cu): terms = 3 (fp32-accurate) or 1 (bf16 mode). */
int mopo_debug_tc3_trace(mopo_ens* h, const float* xn, int64_t B, float* mean, float* logvar, int terms,
                         unsigned long long* trace, void* stream);
/* Pipe-rate microbenchmark: kind 0 tanh.approx.f32, 1 tanh.approx.bf16x2, 2 ex2.approx.f32, 3 tanh.approx.f16x2,
 * 4 fma.bf16x2, 5 fma.f32; cycles_dev = device int64[n_blocks + 1] (per-block clock64 delta; last = scratch). */
int mopo_debug_pipe_bench(int kind, int iters, long long* cycles_dev, int n_blocks, void* stream);
/* tcgen05.mma rate probe: ksteps K=16 bf16 MMAs of (128*cta_group) x N over nacc round-robin accumulators, A from
 * TMEM (1) or shared memory (0); out_dev = device int64[2]: cycles to commit arrival, cycles to last issue. */
int mopo_debug_mma_rate(int cta_group, int N, int nacc, int a_tmem, int ksteps, long long* out_dev, void* stream);

/* Interference probe: the cta_group::2 MMA loop (N = 112 | 96, ksteps K=16 steps into one accumulator) while n_warps
 * other warps per CTA run an epilogue-like loop: imode 0 idle, 1 ex2 + rcp in registers, 2 tcgen05.ld/st x16, 3 the
 * fp32-accurate kernel's group body (ld, swish, fp16 hi/lo split, st).  with_mma = 0 runs the interferers alone for the
 * same nominal duration.  out_dev = device int64[8]: [0] MMA cycles to commit, [1] issue cycles, [2] interferer group
 * iterations completed by the leader CTA's warps. */
int mopo_debug_mma_interference(int N, int a_tmem, int ksteps, int imode, int n_warps, int with_mma, int icol, long long* out_dev, void* stream);   /* icol: first TMEM column of the interferers' groups (the accumulator is [0, N), the A operand [408, 512)) */

/* Stage timeline of the one-kernel SAC update (csrc/sac.cu): once set, every update writes CTA 0's %globaltimer at kernel
 * start, after each stage's grid barrier and at the kernel's end into buf (device, >= 64 entries); NULL switches it off. */
int mopo_debug_sac_trace(mopo_sac* s, unsigned long long* buf);

#ifdef __cplusplus
}
#endif
#endif
