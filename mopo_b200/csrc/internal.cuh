// Internal declarations shared by the translation units of libmopo_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/mopo_b200.h"
#ifdef MOPO_DEBUG_PROBES
#include "../../include/mopo_b200_debug.h"
#endif

#define MOPO_MAX_LAYERS 8
#define MOPO_MAX_D1 64
#define MOPO_MAX_E 16

void mopo_set_error(const char* fmt, ...);
void mopo_count_launch();
#define MOPO_COUNT_LAUNCH mopo_count_launch()

#define CUDA_TRY(expr)                                                                  \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      (void)cudaGetLastError(); /* clear the (non-sticky) error so that later calls do not report it again */ \
      mopo_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return MOPO_ECUDA;                                                                \
    }                                                                                   \
  } while (0)

#define MOPO_API extern "C" __attribute__((visibility("default")))

#define REQUIRE(cond, code, ...)        \
  do {                                  \
    if (!(cond)) {                      \
      mopo_set_error(__VA_ARGS__);      \
      return code;                      \
    }                                   \
  } while (0)

#define TRY(expr)              \
  do {                         \
    int _r = (expr);           \
    if (_r != MOPO_OK) return _r; \
  } while (0)

// Device buffer that only grows.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return MOPO_OK;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 4;
    CUDA_TRY(cudaMalloc(&p, want));
    cap = want;
    return MOPO_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Inputs that are still arriving over PCIe while the kernel runs (mopo_rollout_host): chunk c covers the 256-row tiles
// [end_tile[c-1], end_tile[c]) and is complete once flags[c] == value (a stream-ordered memset behind the chunk's copies).
struct InputReady { const uint32_t* flags = nullptr; uint32_t value = 0; int n = 0; int end_tile[8] = {}; };

struct mopo_train;   // optimiser state while BNN.train runs (train.cu)

struct mopo_ens {
  mopo_ens_cfg cfg;
  int E, in, D1, H, L, HO;   // HO = 2*D1: width of the fused [mean | raw logvar] head
  bool params_set = false, elites_set = false;
  int sm_count = 148;

  // fp32 images (reference layout): hidden layers l = 0..L-1, fused head at index L.
  float* W[MOPO_MAX_LAYERS] = {};   // [E, K_l, N_l]
  float* b[MOPO_MAX_LAYERS] = {};   // [E, N_l]
  float *mu = nullptr, *sigma = nullptr, *max_lv = nullptr, *min_lv = nullptr;
  int32_t* elites = nullptr;        // device [num_elites]
  int n_elites = 0;

  // bf16 UMMA-ready images for the tcgen05 path (built by tc_build_images).
  void* tc_img = nullptr;
  size_t tc_img_bytes = 0;
  // (hi, lo) bf16 images for the fp32-accurate tcgen05 path (built by tc3_build_images).
  void* tc3_img = nullptr;
  size_t tc3_img_bytes = 0;
  // per-launch CUDA-event timing of the forward kernel (mopo_kernel_timing): events on the launching stream
  bool timing = false, timing_tail = false;
  std::vector<cudaEvent_t> ev;      // start/stop pairs of the launches since the last query
  size_t ev_used = 0;
  void* tc_trace = nullptr;         // device uint64[4][2048] timeline buffer while mopo_debug_tc_trace runs

  // scratch
  DevBuf xn, xtiles, actA, actB, mean, lv, blk_counts, misc;
  DevBuf sel, selmv, pen2;         // fused-head step outputs
  DevBuf ro_obs[2], ro_origin[2], ro_counts, ro_act;
  mopo_train* tr = nullptr;
  // mopo_rollout_host: device staging of host inputs, the upload stream and its per-step events
  DevBuf st_obs, st_act;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_start = nullptr;
  std::vector<cudaEvent_t> act_ev;
  // mopo_rollout_host on the streamed-weights kernel: inputs arrive in chunks behind flags the forward polls (ens_tc3.cu)
  DevBuf ready_flags; uint32_t ready_epoch = 0; std::vector<InputReady> host_ready; bool host_ready_on = false;
  uint32_t* flag_src = nullptr;
  int (*after_first_forward)(void*) = nullptr; void* after_first_forward_ctx = nullptr;   // the rest of the uploads, enqueued once step 0's forward is launched
};
void mopo_train_release(mopo_ens* h);

struct mopo_pool {
  int obs_dim, act_dim, device;
  int64_t max_size, size, pointer;
  float *actions, *rewards, *obs, *next_obs;
  uint8_t* terminals;
};

// ---- fp32 FFMA forward (ens_fp32.cu) --------------------------------------------------------
// xn [n_max,in] -> mean [E,ldb,D1], lv [E,ldb,D1]; rows >= *d_n (if d_n) are skipped.
// x_net_stride (floats): 0 = all nets read the same xn rows; n_max * in = per-net inputs xn [E, n_max, in].
int fp32_forward(mopo_ens* h, const float* xn, int64_t n_max, const int* d_n, float* mean, float* lv,
                 int64_t ldb, cudaStream_t st, int64_t x_net_stride = 0);

// ---- tcgen05 bf16 forward (ens_tc.cu) ---------------------------------------------------------
int tc_supported(const mopo_ens* h);
int tc_build_images(mopo_ens* h, cudaStream_t st);
int tc_forward(mopo_ens* h, const float* xn, int64_t n_max, const int* d_n, float* mean, float* lv,
               int64_t ldb, cudaStream_t st);
// Pieces of the tensor-core step: input tiling (+ scaler + elite pick) into h->xtiles, then the forward kernel with
// either the full [E,B,D1] outputs or the fused-head outputs (chosen net's mean/var + per-net variance sums).
struct TcPrep {
  const float *obs, *act; int64_t n_max; const int* d_n;
  int32_t* sel; const int32_t* model_inds; uint64_t seed, step_id; const int32_t* origin; int64_t origin_base;
};
struct TcOut {
  float *mean, *lv; int64_t ldb;                                   // full
  const int32_t* sel; float2* selmv; float* pen2;                  // fused (sel != nullptr): selmv [D1, ldb], pen2 [2, E, ldb]
  bool pdl;                                                        // launch programmatically behind the previous grid (rollout chain)
};
int tc_reserve_xtiles(mopo_ens* h, int64_t n_max, int* K0);
int tc_tiles_from_xn(mopo_ens* h, const float* xn, int64_t n_max, const int* d_n, cudaStream_t st);
int tc_tiles_from_step(mopo_ens* h, const TcPrep& q, cudaStream_t st);
int tc_launch(mopo_ens* h, int64_t n_max, const int* d_n, const TcOut& o, cudaStream_t st);

// ---- tcgen05 split-bf16 forward at fp32 accuracy (ens_tc3.cu) -------------------------------------
// terms = 3: every operand as a bf16 (hi, lo) pair, three MMAs per K step (MOPO_PREC_FP32); terms = 1: plain bf16 with
// streamed weights (MOPO_PREC_BF16 for the shapes tc_supported rejects).  The kernel reads the raw rows itself.
int tc3_supported(const mopo_ens* h);
int tc3_build_images(mopo_ens* h, cudaStream_t st);
struct Tc3In {
  const float *obs, *act; int od, ad;     // x = concat(obs [n, od], act [n, ad]); ad == 0: obs is the whole input row
  bool scale;                             // apply the handle's scaler (false: x is already scaled)
  int64_t net_stride = 0;                 // rows between consecutive nets' inputs (0: every net reads the same rows)
  InputReady ready;                       // flags == nullptr: the inputs are complete when the kernel starts
};
int tc3_pick(mopo_ens* h, int32_t* sel, const int32_t* model_inds, int64_t n_max, const int* d_n, uint64_t seed,
             uint64_t step_id, const int32_t* origin, int64_t origin_base, cudaStream_t st);
int tc3_launch(mopo_ens* h, int terms, const Tc3In& x, int64_t n_max, const int* d_n, const TcOut& o, cudaStream_t st);

// ---- step tail / pool / compaction (tail.cu) ---------------------------------------------------
struct TailArgs {
  const float *mean, *lv;      // [E,ldb,D1]
  int64_t ldb;
  const float *obs, *act;      // [n,obs] [n,act]
  int64_t n_max; const int* d_n;
  int E, D1, obs_dim, act_dim;
  int term_kind, penalty_kind; float coeff; int deterministic;
  const float* noise; const int32_t* model_inds; const int32_t* elites; int n_elites;
  uint64_t seed, step_id; const int32_t* origin; int64_t origin_base;
  float *next_obs, *rewards; uint8_t* terminals;
  float *penalty, *unpenalized, *model_means, *model_stds, *log_prob, *dev; int32_t* model_inds_out;
  // fused-head inputs (instead of mean / lv): chosen net, its (mean, variance) [D1,ldb], per-net variance sums [2,E,ldb]
  const int32_t* sel; const float2* selmv; const float* pen2;
  // optional (fused tail, static rows): also write the NEXT step's input tiles + elite picks (see step_tail_fused_kernel)
  uint4* xt_next; const float* act_next; const float* mu; const float* sigma; int K0;
  int32_t* sel_next; const int32_t* model_inds_next; uint64_t step_id_next;
  bool pdl;                    // launch programmatically behind the forward kernel (rollout chain): see step_tail_fused_kernel
  // optional pool append: row j -> (pool_base + *d_pool_off + j) % pool_max
  float *p_act, *p_rew, *p_obs, *p_nobs; uint8_t* p_term; int64_t pool_max, pool_base; const int64_t* d_pool_off;
  // optional compaction bookkeeping
  int* blk_counts;             // [ceil(n_max/TAIL_ROWS)] survivors per block
  InputReady ready;            // obs / act may still be in flight when a programmatic launch starts this grid (flags == nullptr: no)
  bool obs_behind;             // obs is the previous step's tail output and this grid is launched programmatically: read it behind the wait
};
int launch_prep(const float* obs, const float* act, const float* mu, const float* sigma, float* xn,
                int64_t n_max, const int* d_n, int obs_dim, int act_dim, cudaStream_t st);
int launch_tail(const TailArgs& a, cudaStream_t st);
int tail_rows_per_block();
// survivors of step i -> dense obs/origin for step i+1; writes d_n_next and advances d_pool_off by n.
// optional second output of the compaction: the next step's input tiles + elite picks (xt == nullptr: off)
struct CompactTiles {
  uint4* xt; const float* act_next; const float *mu, *sigma; int act_dim, K0;
  int32_t* sel; const int32_t* model_inds; const int32_t* elites; int n_elites; uint64_t seed, step_id;
};
int launch_compact(const float* next_obs, const uint8_t* term, const int32_t* origin_in, int64_t origin_base,
                   int64_t n_max, const int* d_n, const int* blk_counts, int obs_dim,
                   float* obs_out, int32_t* origin_out, int* d_n_next, int64_t* d_pool_off, const CompactTiles& ta,
                   cudaStream_t st);
int launch_term_only(int kind, const float* next_obs, int64_t n, int obs_dim, uint8_t* out, cudaStream_t st);
int launch_exp(const float* lv, float* var, int64_t n, cudaStream_t st);

// ---- policy (policy.cu): the SAC actor inside the rollout (mopo.py:292-312, 468-485) --------------------------
struct mopo_policy {
  int obs_dim = 0, act_dim = 0, H = 0, device = 0, sm_count = 148;
  bool params_set = false;
  float* W[4] = {};            // fp32, tf.layers.dense layout [in,out]: hidden 0, hidden 1, mu, log_std
  float* b[4] = {};
  void* tc_img = nullptr;      // bf16 UMMA-ready image (policy_tc_kernel)
  size_t tc_img_bytes = 0;
  DevBuf h0, h1;               // fp32 path activations
};
struct PolicyRun {
  const float* obs; int64_t n_max; const int* d_n;
  int deterministic, random_uniform, precision;
  const float* noise; uint64_t seed, step_id; const int32_t* origin; int64_t origin_base;
  float *actions, *mu_out, *std_out;
  // optional (tensor-core path): also be the ensemble step's front end -- concat(obs, act), scaler, bf16 K-panel input
  // tiles and the per-row elite pick, exactly as prep_tiles_kernel (then the step needs no tiling pass)
  uint4* xt; const float *ens_mu, *ens_sigma; int ens_K0;
  int32_t* sel; const int32_t* model_inds; const int32_t* elites; int n_elites; uint64_t ens_step_id;
  bool pdl;                    // launch programmatically behind the previous step's tail (rollout chain)
};
int policy_run(mopo_policy* p, const PolicyRun& r, cudaStream_t st);
int policy_tc_supported(const mopo_policy* p);
int policy_tc_build_image(mopo_policy* p, cudaStream_t st);
int policy_tc_launch(mopo_policy* p, const PolicyRun& r, cudaStream_t st);
