// tcgen05 bf16 ensemble forward (placeholder until the kernel lands).
#include "internal.cuh"
int tc_supported(const mopo_ens* h) { (void)h; return 0; }
int tc_build_images(mopo_ens* h, cudaStream_t st) { (void)h; (void)st; return MOPO_OK; }
int tc_forward(mopo_ens* h, const float*, int64_t, const int*, float*, float*, int64_t, cudaStream_t) {
  (void)h; mopo_set_error("tcgen05 path not built"); return MOPO_EUNSUPPORTED;
}
