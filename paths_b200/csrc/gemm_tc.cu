// placeholder until the tcgen05 engine lands: reports "unsupported" so every contraction takes the fp32 SIMT path
#include "common.cuh"
namespace paths {
bool gemm_tc_supported_nt(const RowSrc&, int, int, int, int, int) { return false; }
bool gemm_tc_supported_tn(int, const RowSrc&, int, int, int) { return false; }
int gemm_nt_tc(int, const RowSrc&, const float*, int, int, float*, int, const float*, int, const int*, int, int, int, cudaStream_t) {
  set_error("tcgen05 path not built"); return -1;
}
int gemm_tn_tc(int, const float*, int, const RowSrc&, float*, int, int, const int*, int, int, cudaStream_t) {
  set_error("tcgen05 path not built"); return -1;
}
}  // namespace paths
