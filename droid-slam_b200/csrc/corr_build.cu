// All-pairs correlation volume + 4-level pyramid (CorrBlock build) -- tcgen05/TMEM GEMM.
// Placeholder until the tcgen05 kernel lands: reports NOT_IMPLEMENTED so that callers fail loudly.
#include "common.cuh"

extern "C" size_t droid_corr_build_workspace_bytes(int B, int C, int H, int W) {
  (void)B; (void)C; (void)H; (void)W;
  return 256;
}

extern "C" int droid_corr_build(const void* fmap1, const void* fmap2, void* lvl0, void* lvl1,
                                void* lvl2, void* lvl3, int B, int C, int H, int W, void* workspace,
                                size_t workspace_bytes, void* stream) {
  (void)fmap1; (void)fmap2; (void)lvl0; (void)lvl1; (void)lvl2; (void)lvl3; (void)B; (void)C;
  (void)H; (void)W; (void)workspace; (void)workspace_bytes; (void)stream;
  return DROID_ERR_NOT_IMPLEMENTED;
}
