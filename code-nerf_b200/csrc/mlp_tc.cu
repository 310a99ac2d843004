// placeholder replaced by the tcgen05 chain
#include "common.cuh"
#include "internal.h"
struct TcChainArgs;
size_t cnrf_tc_saved_bytes(int64_t) { return 0; }
size_t cnrf_tc_workspace_bytes(int64_t, int64_t, int) { return 0; }
int cnrf_tc_chain_fwd(const TcChainArgs&, float*, void*, size_t, void*, size_t, cudaStream_t) { return CNRF_E_UNSUPPORTED; }
int cnrf_tc_chain_bwd(const TcChainArgs&, const float*, const void*, size_t, const cnrf_codenerf_grads*, float*, float*,
                      float*, float*, float*, void*, size_t, cudaStream_t) { return CNRF_E_UNSUPPORTED; }
