// sim_tc.cu — placeholder until the tcgen05 kernels land (eligible() == false keeps every
// call on the SIMT path).
#include "common.cuh"
#include "sim_tc.h"
namespace mm25 { namespace tc {
bool eligible(int64_t, int64_t, int64_t, int64_t, int, int) { return false; }
size_t fwd_workspace_bytes(int64_t, int64_t, int64_t, int64_t) { return 0; }
int launch_fwd(const void*, const void*, int64_t, int64_t, int64_t, int64_t, const float*, int, float, int64_t, float*, float*, float*, float*, void*, size_t, cudaStream_t) { return MM25_ERR_SHAPE; }
size_t bwd_workspace_bytes(int64_t, int64_t, int64_t, int64_t) { return 0; }
int64_t bwd_nparts(int64_t, int64_t, int64_t, int64_t) { return 0; }
int launch_bwd(const void*, const void*, int64_t, int64_t, int64_t, int64_t, const float*, int, float, int64_t, const float*, const float*, float, float, float, float, const float*, float*, float*, float*, void*, size_t, cudaStream_t) { return MM25_ERR_SHAPE; }
}}
