// assign_tc.cu — placeholder until the tcgen05 kernel lands (next commit).
#include "common.cuh"
namespace c3vq {
size_t tc_workspace_bytes(int64_t, int, int) { return 256; }
int launch_assign_tc(const float *, const float *, const int64_t *, int64_t, int, const float *, int, float *,
                     int64_t *, float *, float *, float *, float *, void *, size_t, cudaStream_t)
{
    set_error("tensor-core assignment is not built in this revision");
    return C3VQ_EINVAL;
}
int tc_stats(void *, int64_t out[3], cudaStream_t) { out[0] = out[1] = out[2] = 0; return C3VQ_OK; }
}  // namespace c3vq
