// placeholder until the tcgen05 backward lands
#include "common.cuh"
namespace b200 {
bool tc_bwd_supported(int, int, int, int) { return false; }
size_t tc_bwd_workspace_bytes(int, int, int, int, int) { return 0; }
int tc_bwd(const void*, const void*, const void*, const void*, const float*, const float*, void*, void*, void*, int, int, int, int,
           float, int, void*, size_t, cudaStream_t) { return B200_ERR_UNSUPPORTED; }
}
