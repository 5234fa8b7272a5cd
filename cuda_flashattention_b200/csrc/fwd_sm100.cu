// placeholder until the tcgen05 forward lands
#include "common.cuh"
namespace b200 {
bool tc_fwd_supported(int, int, int, int) { return false; }
int tc_fwd(const void*, const void*, const void*, void*, float*, int, int, int, int, float, int, cudaStream_t) { return B200_ERR_UNSUPPORTED; }
}
