// placeholder until the tcgen05 kernel lands (replaced in the next commit)
#include "common.cuh"
#include "handle.h"
namespace nvllm {
bool tc5_supported(int, int, int) { return false; }
int tc5_splits(int, int, int, int) { return 1; }
int tc5_box_rows_act(int) { return 128; }
int tc5_box_rows_w(int) { return 128; }
int make_tma_2d_bf16(TmaDesc*, const void*, uint64_t, uint64_t, uint64_t, uint32_t, std::string*) { return 0; }
cudaError_t launch_gemm_tc5(cudaStream_t, const Tc5Args&, int) { return cudaErrorNotSupported; }
int build_weight_maps(nvllm_handle*) { return 0; }
int build_activation_maps(nvllm_handle*) { return 0; }
}
