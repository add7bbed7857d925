// placeholder until the tcgen05 kernels land
#include "avt_common.cuh"
namespace avt {
int launch_sim_bf16(const __nv_bfloat16*, const __nv_bfloat16*, const int32_t*, int, int, int, int, int,
                    HeadParams, float*, float*, float*, cudaStream_t) {
    set_error("bf16 path not built"); return AVT_ERR_UNSUPPORTED;
}
int launch_bwd_bf16(const __nv_bfloat16*, const __nv_bfloat16*, __nv_bfloat16*, __nv_bfloat16*, const float*,
                    const float*, const int32_t*, int, int, int, int, int, int, HeadParams, const float*,
                    const float*, const float*, const float*, float*, float*, cudaStream_t) {
    set_error("bf16 path not built"); return AVT_ERR_UNSUPPORTED;
}
}
