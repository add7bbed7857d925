// Shared helpers for the libavtubes kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/avtubes.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libavtubes targets sm_100a (B200) only"
#endif

namespace avt {

void set_error(const char* fmt, ...);

// Launch counter and optional per-stage device timing (avt_profile_* in avtubes.h).
// Stage slots; the events are recorded on the stream the stage is launched on.
enum ProfSlot {
    kProfPrep = 0, kProfSim = 1, kProfFinish = 2, kProfRowGrad = 3, kProfBwdG = 4,
    kProfBwdDv = 5, kProfBwdDa = 6, kProfDvFinish = 7, kProfEval = 8, kProfSlots = 9
};
void count_launch();
bool prof_enabled();
void prof_begin(int slot, cudaStream_t st);
void prof_end(int slot, cudaStream_t st);
struct ProfScope {
    int slot; cudaStream_t st;
    ProfScope(int s, cudaStream_t t) : slot(s), st(t) { prof_begin(slot, st); }
    ~ProfScope() { prof_end(slot, st); }
};

#define AVT_CHECK_CUDA(expr)                                                        \
    do {                                                                            \
        cudaError_t _e = (expr);                                                    \
        if (_e != cudaSuccess) {                                                    \
            avt::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),  \
                           __FILE__, __LINE__);                                     \
            return AVT_ERR_CUDA;                                                    \
        }                                                                           \
    } while (0)

#define AVT_REQUIRE(cond, ...)                                                      \
    do {                                                                            \
        if (!(cond)) {                                                              \
            avt::set_error(__VA_ARGS__);                                            \
            return AVT_ERR_INVALID;                                                 \
        }                                                                           \
    } while (0)

#define AVT_LAUNCH_CHECK()                                                          \
    do {                                                                            \
        avt::count_launch();                                                        \
        cudaError_t _e = cudaGetLastError();                                        \
        if (_e != cudaSuccess) {                                                    \
            avt::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), \
                           __FILE__, __LINE__);                                     \
            return AVT_ERR_CUDA;                                                    \
        }                                                                           \
    } while (0)

constexpr float kNormEps = 1e-12f;          // torch.nn.functional.normalize default eps
constexpr float kLog2e = 1.4426950408889634f;

// audio rows are padded (zero rows) to whole 128-row MMA tiles; above one tile to whole
// 256-row CTA-pair tiles so the cta_group::2 kernels always see an even tile count
inline int n_a_padded(int N_a) { return N_a <= 128 ? 128 : (N_a + 255) / 256 * 256; }

struct HeadParams {
    float eps, eps2, tau, temp, mask_scale;
    int trimap, use_neg;
};

__device__ __forceinline__ float warp_sum(float x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
}
__device__ __forceinline__ float warp_max(float x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, o));
    return x;
}
__device__ __forceinline__ int warp_sum_i(int x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
}

// sigma((s - eps)/tau) with RELATIVE accuracy for tiny values (they are weights of a
// ratio): 1 / (1 + 2^(-z*log2e)), ex2.approx + rcp.approx.
__device__ __forceinline__ float fast_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_rcp(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// accurate variant for the fp32 path (expf-based, keeps relative accuracy too)
__device__ __forceinline__ float sigmoid_acc(float z) {
    return 1.0f / (1.0f + expf(-z));
}

// layout of the small per-call scratch at the head of every head workspace
struct HeadScratch {
    unsigned int rows_done;      // last-block-done counter of the row-finish kernel
    unsigned int gmax_bits;      // backward: bits of max |cell gradient| bound (row_grad_kernel)
    unsigned int pad[62];
};

// Power-of-two scale of the fp16 cell gradients G: bound * scale < 2^14 (the positive column adds
// a second term of at most the same size, so |G * scale| < 2^15 < 65504), exact to undo.
__host__ __device__ __forceinline__ int g_scale_exp(unsigned int bound_bits) {
    if (bound_bits == 0u) return 0;
    int se = 13 - ((int)(bound_bits >> 23) - 127);
    return se < -100 ? -100 : (se > 100 ? 100 : se);
}
__device__ __forceinline__ float pow2i(int e) { return __int_as_float((e + 127) << 23); }

// ---- kernel launchers implemented in the .cu files -------------------------------
// prep: rnorm_v, A (positive map), msum (mean_c v_hat), optional fp16 K-major copy
int launch_prep_visual(const void* v, int v_dtype, const float* a_hat, const int32_t* pos_idx, int col_offset,
                       int N_v, int N_a, int C, int P, float* rnorm_v, float* A, float* msum,
                       __half* vt_f16, HeadScratch* scratch, cudaStream_t st);
int launch_finish_rows(const float* A, const float* msum, int N_v, int N_a, int P, HeadParams hp,
                       float* Pos, float* Neg, float* weighted_A, float* logits,
                       float* loss_rows, float* loss_sum, HeadScratch* scratch, const int32_t* pos_idx,
                       int col_offset, const float* sim, int col_mult, float fuse_scale, float* gsim,
                       float* GA, const float* den, unsigned int* gmax_bits, cudaStream_t st);
int launch_sim_fp32(const float* v, const float* a_hat, const float* rnorm_v,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                    HeadParams hp, float* logits, float* sim, float* den, cudaStream_t st);
int launch_row_grad(const float* A, const float* logits, const float* loss_rows,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int P, HeadParams hp,
                    float dloss_scale, const float* dlogits, const float* dA, float* gsim, float* GA,
                    const float* sim, int col_mult, const float* dloss_rows, const float* den,
                    unsigned int* gmax_bits, cudaStream_t st);
int launch_bwd_fp32(const float* v, const float* a_hat, const float* rnorm_v,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                    HeadParams hp, const float* gsim,
                    const float* GA, float* dvh, float* dah, cudaStream_t st);
// backward of weighted_A (avt_head_weighted.cu)
int launch_weighted_row_grad(const float* A, const float* msum, const float* gW, const float* dA_in,
                             int N_v, int P, int C, HeadParams hp, float* dA_out, float* u_out,
                             cudaStream_t st);
int launch_weighted_dv(const float* v, const float* rnorm_v, const float* msum, const float* u,
                       int N_v, int C, int P, float* dv, cudaStream_t st);
int launch_dv_finish(const float* v, const float* rnorm_v, int N_v, int C, int P, float* dv,
                     cudaStream_t st);
// tcgen05 path (avt_head_tc.cu): vt = normalised maps, fp16 [N_v*P (+slack)][C]; a_h = audio
// rows, fp16 [n_a_pad][C] (rows >= N_a zero)
int launch_sim_f16(const __half* vt, const __half* a_h,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                    HeadParams hp, float* logits, float* sim, float* den, cudaStream_t st);
int launch_bwd_f16(const float* v, const __half* vt, const __half* a_h, __half* at_h,
                    __half* g_h, const float* a_hat, const float* rnorm_v,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int n_a_pad, int C,
                    int P, HeadParams hp, const float* sim, const float* den, const float* gsim,
                    const float* GA, float* dv, float* d_ahat, const unsigned int* gmax_bits,
                    cudaStream_t dv_stream, cudaStream_t st);
int launch_audio_to_f16(const float* a_hat, int N_a, int N_a_pad, int C, __half* out,
                         unsigned int* zero_word, cudaStream_t st);

}  // namespace avt
