// C ABI of the similarity + hard-way loss head: argument checks, workspace carving and
// the launch sequence of both precisions.  Nothing here synchronises or allocates.
#include "avt_common.cuh"

namespace avt {

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Workspace layout (all offsets 1024-byte aligned).  The forward leaves the fp16 operand
// copies in place; a backward called with the same workspace reuses them.
struct HeadWorkspace {
    size_t scratch, msum, gsim, ga, da_w, u_w, vt, a_h, at_h, g_h, total;
};

HeadWorkspace head_workspace_layout(int N_v, int N_a, int C, int P, int precision, int backward) {
    HeadWorkspace w{};
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 1024); return o; };
    w.scratch = take(sizeof(HeadScratch));
    w.msum = take((size_t)N_v * P * sizeof(float));
    w.gsim = w.ga = w.da_w = w.u_w = w.vt = w.a_h = w.at_h = w.g_h = 0;
    if (precision == AVT_PRECISION_F16) {
        const size_t n_a_pad = (size_t)n_a_padded(N_a);
        // + one tile of slack rows so the last image's 208-row TMA box stays in bounds
        w.vt = take(((size_t)N_v * P + 64) * C * sizeof(__half));
        w.a_h = take(n_a_pad * C * sizeof(__half));
        if (backward) {
            w.at_h = take((size_t)C * n_a_pad * sizeof(__half));
            w.g_h = take(((size_t)N_v * P + 128) * n_a_pad * sizeof(__half));
        }
    }
    if (backward) {
        w.gsim = take((size_t)N_v * N_a * sizeof(float));
        w.ga = take((size_t)N_v * P * sizeof(float));
        w.da_w = take((size_t)N_v * P * sizeof(float));      // weighted_A backward: dA total, u
        w.u_w = take((size_t)N_v * P * sizeof(float));
    }
    w.total = off;
    return w;
}

static int check_shapes(const char* fn, int N_v, int N_a, int C, int P, int precision) {
    AVT_REQUIRE(N_v > 0 && N_a > 0, "%s: N_v=%d N_a=%d must be positive", fn, N_v, N_a);
    AVT_REQUIRE(C > 0 && C % 64 == 0, "%s: C=%d must be a positive multiple of 64", fn, C);
    AVT_REQUIRE(P > 0 && P <= 256, "%s: P=%d must be in 1..256", fn, P);
    AVT_REQUIRE(precision == AVT_PRECISION_FP32 || precision == AVT_PRECISION_F16,
                "%s: unknown precision %d", fn, precision);
    if (precision == AVT_PRECISION_F16) {
        AVT_REQUIRE(P <= 208 && P % 4 == 0, "%s: fp16 path needs P <= 208 and P %% 4 == 0 (P=%d)", fn, P);
        AVT_REQUIRE(C % 64 == 0 && C <= 4096, "%s: fp16 path needs C %% 64 == 0", fn);
    }
    return AVT_OK;
}

}  // namespace avt

using namespace avt;

extern "C" size_t avt_head_workspace_bytes(int N_v, int N_a, int C, int P, int precision,
                                           int backward) {
    if (N_v <= 0 || N_a <= 0 || C <= 0 || P <= 0) return 0;
    return head_workspace_layout(N_v, N_a, C, P, precision, backward).total;
}

extern "C" int avt_head_forward(const void* v_any, int v_dtype, const float* a_hat, int N_v, int N_a, int C, int P,
                                const int32_t* pos_idx, int col_offset, float eps, float eps2,
                                float tau, float temp, float mask_scale, int trimap, int use_neg,
                                int precision, int col_mult, float* A, float* Pos, float* Neg, float* logits,
                                float* loss_rows, float* loss_sum, float* weighted_A,
                                float* rnorm_v, float* sim, float* den, int stages,
                                float fuse_grad_scale, void* workspace,
                                size_t workspace_bytes, void* stream) {
    int rc = check_shapes("avt_head_forward", N_v, N_a, C, P, precision);
    if (rc != AVT_OK) return rc;
    AVT_REQUIRE(v_any && a_hat && A && Pos && Neg && logits && loss_rows && rnorm_v && sim && den,
                "avt_head_forward: null pointer among required arguments");
    AVT_REQUIRE(v_dtype == AVT_DTYPE_F32 || ((v_dtype == AVT_DTYPE_F16 || v_dtype == AVT_DTYPE_BF16) &&
                                             precision == AVT_PRECISION_F16),
                "avt_head_forward: v_dtype=%d (fp16 / bf16 maps need AVT_PRECISION_F16)", v_dtype);
    const float* v = static_cast<const float*>(v_any);      // only dereferenced when v_dtype is fp32
    AVT_REQUIRE(tau > 0.f && temp > 0.f, "avt_head_forward: tau and temp must be positive");
    AVT_REQUIRE(col_mult >= 1, "avt_head_forward: col_mult=%d must be >= 1", col_mult);
    AVT_REQUIRE(col_mult == 1 || pos_idx != nullptr,
                "avt_head_forward: tube mode (col_mult > 1) needs pos_idx (the audio row of every frame)");
    AVT_REQUIRE(workspace != nullptr, "avt_head_forward: workspace is null");
    AVT_REQUIRE(stages >= 0 && stages <= 3, "avt_head_forward: stages=%d", stages);
    if (stages == 0) stages = AVT_STAGE_PREP | AVT_STAGE_SIM;
    const bool fuse = fuse_grad_scale != 0.0f && (stages & AVT_STAGE_SIM);
    // the forward regions lead both layouts, so a workspace sized for backward works too
    HeadWorkspace w = head_workspace_layout(N_v, N_a, C, P, precision, fuse ? 1 : 0);
    if (workspace_bytes < w.total) {
        set_error("avt_head_forward: workspace %zu B < required %zu B", workspace_bytes, w.total);
        return AVT_ERR_WORKSPACE;
    }
    char* ws = static_cast<char*>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    HeadParams hp{eps, eps2, tau, temp, mask_scale, trimap, use_neg};
    HeadScratch* scratch = reinterpret_cast<HeadScratch*>(ws + w.scratch);
    float* msum = reinterpret_cast<float*>(ws + w.msum);

    if (precision == AVT_PRECISION_F16 && !avt_device_ok()) {
        set_error("avt_head_forward: the fp16 tcgen05 path needs an sm_100 device");
        return AVT_ERR_UNSUPPORTED;
    }
    __half* vt = precision == AVT_PRECISION_F16 ? reinterpret_cast<__half*>(ws + w.vt) : nullptr;
    if (stages & AVT_STAGE_PREP) {
        rc = launch_prep_visual(v_any, v_dtype, a_hat, pos_idx, col_offset, N_v, N_a, C, P, rnorm_v, A, msum, vt,
                                scratch, st);
        if (rc != AVT_OK) return rc;
    }
    if (!(stages & AVT_STAGE_SIM)) return AVT_OK;
    if (precision == AVT_PRECISION_FP32) {
        rc = launch_sim_fp32(v, a_hat, rnorm_v, pos_idx, col_offset, N_v, N_a, C, P, hp, logits, sim,
                             den, st);
        if (rc != AVT_OK) return rc;
    } else {
        __half* a_h = reinterpret_cast<__half*>(ws + w.a_h);
        const int n_a_pad = n_a_padded(N_a);
        rc = launch_audio_to_f16(a_hat, N_a, n_a_pad, C, a_h, fuse ? &scratch->gmax_bits : nullptr, st);
        if (rc != AVT_OK) return rc;
        rc = launch_sim_f16(vt, a_h, pos_idx, col_offset, N_v, N_a, C, P, hp, logits, sim, den, st);
        if (rc != AVT_OK) return rc;
    }
    float* gsim = nullptr; float* GA = nullptr; unsigned int* gmax_bits = nullptr;
    if (fuse) {             // fused training step: the row gradients come out of the finish kernel
        gsim = reinterpret_cast<float*>(ws + w.gsim);
        GA = reinterpret_cast<float*>(ws + w.ga);
        if (precision == AVT_PRECISION_F16) gmax_bits = &scratch->gmax_bits;    // zeroed by the audio operand kernel
    }
    return launch_finish_rows(A, msum, N_v, N_a, P, hp, Pos, Neg, weighted_A, logits, loss_rows,
                              loss_sum, scratch, pos_idx, col_offset, sim, col_mult, fuse_grad_scale, gsim, GA,
                              den, gmax_bits, st);
}

extern "C" int avt_head_backward(const void* v_any, int v_dtype, const float* a_hat, int N_v, int N_a, int C, int P,
                                 const int32_t* pos_idx, int col_offset, float eps, float eps2,
                                 float tau, float temp, float mask_scale, int trimap, int use_neg,
                                 int precision, int col_mult, const float* A, const float* logits,
                                 const float* loss_rows, const float* rnorm_v, const float* sim,
                                 const float* den, float dloss_scale, const float* dloss_rows,
                                 const float* dlogits,
                                 const float* dA, const float* dWeighted, float* dv,
                                 float* d_ahat, int flags, void* dv_stream, void* workspace,
                                 size_t workspace_bytes, void* stream) {
    int rc = check_shapes("avt_head_backward", N_v, N_a, C, P, precision);
    if (rc != AVT_OK) return rc;
    AVT_REQUIRE(v_any && a_hat && A && logits && loss_rows && rnorm_v && sim && den && dv && d_ahat,
                "avt_head_backward: null pointer among required arguments");
    AVT_REQUIRE(v_dtype == AVT_DTYPE_F32 || ((v_dtype == AVT_DTYPE_F16 || v_dtype == AVT_DTYPE_BF16) &&
                                             precision == AVT_PRECISION_F16 && dWeighted == nullptr),
                "avt_head_backward: v_dtype=%d (fp16 / bf16 maps need AVT_PRECISION_F16 and no dWeighted)", v_dtype);
    const float* v = static_cast<const float*>(v_any);      // the fp16 path never reads it (uses its own copy)
    AVT_REQUIRE(workspace != nullptr, "avt_head_backward: workspace is null");
    AVT_REQUIRE(col_mult >= 1 && (col_mult == 1 || pos_idx != nullptr),
                "avt_head_backward: col_mult=%d (tube mode needs pos_idx)", col_mult);
    HeadWorkspace w = head_workspace_layout(N_v, N_a, C, P, precision, 1);
    if (workspace_bytes < w.total) {
        set_error("avt_head_backward: workspace %zu B < required %zu B", workspace_bytes, w.total);
        return AVT_ERR_WORKSPACE;
    }
    char* ws = static_cast<char*>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    HeadParams hp{eps, eps2, tau, temp, mask_scale, trimap, use_neg};
    float* gsim = reinterpret_cast<float*>(ws + w.gsim);
    float* GA = reinterpret_cast<float*>(ws + w.ga);

    const float* msum = reinterpret_cast<const float*>(ws + w.msum);   // left by the forward
    float* u_w = reinterpret_cast<float*>(ws + w.u_w);
    if (dWeighted) {            // model.py:148-152: fold the map part into dA, keep u for dv
        float* da_w = reinterpret_cast<float*>(ws + w.da_w);
        rc = launch_weighted_row_grad(A, msum, dWeighted, dA, N_v, P, C, hp, da_w, u_w, st);
        if (rc != AVT_OK) return rc;
        dA = da_w;
    }
    HeadScratch* scratch = reinterpret_cast<HeadScratch*>(ws + w.scratch);
    unsigned int* gmax_bits = precision == AVT_PRECISION_F16 ? &scratch->gmax_bits : nullptr;
    if (flags & AVT_BWD_ROWGRAD_DONE) {
        // the forward ran with fuse_grad_scale: gsim / GA / the bound of G are in the workspace
        AVT_REQUIRE(!dlogits && !dA && !dWeighted && !dloss_rows,
                    "avt_head_backward: AVT_BWD_ROWGRAD_DONE takes no upstream tensors (the fused forward "
                    "already applied fuse_grad_scale)");
    } else {
        if (gmax_bits)                             // bound of the cell gradients -> fp16 scale of G
            AVT_CHECK_CUDA(cudaMemsetAsync(gmax_bits, 0, sizeof(unsigned int), st));
        rc = launch_row_grad(A, logits, loss_rows, pos_idx, col_offset, N_v, N_a, P, hp, dloss_scale,
                             dlogits, dA, gsim, GA, sim, col_mult, dloss_rows, den, gmax_bits, st);
        if (rc != AVT_OK) return rc;
    }
    AVT_CHECK_CUDA(cudaMemsetAsync(d_ahat, 0, (size_t)N_a * C * sizeof(float), st));
    if (precision == AVT_PRECISION_FP32) {
        AVT_CHECK_CUDA(cudaMemsetAsync(dv, 0, (size_t)N_v * C * P * sizeof(float), st));
        rc = launch_bwd_fp32(v, a_hat, rnorm_v, pos_idx, col_offset, N_v, N_a, C, P, hp, gsim, GA, dv,
                             d_ahat, st);
        if (rc != AVT_OK) return rc;
        rc = launch_dv_finish(v, rnorm_v, N_v, C, P, dv, st);
        if (rc != AVT_OK || !dWeighted) return rc;
        return launch_weighted_dv(v, rnorm_v, msum, u_w, N_v, C, P, dv, st);
    }
    if (!avt_device_ok()) {
        set_error("avt_head_backward: the fp16 tcgen05 path needs an sm_100 device");
        return AVT_ERR_UNSUPPORTED;
    }
    const int n_a_pad = n_a_padded(N_a);
    __half* vt = reinterpret_cast<__half*>(ws + w.vt);
    __half* a_h = reinterpret_cast<__half*>(ws + w.a_h);
    __half* at_h = reinterpret_cast<__half*>(ws + w.at_h);
    __half* g_h = reinterpret_cast<__half*>(ws + w.g_h);
    rc = launch_bwd_f16(v, vt, a_h, at_h, g_h, a_hat, rnorm_v, pos_idx, col_offset, N_v, N_a,
                         n_a_pad, C, P, hp, sim, den, gsim, GA, dv, d_ahat, gmax_bits,
                         static_cast<cudaStream_t>(dv_stream), st);
    if (rc != AVT_OK || !dWeighted) return rc;
    return launch_weighted_dv(v, rnorm_v, msum, u_w, N_v, C, P, dv, st);
}
