// Per-row gradient stage of the head, shared by row_grad_kernel (stand-alone backward) and
// finish_rows_kernel (fused training step: the row's softmax is at hand the moment its
// log-sum-exp is, so the two latency-bound row kernels become one launch).
#pragma once
#include "avt_common.cuh"

namespace avt {

// ---------------------------------------------------------------- per-row gradients
// Upstream gradient of every pooled similarity and of the positive map:
//   gl[j]     = dloss_scale * (softmax(logits[n])[j] - [j == 0]) + dlogits[n, j]
//   gsim[n,k] = gl[1 + k] * mask[n,k] / temp
//   GA[n,p]   = d/dA[n,p] of (gl[0] * sim1 + gl[L-1] * sim2) / temp  +  dA[n,p]
// One block of kRowGradThreads per row.  lse = log-sum-exp of the row, lg0 / lgL = its first / last
// logit: passed by VALUE because the fused caller has just produced them (columns 1..N_a of
// `logits` are read from memory: the similarity kernel wrote them in an earlier launch).
constexpr int kRowGradThreads = 256;

__device__ __forceinline__ float block_sum_rg(float x, float* s_tmp) {
    x = warp_sum(x);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) s_tmp[w] = x;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kRowGradThreads / 32; ++i) t += s_tmp[i];
    return t;
}

__device__ __forceinline__ void
row_grad_row(const int n, const float lse, const float lg0, const float lgL,
             const float* __restrict__ A, const float* __restrict__ logits,
                const int32_t* __restrict__ pos_idx,
                int col_offset, int N_a, int P, HeadParams hp, float dloss_scale_all,
                const float* __restrict__ dlogits, const float* __restrict__ dA,
                float* __restrict__ gsim, float* __restrict__ GA, const float* __restrict__ sim,
                int col_mult, const float* __restrict__ dloss_rows, const float* __restrict__ den,
                unsigned int* __restrict__ gmax_bits) {
    __shared__ float s_tmp[kRowGradThreads / 32];
    // upstream gradient of this row's loss: the uniform scalar plus an optional per-row term
    // (loss_rows.mean().backward() and friends never leave the library for a softmax)
    const float dloss_scale = dloss_scale_all + (dloss_rows ? dloss_rows[n] : 0.f);
    // bound of |cell gradient| for the fp16 scale of G:  G = gsim/den * w (1 + (S - sim)(1 - w)/tau)
    // with w <= min(1, den) (den is the sum of the weights) and |S - sim| <= 2, so
    // |G| <= |gsim| min(1, 1/den) (1 + 0.5/tau)  (+ |GA| on the positive column).  The bound is
    // within P (1 + 0.5/tau) < 2^12 of the largest cell even when every weight is tiny
    // (S far below eps): the scale then still leaves 16 binades of normal fp16 range under it.
    const float gfac = 1.0f + 0.5f / hp.tau;
    float mx = 0.f;
    const int L = N_a + 1 + (hp.use_neg ? 1 : 0);
    const float* lg = logits + (size_t)n * L;
    const int pos = pos_idx ? pos_idx[n] : col_offset + n;
    const float itemp = 1.0f / hp.temp, itau = 1.0f / hp.tau;

    if (col_mult > 1) {
        // tube mode: audio row k stands for T identical columns; d loss / d sim[n,k] sums their
        // softmax weights, the masked copy (positive audio only) with its factor 1 - mask_scale.
        // dlogits (if any) arrives already reduced over the copies WITH the mask factor applied.
        const float tf = (float)col_mult;
        for (int k = threadIdx.x; k < N_a; k += kRowGradThreads) {
            float g;
            if (k == pos) {
                const float un = expf(sim[(size_t)n * N_a + k] * itemp - lse);
                g = dloss_scale * ((tf - 1.0f) * un + (1.0f - hp.mask_scale) * expf(lg[1 + k] - lse));
            } else {
                g = dloss_scale * tf * expf(lg[1 + k] - lse);
            }
            if (dlogits) g += dlogits[(size_t)n * L + 1 + k];
            gsim[(size_t)n * N_a + k] = g * itemp;
            if (den) mx = fmaxf(mx, fabsf(g * itemp) * fminf(1.0f, 1.0f / den[(size_t)n * N_a + k]) * gfac);
        }
    } else {
        for (int k = threadIdx.x; k < N_a; k += kRowGradThreads) {
            float g = dloss_scale * expf(lg[1 + k] - lse);
            if (dlogits) g += dlogits[(size_t)n * L + 1 + k];
            float m = (k == pos) ? (1.0f - hp.mask_scale) : 1.0f;
            gsim[(size_t)n * N_a + k] = g * m * itemp;
            if (den) mx = fmaxf(mx, fabsf(g * m * itemp) * fminf(1.0f, 1.0f / den[(size_t)n * N_a + k]) * gfac);
        }
    }
    float g1 = dloss_scale * (expf(lg0 - lse) - 1.0f);
    if (dlogits) g1 += dlogits[(size_t)n * L];
    g1 *= itemp;
    float g2 = 0.f;
    if (hp.use_neg) {
        g2 = dloss_scale * expf(lgL - lse);
        if (dlogits) g2 += dlogits[(size_t)n * L + L - 1];
        g2 *= itemp;
    }
    // recompute the two masked means from A (model.py:128-133,138,141)
    float sp = 0.f, spa = 0.f, sn = 0.f, sna = 0.f;
    for (int p = threadIdx.x; p < P; p += kRowGradThreads) {
        float a = A[(size_t)n * P + p];
        float z1 = (a - hp.eps) * itau;
        float w1 = 1.0f / (1.0f + expf(-z1));
        float z2 = hp.trimap ? (a - hp.eps2) * itau : z1;
        float w2 = 1.0f / (1.0f + expf(z2));
        sp += w1; spa = fmaf(w1, a, spa);
        sn += w2; sna = fmaf(w2, a, sna);
    }
    {   // four sums behind one pair of barriers
        __shared__ float s_multi[kRowGradThreads / 32][4];
        sp = warp_sum(sp); spa = warp_sum(spa); sn = warp_sum(sn); sna = warp_sum(sna);
        if ((threadIdx.x & 31) == 0) {
            const int w = threadIdx.x >> 5;
            s_multi[w][0] = sp; s_multi[w][1] = spa; s_multi[w][2] = sn; s_multi[w][3] = sna;
        }
        __syncthreads();
        sp = spa = sn = sna = 0.f;
#pragma unroll
        for (int i = 0; i < kRowGradThreads / 32; ++i) {
            sp += s_multi[i][0]; spa += s_multi[i][1]; sn += s_multi[i][2]; sna += s_multi[i][3];
        }
    }
    const float sim1 = spa / sp, sim2 = sna / sn;
    for (int p = threadIdx.x; p < P; p += kRowGradThreads) {
        float a = A[(size_t)n * P + p];
        float z1 = (a - hp.eps) * itau;
        float w1 = 1.0f / (1.0f + expf(-z1));
        float z2 = hp.trimap ? (a - hp.eps2) * itau : z1;
        float w2 = 1.0f / (1.0f + expf(z2));            // Neg; dNeg/dA = -Neg (1 - Neg) / tau
        float g = g1 * (w1 + (a - sim1) * w1 * (1.0f - w1) * itau) / sp;
        g += g2 * (w2 - (a - sim2) * w2 * (1.0f - w2) * itau) / sn;
        if (dA) g += dA[(size_t)n * P + p];
        GA[(size_t)n * P + p] = g;
        mx = fmaxf(mx, fabsf(g));
    }
    if (gmax_bits) {                                     // non-negative floats order like their bits
        mx = warp_max(mx);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) s_tmp[threadIdx.x >> 5] = mx;
        __syncthreads();
        if (threadIdx.x == 0) {
            float t = s_tmp[0];
#pragma unroll
            for (int i = 1; i < kRowGradThreads / 32; ++i) t = fmaxf(t, s_tmp[i]);
            if (!(t <= 3.0e38f)) t = 3.0e38f;            // NaN / Inf upstream: keep the scale finite
            atomicMax(gmax_bits, __float_as_uint(t));
        }
    }
}


}  // namespace avt
