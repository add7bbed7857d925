// AVT_PRECISION_FP32: the all-pairs contraction on the CUDA cores in the reference's own
// arithmetic (fp32 FMA, expf sigmoid).  It is the exact mode of the library and the
// on-device cross-check for the tcgen05 path; the similarity tensor is still never
// written to memory.  Also here: the per-row gradient kernel shared by both precisions.
#include "avt_common.cuh"
#include "avt_rows.cuh"

namespace avt {

constexpr int kSimThreads = 256;   // one thread per spatial position (P <= 256)
constexpr int kKT = 16;            // audio rows per block

// smem layout helper: audio tile [C][kKT] so that one position's 16 dot products read
// four broadcast float4 per channel
__device__ __forceinline__ void load_audio_tile(float* a_s, const float* __restrict__ a_hat,
                                                int k0, int N_a, int C) {
    for (int i = threadIdx.x; i < C * kKT; i += blockDim.x) {
        int k = i / C, c = i - k * C;                      // coalesced along c
        a_s[c * kKT + k] = (k0 + k < N_a) ? a_hat[(size_t)(k0 + k) * C + c] : 0.f;
    }
}

// S[k] = rn * sum_c v[n,c,p] * a_hat[k0+k,c]      model.py:117,125
__device__ __forceinline__ void dot_rows(const float* __restrict__ vp, int C, int P,
                                         const float* a_s, float rn, float (&S)[kKT]) {
#pragma unroll
    for (int k = 0; k < kKT; ++k) S[k] = 0.f;
#pragma unroll 4
    for (int c = 0; c < C; ++c) {
        float x = __ldg(vp + (size_t)c * P);
        const float4* ar = reinterpret_cast<const float4*>(a_s + c * kKT);
#pragma unroll
        for (int q = 0; q < kKT / 4; ++q) {
            float4 w = ar[q];
            S[4 * q + 0] = fmaf(x, w.x, S[4 * q + 0]);
            S[4 * q + 1] = fmaf(x, w.y, S[4 * q + 1]);
            S[4 * q + 2] = fmaf(x, w.z, S[4 * q + 2]);
            S[4 * q + 3] = fmaf(x, w.w, S[4 * q + 3]);
        }
    }
#pragma unroll
    for (int k = 0; k < kKT; ++k) S[k] *= rn;
}

// ---------------------------------------------------------------- forward, fp32
// grid (ceil(N_a / 16), N_v).  model.py:125,135,140 + the diagonal mask and /0.07 of
// :115,143-146 for columns 1..N_a of the logits.
__global__ void __launch_bounds__(kSimThreads)
sim_fp32_kernel(const float* __restrict__ v, const float* __restrict__ a_hat,
                const float* __restrict__ rnorm_v, const int32_t* __restrict__ pos_idx,
                int col_offset, int N_a, int C, int P, HeadParams hp, int L,
                float* __restrict__ logits, float* __restrict__ sim, float* __restrict__ den) {
    extern __shared__ float smem[];
    float* a_s = smem;                                  // [C][kKT]
    float* red = smem + (size_t)C * kKT;                // [2][kSimThreads/32][kKT]
    const int n = blockIdx.y, k0 = blockIdx.x * kKT;
    const int p = threadIdx.x;
    load_audio_tile(a_s, a_hat, k0, N_a, C);
    __syncthreads();
    float S[kKT];
    float num[kKT], dn[kKT];
    if (p < P) {
        dot_rows(v + (size_t)n * C * P + p, C, P, a_s, rnorm_v[(size_t)n * P + p], S);
        const float itau = 1.0f / hp.tau;
#pragma unroll
        for (int k = 0; k < kKT; ++k) {
            float w = sigmoid_acc((S[k] - hp.eps) * itau);
            dn[k] = w;
            num[k] = w * S[k];
        }
    } else {
#pragma unroll
        for (int k = 0; k < kKT; ++k) { dn[k] = 0.f; num[k] = 0.f; }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < kKT; ++k) {
        float a = warp_sum(num[k]), b = warp_sum(dn[k]);
        if (lane == 0) {
            red[(0 * (kSimThreads / 32) + warp) * kKT + k] = a;
            red[(1 * (kSimThreads / 32) + warp) * kKT + k] = b;
        }
    }
    __syncthreads();
    if (threadIdx.x < kKT && k0 + threadIdx.x < N_a) {
        const int k = threadIdx.x;
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int w = 0; w < kSimThreads / 32; ++w) {
            a += red[(0 * (kSimThreads / 32) + w) * kKT + k];
            b += red[(1 * (kSimThreads / 32) + w) * kKT + k];
        }
        const int pos = pos_idx ? pos_idx[n] : col_offset + n;
        const float s = a / b;
        const float m = (k0 + k == pos) ? (1.0f - hp.mask_scale) : 1.0f;
        sim[(size_t)n * N_a + k0 + k] = s;
        den[(size_t)n * N_a + k0 + k] = b;
        logits[(size_t)n * L + 1 + k0 + k] = s * m / hp.temp;
    }
}

int launch_sim_fp32(const float* v, const float* a_hat, const float* rnorm_v,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                    HeadParams hp, float* logits, float* sim, float* den, cudaStream_t st) {
    ProfScope _ps(kProfSim, st);
    const int L = N_a + 1 + (hp.use_neg ? 1 : 0);
    size_t smem = ((size_t)C * kKT + 2 * (kSimThreads / 32) * kKT) * sizeof(float);
    AVT_CHECK_CUDA(cudaFuncSetAttribute(sim_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)smem));
    dim3 grid((N_a + kKT - 1) / kKT, N_v);
    sim_fp32_kernel<<<grid, kSimThreads, smem, st>>>(v, a_hat, rnorm_v, pos_idx, col_offset, N_a, C,
                                                    P, hp, L, logits, sim, den);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

// ---------------------------------------------------------------- per-row gradients (avt_rows.cuh)
__global__ void __launch_bounds__(kRowGradThreads)
row_grad_kernel(const float* __restrict__ A, const float* __restrict__ logits,
                const float* __restrict__ loss_rows, const int32_t* __restrict__ pos_idx,
                int col_offset, int N_a, int P, HeadParams hp, float dloss_scale_all,
                const float* __restrict__ dlogits, const float* __restrict__ dA,
                float* __restrict__ gsim, float* __restrict__ GA, const float* __restrict__ sim,
                int col_mult, const float* __restrict__ dloss_rows, const float* __restrict__ den,
                unsigned int* __restrict__ gmax_bits) {
    const int n = blockIdx.x;
    const int L = N_a + 1 + (hp.use_neg ? 1 : 0);
    const float lg0 = logits[(size_t)n * L], lgL = logits[(size_t)n * L + L - 1];
    row_grad_row(n, loss_rows[n] + lg0, lg0, lgL, A, logits, pos_idx, col_offset, N_a, P, hp, dloss_scale_all, dlogits, dA,
                 gsim, GA, sim, col_mult, dloss_rows, den, gmax_bits);
}

int launch_row_grad(const float* A, const float* logits, const float* loss_rows,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int P, HeadParams hp,
                    float dloss_scale, const float* dlogits, const float* dA, float* gsim, float* GA,
                    const float* sim, int col_mult, const float* dloss_rows, const float* den,
                    unsigned int* gmax_bits, cudaStream_t st) {
    ProfScope _ps(kProfRowGrad, st);
    row_grad_kernel<<<N_v, kRowGradThreads, 0, st>>>(A, logits, loss_rows, pos_idx, col_offset, N_a,
                                                     P, hp, dloss_scale, dlogits, dA, gsim, GA, sim,
                                                     col_mult, dloss_rows, den, gmax_bits);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

// ---------------------------------------------------------------- backward, fp32
// grid (ceil(N_a / 16), N_v), 256 threads.  Recomputes S for the block's 16 audio rows,
// forms the cell gradient G in registers, then sweeps the channels once more:
//   dv_hat[n,c,p] += sum_k G[k,p] a_hat[k,c]        (thread = position)
//   d_ahat[k,c]   += sum_p G[k,p] v_hat[n,c,p]      (warp = two audio rows, lane = channel)
// Both are accumulated with fp32 atomics (dv / d_ahat zeroed by the caller).
constexpr int kBwdChunk = 32;                        // channels staged per step

__global__ void __launch_bounds__(kSimThreads)
bwd_fp32_kernel(const float* __restrict__ v, const float* __restrict__ a_hat,
                const float* __restrict__ rnorm_v, const int32_t* __restrict__ pos_idx,
                int col_offset, int N_a, int C, int P, HeadParams hp,
                const float* __restrict__ gsim, const float* __restrict__ GA,
                float* __restrict__ dvh, float* __restrict__ dah) {
    extern __shared__ float smem[];
    const int PS = P + 1;                                // padded row of the staging tiles
    float* a_s = smem;                                   // [C][kKT]
    float* G_s = a_s + (size_t)C * kKT;                  // [kKT][PS]
    float* v_s = G_s + (size_t)kKT * PS;                 // [kBwdChunk][PS]
    float* rn_s = v_s + (size_t)kBwdChunk * PS;          // [P]
    float* red_s = rn_s + P;                             // [2][warps][kKT]
    float* sd_s = red_s + 2 * (kSimThreads / 32) * kKT;   // [2][kKT] pooled sim, den
    const int n = blockIdx.y, k0 = blockIdx.x * kKT;
    const int p = threadIdx.x;
    const int pos = pos_idx ? pos_idx[n] : col_offset + n;
    load_audio_tile(a_s, a_hat, k0, N_a, C);
    for (int i = threadIdx.x; i < P; i += blockDim.x) rn_s[i] = rnorm_v[(size_t)n * P + i];
    __syncthreads();

    // S for this block's 16 audio rows, then the pooled sim / den recomputed in fp32 here
    // (self-consistent gradient; the saved forward values may come from the fp16 path)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float itau = 1.0f / hp.tau;
    float S[kKT], W[kKT];
    if (p < P) {
        dot_rows(v + (size_t)n * C * P + p, C, P, a_s, rn_s[p], S);
#pragma unroll
        for (int k = 0; k < kKT; ++k) W[k] = sigmoid_acc((S[k] - hp.eps) * itau);
    } else {
#pragma unroll
        for (int k = 0; k < kKT; ++k) { S[k] = 0.f; W[k] = 0.f; }
    }
#pragma unroll
    for (int k = 0; k < kKT; ++k) {
        float a = warp_sum(W[k] * S[k]), b = warp_sum(W[k]);
        if (lane == 0) {
            red_s[(0 * (kSimThreads / 32) + warp) * kKT + k] = a;
            red_s[(1 * (kSimThreads / 32) + warp) * kKT + k] = b;
        }
    }
    __syncthreads();
    if (threadIdx.x < kKT) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int w = 0; w < kSimThreads / 32; ++w) {
            a += red_s[(0 * (kSimThreads / 32) + w) * kKT + threadIdx.x];
            b += red_s[(1 * (kSimThreads / 32) + w) * kKT + threadIdx.x];
        }
        sd_s[threadIdx.x] = a / b;
        sd_s[kKT + threadIdx.x] = b;
    }
    __syncthreads();
    float G[kKT];
    if (p < P) {
#pragma unroll
        for (int k = 0; k < kKT; ++k) {
            float g = 0.f;
            if (k0 + k < N_a) {
                const float w = W[k], sm = sd_s[k], dn = sd_s[kKT + k];
                g = gsim[(size_t)n * N_a + k0 + k] * (w + (S[k] - sm) * w * (1.0f - w) * itau) / dn;
                if (k0 + k == pos) g += GA[(size_t)n * P + p];
            }
            G[k] = g;
            G_s[k * PS + p] = g;
        }
    } else {
#pragma unroll
        for (int k = 0; k < kKT; ++k) G[k] = 0.f;
    }
    for (int c0 = 0; c0 < C; c0 += kBwdChunk) {
        __syncthreads();
        for (int i = threadIdx.x; i < kBwdChunk * P; i += blockDim.x) {
            int cc = i / P, pp = i - cc * P;
            v_s[cc * PS + pp] = __ldg(v + ((size_t)n * C + c0 + cc) * P + pp) * rn_s[pp];
        }
        __syncthreads();
        if (p < P) {
#pragma unroll 4
            for (int cc = 0; cc < kBwdChunk; ++cc) {
                const float4* ar = reinterpret_cast<const float4*>(a_s + (c0 + cc) * kKT);
                float acc = 0.f;
#pragma unroll
                for (int q = 0; q < kKT / 4; ++q) {
                    float4 w = ar[q];
                    acc = fmaf(G[4 * q + 0], w.x, acc);
                    acc = fmaf(G[4 * q + 1], w.y, acc);
                    acc = fmaf(G[4 * q + 2], w.z, acc);
                    acc = fmaf(G[4 * q + 3], w.w, acc);
                }
                atomicAdd(dvh + ((size_t)n * C + c0 + cc) * P + p, acc);
            }
        }
        {   // d_ahat: warp -> audio rows 2w, 2w+1; lane -> channel c0 + lane
            const int ka = 2 * warp, kb = 2 * warp + 1;
            const float* ga = G_s + ka * PS;
            const float* gb = G_s + kb * PS;
            const float* vr = v_s + lane * PS;
            float acc0 = 0.f, acc1 = 0.f;
            for (int pp = 0; pp < P; ++pp) {
                float x = vr[pp];
                acc0 = fmaf(ga[pp], x, acc0);
                acc1 = fmaf(gb[pp], x, acc1);
            }
            if (k0 + ka < N_a) atomicAdd(dah + (size_t)(k0 + ka) * C + c0 + lane, acc0);
            if (k0 + kb < N_a) atomicAdd(dah + (size_t)(k0 + kb) * C + c0 + lane, acc1);
        }
    }
}

int launch_bwd_fp32(const float* v, const float* a_hat, const float* rnorm_v,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                    HeadParams hp, const float* gsim,
                    const float* GA, float* dvh, float* dah, cudaStream_t st) {
    ProfScope _ps(kProfBwdG, st);
    const int PS = P + 1;
    size_t smem = ((size_t)C * kKT + (size_t)kKT * PS + (size_t)kBwdChunk * PS + P +
                   2 * (kSimThreads / 32) * kKT + 2 * kKT) * sizeof(float);
    AVT_CHECK_CUDA(cudaFuncSetAttribute(bwd_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)smem));
    dim3 grid((N_a + kKT - 1) / kKT, N_v);
    bwd_fp32_kernel<<<grid, kSimThreads, smem, st>>>(v, a_hat, rnorm_v, pos_idx, col_offset, N_a, C, P,
                                                    hp, gsim, GA, dvh, dah);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

// ---------------------------------------------------------------- F.normalize Jacobian
// In place on dv (which holds dv_hat):  dv = rn * (dv_hat - v_hat * <v_hat, dv_hat>_c)
// Same block shape as the prep pass: grid (2, N_v), 4 channel groups x 128 lanes.
constexpr int kFinThreads = 512;
constexpr int kFinGroups = 4;

__global__ void __launch_bounds__(kFinThreads)
dv_finish_kernel(const float* __restrict__ v, const float* __restrict__ rnorm_v, int C, int P,
                 float* __restrict__ dv) {
    __shared__ float s_red[kFinGroups][128];
    const int n = blockIdx.y;
    const int PT = (P + 1) / 2;
    const int lane = threadIdx.x & 127, g = threadIdx.x >> 7;
    const int p = blockIdx.x * PT + lane;
    const bool active = (lane < PT) && (p < P);
    const float rn = active ? rnorm_v[(size_t)n * P + p] : 0.f;
    const float* vp = v + (size_t)n * C * P + p;
    float* dp = dv + (size_t)n * C * P + p;
    float t = 0.f;
    if (active) {
        for (int c0 = g * 8; c0 < C; c0 += 8 * kFinGroups) {
            float x[8], y[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) { x[i] = __ldg(vp + (size_t)(c0 + i) * P); y[i] = dp[(size_t)(c0 + i) * P]; }
#pragma unroll
            for (int i = 0; i < 8; ++i) t = fmaf(x[i] * rn, y[i], t);
        }
    }
    s_red[g][lane] = t;
    __syncthreads();
    if (active) {
        t = 0.f;
#pragma unroll
        for (int j = 0; j < kFinGroups; ++j) t += s_red[j][lane];
        for (int c0 = g * 8; c0 < C; c0 += 8 * kFinGroups) {
            float x[8], y[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) { x[i] = __ldg(vp + (size_t)(c0 + i) * P); y[i] = dp[(size_t)(c0 + i) * P]; }
#pragma unroll
            for (int i = 0; i < 8; ++i) dp[(size_t)(c0 + i) * P] = rn * (y[i] - x[i] * rn * t);
        }
    }
}

int launch_dv_finish(const float* v, const float* rnorm_v, int N_v, int C, int P, float* dv,
                     cudaStream_t st) {
    ProfScope _ps(kProfDvFinish, st);
    dim3 grid(2, N_v);
    dv_finish_kernel<<<grid, kFinThreads, 0, st>>>(v, rnorm_v, C, P, dv);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

}  // namespace avt
