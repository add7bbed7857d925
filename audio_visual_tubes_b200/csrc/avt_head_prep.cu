// Row-wise pieces of the head: audio normalisation (+ adjoint), the single pass over
// the visual maps that yields 1/||v||, the positive map A and the bf16 K-major copy
// for the tensor-core contraction, and the per-row finish (tri-map, sim1/sim2,
// weighted_A, cross-entropy).  HBM-bound kernels: coalesced along P, 16 independent
// loads in flight per thread.
#include "avt_common.cuh"

namespace avt {

// ------------------------------------------------------------------ audio normalise
// model.py:122  aud = F.normalize(aud, dim=1)        one warp per audio row
__global__ void audio_normalize_kernel(const float* __restrict__ a, int N_a, int C,
                                       float* __restrict__ a_hat, float* __restrict__ rnorm_a) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= N_a) return;
    const float* r = a + (size_t)row * C;
    float ss = 0.f;
    for (int c = lane; c < C; c += 32) { float x = r[c]; ss = fmaf(x, x, ss); }
    ss = warp_sum(ss);
    float rn = 1.0f / fmaxf(sqrtf(ss), kNormEps);
    for (int c = lane; c < C; c += 32) a_hat[(size_t)row * C + c] = r[c] * rn;
    if (lane == 0) rnorm_a[row] = rn;
}

__global__ void audio_normalize_bwd_kernel(const float* __restrict__ a_hat,
                                           const float* __restrict__ rnorm_a,
                                           const float* __restrict__ d_ahat, int N_a, int C,
                                           float* __restrict__ da) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= N_a) return;
    const float* h = a_hat + (size_t)row * C;
    const float* g = d_ahat + (size_t)row * C;
    float dot = 0.f;
    for (int c = lane; c < C; c += 32) dot = fmaf(h[c], g[c], dot);
    dot = warp_sum(dot);
    float rn = rnorm_a[row];
    for (int c = lane; c < C; c += 32) da[(size_t)row * C + c] = rn * (g[c] - h[c] * dot);
}

__global__ void audio_to_bf16_kernel(const float* __restrict__ a_hat, int N_a, int N_a_pad, int C,
                                     __nv_bfloat16* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t total = (size_t)N_a_pad * C;
    if (i >= total) return;
    size_t row = i / C;
    out[i] = __float2bfloat16_rn(row < (size_t)N_a ? a_hat[i] : 0.f);
}

int launch_audio_to_bf16(const float* a_hat, int N_a, int N_a_pad, int C, __nv_bfloat16* out,
                         cudaStream_t st) {
    size_t total = (size_t)N_a_pad * C;
    audio_to_bf16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(a_hat, N_a, N_a_pad, C, out);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

// ------------------------------------------------------------------ visual prep pass
// model.py:117 (norm over C per position), :124 (positive map), :150 (mean_c of the
// normalised map).  grid (2, N_v): each block owns one half of the positions of one
// image; 4 channel groups x 128 lanes; every thread streams 16 channels per step.
// With kWriteBf16 a second sweep emits v_hat = v * rnorm as bf16, transposed to [n][p][c].
constexpr int kPrepThreads = 512;
constexpr int kPrepGroups = 4;

template <bool kWriteBf16>
__global__ void __launch_bounds__(kPrepThreads)
prep_visual_kernel(const float* __restrict__ v, const float* __restrict__ a_hat,
                   const int32_t* __restrict__ pos_idx, int col_offset, int N_a, int C, int P,
                   float* __restrict__ rnorm_v, float* __restrict__ A, float* __restrict__ msum,
                   __nv_bfloat16* __restrict__ vt, HeadScratch* scratch) {
    extern __shared__ float smem[];
    float* s_a = smem;                       // [C]  positive audio row
    float* s_red = smem + C;                 // [3][kPrepGroups][128]
    const int n = blockIdx.y;
    const int PT = (P + 1) / 2;
    const int lane = threadIdx.x & 127;
    const int g = threadIdx.x >> 7;
    const int p = blockIdx.x * PT + lane;
    const bool active = (lane < PT) && (p < P);
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) scratch->rows_done = 0u;

    int pos = pos_idx ? pos_idx[n] : col_offset + n;
    const bool has_pos = (pos >= 0 && pos < N_a);
    for (int c = threadIdx.x; c < C; c += kPrepThreads)
        s_a[c] = has_pos ? a_hat[(size_t)pos * C + c] : 0.f;
    __syncthreads();

    float ssq = 0.f, dot = 0.f, sum = 0.f;
    const float* vp = v + (size_t)n * C * P + p;
    if (active) {
        for (int c0 = g * 16; c0 < C; c0 += 16 * kPrepGroups) {
            float x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = __ldg(vp + (size_t)(c0 + i) * P);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                ssq = fmaf(x[i], x[i], ssq);
                dot = fmaf(x[i], s_a[c0 + i], dot);
                sum += x[i];
            }
        }
    }
    s_red[(0 * kPrepGroups + g) * 128 + lane] = ssq;
    s_red[(1 * kPrepGroups + g) * 128 + lane] = dot;
    s_red[(2 * kPrepGroups + g) * 128 + lane] = sum;
    __syncthreads();
    float rn = 0.f;
    if (active) {
        float t0 = 0.f, t1 = 0.f, t2 = 0.f;
#pragma unroll
        for (int j = 0; j < kPrepGroups; ++j) {
            t0 += s_red[(0 * kPrepGroups + j) * 128 + lane];
            t1 += s_red[(1 * kPrepGroups + j) * 128 + lane];
            t2 += s_red[(2 * kPrepGroups + j) * 128 + lane];
        }
        rn = 1.0f / fmaxf(sqrtf(t0), kNormEps);
        if (g == 0) {
            size_t o = (size_t)n * P + p;
            rnorm_v[o] = rn;
            A[o] = t1 * rn;
            msum[o] = t2 * rn / (float)C;
        }
    }
    if (kWriteBf16 && active) {
        // second sweep (L2-resident: this block just streamed these lines): the
        // normalised map as bf16, K-major [n][p][c] for the tensor-core contraction
        __nv_bfloat16* op = vt + ((size_t)n * P + p) * C;
        for (int c0 = g * 16; c0 < C; c0 += 16 * kPrepGroups) {
            float x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = __ldg(vp + (size_t)(c0 + i) * P) * rn;
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                __nv_bfloat162 h = __floats2bfloat162_rn(x[2 * i], x[2 * i + 1]);
                w[i] = *reinterpret_cast<uint32_t*>(&h);
            }
            uint4* o = reinterpret_cast<uint4*>(op + c0);
            o[0] = make_uint4(w[0], w[1], w[2], w[3]);
            o[1] = make_uint4(w[4], w[5], w[6], w[7]);
        }
    }
}

int launch_prep_visual(const float* v, const float* a_hat, const int32_t* pos_idx, int col_offset,
                       int N_v, int N_a, int C, int P, float* rnorm_v, float* A, float* msum,
                       __nv_bfloat16* vt_bf16, HeadScratch* scratch, cudaStream_t st) {
    ProfScope _ps(kProfPrep, st);
    dim3 grid(2, N_v);
    size_t smem = (size_t)(C + 3 * kPrepGroups * 128) * sizeof(float);
    if (vt_bf16)
        prep_visual_kernel<true><<<grid, kPrepThreads, smem, st>>>(v, a_hat, pos_idx, col_offset, N_a,
                                                                  C, P, rnorm_v, A, msum, vt_bf16, scratch);
    else
        prep_visual_kernel<false><<<grid, kPrepThreads, smem, st>>>(v, a_hat, pos_idx, col_offset, N_a,
                                                                   C, P, rnorm_v, A, msum, nullptr, scratch);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

// ------------------------------------------------------------------ per-row finish
// model.py:128-133 (Pos / Neg), :138,:141 (sim1, sim2), :143-146 (first / last logit),
// :148-152 (weighted_A), CrossEntropyLoss with target 0.  One block per visual row.
constexpr int kRowThreads = 256;

__device__ __forceinline__ float block_sum(float x, float* s_tmp) {
    x = warp_sum(x);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) s_tmp[w] = x;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kRowThreads / 32; ++i) t += s_tmp[i];
    return t;
}
__device__ __forceinline__ float block_max(float x, float* s_tmp) {
    x = warp_max(x);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) s_tmp[w] = x;
    __syncthreads();
    float t = s_tmp[0];
#pragma unroll
    for (int i = 1; i < kRowThreads / 32; ++i) t = fmaxf(t, s_tmp[i]);
    return t;
}

__global__ void __launch_bounds__(kRowThreads)
finish_rows_kernel(const float* __restrict__ A, const float* __restrict__ msum, int N_v, int N_a,
                   int P, HeadParams hp, float* __restrict__ Pos, float* __restrict__ Neg,
                   float* __restrict__ weighted_A, float* __restrict__ logits,
                   float* __restrict__ loss_rows, float* __restrict__ loss_sum,
                   HeadScratch* scratch) {
    __shared__ float s_tmp[kRowThreads / 32];
    __shared__ bool s_last;
    const int n = blockIdx.x;
    const int L = N_a + 1 + (hp.use_neg ? 1 : 0);
    const float itau = 1.0f / hp.tau;

    float sp = 0.f, spa = 0.f, sn = 0.f, sna = 0.f, spp = 0.f;
    for (int p = threadIdx.x; p < P; p += kRowThreads) {
        float a = A[(size_t)n * P + p];
        float z1 = (a - hp.eps) * itau;
        float pos = 1.0f / (1.0f + expf(-z1));
        float neg;
        if (hp.trimap) {
            float z2 = (a - hp.eps2) * itau;
            neg = 1.0f / (1.0f + expf(z2));           // 1 - sigmoid(z2) without cancellation
        } else {
            neg = 1.0f / (1.0f + expf(z1));
        }
        Pos[(size_t)n * P + p] = pos;
        Neg[(size_t)n * P + p] = neg;
        sp += pos; spa = fmaf(pos, a, spa);
        sn += neg; sna = fmaf(neg, a, sna);
        spp = fmaf(pos, pos, spp);
    }
    sp = block_sum(sp, s_tmp);
    spa = block_sum(spa, s_tmp);
    sn = block_sum(sn, s_tmp);
    sna = block_sum(sna, s_tmp);
    spp = block_sum(spp, s_tmp);
    const float l_first = (spa / sp) / hp.temp;
    const float l_last = (sna / sn) / hp.temp;
    if (weighted_A) {
        float rpn = 1.0f / fmaxf(sqrtf(spp), kNormEps);
        for (int p = threadIdx.x; p < P; p += kRowThreads) {
            size_t o = (size_t)n * P + p;
            weighted_A[o] = Pos[o] * rpn * msum[o];
        }
    }
    float* lg = logits + (size_t)n * L;
    if (threadIdx.x == 0) {
        lg[0] = l_first;
        if (hp.use_neg) lg[L - 1] = l_last;
    }
    // log-sum-exp over the row; columns 1..N_a were written by the similarity kernel
    float m = l_first;
    if (hp.use_neg) m = fmaxf(m, l_last);
    for (int j = 1 + threadIdx.x; j <= N_a; j += kRowThreads) m = fmaxf(m, lg[j]);
    m = block_max(m, s_tmp);
    float s = 0.f;
    for (int j = 1 + threadIdx.x; j <= N_a; j += kRowThreads) s += expf(lg[j] - m);
    if (threadIdx.x == 0) {
        s += expf(l_first - m);
        if (hp.use_neg) s += expf(l_last - m);
    }
    s = block_sum(s, s_tmp);
    if (threadIdx.x == 0) loss_rows[n] = (m + logf(s)) - l_first;

    if (loss_sum == nullptr) return;
    // deterministic total: the last block to finish adds the rows in index order
    __threadfence();
    if (threadIdx.x == 0) {
        unsigned int done = atomicAdd(&scratch->rows_done, 1u);
        s_last = (done == (unsigned)N_v - 1u);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    float t = 0.f;
    for (int i = threadIdx.x; i < N_v; i += kRowThreads) t += __ldcg(loss_rows + i);
    t = block_sum(t, s_tmp);
    if (threadIdx.x == 0) { *loss_sum = t; scratch->rows_done = 0u; }
}

int launch_finish_rows(const float* A, const float* msum, int N_v, int N_a, int P, HeadParams hp,
                       float* Pos, float* Neg, float* weighted_A, float* logits,
                       float* loss_rows, float* loss_sum, HeadScratch* scratch, cudaStream_t st) {
    ProfScope _ps(kProfFinish, st);
    finish_rows_kernel<<<N_v, kRowThreads, 0, st>>>(A, msum, N_v, N_a, P, hp, Pos, Neg, weighted_A,
                                                    logits, loss_rows, loss_sum, scratch);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

}  // namespace avt

// ------------------------------------------------------------------ C ABI
extern "C" int avt_audio_normalize(const float* a, int N_a, int C, float* a_hat, float* rnorm_a,
                                   void* stream) {
    AVT_REQUIRE(a && a_hat && rnorm_a, "avt_audio_normalize: null pointer");
    AVT_REQUIRE(N_a > 0 && C > 0, "avt_audio_normalize: bad shape N_a=%d C=%d", N_a, C);
    avt::audio_normalize_kernel<<<(N_a + 7) / 8, 256, 0, (cudaStream_t)stream>>>(a, N_a, C, a_hat, rnorm_a);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

extern "C" int avt_audio_normalize_backward(const float* a_hat, const float* rnorm_a,
                                            const float* d_ahat, int N_a, int C, float* da,
                                            void* stream) {
    AVT_REQUIRE(a_hat && rnorm_a && d_ahat && da, "avt_audio_normalize_backward: null pointer");
    AVT_REQUIRE(N_a > 0 && C > 0, "avt_audio_normalize_backward: bad shape");
    avt::audio_normalize_bwd_kernel<<<(N_a + 7) / 8, 256, 0, (cudaStream_t)stream>>>(a_hat, rnorm_a,
                                                                                  d_ahat, N_a, C, da);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}
