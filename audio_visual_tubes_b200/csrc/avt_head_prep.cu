// Row-wise pieces of the head: audio normalisation (+ adjoint), the single pass over
// the visual maps that yields 1/||v||, the positive map A and the fp16 K-major copy
// for the tensor-core contraction, and the per-row finish (tri-map, sim1/sim2,
// weighted_A, cross-entropy).  HBM-bound kernels: coalesced along P, 16 independent
// loads in flight per thread.
#include "avt_common.cuh"
#include "avt_rows.cuh"

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

// also zeroes *zero_word (the bound of the cell gradients that the fused row-finish kernel max-reduces
// into): one memset node less between the contraction and the row kernel
__global__ void audio_to_f16_kernel(const float* __restrict__ a_hat, int N_a, int N_a_pad, int C,
                                     __half* __restrict__ out, unsigned int* __restrict__ zero_word) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 && zero_word) *zero_word = 0u;
    size_t total = (size_t)N_a_pad * C;
    if (i >= total) return;
    size_t row = i / C;
    out[i] = __float2half_rn(row < (size_t)N_a ? a_hat[i] : 0.f);
}

int launch_audio_to_f16(const float* a_hat, int N_a, int N_a_pad, int C, __half* out,
                         unsigned int* zero_word, cudaStream_t st) {
    size_t total = (size_t)N_a_pad * C;
    audio_to_f16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(a_hat, N_a, N_a_pad, C, out, zero_word);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

// ------------------------------------------------------------------ visual prep pass
// model.py:117 (norm over C per position), :124 (positive map), :150 (mean_c of the
// normalised map) and, for the tensor-core path, v_hat as fp16 K-major [n][p][c].
// ONE read of v: a block owns 32 positions of one image, its 16 warps own C/16 channels
// each, and every thread keeps its <= 32 values in registers between the reduction over
// the channels and the normalised write-out (the earlier two-sweep version re-read v from
// DRAM: ncu showed 750 MB of reads for a 411 MB input).  Loads are 128-byte rows of
// consecutive positions; stores are 64 contiguous bytes per thread.
constexpr int kPrepThreads = 512;
constexpr int kPrepWarps = 16;
constexpr int kPrepMaxCpw = 32;                  // channels per warp, C <= 512

// kC / kP: compile-time C and P of the reference's shapes (512 channels, 14 x 14 positions) so that
// every guard folds and every load / store offset is an immediate; 0 = take the run-time value.
// TIn: element type of the caller's maps -- fp32 from the reference's fp32 backbones, or fp16 / bf16
// when the backbone ran under autocast (halves the bytes of this pass and of a host->device copy).
__device__ __forceinline__ float ld_in(const float* p) { return __ldg(p); }
__device__ __forceinline__ float ld_in(const __half* p) { return __half2float(__ldg(p)); }
__device__ __forceinline__ float ld_in(const __nv_bfloat16* p) { return __bfloat162float(__ldg(p)); }

template <bool kWriteF16, int kC, int kP, typename TIn>
__global__ void __launch_bounds__(kPrepThreads, 2)
prep_visual_kernel(const TIn* __restrict__ v, const float* __restrict__ a_hat,
                   const int32_t* __restrict__ pos_idx, int col_offset, int N_a, int C_rt, int P_rt,
                   float* __restrict__ rnorm_v, float* __restrict__ A, float* __restrict__ msum,
                   __half* __restrict__ vt, HeadScratch* scratch) {
    const int C = kC ? kC : C_rt, P = kP ? kP : P_rt;
    extern __shared__ float smem[];
    float* s_a = smem;                       // [C]  positive audio row
    float* s_red = smem + C;                 // [3][kPrepWarps][32]
    const int n = blockIdx.y;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int p = blockIdx.x * 32 + lane;
    const bool active = p < P;
    const int cpw = C / kPrepWarps;          // channels of this warp: [w*cpw, (w+1)*cpw)
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) scratch->rows_done = 0u;

    const int pos = pos_idx ? pos_idx[n] : col_offset + n;
    const bool has_pos = (pos >= 0 && pos < N_a);
    for (int c = threadIdx.x; c < C; c += kPrepThreads)
        s_a[c] = has_pos ? a_hat[(size_t)pos * C + c] : 0.f;

    float x[kPrepMaxCpw];
    const TIn* vp = v + ((size_t)n * C + (size_t)w * cpw) * P + p;
#pragma unroll
    for (int i = 0; i < kPrepMaxCpw; ++i)
        x[i] = (active && i < cpw) ? ld_in(vp + (size_t)i * P) : 0.f;
    __syncthreads();                         // s_a ready
    float ssq = 0.f, dot = 0.f, sum = 0.f;
#pragma unroll
    for (int i = 0; i < kPrepMaxCpw; ++i) {
        if (i < cpw) {
            ssq = fmaf(x[i], x[i], ssq);
            dot = fmaf(x[i], s_a[w * cpw + i], dot);
            sum += x[i];
        }
    }
    s_red[(0 * kPrepWarps + w) * 32 + lane] = ssq;
    s_red[(1 * kPrepWarps + w) * 32 + lane] = dot;
    s_red[(2 * kPrepWarps + w) * 32 + lane] = sum;
    __syncthreads();
    float t0 = 0.f;
#pragma unroll
    for (int j = 0; j < kPrepWarps; ++j) t0 += s_red[(0 * kPrepWarps + j) * 32 + lane];
    const float rn = 1.0f / fmaxf(sqrtf(t0), kNormEps);
    if (w == 0 && active) {
        float t1 = 0.f, t2 = 0.f;
#pragma unroll
        for (int j = 0; j < kPrepWarps; ++j) {
            t1 += s_red[(1 * kPrepWarps + j) * 32 + lane];
            t2 += s_red[(2 * kPrepWarps + j) * 32 + lane];
        }
        const size_t o = (size_t)n * P + p;
        rnorm_v[o] = rn;
        A[o] = t1 * rn;
        msum[o] = t2 * rn / (float)C;
    }
    if (kWriteF16) {
        // transpose through shared memory so the K-major rows leave as 512-byte bursts:
        // tile[p][c] fp16 with a 16-byte pad per row (16-byte accesses stay conflict-free)
        uint8_t* s_tile = reinterpret_cast<uint8_t*>(s_red + 3 * kPrepWarps * 32);
        const int pitch = C * 2 + 16;
#pragma unroll
        for (int i = 0; i < kPrepMaxCpw; i += 8) {
            if (i < cpw) {
                uint32_t wds[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    __half2 h = __floats2half2_rn(x[i + 2 * e] * rn, x[i + 2 * e + 1] * rn);
                    wds[e] = *reinterpret_cast<uint32_t*>(&h);
                }
                if (cpw % 8 == 0) {
                    *reinterpret_cast<uint4*>(s_tile + (size_t)lane * pitch + (w * cpw + i) * 2) =
                        make_uint4(wds[0], wds[1], wds[2], wds[3]);
                } else {                         // cpw a multiple of 4 only: 8-byte pieces
                    uint8_t* o = s_tile + (size_t)lane * pitch + (w * cpw + i) * 2;
                    *reinterpret_cast<uint2*>(o) = make_uint2(wds[0], wds[1]);
                    if (i + 4 < cpw) *reinterpret_cast<uint2*>(o + 8) = make_uint2(wds[2], wds[3]);
                }
            }
        }
        __syncthreads();
        // warp w writes rows w, w+16 of the tile; a lane moves 16 bytes per step
        const int p0 = blockIdx.x * 32;
        for (int r = w; r < 32 && p0 + r < P; r += kPrepWarps) {
            uint8_t* dst = reinterpret_cast<uint8_t*>(vt + ((size_t)n * P + p0 + r) * C);
            for (int b = lane * 16; b < C * 2; b += 32 * 16)
                *reinterpret_cast<uint4*>(dst + b) = *reinterpret_cast<const uint4*>(s_tile + (size_t)r * pitch + b);
        }
    }
}

int launch_prep_visual(const void* v_any, int v_dtype, const float* a_hat, const int32_t* pos_idx, int col_offset,
                       int N_v, int N_a, int C, int P, float* rnorm_v, float* A, float* msum,
                       __half* vt_f16, HeadScratch* scratch, cudaStream_t st) {
    ProfScope _ps(kProfPrep, st);
    AVT_REQUIRE(C % 64 == 0 && C <= kPrepWarps * kPrepMaxCpw, "prep: C=%d must be a multiple of 64, <= 512", C);
    dim3 grid((P + 31) / 32, N_v);
    size_t smem = (size_t)(C + 3 * kPrepWarps * 32) * sizeof(float) + (vt_f16 ? (size_t)32 * (C * 2 + 16) : 0);
    const bool ref_shape = (C == 512 && P == 196);
#define AVT_PREP(BF, KC, KP, T)                                                                     \
    prep_visual_kernel<BF, KC, KP, T><<<grid, kPrepThreads, smem, st>>>(static_cast<const T*>(v_any), a_hat, pos_idx, \
                                                                        col_offset, N_a, C, P, rnorm_v, A, msum, vt_f16, scratch)
    if (v_dtype == AVT_DTYPE_F32) {
        if (vt_f16) { if (ref_shape) AVT_PREP(true, 512, 196, float); else AVT_PREP(true, 0, 0, float); }
        else { if (ref_shape) AVT_PREP(false, 512, 196, float); else AVT_PREP(false, 0, 0, float); }
    } else {
        AVT_REQUIRE(vt_f16 != nullptr, "prep: fp16 / bf16 maps are taken by the tensor-core path only");
        if (v_dtype == AVT_DTYPE_F16) { if (ref_shape) AVT_PREP(true, 512, 196, __half); else AVT_PREP(true, 0, 0, __half); }
        else { if (ref_shape) AVT_PREP(true, 512, 196, __nv_bfloat16); else AVT_PREP(true, 0, 0, __nv_bfloat16); }
    }
#undef AVT_PREP
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
                   HeadScratch* scratch, const int32_t* __restrict__ pos_idx, int col_offset,
                   const float* __restrict__ sim, int col_mult,
                   float fuse_scale, float* __restrict__ gsim, float* __restrict__ GA,
                   const float* __restrict__ den, unsigned int* __restrict__ gmax_bits) {
    __shared__ float s_tmp[kRowThreads / 32];
    __shared__ float s_multi[kRowThreads / 32][5];
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
    {   // five sums behind one pair of barriers (fixed order: deterministic)
        sp = warp_sum(sp); spa = warp_sum(spa); sn = warp_sum(sn); sna = warp_sum(sna); spp = warp_sum(spp);
        const int w = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) {
            s_multi[w][0] = sp; s_multi[w][1] = spa; s_multi[w][2] = sn; s_multi[w][3] = sna; s_multi[w][4] = spp;
        }
        __syncthreads();
        sp = spa = sn = sna = spp = 0.f;
#pragma unroll
        for (int i = 0; i < kRowThreads / 32; ++i) {
            sp += s_multi[i][0]; spa += s_multi[i][1]; sn += s_multi[i][2]; sna += s_multi[i][3]; spp += s_multi[i][4];
        }
    }
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
    // log-sum-exp over the row; columns 1..N_a were written by the similarity kernel.
    // Tube mode (col_mult = T > 1, train_3D.py:129-130): audio row k stands for T identical
    // logit columns, of which ONE -- in the row's positive audio -- carries the -99 mask
    // (model.py:47,57).  Column k then weighs T (log T added to its logit); the positive audio
    // contributes its masked logit once and its unmasked one T - 1 times.
    const float log_t = (col_mult > 1) ? logf((float)col_mult) : 0.f;
    const int posk = (col_mult > 1) ? (pos_idx ? pos_idx[n] : col_offset + n) : -1;
    float y_pos_un = -3.0e38f;                               // unmasked logit of the positive audio
    if (col_mult > 1 && posk >= 0 && posk < N_a)
        y_pos_un = sim[(size_t)n * N_a + posk] / hp.temp + logf((float)(col_mult - 1));
    float m = l_first;
    if (hp.use_neg) m = fmaxf(m, l_last);
    m = fmaxf(m, y_pos_un);
    for (int j = 1 + threadIdx.x; j <= N_a; j += kRowThreads) m = fmaxf(m, lg[j] + ((j - 1 == posk) ? 0.f : log_t));
    m = block_max(m, s_tmp);
    float s = 0.f;
    for (int j = 1 + threadIdx.x; j <= N_a; j += kRowThreads) s += expf(lg[j] + ((j - 1 == posk) ? 0.f : log_t) - m);
    if (threadIdx.x == 0) {
        s += expf(l_first - m);
        if (hp.use_neg) s += expf(l_last - m);
        if (y_pos_un > -1.0e38f) s += expf(y_pos_un - m);
    }
    s = block_sum(s, s_tmp);
    if (threadIdx.x == 0) loss_rows[n] = (m + logf(s)) - l_first;

    // fused training step (gsim != NULL): the row's upstream gradients right here -- the softmax
    // needs nothing but this row's log-sum-exp, so the separate row_grad_kernel launch (12-18 us,
    // fixed, on the critical path of small shards) disappears
    if (gsim != nullptr)
        row_grad_row(n, m + logf(s), l_first, l_last, A, logits, pos_idx, col_offset, N_a, P, hp, fuse_scale,
                     nullptr, nullptr, gsim, GA, sim, col_mult, nullptr, den, gmax_bits);

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
                       float* loss_rows, float* loss_sum, HeadScratch* scratch, const int32_t* pos_idx,
                       int col_offset, const float* sim, int col_mult, float fuse_scale, float* gsim,
                       float* GA, const float* den, unsigned int* gmax_bits, cudaStream_t st) {
    ProfScope _ps(kProfFinish, st);
    static_assert(kRowThreads == kRowGradThreads, "fused row gradients share the block shape");
    finish_rows_kernel<<<N_v, kRowThreads, 0, st>>>(A, msum, N_v, N_a, P, hp, Pos, Neg, weighted_A,
                                                    logits, loss_rows, loss_sum, scratch, pos_idx,
                                                    col_offset, sim, col_mult, fuse_scale, gsim, GA, den,
                                                    gmax_bits);
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
