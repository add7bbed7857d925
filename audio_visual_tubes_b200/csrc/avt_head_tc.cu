// AVT_PRECISION_BF16: the all-pairs contraction S[n,k,p] = <v_hat[n,:,p], a_hat[k,:]>
// (model.py:125) on the 5th-generation tensor cores, with the tri-map sigmoid and the masked
// pooling (model.py:135,140) applied straight out of tensor memory -- S never reaches HBM.
//
// Mapping: UMMA M = 128 audio rows (one TMEM lane each), UMMA N = 208 >=
// P = 196 positions of ONE image, K = C = 512 channels.  A TMEM lane then owns one (n, k)
// pair and the pooling over p is a private loop over TMEM columns: no cross-thread
// reduction anywhere.
//
//   operands   A = a_hat as bf16 [n_a_pad][C]   K-major, 128-byte swizzle, RESIDENT in shared
//                  memory for all images a block visits (128 KB)
//              B = v_hat as bf16 [N_v*P][C]     K-major, streamed by TMA, 3 stages x 26 KB;
//                  rows 196..207 of a box belong to the next image (or are zero-filled past
//                  the end) and land in accumulator columns nobody reads
//   TMEM       2 accumulators x 256 columns: the epilogue of tile i overlaps the MMAs of i+1
//   warps      0: TMA producer   1: MMA issuer + TMEM owner   2-5: epilogue (one lane quarter each)
//   grid       persistent: block = (audio tile, image group); images of a group are visited in
//              step by the blocks of all audio tiles, so each image is read from HBM once
#include "avt_common.cuh"
#include "avt_tc.cuh"

namespace avt {
using namespace tc;

constexpr int kTcThreads = 192;
constexpr int kBM = 128;
constexpr int kBN = 208;
constexpr int kBK = 64;
constexpr int kStagesB = 3;
constexpr int kMaxKC = 8;                               // C <= 512
constexpr int kABytes = kBM * kBK * 2;                  // 16384 per K chunk
constexpr int kBBytes = kBN * kBK * 2;                  // 26624 per stage
constexpr int kAccCols = 256;
constexpr int kTmemCols = 512;
constexpr size_t kTcSmemBytes = 1024 + (size_t)kMaxKC * kABytes + (size_t)kStagesB * kBBytes + 256;

// kGrad = false: forward (writes sim / den / logits).
// kGrad = true : backward stage 1.  The same contraction is recomputed and the epilogue turns
//   every cell into its gradient  G = gsim/den * w (1 + (S - sim)(1 - w)/tau)  (+ GA on the
//   positive column), stored as bf16 in G[n*P + p][k] (row pitch n_a_pad) --
//   the K-major A operand of the dv contraction and the MN-major A operand of the d_ahat one.
template <bool kGrad>
__global__ void __launch_bounds__(kTcThreads, 1)
sim_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                const int32_t* __restrict__ pos_idx, int col_offset, int N_v, int N_a, int n_mt,
                int groups, int KC, int P, HeadParams hp, int L, float* __restrict__ logits,
                float* __restrict__ sim, float* __restrict__ den,
                const float* __restrict__ gsim, const float* __restrict__ GA,
                __nv_bfloat16* __restrict__ G, int n_a_pad) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = base;
    uint8_t* sB = sA + (size_t)kMaxKC * kABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + (size_t)kStagesB * kBBytes);
    uint64_t* full_b = bars;                             // [kStagesB]
    uint64_t* empty_b = bars + kStagesB;                 // [kStagesB]
    uint64_t* a_full = bars + 2 * kStagesB;
    uint64_t* a_empty = a_full + 1;
    uint64_t* tmem_full = a_empty + 1;                   // [2]
    uint64_t* tmem_empty = tmem_full + 2;                // [2]
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mt0 = blockIdx.x / groups, g = blockIdx.x % groups, mt_step = gridDim.x / groups;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&map_a);
        prefetch_tmap(&map_b);
        for (int i = 0; i < kStagesB; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 128); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_holder, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;

    if (warp == 0) {
        if (lane == 0) {                                 // ---------------- TMA producer
            int stage = 0; uint32_t phase = 0, a_par = 0;
            for (int mt = mt0; mt < n_mt; mt += mt_step) {
                if (mt != mt0) { mbar_wait(a_empty, a_par); a_par ^= 1; }
                mbar_arrive_expect_tx(a_full, (uint32_t)(KC * kABytes));
                for (int kc = 0; kc < KC; ++kc)
                    tma_load_2d(sA + (size_t)kc * kABytes, &map_a, a_full, kc * kBK, mt * kBM);
                for (int n = g; n < N_v; n += groups) {
                    for (int kc = 0; kc < KC; ++kc) {
                        mbar_wait(&empty_b[stage], phase ^ 1);
                        mbar_arrive_expect_tx(&full_b[stage], (uint32_t)kBBytes);
                        tma_load_2d(sB + (size_t)stage * kBBytes, &map_b, &full_b[stage], kc * kBK, n * P);
                        if (++stage == kStagesB) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {                                 // ---------------- MMA issuer
            const uint32_t idesc = make_idesc_bf16(kBM, kBN, 0, 0);
            int stage = 0; uint32_t phase = 0, a_par = 0;
            uint32_t tile = 0;
            for (int mt = mt0; mt < n_mt; mt += mt_step) {
                mbar_wait(a_full, a_par); a_par ^= 1;
                tc_fence_after();
                for (int n = g; n < N_v; n += groups, ++tile) {
                    const uint32_t buf = tile & 1u, use = tile >> 1;
                    mbar_wait(&tmem_empty[buf], (use & 1u) ^ 1u);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * kAccCols;
                    for (int kc = 0; kc < KC; ++kc) {
                        mbar_wait(&full_b[stage], phase);
                        tc_fence_after();
                        const uint32_t a_addr = smem_u32(sA + (size_t)kc * kABytes);
                        const uint32_t b_addr = smem_u32(sB + (size_t)stage * kBBytes);
#pragma unroll
                        for (int k = 0; k < kBK / 16; ++k) {
                            const uint64_t ad = make_smem_desc(a_addr + k * 32, 16, 1024);
                            const uint64_t bd = make_smem_desc(b_addr + k * 32, 16, 1024);
                            umma_bf16_ss(d_tmem, ad, bd, idesc, (uint32_t)((kc | k) != 0));
                        }
                        umma_commit(&empty_b[stage]);    // frees the stage once these MMAs retire
                        if (++stage == kStagesB) { stage = 0; phase ^= 1; }
                    }
                    umma_commit(&tmem_full[buf]);
                }
                umma_commit(a_empty);
            }
        }
    } else {                                             // ---------------- epilogue warps 2..5
        const int q = warp & 3;                          // TMEM lane quarter this warp may touch
        const int row = q * 32 + lane;
        const float c1 = -kLog2e / hp.tau, c0 = hp.eps * kLog2e / hp.tau;
        const float itemp = 1.0f / hp.temp, itau = 1.0f / hp.tau;
        uint32_t tile = 0;
        for (int mt = mt0; mt < n_mt; mt += mt_step) {
            const int k = mt * kBM + row;
            for (int n = g; n < N_v; n += groups, ++tile) {
                const uint32_t buf = tile & 1u, use = tile >> 1;
                const int pos = pos_idx ? pos_idx[n] : col_offset + n;
                float sm = 0.f, coef = 0.f;
                if (kGrad && k < N_a) {
                    sm = sim[(size_t)n * N_a + k];
                    coef = gsim[(size_t)n * N_a + k] / den[(size_t)n * N_a + k];
                }
                const bool is_pos = kGrad && (k == pos);
                const float* ga = GA + (size_t)n * P;
                __nv_bfloat16* gout = G + (size_t)n * P * n_a_pad + k;
                mbar_wait(&tmem_full[buf], use & 1u);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * kAccCols;
                float num = 0.f, dn = 0.f;
                int c = 0;
                for (; c + 32 <= P; c += 32) {
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + c, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float s = __uint_as_float(r[i]);
                        const float w = fast_rcp(1.0f + fast_ex2(fmaf(s, c1, c0)));
                        if (kGrad) {
                            const float x = fmaf(-w, itau, itau);            // (1 - w) / tau
                            float gv = coef * w * fmaf(s - sm, x, 1.0f);
                            if (is_pos) gv += ga[c + i];
                            gout[(size_t)(c + i) * n_a_pad] = __float2bfloat16_rn(gv);
                        } else {
                            dn += w;
                            num = fmaf(w, s, num);
                        }
                    }
                }
                for (; c < P; c += 4) {
                    uint32_t r[4];
                    tmem_ld_32x4(taddr + c, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float s = __uint_as_float(r[i]);
                        const float w = fast_rcp(1.0f + fast_ex2(fmaf(s, c1, c0)));
                        if (kGrad) {
                            const float x = fmaf(-w, itau, itau);
                            float gv = coef * w * fmaf(s - sm, x, 1.0f);
                            if (is_pos) gv += ga[c + i];
                            gout[(size_t)(c + i) * n_a_pad] = __float2bfloat16_rn(gv);
                        } else {
                            dn += w;
                            num = fmaf(w, s, num);
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&tmem_empty[buf]);           // accumulator may be overwritten
                if (!kGrad && k < N_a) {
                    const float smv = num / dn;
                    const float m = (k == pos) ? (1.0f - hp.mask_scale) : 1.0f;
                    sim[(size_t)n * N_a + k] = smv;
                    den[(size_t)n * N_a + k] = dn;
                    logits[(size_t)n * L + 1 + k] = smv * m * itemp;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

static int launch_sim_tc(bool grad, const __nv_bfloat16* vt, const __nv_bfloat16* a_bf,
                         const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                         HeadParams hp, float* logits, float* sim, float* den, const float* gsim,
                         const float* GA, __nv_bfloat16* G, int n_a_pad, cudaStream_t st) {
    AVT_REQUIRE(C % kBK == 0 && C <= kMaxKC * kBK, "bf16 head: C=%d must be a multiple of 64, <= 512", C);
    AVT_REQUIRE(P % 4 == 0 && P <= kBN, "bf16 head: P=%d must be a multiple of 4, <= 208", P);
    const int n_mt = (N_a + kBM - 1) / kBM;
    CUtensorMap map_a, map_b;
    if (!make_tmap_bf16_2d(&map_a, a_bf, (uint64_t)n_mt * kBM, (uint64_t)C, (uint64_t)C * 2, kBK, kBM) ||
        !make_tmap_bf16_2d(&map_b, vt, (uint64_t)N_v * P, (uint64_t)C, (uint64_t)C * 2, kBK, kBN)) {
        set_error("bf16 head: cuTensorMapEncodeTiled failed");
        return AVT_ERR_CUDA;
    }
    int dev = 0, sms = 148;
    AVT_CHECK_CUDA(cudaGetDevice(&dev));
    AVT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int mt_par = n_mt < sms ? n_mt : sms;
    int groups = sms / mt_par;
    if (groups > N_v) groups = N_v;
    if (groups < 1) groups = 1;
    const int grid = groups * mt_par;
    const int L = N_a + 1 + (hp.use_neg ? 1 : 0);
    if (grad) {
        AVT_CHECK_CUDA(cudaFuncSetAttribute(sim_bf16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)kTcSmemBytes));
        sim_bf16_kernel<true><<<grid, kTcThreads, kTcSmemBytes, st>>>(
            map_a, map_b, pos_idx, col_offset, N_v, N_a, n_mt, groups, C / kBK, P, hp, L, logits, sim, den,
            gsim, GA, G, n_a_pad);
    } else {
        AVT_CHECK_CUDA(cudaFuncSetAttribute(sim_bf16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)kTcSmemBytes));
        sim_bf16_kernel<false><<<grid, kTcThreads, kTcSmemBytes, st>>>(
            map_a, map_b, pos_idx, col_offset, N_v, N_a, n_mt, groups, C / kBK, P, hp, L, logits, sim, den,
            nullptr, nullptr, nullptr, 0);
    }
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

int launch_sim_bf16(const __nv_bfloat16* vt, const __nv_bfloat16* a_bf, const int32_t* pos_idx,
                    int col_offset, int N_v, int N_a, int C, int P, HeadParams hp, float* logits,
                    float* sim, float* den, cudaStream_t st) {
    ProfScope _ps(kProfSim, st);
    return launch_sim_tc(false, vt, a_bf, pos_idx, col_offset, N_v, N_a, C, P, hp, logits, sim, den,
                         nullptr, nullptr, nullptr, 0, st);
}

// backward stage 1: cell gradients G (bf16, [N_v*P][n_a_pad])
int launch_grad_cells_bf16(const __nv_bfloat16* vt, const __nv_bfloat16* a_bf, const int32_t* pos_idx,
                           int col_offset, int N_v, int N_a, int n_a_pad, int C, int P, HeadParams hp,
                           const float* sim, const float* den, const float* gsim, const float* GA,
                           __nv_bfloat16* G, cudaStream_t st) {
    ProfScope _ps(kProfBwdG, st);
    return launch_sim_tc(true, vt, a_bf, pos_idx, col_offset, N_v, N_a, C, P, hp, nullptr,
                         const_cast<float*>(sim), const_cast<float*>(den), gsim, GA, G, n_a_pad, st);
}

}  // namespace avt
