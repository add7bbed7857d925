// AVT_PRECISION_F16: the all-pairs contraction S[n,k,p] = <v_hat[n,:,p], a_hat[k,:]>
// (model.py:125) on the 5th-generation tensor cores, with the tri-map sigmoid and the masked
// pooling (model.py:135,140) applied straight out of tensor memory -- S never reaches HBM.
//
// Mapping: UMMA M = 128 audio rows (one TMEM lane each), UMMA N = 208 >=
// P = 196 positions of ONE image, K = C = 512 channels.  A TMEM lane then owns one (n, k)
// pair and the pooling over p is a private loop over TMEM columns: no cross-thread
// reduction anywhere.
//
//   operands   A = a_hat as fp16 [n_a_pad][C]   K-major, 128-byte swizzle, RESIDENT in shared
//                  memory for all images a block visits (128 KB)
//              B = v_hat as fp16 [N_v*P][C]     K-major, streamed by TMA, 3 stages x 26 KB;
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

constexpr int kTcThreads = 320;                         // warp 0 TMA, 1 MMA, 2..9 epilogue
constexpr int kBM = 128;
constexpr int kBN = 208;
constexpr int kBK = 64;
constexpr int kStagesB = 3;
constexpr int kMaxKC = 8;                               // C <= 512
constexpr int kABytes = kBM * kBK * 2;                  // 16384 per K chunk
constexpr int kBBytes = kBN * kBK * 2;                  // 26624 per stage
constexpr int kAccCols = 256;
constexpr int kTmemCols = 512;
constexpr size_t kTcSmemBytes = 1024 + (size_t)kMaxKC * kABytes + (size_t)kStagesB * kBBytes + 256 + 2 * kBM * 8;

// One chunk of N accumulator columns of one TMEM lane.  w = sigma((s - eps)/tau) =
// 1 / (1 + 2^(s*c1 + c0)).  The fast variant shares one MUFU.RCP between two cells
// (r = 1/(x0*x1); w0 = r*x1; w1 = r*x0), which needs the exponent clamped so the product
// cannot overflow; clamped cells only matter when EVERY weight of the (n, k) pair is below
// 2^-30, and in that case the caller redoes the pair with kExact = true (one RCP per cell,
// no clamp), so results never depend on the clamp.
struct CellCtx {
    float c1, c0, itau;          // sigmoid constants
    float sm, coef;              // backward: pooled sim and gsim / den of this (n, k)
};
constexpr float kExpClamp = 60.0f;
constexpr float kTinyDen = 9.313225746154785e-10f;      // 2^-30

template <bool kGrad, bool kExact, int N>
__device__ __forceinline__ void cell_chunk(const uint32_t (&r)[N], const CellCtx& cx, float& num,
                                           float& dn, __half* gp, uint32_t pitch) {
#pragma unroll
    for (int i = 0; i < N; i += 2) {
        const float s0 = __uint_as_float(r[i]), s1 = __uint_as_float(r[i + 1]);
        float w0, w1;
        if (kExact) {
            w0 = fast_rcp(1.0f + fast_ex2(fmaf(s0, cx.c1, cx.c0)));
            w1 = fast_rcp(1.0f + fast_ex2(fmaf(s1, cx.c1, cx.c0)));
        } else {
            const float x0 = 1.0f + fast_ex2(fminf(fmaf(s0, cx.c1, cx.c0), kExpClamp));
            const float x1 = 1.0f + fast_ex2(fminf(fmaf(s1, cx.c1, cx.c0), kExpClamp));
            const float rr = fast_rcp(x0 * x1);
            w0 = rr * x1;
            w1 = rr * x0;
        }
        if (kGrad) {
            const float g0 = (cx.coef * w0) * fmaf(s0 - cx.sm, fmaf(-w0, cx.itau, cx.itau), 1.0f);
            const float g1 = (cx.coef * w1) * fmaf(s1 - cx.sm, fmaf(-w1, cx.itau, cx.itau), 1.0f);
            gp[(uint32_t)i * pitch] = __float2half_rn(g0);
            gp[(uint32_t)(i + 1) * pitch] = __float2half_rn(g1);
        } else {
            dn += w0 + w1;
            num = fmaf(w0, s0, num);
            num = fmaf(w1, s1, num);
        }
    }
}

// all columns [c_begin, c_end) of one TMEM lane (c_end - c_begin a multiple of 4)
template <bool kGrad, bool kExact>
__device__ __forceinline__ void cell_sweep(uint32_t taddr, int c_begin, int c_end, const CellCtx& cx,
                                           float& num, float& dn, __half* gout, uint32_t pitch) {
    int c = c_begin;
    for (; c + 32 <= c_end; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c, r);
        tmem_ld_wait();
        cell_chunk<kGrad, kExact, 32>(r, cx, num, dn, gout + (uint32_t)c * pitch, pitch);
    }
    for (; c < c_end; c += 4) {
        uint32_t r[4];
        tmem_ld_32x4(taddr + c, r);
        tmem_ld_wait();
        cell_chunk<kGrad, kExact, 4>(r, cx, num, dn, gout + (uint32_t)c * pitch, pitch);
    }
}

// kGrad = false: forward (writes sim / den / logits).
// kGrad = true : backward stage 1.  The same contraction is recomputed and the epilogue turns
//   every cell into its gradient  G = gsim/den * w (1 + (S - sim)(1 - w)/tau)  (+ GA on the
//   positive column), stored as fp16 in G[n*P + p][k] (row pitch n_a_pad) --
//   the K-major A operand of the dv contraction and the MN-major A operand of the d_ahat one.
template <bool kGrad>
__global__ void __launch_bounds__(kTcThreads, 1)
sim_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                const int32_t* __restrict__ pos_idx, int col_offset, int N_v, int N_a, int n_mt,
                int groups, int KC, int P, HeadParams hp, int L, float* __restrict__ logits,
                float* __restrict__ sim, float* __restrict__ den,
                const float* __restrict__ gsim, const float* __restrict__ GA,
                __half* __restrict__ G, int n_a_pad, const unsigned int* __restrict__ gmax_bits) {
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
    float2* part = reinterpret_cast<float2*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2][kBM]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mt0 = blockIdx.x / groups, g = blockIdx.x % groups, mt_step = gridDim.x / groups;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&map_a);
        prefetch_tmap(&map_b);
        for (int i = 0; i < kStagesB; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 256); }
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
            const uint32_t idesc = make_idesc_f16(kBM, kBN, 0, 0);
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
                            umma_f16_ss(d_tmem, ad, bd, idesc, (uint32_t)((kc | k) != 0));
                        }
                        umma_commit(&empty_b[stage]);    // frees the stage once these MMAs retire
                        if (++stage == kStagesB) { stage = 0; phase ^= 1; }
                    }
                    umma_commit(&tmem_full[buf]);
                }
                umma_commit(a_empty);
            }
        }
    } else {                                             // ---------------- epilogue warps 2..9
        // two warps per TMEM lane quarter (one per scheduler pair): each takes half of the
        // image's positions, so every SM sub-partition has two warps to hide MUFU / TMEM latency
        const int q = warp & 3;                          // TMEM lane quarter this warp may touch
        const int half = (warp - 2) >> 2;
        const int split = ((P >> 1) >> 5) << 5;          // 96 for P = 196
        const int c_begin = half ? split : 0, c_end = half ? P : split;
        const int row = q * 32 + lane;
        const float c1 = -kLog2e / hp.tau, c0 = hp.eps * kLog2e / hp.tau;
        const float itemp = 1.0f / hp.temp, itau = 1.0f / hp.tau;
        // fp16 keeps 11 significant bits but only 2^-14..2^15 of range: the cell gradients are
        // stored times a power of two derived from their bound (row_grad_kernel), undone exactly
        // by the two contractions that consume them
        const float gscale = kGrad ? pow2i(g_scale_exp(__ldg(gmax_bits))) : 1.0f;
        uint32_t tile = 0;
        for (int mt = mt0; mt < n_mt; mt += mt_step) {
            const int k = mt * kBM + row;
            // per-(n, k) coefficients of the NEXT tile are requested while this one is swept, so
            // their global-load latency never sits between the accumulator and the sweep
            float nx_den = 1.0f, nx_sim = 0.f, nx_gs = 0.f;
            auto fetch = [&](int n) {
                if (kGrad && k < N_a && n < N_v) {
                    nx_den = den[(size_t)n * N_a + k];
                    nx_sim = sim[(size_t)n * N_a + k];
                    nx_gs = gsim[(size_t)n * N_a + k];
                }
            };
            fetch(g);
            for (int n = g; n < N_v; n += groups, ++tile) {
                const uint32_t buf = tile & 1u, use = tile >> 1;
                const int pos = pos_idx ? pos_idx[n] : col_offset + n;
                CellCtx cx{c1, c0, itau, 0.f, 0.f};
                bool exact = false;
                if (kGrad) {
                    float dnk = 1.0f;
                    if (k < N_a) {
                        dnk = nx_den;
                        cx.sm = nx_sim;
                        cx.coef = nx_gs / dnk * gscale;
                    }
                    exact = __any_sync(0xffffffffu, dnk < kTinyDen);
                    fetch(n + groups);
                }
                __half* gout = G + (size_t)n * P * n_a_pad + k;
                const uint32_t pitch = (uint32_t)n_a_pad;
                mbar_wait(&tmem_full[buf], use & 1u);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * kAccCols;
                float num = 0.f, dn = 0.f;
                // forward: MMA-bound, one RCP per cell is free; backward: MUFU-bound, share RCPs
                // unless some pair of this warp has every weight below 2^-30 (rare)
                if (!kGrad || exact) cell_sweep<kGrad, true>(taddr, c_begin, c_end, cx, num, dn, gout, pitch);
                else cell_sweep<kGrad, false>(taddr, c_begin, c_end, cx, num, dn, gout, pitch);
                tc_fence_before();
                mbar_arrive(&tmem_empty[buf]);           // accumulator may be overwritten
                if (kGrad) {
                    // positive column: G[n, pos, p] += GA[n, p] (gradient through A, sim1, sim2).
                    // The whole warp shares the fix-up of its one positive lane.
                    const uint32_t m = __ballot_sync(0xffffffffu, k == pos);
                    if (m) {
                        __syncwarp();
                        const int kp = mt * kBM + q * 32 + (__ffs(m) - 1);
                        __half* gcol = G + (size_t)n * P * n_a_pad + kp;
                        for (int pp = c_begin + lane; pp < c_end; pp += 32) {
                            __half* a = gcol + (uint32_t)pp * pitch;
                            *a = __float2half_rn(fmaf(GA[(size_t)n * P + pp], gscale, __half2float(*a)));
                        }
                    }
                }
                if (!kGrad) {                            // combine the two column halves
                    float2* pp = part + (tile & 1u) * kBM;
                    if (half) pp[row] = make_float2(num, dn);
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    if (!half) { const float2 o = pp[row]; num += o.x; dn += o.y; }
                }
                if (!kGrad && !half && k < N_a) {
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

static int launch_sim_tc(bool grad, const __half* vt, const __half* a_h,
                         const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                         HeadParams hp, float* logits, float* sim, float* den, const float* gsim,
                         const float* GA, __half* G, int n_a_pad, const unsigned int* gmax_bits,
                         cudaStream_t st) {
    AVT_REQUIRE(C % kBK == 0 && C <= kMaxKC * kBK, "fp16 head: C=%d must be a multiple of 64, <= 512", C);
    AVT_REQUIRE(P % 4 == 0 && P <= kBN, "fp16 head: P=%d must be a multiple of 4, <= 208", P);
    // backward: every column of G up to n_a_pad must be written (pad columns get zeros)
    const int n_mt = grad ? n_a_pad / kBM : (N_a + kBM - 1) / kBM;
    CUtensorMap map_a, map_b;
    if (!make_tmap_f16_2d(&map_a, a_h, (uint64_t)n_mt * kBM, (uint64_t)C, (uint64_t)C * 2, kBK, kBM) ||
        !make_tmap_f16_2d(&map_b, vt, (uint64_t)N_v * P, (uint64_t)C, (uint64_t)C * 2, kBK, kBN)) {
        set_error("fp16 head: cuTensorMapEncodeTiled failed");
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
        AVT_CHECK_CUDA(cudaFuncSetAttribute(sim_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)kTcSmemBytes));
        sim_tc_kernel<true><<<grid, kTcThreads, kTcSmemBytes, st>>>(
            map_a, map_b, pos_idx, col_offset, N_v, N_a, n_mt, groups, C / kBK, P, hp, L, logits, sim, den,
            gsim, GA, G, n_a_pad, gmax_bits);
    } else {
        AVT_CHECK_CUDA(cudaFuncSetAttribute(sim_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)kTcSmemBytes));
        sim_tc_kernel<false><<<grid, kTcThreads, kTcSmemBytes, st>>>(
            map_a, map_b, pos_idx, col_offset, N_v, N_a, n_mt, groups, C / kBK, P, hp, L, logits, sim, den,
            nullptr, nullptr, nullptr, 0, nullptr);
    }
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

int launch_sim_f16(const __half* vt, const __half* a_h, const int32_t* pos_idx,
                    int col_offset, int N_v, int N_a, int C, int P, HeadParams hp, float* logits,
                    float* sim, float* den, cudaStream_t st) {
    ProfScope _ps(kProfSim, st);
    return launch_sim_tc(false, vt, a_h, pos_idx, col_offset, N_v, N_a, C, P, hp, logits, sim, den,
                         nullptr, nullptr, nullptr, 0, nullptr, st);
}

// backward stage 1: cell gradients G (fp16, [N_v*P][n_a_pad])
int launch_grad_cells_f16(const __half* vt, const __half* a_h, const int32_t* pos_idx,
                           int col_offset, int N_v, int N_a, int n_a_pad, int C, int P, HeadParams hp,
                           const float* sim, const float* den, const float* gsim, const float* GA,
                           __half* G, const unsigned int* gmax_bits, cudaStream_t st) {
    ProfScope _ps(kProfBwdG, st);
    return launch_sim_tc(true, vt, a_h, pos_idx, col_offset, N_v, N_a, C, P, hp, nullptr,
                         const_cast<float*>(sim), const_cast<float*>(den), gsim, GA, G, n_a_pad, gmax_bits, st);
}

}  // namespace avt
