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
//   warps      0-15: epilogue (lane quarter = warp % 4, four column groups)   16: TMA producer   17: MMA issuer + TMEM owner
//   grid       persistent: block = (audio tile, image group); images of a group are visited in
//              step by the blocks of all audio tiles, so each image is read from HBM once
#include "avt_common.cuh"
#include "avt_tc.cuh"
#include <stdlib.h>

namespace avt {
using namespace tc;

constexpr int kColGroups = 4;                           // epilogue warps per TMEM lane quarter
constexpr int kEpiWarps = 4 * kColGroups;
constexpr int kTcThreads = 64 + 32 * kEpiWarps;         // warps 0..15 epilogue, 16 TMA, 17 MMA
// The two single-thread roles sit at the HIGHEST warp ids: the sub-partition arbiter prefers the
// highest warp id among eligible warps, and the MMA issuer shares its scheduler with four epilogue
// warps -- as warp 1 it only issued when all four were stalled (cycle trace, round 2: the issuer spent
// 85 % of its time in its own ~70 instructions per K chunk, 144 clocks per MMA against the 104 the
// tensor pipe needs).
constexpr int kWarpTma = kEpiWarps, kWarpMma = kEpiWarps + 1;
constexpr int kBM = 128;
constexpr int kBN = 208;
constexpr int kBK = 64;
constexpr int kStagesB = 3;
constexpr int kMaxKC = 8;                               // C <= 512
constexpr int kABytes = kBM * kBK * 2;                  // 16384 per K chunk
constexpr int kBBytes = kBN * kBK * 2;                  // 26624 per stage
constexpr int kAccCols = 256;
constexpr int kTmemCols = 512;
constexpr int kStagesBPair = 7;                         // cta_group::2: half a B tile (13 KB) per CTA and stage
constexpr int kPartBytes = 2 * (kColGroups - 1) * kBM * 8;
constexpr size_t kTcSmemBytes = 1024 + (size_t)kMaxKC * kABytes + (size_t)kStagesB * kBBytes + 256 + kPartBytes;
constexpr size_t kTcSmemBytesPair = 1024 + (size_t)kMaxKC * kABytes + (size_t)kStagesBPair * (kBBytes / 2) + 256 + kPartBytes;
static_assert(kTcSmemBytesPair <= 232448 && kTcSmemBytes <= 232448, "shared memory budget (227 KB)");
static_assert((kBBytes / 2) % 1024 == 0, "half a B tile must keep the 1024-byte swizzle alignment");

// ---- packed fp32 pairs (sm_100: FFMA2 / FMUL2 / FADD2 issue one instruction for two lanes' worth
// of fp32 work).  The epilogues below are instruction-issue bound (ncu, round 2: 18 instructions per
// cell in the gradient epilogue, issue slots 50 % busy at 2 warps per scheduler), so every
// elementwise step is written on float2.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("{.reg .b64 ra, rb, rc, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
        "mov.b64 {%0, %1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    float2 d;
    asm("{.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
        "mul.rn.f32x2 rd, ra, rb;\n\tmov.b64 {%0, %1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    float2 d;
    asm("{.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
        "add.rn.f32x2 rd, ra, rb;\n\tmov.b64 {%0, %1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return d;
}
__device__ __forceinline__ float2 splat(float x) { return make_float2(x, x); }

// One chunk of N accumulator columns of one TMEM lane.  w = sigma((s - eps)/tau) =
// 1 / (1 + 2^(s*c1 + c0)).  Fast variant: two cells share one MUFU.RCP (r = 1/(x0*x1); w0 = r*x1;
// w1 = r*x0) -- 1.5 MUFU operations per cell instead of 2 (the forward epilogue at 2 per cell was
// MUFU-bound: 3136 of the 3328 clocks an image tile's MMAs take).  The product x0*x1 overflows only
// when both weights are below 2^-64; such cells then read 0, which matters only when EVERY weight
// of the (n, k) pair is tiny -- the caller detects that (den below 2^-30) and redoes the pair with
// kExact = true (one RCP per cell), so results never depend on it.
struct CellCtx {
    float c1, c0, itau;          // sigmoid constants
    float sm, coef;              // backward: pooled sim and gsim / den of this (n, k)
};
constexpr float kTinyDen = 9.313225746154785e-10f;      // 2^-30
constexpr float kTinyFwd = 8.8817841970012523e-16f;     // 2^-50 per column: forward redo threshold

// kPitch: compile-time row pitch of G in elements (0 = run-time `pitch`), so that the 32 stores of
// a chunk are STG [base + immediate] instead of one 64-bit address add each
template <bool kGrad, bool kExact, int N, int kPitch>
__device__ __forceinline__ void cell_chunk(const uint32_t (&r)[N], const CellCtx& cx, float2& num,
                                           float2& dn, __half* gp, uint32_t pitch_rt) {
    const uint32_t pitch = kPitch ? (uint32_t)kPitch : pitch_rt;
#pragma unroll
    for (int i = 0; i < N; i += 2) {
        const float2 s = make_float2(__uint_as_float(r[i]), __uint_as_float(r[i + 1]));
        const float2 t = ffma2(s, splat(cx.c1), splat(cx.c0));
        const float2 x = fadd2(make_float2(fast_ex2(t.x), fast_ex2(t.y)), splat(1.0f));
        float2 w;
        if (kExact) {
            w = make_float2(fast_rcp(x.x), fast_rcp(x.y));
        } else {
            const float rr = fast_rcp(x.x * x.y);
            w = make_float2(rr * x.y, rr * x.x);
        }
        if (kGrad) {
            // G = coef w (1 + (s - sim)(1 - w)/tau)
            const float2 u = ffma2(w, splat(-cx.itau), splat(cx.itau));
            const float2 f = ffma2(fadd2(s, splat(-cx.sm)), u, splat(1.0f));
            const float2 gv = fmul2(fmul2(w, splat(cx.coef)), f);
            const __half2 h = __floats2half2_rn(gv.x, gv.y);
#ifndef AVT_G_SCALAR_ST
            // Lanes 2j / 2j+1 hold audio columns k, k+1 of the same two positions: they swap halves so
            // that the even lane stores (k, k+1) of position i and the odd lane (k, k+1) of position i+1
            // as ONE 4-byte word each -- half the store instructions (the epilogue sits in LG throttle
            // 10 % of the time: profiles/r2_ncu_head.txt).  `gp` was moved to this lane's word by the caller.
            const uint32_t mine = *reinterpret_cast<const uint32_t*>(&h);
            const uint32_t other = __shfl_xor_sync(0xffffffffu, mine, 1);
            const uint32_t word = (threadIdx.x & 1u) ? __byte_perm(other, mine, 0x7632) : __byte_perm(mine, other, 0x5410);
            *reinterpret_cast<uint32_t*>(gp + (uint32_t)i * pitch) = word;
#else
            gp[(uint32_t)i * pitch] = __low2half(h);
            gp[(uint32_t)(i + 1) * pitch] = __high2half(h);
#endif
        } else {
            dn = fadd2(dn, w);
            num = ffma2(w, s, num);
        }
    }
}

// all columns [c_begin, c_end) of one TMEM lane (c_end - c_begin a multiple of 4)
template <bool kGrad, bool kExact, int kPitch>
__device__ __forceinline__ void cell_sweep(uint32_t taddr, int c_begin, int c_end, const CellCtx& cx,
                                           float& num_out, float& dn_out, __half* gout, uint32_t pitch_rt) {
    const uint32_t pitch = kPitch ? (uint32_t)kPitch : pitch_rt;
    float2 num = make_float2(0.f, 0.f), dn = make_float2(0.f, 0.f);
    int c = c_begin;
    for (; c + 32 <= c_end; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c, r);
        tmem_ld_wait();
        cell_chunk<kGrad, kExact, 32, kPitch>(r, cx, num, dn, gout + (uint32_t)c * pitch, pitch_rt);
    }
    if (c + 16 <= c_end) {
        uint32_t r[16];
        tmem_ld_32x16(taddr + c, r);
        tmem_ld_wait();
        cell_chunk<kGrad, kExact, 16, kPitch>(r, cx, num, dn, gout + (uint32_t)c * pitch, pitch_rt);
        c += 16;
    }
    for (; c < c_end; c += 4) {
        uint32_t r[4];
        tmem_ld_32x4(taddr + c, r);
        tmem_ld_wait();
        cell_chunk<kGrad, kExact, 4, kPitch>(r, cx, num, dn, gout + (uint32_t)c * pitch, pitch_rt);
    }
    num_out = num.x + num.y;
    dn_out = dn.x + dn.y;
}

// kGrad = false: forward (writes sim / den / logits).
// kGrad = true : backward stage 1.  The same contraction is recomputed and the epilogue turns
//   every cell into its gradient  G = gsim/den * w (1 + (S - sim)(1 - w)/tau)  (+ GA on the
//   positive column), stored as fp16 in G[n*P + p][k] (row pitch n_a_pad) --
//   the K-major A operand of the dv contraction and the MN-major A operand of the d_ahat one.
// kPair = true : tcgen05.mma.cta_group::2.  A cluster of two CTAs owns TWO audio tiles (M = 256) of
//   the same image: each CTA keeps its own 128 audio rows resident and receives its own 128
//   accumulator lanes, but streams only HALF of the image tile (104 of the 208 rows of B) -- the
//   hardware reads both halves across the pair.  ncu of the 1-CTA kernel (profiles/r2_ncu_head_a):
//   tensor pipe 61 % with the L2 and the crossbar at ~30 %: with A resident only 3 x 26 KB of B
//   fit in flight per SM against 64 B/clk needed -- latency-bound on the operand feed.  In pair
//   mode a CTA needs 32 B/clk and holds 7 stages of 13 KB.
#ifdef AVT_TC_TRACE
// development build only (-DAVT_TC_TRACE): per-block cycle counters of the MMA issuer and of one
// epilogue warp -- [0] issuer waiting for operands, [1] issuer waiting for a free accumulator,
// [2] issuer total, [3] epilogue waiting for the accumulator, [4] epilogue sweeping, [5] epilogue total
__device__ unsigned long long g_tc_trace[1024 * 8];
#define AVT_TRACE(x) x
#else
#define AVT_TRACE(x)
#endif

template <bool kGrad, bool kPair, int kPitch>
__global__ void __launch_bounds__(kTcThreads, 1)
sim_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
              const int32_t* __restrict__ pos_idx, int col_offset, int N_v, int N_a, int n_mt,
              int groups, int KC, int P, HeadParams hp, int L, float* __restrict__ logits,
              float* __restrict__ sim, float* __restrict__ den,
              const float* __restrict__ gsim, const float* __restrict__ GA,
              __half* __restrict__ G, int n_a_pad, const unsigned int* __restrict__ gmax_bits) {
    constexpr int kStages = kPair ? kStagesBPair : kStagesB;
    constexpr int kStageBytes = kPair ? kBBytes / 2 : kBBytes;
    constexpr uint32_t kEpiArrivals = (uint32_t)(kEpiWarps * 32) * (kPair ? 2u : 1u);
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = base;
    uint8_t* sB = sA + (size_t)kMaxKC * kABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + (size_t)kStages * kStageBytes);
    uint64_t* full_b = bars;                             // [kStages]  (pair: the leader's copy counts both CTAs' bytes)
    uint64_t* empty_b = bars + kStages;                  // [kStages]
    uint64_t* a_full = bars + 2 * kStages;
    uint64_t* a_empty = a_full + 1;
    uint64_t* tmem_full = a_empty + 1;                   // [2]
    uint64_t* tmem_empty = tmem_full + 2;                // [2]  (pair: leader's copies, arrivals of both CTAs)
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    float2* part = reinterpret_cast<float2*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2][kColGroups-1][kBM]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = kPair ? cluster_ctarank() : 0u;
    // work item of a block (pair): audio tile (pair) u0, u0 + u_step, ...; images g, g + groups, ...
    const int unit = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int n_units = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int n_ut = kPair ? n_mt / 2 : n_mt;            // audio tiles (pairs) in total
    const int u0 = unit / groups, g = unit % groups, u_step = n_units / groups;

    if (warp == kWarpTma && lane == 0) {
        prefetch_tmap(&map_a);
        prefetch_tmap(&map_b);
        for (int i = 0; i < kStages; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], kEpiArrivals); }
        fence_barrier_init();
    }
    if (warp == kWarpMma) { if (kPair) tmem_alloc2(tmem_holder, kTmemCols); else tmem_alloc(tmem_holder, kTmemCols); }
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync();                           // both CTAs' barriers exist before any remote arrive
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;

    if (warp == kWarpTma) {
        if (lane == 0) {                                 // ---------------- TMA producer (both CTAs of a pair)
            int stage = 0; uint32_t phase = 0, a_par = 0;
            for (int u = u0; u < n_ut; u += u_step) {
                const int mt = kPair ? 2 * u + (int)rank : u;
                if (u != u0) { mbar_wait(a_empty, a_par); a_par ^= 1; }
                if (!kPair || rank == 0) mbar_arrive_expect_tx(a_full, (uint32_t)(KC * kABytes) * (kPair ? 2u : 1u));
                for (int kc = 0; kc < KC; ++kc) {
                    if (kPair) tma_load_2d_pair(sA + (size_t)kc * kABytes, &map_a, a_full, kc * kBK, mt * kBM);
                    else tma_load_2d(sA + (size_t)kc * kABytes, &map_a, a_full, kc * kBK, mt * kBM);
                }
                for (int n = g; n < N_v; n += groups) {
                    for (int kc = 0; kc < KC; ++kc) {
                        mbar_wait(&empty_b[stage], phase ^ 1);
                        if (kPair) {
                            if (rank == 0) mbar_arrive_expect_tx(&full_b[stage], (uint32_t)kBBytes);
                            tma_load_2d_pair(sB + (size_t)stage * kStageBytes, &map_b, &full_b[stage], kc * kBK,
                                             n * P + (int)rank * (kBN / 2));
                        } else {
                            mbar_arrive_expect_tx(&full_b[stage], (uint32_t)kBBytes);
                            tma_load_2d(sB + (size_t)stage * kStageBytes, &map_b, &full_b[stage], kc * kBK, n * P);
                        }
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == kWarpMma) {
        if (rank == 0) {                                 // ---------------- MMA issuer (pair: the leader only)
            // the whole warp walks the loops (uniform control flow, uniform registers); one elected
            // lane issues the MMAs and their commits
            const uint32_t idesc = make_idesc_f16(kPair ? 2 * kBM : kBM, kBN, 0, 0);
            const uint64_t a_desc0 = make_smem_desc(smem_u32(sA), 16, 1024);
            const uint64_t b_desc0 = make_smem_desc(smem_u32(sB), 16, 1024);
            int stage = 0; uint32_t phase = 0, a_par = 0;
            uint32_t tile = 0;
            AVT_TRACE(long long tw_full = 0; long long tw_empty = 0; const long long t_begin = clock64(); long long t0;)
            for (int u = u0; u < n_ut; u += u_step) {
                mbar_wait(a_full, a_par); a_par ^= 1;
                tc_fence_after();
                for (int n = g; n < N_v; n += groups, ++tile) {
                    const uint32_t buf = tile & 1u, use = tile >> 1;
                    AVT_TRACE(t0 = clock64();)
                    mbar_wait(&tmem_empty[buf], (use & 1u) ^ 1u);
                    AVT_TRACE(tw_empty += clock64() - t0;)
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * kAccCols;
                    // descriptors differ only in their 14-bit address field (bytes >> 4): one add each
                    uint64_t a_desc = a_desc0;
                    for (int kc = 0; kc < KC; ++kc, a_desc += (uint64_t)(kABytes >> 4)) {
                        AVT_TRACE(t0 = clock64();)
                        mbar_wait(&full_b[stage], phase);
                        AVT_TRACE(tw_full += clock64() - t0;)
                        tc_fence_after();
                        const uint64_t b_desc = b_desc0 + (uint64_t)((uint32_t)stage * (uint32_t)(kStageBytes >> 4));
                        if (elect_one()) {
#pragma unroll
                            for (int k = 0; k < kBK / 16; ++k) {
                                if (kPair) umma2_f16_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (uint32_t)((kc | k) != 0));
                                else umma_f16_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (uint32_t)((kc | k) != 0));
                            }
                            // frees the stage (in both CTAs) once these MMAs retire
                            if (kPair) umma2_commit_mc(&empty_b[stage], 3); else umma_commit(&empty_b[stage]);
                        }
                        __syncwarp();
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                    if (elect_one()) { if (kPair) umma2_commit_mc(&tmem_full[buf], 3); else umma_commit(&tmem_full[buf]); }
                    __syncwarp();
                }
                if (elect_one()) { if (kPair) umma2_commit_mc(a_empty, 3); else umma_commit(a_empty); }
                __syncwarp();
            }
            AVT_TRACE(if (lane == 0) { g_tc_trace[blockIdx.x * 8 + 0] = tw_full; g_tc_trace[blockIdx.x * 8 + 1] = tw_empty;
                      g_tc_trace[blockIdx.x * 8 + 2] = clock64() - t_begin; g_tc_trace[blockIdx.x * 8 + 6] = tile; })
        }
    } else {                                             // ---------------- epilogue warps 0 .. kEpiWarps-1
        // kColGroups warps per TMEM lane quarter (one per scheduler each): every warp takes a slice of
        // the image's positions, so each SM sub-partition has four warps to hide MUFU / TMEM latency
        // (ncu of the 2-group version: issue slots 50 % busy, tensor pipe 46 % -- latency-bound)
        const int q = warp & 3;                          // TMEM lane quarter this warp may touch
        const int grp = warp >> 2;                 // column group 0 .. kColGroups-1
        const int cw = ((P / kColGroups) >> 4) << 4;     // 48 for P = 196; the last group takes the rest
        const int c_begin = grp * cw, c_end = (grp == kColGroups - 1) ? P : (grp + 1) * cw;
        const int row = q * 32 + lane;
        const float c1 = -kLog2e / hp.tau, c0 = hp.eps * kLog2e / hp.tau;
        const float itemp = 1.0f / hp.temp, itau = 1.0f / hp.tau;
        // fp16 keeps 11 significant bits but only 2^-14..2^15 of range: the cell gradients are
        // stored times a power of two derived from their bound (row_grad_kernel), undone exactly
        // by the two contractions that consume them
        const float gscale = kGrad ? pow2i(g_scale_exp(__ldg(gmax_bits))) : 1.0f;
        uint32_t tile = 0;
        AVT_TRACE(long long te_wait = 0; long long te_sweep = 0; const long long te_begin = clock64(); long long e0;)
        for (int u = u0; u < n_ut; u += u_step) {
            const int mt = kPair ? 2 * u + (int)rank : u;
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
#ifndef AVT_G_SCALAR_ST
                // this lane's 4-byte word: columns (k, k+1) of position i (even lane) or (k-1, k) of i+1 (odd)
                __half* gout = G + (size_t)n * P * n_a_pad + k + ((lane & 1) ? (int)n_a_pad - 1 : 0);
#else
                __half* gout = G + (size_t)n * P * n_a_pad + k;
#endif
                const uint32_t pitch = (uint32_t)n_a_pad;
                AVT_TRACE(e0 = clock64();)
                mbar_wait(&tmem_full[buf], use & 1u);
                AVT_TRACE(te_wait += clock64() - e0; e0 = clock64();)
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * kAccCols;
                float num = 0.f, dn = 0.f;
                // two cells share one RCP unless some pair of this warp has every weight tiny (rare): the
                // backward knows from the saved den, the forward from the sum it has just formed (the
                // accumulator is still there, so the slice is simply swept again)
                if (kGrad && exact) cell_sweep<kGrad, true, kPitch>(taddr, c_begin, c_end, cx, num, dn, gout, pitch);
                else cell_sweep<kGrad, false, kPitch>(taddr, c_begin, c_end, cx, num, dn, gout, pitch);
                if (!kGrad && __any_sync(0xffffffffu, !(dn >= kTinyFwd * (float)(c_end - c_begin))))
                    cell_sweep<false, true, 0>(taddr, c_begin, c_end, cx, num, dn, gout, pitch);
                AVT_TRACE(te_sweep += clock64() - e0;)
                tc_fence_before();
                // accumulator may be overwritten (pair: the leader's barrier collects both CTAs)
                if (!kPair || rank == 0) mbar_arrive(&tmem_empty[buf]); else mbar_arrive_leader(&tmem_empty[buf]);
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
                if (!kGrad) {                            // combine the column groups
                    float2* pp = part + (tile & 1u) * (kColGroups - 1) * kBM;
                    if (grp) pp[(grp - 1) * kBM + row] = make_float2(num, dn);
                    asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
                    if (!grp) {
#pragma unroll
                        for (int j = 0; j < kColGroups - 1; ++j) { const float2 o = pp[j * kBM + row]; num += o.x; dn += o.y; }
                    }
                }
                if (!kGrad && !grp && k < N_a) {
                    const float smv = num / dn;
                    const float m = (k == pos) ? (1.0f - hp.mask_scale) : 1.0f;
                    sim[(size_t)n * N_a + k] = smv;
                    den[(size_t)n * N_a + k] = dn;
                    logits[(size_t)n * L + 1 + k] = smv * m * itemp;
                }
            }
        }
        AVT_TRACE(if (warp == 0 && lane == 0) { g_tc_trace[blockIdx.x * 8 + 3] = te_wait; g_tc_trace[blockIdx.x * 8 + 4] = te_sweep;
                                                g_tc_trace[blockIdx.x * 8 + 5] = clock64() - te_begin; })
    }
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync();                           // nobody leaves while the pair may still touch its smem / TMEM
    if (warp == kWarpMma) {
        tc_fence_after();
        if (kPair) tmem_dealloc2(tmem_base, kTmemCols); else tmem_dealloc(tmem_base, kTmemCols);
    }
}

template <bool kGrad, bool kPair, int kPitch>
static int launch_sim_variant(const CUtensorMap& map_a, const CUtensorMap& map_b, int grid, size_t smem,
                              const int32_t* pos_idx, int col_offset, int N_v, int N_a, int n_mt, int groups,
                              int KC, int P, HeadParams hp, int L, float* logits, float* sim, float* den,
                              const float* gsim, const float* GA, __half* G, int n_a_pad,
                              const unsigned int* gmax_bits, cudaStream_t st) {
    auto kernel = sim_tc_kernel<kGrad, kPair, kPitch>;
    AVT_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kPair ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    AVT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kernel, map_a, map_b, pos_idx, col_offset, N_v, N_a, n_mt, groups, KC,
                                      P, hp, L, logits, sim, den, gsim, GA, G, n_a_pad, gmax_bits));
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

static int launch_sim_tc(bool grad, const __half* vt, const __half* a_h,
                         const int32_t* pos_idx, int col_offset, int N_v, int N_a, int C, int P,
                         HeadParams hp, float* logits, float* sim, float* den, const float* gsim,
                         const float* GA, __half* G, int n_a_pad_in, const unsigned int* gmax_bits,
                         cudaStream_t st) {
    const int n_a_pad = n_a_pad_in ? n_a_pad_in : n_a_padded(N_a);      // rows of a_h (zero rows past N_a)
    AVT_REQUIRE(C % kBK == 0 && C <= kMaxKC * kBK, "fp16 head: C=%d must be a multiple of 64, <= 512", C);
    AVT_REQUIRE(P % 4 == 0 && P <= kBN, "fp16 head: P=%d must be a multiple of 4, <= 208", P);
    // backward: every column of G up to n_a_pad must be written (pad columns get zeros)
    int n_mt = grad ? n_a_pad / kBM : (N_a + kBM - 1) / kBM;
    static const bool no_pair = getenv("AVT_NO_PAIR") != nullptr;           // development switch, read once
    const bool pair = !no_pair && n_a_pad >= 2 * kBM;                        // n_a_pad is then a multiple of 256
    if (pair) n_mt = (n_mt + 1) / 2 * 2;                                     // whole tile pairs (a_h rows are padded)
    CUtensorMap map_a, map_b;
    if (!make_tmap_f16_2d(&map_a, a_h, (uint64_t)n_a_pad, (uint64_t)C, (uint64_t)C * 2, kBK, kBM) ||
        !make_tmap_f16_2d(&map_b, vt, (uint64_t)N_v * P, (uint64_t)C, (uint64_t)C * 2, kBK, pair ? kBN / 2 : kBN)) {
        set_error("fp16 head: cuTensorMapEncodeTiled failed");
        return AVT_ERR_CUDA;
    }
    int dev = 0, sms = 148;
    AVT_CHECK_CUDA(cudaGetDevice(&dev));
    AVT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int slots = pair ? sms / 2 : sms;              // CTAs (or CTA pairs) resident at once
    const int n_ut = pair ? n_mt / 2 : n_mt;
    const int ut_par = n_ut < slots ? n_ut : slots;
    int groups = slots / ut_par;
    if (groups > N_v) groups = N_v;
    if (groups < 1) groups = 1;
    const int grid = groups * ut_par * (pair ? 2 : 1);
    const int L = N_a + 1 + (hp.use_neg ? 1 : 0);
    const int KC = C / kBK;
#define AVT_SIM(GR, PR, PI)                                                                               \
    launch_sim_variant<GR, PR, PI>(map_a, map_b, grid, PR ? kTcSmemBytesPair : kTcSmemBytes, pos_idx, col_offset, N_v, \
                                   N_a, n_mt, groups, KC, P, hp, L, logits, sim, den, gsim, GA, G, n_a_pad, gmax_bits, st)
    if (!grad) return pair ? AVT_SIM(false, true, 0) : AVT_SIM(false, false, 0);
    if (!pair) return n_a_pad == 128 ? AVT_SIM(true, false, 128) : AVT_SIM(true, false, 0);
    switch (n_a_pad) {                                   // common audio counts get immediate store offsets
        case 256: return AVT_SIM(true, true, 256);
        case 512: return AVT_SIM(true, true, 512);
        case 1024: return AVT_SIM(true, true, 1024);
        default: return AVT_SIM(true, true, 0);
    }
#undef AVT_SIM
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

#ifdef AVT_TC_TRACE
extern "C" int avt_debug_tc_trace(unsigned long long* host_out, int n_blocks) {
    return cudaMemcpyFromSymbol(host_out, avt::g_tc_trace, sizeof(unsigned long long) * 8 * (size_t)n_blocks) == cudaSuccess ? 0 : -2;
}
#endif
