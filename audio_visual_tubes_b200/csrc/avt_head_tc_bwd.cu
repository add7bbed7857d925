// AVT_PRECISION_F16 backward on the tensor cores.  Three stages:
//
//   1. sim_tc_kernel<true> (avt_head_tc.cu) recomputes S tile by tile and stores the cell
//      gradients G[n*P + p][k] as fp16 (row pitch n_a_pad).
//   2. bwd_dv_kernel:  dv_hat[np, c] = sum_k G[np, k] a_hat[k, c]   (M = 128 rows of np,
//      N = C in two halves of <= 256, K = n_a_pad), epilogue = the F.normalize Jacobian
//      dv = rn (dv_hat - v_hat <v_hat, dv_hat>) (model.py:117) written straight into the
//      NCHW fp32 gradient.  Both operands K-major (G rows, a_hat^T rows).
//   3. bwd_da_kernel:  d_ahat[k, c] = sum_np G[np, k] v_hat[np, c]  (M = 128 audio rows,
//      N = C, K = N_v*P split over the grid).  Both operands are MN-major views of the
//      same row-major buffers stage 1 and the forward left behind, so nothing is
//      transposed in memory; partial sums are reduced with red.global.add.
//
// All three are TMA-fed tcgen05 kernels with the accumulators in tensor memory.
#include "avt_common.cuh"
#include <mutex>
#include "avt_tc.cuh"
#include <stdlib.h>
#ifdef AVT_DV_PLAIN_STORE
#define AVT_DV_ST(p, x) (*(p) = (x))
#else
#define AVT_DV_ST(p, x) __stcs((p), (x))
#endif

namespace avt {
using namespace tc;

#ifdef AVT_TC_TRACE
// development build only: per-block cycle counters of the dv kernel -- [0] issuer waiting for operands,
// [1] issuer waiting for a free accumulator half, [2] issuer total, [3] epilogue waiting for accumulators,
// [4] epilogue pass 1 (t), [5] epilogue total, [6] tiles, [7] epilogue pass 2 (stores)
__device__ unsigned long long g_dv_trace[1024 * 8];
#define AVT_TRACE(x) x
#else
#define AVT_TRACE(x)
#endif

// ------------------------------------------------------------------ a_hat^T (fp16)
__global__ void transpose_audio_kernel(const __half* __restrict__ a_h, int n_a_pad, int C,
                                       __half* __restrict__ at_h) {
    __shared__ __half tile[32][33];
    const int k0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y)
        tile[r][threadIdx.x] = a_h[(size_t)(k0 + r) * C + c0 + threadIdx.x];
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y)
        at_h[(size_t)(c0 + r) * n_a_pad + k0 + threadIdx.x] = tile[threadIdx.x][r];
}

// ------------------------------------------------------------------ stage 2: dv
constexpr int kDvThreads = 320;          // warps 0..7 epilogue, warp 8 TMA, warp 9 MMA + TMEM owner (highest ids:
                                         // the sub-partition arbiter prefers them, see avt_head_tc.cu)
constexpr int kDvWarpTma = 8, kDvWarpMma = 9;
constexpr int kDvBM = 128;
constexpr int kDvBK = 64;
constexpr int kDvStages = 4;
constexpr int kDvABytes = kDvBM * kDvBK * 2;          // 16384
constexpr int kDvBBytes = 256 * kDvBK * 2;            // 32768 (one half of C)
constexpr int kDvStageBytes = kDvABytes + kDvBBytes;
constexpr size_t kDvSmemBytes = 1024 + (size_t)kDvStages * kDvStageBytes + 256 + 2 * 2 * kDvBM * 4;

__device__ __forceinline__ float h_lo(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w & 0xffffu))); }
__device__ __forceinline__ float h_hi(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w >> 16))); }
// 16 consecutive fp16 (32 bytes, 32-byte aligned) through the read-only path in one request
__device__ __forceinline__ void ld_nc_v8(const __half* p, uint32_t* r) {
    asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}

__global__ void __launch_bounds__(kDvThreads, 1)
bwd_dv_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_at,
              const __half* __restrict__ vt,
              const float* __restrict__ rnorm_v, float* __restrict__ dv, int n_rows, int n_tiles,
              int KC, int C, int P, const unsigned int* __restrict__ gmax_bits) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* stages = base;
    uint64_t* bars = reinterpret_cast<uint64_t*>(stages + (size_t)kDvStages * kDvStageBytes);
    uint64_t* full_b = bars;                             // [kDvStages]
    uint64_t* empty_b = bars + kDvStages;                // [kDvStages]
    uint64_t* tmem_full = bars + 2 * kDvStages;          // [2] one per half of C
    uint64_t* tmem_empty = tmem_full + 2;                // [2]
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    float* t_part = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2][2][128]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nh = C > 256 ? 2 : 1;
    const int hN0 = C > 256 ? 256 : C, hN1 = C - hN0;

    if (warp == kDvWarpTma && lane == 0) {
        prefetch_tmap(&map_g);
        prefetch_tmap(&map_at);
        for (int i = 0; i < kDvStages; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 256); }
        fence_barrier_init();
    }
    if (warp == kDvWarpMma) tmem_alloc(tmem_holder, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;

    if (warp == kDvWarpTma) {
        if (lane == 0) {                                 // ---------------- TMA producer
            int stage = 0; uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int h = 0; h < nh; ++h) {
                    for (int kc = 0; kc < KC; ++kc) {
                        mbar_wait(&empty_b[stage], phase ^ 1);
                        mbar_arrive_expect_tx(&full_b[stage], (uint32_t)kDvStageBytes);
                        uint8_t* sA = stages + (size_t)stage * kDvStageBytes;
                        tma_load_2d(sA, &map_g, &full_b[stage], kc * kDvBK, tile * kDvBM);
                        tma_load_2d(sA + kDvABytes, &map_at, &full_b[stage], kc * kDvBK, h * 256);
                        if (++stage == kDvStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == kDvWarpMma) {
        {                                                // ---------------- MMA issuer (warp-uniform loop,
            const uint64_t desc0 = make_smem_desc(smem_u32(stages), 16, 1024);   // one elected lane issues)
            int stage = 0; uint32_t phase = 0, it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
                for (int h = 0; h < nh; ++h) {
                    const uint32_t idesc = make_idesc_f16(kDvBM, h ? hN1 : hN0, 0, 0);
                    mbar_wait(&tmem_empty[h], (it & 1u) ^ 1u);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + (uint32_t)h * 256u;
                    for (int kc = 0; kc < KC; ++kc) {
                        mbar_wait(&full_b[stage], phase);
                        tc_fence_after();
                        const uint64_t ad = desc0 + (uint64_t)((uint32_t)stage * (uint32_t)(kDvStageBytes >> 4));
                        const uint64_t bd = ad + (uint64_t)(kDvABytes >> 4);
                        if (elect_one()) {
#pragma unroll
                            for (int k = 0; k < kDvBK / 16; ++k)
                                umma_f16_ss(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (uint32_t)((kc | k) != 0));
                            umma_commit(&empty_b[stage]);
                        }
                        __syncwarp();
                        if (++stage == kDvStages) { stage = 0; phase ^= 1; }
                    }
                    if (elect_one()) umma_commit(&tmem_full[h]);
                    __syncwarp();
                }
            }
        }
    } else {                                             // ---------------- epilogue warps 0..7
        const int q = warp & 3;                          // TMEM lane quarter
        const int grp = warp >> 2;                 // column group inside a half
        const int r_in = q * 32 + lane;
        const uint32_t Pu = (uint32_t)P;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int row = tile * kDvBM + r_in;
            const bool valid = row < n_rows;
            const int rowc = valid ? row : n_rows - 1;   // loads of a ragged last tile stay in bounds
            const int n = rowc / P;
            const int p = rowc - n * P;
            const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
            const __half* vrow = vt + (size_t)rowc * C;
            // pass 1: t = <v_hat[row, :], dv_hat[row, :]> over this thread's columns
            float t = 0.f;
            for (int h = 0; h < nh; ++h) {
                const int hN = h ? hN1 : hN0, cw = hN >> 1;
                const int cbase = h * 256 + grp * cw;
                mbar_wait(&tmem_full[h], it & 1u);
                tc_fence_after();
                for (int c = 0; c < cw; c += 32) {
                    uint32_t r[32], x[16];
                    tmem_ld_32x32(lane_addr + (uint32_t)(cbase + c), r);
                    ld_nc_v8(vrow + cbase + c, x);
                    ld_nc_v8(vrow + cbase + c + 16, x + 8);
                    tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        t = fmaf(h_lo(x[e]), __uint_as_float(r[2 * e]), t);
                        t = fmaf(h_hi(x[e]), __uint_as_float(r[2 * e + 1]), t);
                    }
                }
            }
            float* tp = t_part + (it & 1u) * 2 * kDvBM;
            tp[grp * kDvBM + r_in] = t;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            t = tp[r_in] + tp[kDvBM + r_in];
            // pass 2: dv = rn * (dv_hat - v_hat * t), NCHW fp32.  v_hat comes from the fp16
            // row this thread just read (L1/L2-resident), not from a second sweep over v.
            const float rn = rnorm_v[rowc] * pow2i(-g_scale_exp(__ldg(gmax_bits)));   // undo the scale of G
            const float rt = rn * t;
            for (int h = 0; h < nh; ++h) {
                const int hN = h ? hN1 : hN0, cw = hN >> 1;
                const int cbase = h * 256 + grp * cw;
                float* obase = dv + ((size_t)n * C + cbase) * P + p;
                for (int c = 0; c < cw; c += 32) {
                    uint32_t r[32], x[16];
                    tmem_ld_32x32(lane_addr + (uint32_t)(cbase + c), r);
                    ld_nc_v8(vrow + cbase + c, x);
                    ld_nc_v8(vrow + cbase + c + 16, x + 8);
                    tmem_ld_wait();
                    float* o = obase + (uint32_t)c * Pu;
                    if (valid) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            o[(uint32_t)(2 * e) * Pu] = fmaf(-h_lo(x[e]), rt, rn * __uint_as_float(r[2 * e]));
                            o[(uint32_t)(2 * e + 1) * Pu] = fmaf(-h_hi(x[e]), rt, rn * __uint_as_float(r[2 * e + 1]));
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&tmem_empty[h]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kDvWarpMma) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------------ stage 2, CTA pairs
// cta_group::2 version: a cluster of two CTAs owns 256 consecutive rows of np (M = 256).  As in
// the 1-CTA kernel the two halves of C are two passes over K into two 256-column accumulators,
// so the epilogue of one half overlaps the MMAs of the other (a single-pass N = 512 variant
// that reads G once was measured slower: with all 512 TMEM columns busy the write-heavy
// epilogue serialises with the MMAs on every SM at the same time).  Per stage a CTA pulls its
// own 128 rows of G (16 KB) and HALF of the a_hat^T chunk (16 KB): 32 KB for an M = 256 step
// against 48 KB for an M = 128 step before.
constexpr int kDv2BHalfBytes = 128 * kDvBK * 2;       // this CTA's 128 channel rows of one N = 256 MMA
constexpr int kDv2StageBytes = kDvABytes + kDv2BHalfBytes;         // 32 KB
constexpr int kDv2Stages = 6;
constexpr size_t kDv2SmemBytes = 1024 + (size_t)kDv2Stages * kDv2StageBytes + 256 + 2 * 4 * kDvBM * 4;

// kG = column groups per half of C (epilogue warps = 4 * kG): 2 -> 8 warps of 128 columns per half,
// 4 -> 16 warps of 64 columns (twice the warps per scheduler to hide TMEM / store latency)
// kP: compile-time P (196 = the reference's 14 x 14 maps; 0 = run time) so that the 256 strided NCHW
// stores of a thread are STG [base + immediate] (with a run-time P every store paid an address
// computation and the kernel spilled 400 bytes)
// kBmn: the B operand is a_hat itself, [k][c] row-major = MN-major (N = c contiguous), fetched as two
// [64 k][64 c] SW128 blocks per CTA exactly like the d_ahat kernel's operands -- no transposed copy of
// the audio rows, no transpose launch.  false: K-major rows of a_hat^T (the first version; A/B switch).
template <int kG, int kP, bool kBmn>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64 + 128 * kG, 1)
bwd_dv_pair_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_at,
                   const __half* __restrict__ vt, const float* __restrict__ rnorm_v,
                   float* __restrict__ dv, int n_rows, int n_tile_pairs, int KC, int C, int P_rt, int reverse,
                   const unsigned int* __restrict__ gmax_bits) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* stages = base;
    uint64_t* bars = reinterpret_cast<uint64_t*>(stages + (size_t)kDv2Stages * kDv2StageBytes);
    uint64_t* full_b = bars;                             // leader's copy counts both CTAs' bytes
    uint64_t* empty_b = bars + kDv2Stages;
    uint64_t* tmem_full = bars + 2 * kDv2Stages;         // [2] one per half of C
    uint64_t* tmem_empty = tmem_full + 2;                // [2] leader's copies: 256 local + 256 peer arrivals
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    float* t_part = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2][kG][128]
    constexpr int kWTma = 4 * kG, kWMma = 4 * kG + 1;     // the two single-thread roles keep the highest warp ids
    constexpr int kCW = 256 / kG;                          // columns of one thread per half
    constexpr int kJ = kCW / 32;                           // 32-column chunks per thread and half
    constexpr uint32_t kEpiThreads = 128u * kG;

    const int P = kP ? kP : P_rt;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair0 = blockIdx.x >> 1, pair_step = gridDim.x >> 1;
    const int nh = C / 256;                              // C is 256 or 512

    if (warp == kWTma && lane == 0) {
        prefetch_tmap(&map_g);
        prefetch_tmap(&map_at);
        for (int i = 0; i < kDv2Stages; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 2 * kEpiThreads); }
        fence_barrier_init();
    }
    if (warp == kWMma) tmem_alloc2(tmem_holder, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;

    if (warp == kWTma) {
        if (lane == 0) {                                 // ---------------- TMA producer (both CTAs)
            int stage = 0; uint32_t phase = 0;
            for (int ti = pair0; ti < n_tile_pairs; ti += pair_step) {
                const int tp = reverse ? n_tile_pairs - 1 - ti : ti;
                const int tile = tp * 2 + (int)rank;
                for (int h = 0; h < nh; ++h) {
                    for (int kc = 0; kc < KC; ++kc) {
                        mbar_wait(&empty_b[stage], phase ^ 1);
                        if (rank == 0) mbar_arrive_expect_tx(&full_b[stage], (uint32_t)(2 * kDv2StageBytes));
                        uint8_t* sA = stages + (size_t)stage * kDv2StageBytes;
                        tma_load_2d_pair(sA, &map_g, &full_b[stage], kc * kDvBK, tile * kDvBM);
                        if (kBmn) tma_load_3d_pair(sA + kDvABytes, &map_at, &full_b[stage], 0, kc * kDvBK,
                                                   h * 4 + (int)rank * 2);
                        else tma_load_2d_pair(sA + kDvABytes, &map_at, &full_b[stage], kc * kDvBK,
                                              h * 256 + (int)rank * 128);
                        if (++stage == kDv2Stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == kWMma) {
        if (rank == 0) {                                 // ---------------- MMA issuer (leader only; warp-uniform
            const uint32_t idesc = make_idesc_f16(256, 256, 0, kBmn ? 1 : 0);   // loop, one elected lane issues)
            const uint64_t desc0 = make_smem_desc(smem_u32(stages), 16, 1024);
            // MN-major B: 64-channel blocks one LBO (8 KB) apart, 8 k rows per 1024 B; 16 k rows = 2048 B
            const uint64_t bdesc0 = kBmn ? make_smem_desc(smem_u32(stages) + kDvABytes, 64 * 128, 1024)
                                         : desc0 + (uint64_t)(kDvABytes >> 4);
            constexpr uint32_t kBStep = kBmn ? (2048u >> 4) : 2u;
            int stage = 0; uint32_t phase = 0, it = 0;
            AVT_TRACE(long long tw_full = 0; long long tw_empty = 0; const long long t_begin = clock64(); long long t0;)
            for (int ti = pair0; ti < n_tile_pairs; ti += pair_step, ++it) {
                for (int h = 0; h < nh; ++h) {
                    AVT_TRACE(t0 = clock64();)
                    mbar_wait(&tmem_empty[h], (it & 1u) ^ 1u);
                    AVT_TRACE(tw_empty += clock64() - t0;)
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + (uint32_t)h * 256u;
                    for (int kc = 0; kc < KC; ++kc) {
                        AVT_TRACE(t0 = clock64();)
                        mbar_wait(&full_b[stage], phase);
                        AVT_TRACE(tw_full += clock64() - t0;)
                        tc_fence_after();
                        const uint64_t ad = desc0 + (uint64_t)((uint32_t)stage * (uint32_t)(kDv2StageBytes >> 4));
                        const uint64_t bd = bdesc0 + (uint64_t)((uint32_t)stage * (uint32_t)(kDv2StageBytes >> 4));
                        if (elect_one()) {
#pragma unroll
                            for (int k = 0; k < kDvBK / 16; ++k)
                                umma2_f16_ss(d_tmem, ad + 2 * k, bd + kBStep * k, idesc, (uint32_t)((kc | k) != 0));
                            umma2_commit_mc(&empty_b[stage], 3);
                        }
                        __syncwarp();
                        if (++stage == kDv2Stages) { stage = 0; phase ^= 1; }
                    }
                    if (elect_one()) umma2_commit_mc(&tmem_full[h], 3);
                    __syncwarp();
                }
            }
            AVT_TRACE(if (lane == 0) { g_dv_trace[blockIdx.x * 8 + 0] = tw_full; g_dv_trace[blockIdx.x * 8 + 1] = tw_empty;
                      g_dv_trace[blockIdx.x * 8 + 2] = clock64() - t_begin; g_dv_trace[blockIdx.x * 8 + 6] = it; })
        }
    } else {                                             // ---------------- epilogue warps 0 .. 4*kG-1
        const int q = warp & 3;
        const int grp = warp >> 2;
        const int r_in = q * 32 + lane;
        const uint32_t Pu = (uint32_t)P;
        uint32_t it = 0;
        AVT_TRACE(long long te_wait = 0; long long te_p1 = 0; long long te_p2 = 0; const long long te_begin = clock64(); long long e0;)
        for (int ti = pair0; ti < n_tile_pairs; ti += pair_step, ++it) {
            const int tp = reverse ? n_tile_pairs - 1 - ti : ti;
            const int row = (tp * 2 + (int)rank) * kDvBM + r_in;
            const bool valid = row < n_rows;
            const int rowc = row < n_rows ? row : n_rows - 1;
            const int n = rowc / P;
            const int p = rowc - n * P;
            const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
            const __half* vrow = vt + (size_t)rowc * C;
            // v_hat row pieces are fetched ahead of their use: the first half before the
            // accumulator is even ready, the second while the first is being reduced, and the
            // second pass starts with the half that is still in registers.
            uint32_t xa[kJ][16], xb[kJ][16];
            const int ca = grp * kCW, cb = 256 + grp * kCW;        // this thread's columns of h = 0 / 1
#pragma unroll
            for (int j = 0; j < kJ; ++j) { ld_nc_v8(vrow + ca + 32 * j, xa[j]); ld_nc_v8(vrow + ca + 32 * j + 16, xa[j] + 8); }
            if (nh == 2) {
#pragma unroll
                for (int j = 0; j < kJ; ++j) { ld_nc_v8(vrow + cb + 32 * j, xb[j]); ld_nc_v8(vrow + cb + 32 * j + 16, xb[j] + 8); }
            }
            AVT_TRACE(e0 = clock64();)
            mbar_wait(&tmem_full[0], it & 1u);
            AVT_TRACE(te_wait += clock64() - e0; e0 = clock64();)
            tc_fence_after();
            float t = 0.f;
#pragma unroll
            for (int j = 0; j < kJ; ++j) {
                uint32_t r[32];
                tmem_ld_32x32(lane_addr + (uint32_t)(ca + 32 * j), r);
                tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    t = fmaf(h_lo(xa[j][e]), __uint_as_float(r[2 * e]), t);
                    t = fmaf(h_hi(xa[j][e]), __uint_as_float(r[2 * e + 1]), t);
                }
            }
            if (nh == 2) {
                AVT_TRACE(te_p1 += clock64() - e0; e0 = clock64();)
                mbar_wait(&tmem_full[1], it & 1u);
                AVT_TRACE(te_wait += clock64() - e0; e0 = clock64();)
                tc_fence_after();
#pragma unroll
                for (int j = 0; j < kJ; ++j) {
                    uint32_t r[32];
                    tmem_ld_32x32(lane_addr + (uint32_t)(cb + 32 * j), r);
                    tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        t = fmaf(h_lo(xb[j][e]), __uint_as_float(r[2 * e]), t);
                        t = fmaf(h_hi(xb[j][e]), __uint_as_float(r[2 * e + 1]), t);
                    }
                }
            }
            float* tpart = t_part + (it & 1u) * kG * kDvBM;
            tpart[grp * kDvBM + r_in] = t;
            asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
            t = 0.f;
#pragma unroll
            for (int j = 0; j < kG; ++j) t += tpart[j * kDvBM + r_in];
            const float rn = rnorm_v[rowc] * pow2i(-g_scale_exp(__ldg(gmax_bits)));   // undo the scale of G
            const float rt = rn * t;
            AVT_TRACE(te_p1 += clock64() - e0; e0 = clock64();)
            // pass 2: dv = rn * dv_hat - (rn t) * v_hat, NCHW fp32; each half is released as soon
            // as it has been written so the next tile's MMAs can start on it
            {
                float* obase = dv + ((size_t)n * C + ca) * P + p;
#pragma unroll
                for (int j = 0; j < kJ; ++j) {
                    uint32_t r[32];
                    tmem_ld_32x32(lane_addr + (uint32_t)(ca + 32 * j), r);
                    tmem_ld_wait();
                    float* o = obase + (uint32_t)(32 * j) * Pu;
                    if (valid) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            AVT_DV_ST(o + (uint32_t)(2 * e) * Pu, fmaf(-h_lo(xa[j][e]), rt, rn * __uint_as_float(r[2 * e])));
                            AVT_DV_ST(o + (uint32_t)(2 * e + 1) * Pu, fmaf(-h_hi(xa[j][e]), rt, rn * __uint_as_float(r[2 * e + 1])));
                        }
                    }
                }
                tc_fence_before();
                if (rank == 0) mbar_arrive(&tmem_empty[0]); else mbar_arrive_leader(&tmem_empty[0]);
            }
            if (nh == 2) {
                float* obase = dv + ((size_t)n * C + cb) * P + p;
#pragma unroll
                for (int j = 0; j < kJ; ++j) {
                    uint32_t r[32];
                    tmem_ld_32x32(lane_addr + (uint32_t)(cb + 32 * j), r);
                    tmem_ld_wait();
                    float* o = obase + (uint32_t)(32 * j) * Pu;
                    if (valid) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            AVT_DV_ST(o + (uint32_t)(2 * e) * Pu, fmaf(-h_lo(xb[j][e]), rt, rn * __uint_as_float(r[2 * e])));
                            AVT_DV_ST(o + (uint32_t)(2 * e + 1) * Pu, fmaf(-h_hi(xb[j][e]), rt, rn * __uint_as_float(r[2 * e + 1])));
                        }
                    }
                }
                tc_fence_before();
                if (rank == 0) mbar_arrive(&tmem_empty[1]); else mbar_arrive_leader(&tmem_empty[1]);
            }
            AVT_TRACE(te_p2 += clock64() - e0;)
        }
        AVT_TRACE(if (warp == 0 && lane == 0) { g_dv_trace[blockIdx.x * 8 + 3] = te_wait; g_dv_trace[blockIdx.x * 8 + 4] = te_p1;
                                                g_dv_trace[blockIdx.x * 8 + 5] = clock64() - te_begin; g_dv_trace[blockIdx.x * 8 + 7] = te_p2; })
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (warp == kWMma) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512);
    }
}

// ------------------------------------------------------------------ stage 3: d_ahat
constexpr int kDaThreads = 192;          // warps 0..3 epilogue, warp 4 TMA, warp 5 MMA + TMEM owner
constexpr int kDaWarpTma = 4, kDaWarpMma = 5;
constexpr int kDaBM = 128;
constexpr int kDaBK = 32;                // rows of np per stage
constexpr int kDaStages = 4;
constexpr int kDaBoxBytes = kDaBK * 128; // one [32 rows][64 elements] box
constexpr int kDaABytes = 2 * kDaBoxBytes;            // 128 audio columns
constexpr int kDaMaxStageBytes = kDaABytes + 8 * kDaBoxBytes;   // C <= 512
constexpr size_t kDaSmemBytes = 1024 + (size_t)kDaStages * kDaMaxStageBytes + 256;

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

__global__ void __launch_bounds__(kDaThreads, 1)
bwd_da_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_vt,
              float* __restrict__ d_ahat, int N_a, int n_mt, int splits, int n_chunks, int C,
                   const unsigned int* __restrict__ gmax_bits) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* stages = base;
    uint64_t* bars = reinterpret_cast<uint64_t*>(stages + (size_t)kDaStages * kDaMaxStageBytes);
    uint64_t* full_b = bars;
    uint64_t* empty_b = bars + kDaStages;
    uint64_t* tmem_full = bars + 2 * kDaStages;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mt = blockIdx.x % n_mt, split = blockIdx.x / n_mt;
    const int ch0 = (int)((long long)split * n_chunks / splits);
    const int ch1 = (int)((long long)(split + 1) * n_chunks / splits);
    const int cboxes = C >> 6;
    const uint32_t stage_bytes = (uint32_t)(kDaABytes + cboxes * kDaBoxBytes);
    const int nh = C > 256 ? 2 : 1;
    const int hN0 = C > 256 ? 256 : C, hN1 = C - hN0;

    if (warp == kDaWarpTma && lane == 0) {
        prefetch_tmap(&map_g);
        prefetch_tmap(&map_vt);
        for (int i = 0; i < kDaStages; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
    }
    if (warp == kDaWarpMma) tmem_alloc(tmem_holder, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;

    if (warp == kDaWarpTma) {
        if (lane == 0) {                                 // ---------------- TMA producer
            int stage = 0; uint32_t phase = 0;
            for (int ch = ch0; ch < ch1; ++ch) {
                mbar_wait(&empty_b[stage], phase ^ 1);
                mbar_arrive_expect_tx(&full_b[stage], stage_bytes);
                uint8_t* sA = stages + (size_t)stage * kDaMaxStageBytes;
                uint8_t* sB = sA + kDaABytes;
                for (int j = 0; j < 2; ++j)
                    tma_load_2d(sA + j * kDaBoxBytes, &map_g, &full_b[stage], mt * kDaBM + j * 64, ch * kDaBK);
                for (int j = 0; j < cboxes; ++j)
                    tma_load_2d(sB + j * kDaBoxBytes, &map_vt, &full_b[stage], j * 64, ch * kDaBK);
                if (++stage == kDaStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == kDaWarpMma) {
        {                                                // ---------------- MMA issuer (warp-uniform loop)
            // MN-major, 128-byte swizzle: 16 K rows = 2 groups of 8 rows (SBO 1024 B);
            // consecutive 64-element MN blocks are one box apart (LBO)
            const uint64_t desc0 = make_smem_desc(smem_u32(stages), kDaBoxBytes, 1024);
            int stage = 0; uint32_t phase = 0;
            for (int ch = ch0; ch < ch1; ++ch) {
                mbar_wait(&full_b[stage], phase);
                tc_fence_after();
                const uint64_t ad0 = desc0 + (uint64_t)((uint32_t)stage * (uint32_t)(kDaMaxStageBytes >> 4));
                const uint64_t bd0 = ad0 + (uint64_t)(kDaABytes >> 4);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < kDaBK / 16; ++k) {
                        for (int h = 0; h < nh; ++h) {
                            const uint32_t idesc = make_idesc_f16(kDaBM, h ? hN1 : hN0, 1, 1);
                            umma_f16_ss(tmem_base + (uint32_t)h * 256u, ad0 + (uint64_t)(k * (2048 >> 4)),
                                        bd0 + (uint64_t)((h * 4 * kDaBoxBytes + k * 2048) >> 4), idesc,
                                        (uint32_t)((ch != ch0) || (k != 0)));
                        }
                    }
                    umma_commit(&empty_b[stage]);
                }
                __syncwarp();
                if (++stage == kDaStages) { stage = 0; phase ^= 1; }
            }
            if (elect_one()) umma_commit(tmem_full);
            __syncwarp();
        }
    } else {                                             // ---------------- epilogue warps 0..3
        const int q = warp & 3;
        const int k = mt * kDaBM + q * 32 + lane;
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
        const float inv = pow2i(-g_scale_exp(__ldg(gmax_bits)));          // undo the scale of G
        for (int c = 0; c < C; c += 32) {
            uint32_t r[32];
            tmem_ld_32x32(lane_addr + (uint32_t)c, r);
            tmem_ld_wait();
            if (k < N_a) {
                float* o = d_ahat + (size_t)k * C + c;
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                    red_add_v4(o + i, inv * __uint_as_float(r[i]), inv * __uint_as_float(r[i + 1]),
                               inv * __uint_as_float(r[i + 2]), inv * __uint_as_float(r[i + 3]));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kDaWarpMma) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------------ stage 3, CTA pairs
// Same contraction with tcgen05.mma.cta_group::2: a cluster of two CTAs owns 256 audio rows
// (M = 256) of one K split.  Per 64-row K chunk each CTA pulls its own 128 columns of G
// (16 KB) but only HALF of the v_hat chunk (2 x 128 of the 512 channels, 32 KB instead of
// 64 KB); the hardware reads both halves of B across the pair.  Three TMA instructions per
// stage per CTA (3-D boxes that land as consecutive MN-major SW128 blocks).
constexpr int kDa2Threads = 192;
constexpr int kDa2BK = 64;
constexpr int kDa2Stages = 4;
constexpr int kDa2BlockBytes = kDa2BK * 128;          // one [64 rows][64 elements] block, 8 KB
constexpr int kDa2ABytes = 2 * kDa2BlockBytes;        // this CTA's 128 audio columns
constexpr int kDa2BHalfBytes = 2 * kDa2BlockBytes;    // this CTA's 128 channels of one N = 256 MMA
constexpr int kDa2StageBytes = kDa2ABytes + 2 * kDa2BHalfBytes;   // 48 KB
constexpr size_t kDa2SmemBytes = 1024 + (size_t)kDa2Stages * kDa2StageBytes + 256;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kDa2Threads, 1)
bwd_da_pair_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_vt,
                   float* __restrict__ d_ahat, int N_a, int n_mt, int splits, int n_chunks, int C,
                   const unsigned int* __restrict__ gmax_bits, int order) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* stages = base;
    uint64_t* bars = reinterpret_cast<uint64_t*>(stages + (size_t)kDa2Stages * kDa2StageBytes);
    uint64_t* full_b = bars;                 // used on the leader; both CTAs' TMA bytes land here
    uint64_t* empty_b = bars + kDa2Stages;   // one per CTA, released by the leader's multicast commit
    uint64_t* tmem_full = bars + 2 * kDa2Stages;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int mt = blockIdx.x % n_mt, split = blockIdx.x / n_mt;      // n_mt even: a pair shares `split`
    // order bit 0: chunks dealt round robin over the splits (all CTAs sweep the rows of G together, in step
    // with a dv kernel running beside this one, so G and v_hat come from HBM once for both); bit 1: last rows first
    const bool rr = (order & 1) != 0, rev = (order & 2) != 0;
    const int ch0 = rr ? 0 : (int)((long long)split * n_chunks / splits);
    const int ch1 = rr ? (n_chunks - split + splits - 1) / splits : (int)((long long)(split + 1) * n_chunks / splits);
    auto chunk_of = [&](int i) -> int { const int c = rr ? split + i * splits : i; return rev ? n_chunks - 1 - c : c; };
    const int nh = C / 256;                  // C is 256 or 512

    if (warp == kDaWarpTma && lane == 0) {
        prefetch_tmap(&map_g);
        prefetch_tmap(&map_vt);
        for (int i = 0; i < kDa2Stages; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
    }
    if (warp == kDaWarpMma) tmem_alloc2(tmem_holder, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync();                          // both CTAs' barriers are initialised before any remote arrive
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;

    if (warp == kDaWarpTma) {
        if (lane == 0) {                                 // ---------------- TMA producer (both CTAs)
            int stage = 0; uint32_t phase = 0;
            for (int ch = ch0; ch < ch1; ++ch) {
                mbar_wait(&empty_b[stage], phase ^ 1);
                if (rank == 0)                           // the leader's barrier counts both CTAs' bytes
                    mbar_arrive_expect_tx(&full_b[stage], (uint32_t)(2 * (kDa2ABytes + nh * kDa2BHalfBytes)));
                uint8_t* sA = stages + (size_t)stage * kDa2StageBytes;
                const int row0 = chunk_of(ch) * kDa2BK;
                tma_load_3d_pair(sA, &map_g, &full_b[stage], 0, row0, mt * 2);
                for (int h = 0; h < nh; ++h)
                    tma_load_3d_pair(sA + kDa2ABytes + h * kDa2BHalfBytes, &map_vt, &full_b[stage], 0,
                                     row0, h * 4 + (int)rank * 2);
                if (++stage == kDa2Stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == kDaWarpMma) {
#ifdef AVT_DA_LANE0
        if (rank == 0 && lane == 0) {
#else
        if (rank == 0) {                                 // ---------------- MMA issuer (leader only; warp-uniform loop)
#endif
            const uint32_t idesc = make_idesc_f16(256, 256, 1, 1);
            const uint64_t desc0 = make_smem_desc(smem_u32(stages), kDa2BlockBytes, 1024);
            int stage = 0; uint32_t phase = 0;
            for (int ch = ch0; ch < ch1; ++ch) {
                mbar_wait(&full_b[stage], phase);
                tc_fence_after();
                const uint64_t ad0 = desc0 + (uint64_t)((uint32_t)stage * (uint32_t)(kDa2StageBytes >> 4));
                const uint64_t bd0 = ad0 + (uint64_t)(kDa2ABytes >> 4);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < kDa2BK / 16; ++k) {
                        for (int h = 0; h < nh; ++h)
                            umma2_f16_ss(tmem_base + (uint32_t)h * 256u, ad0 + (uint64_t)(k * (2048 >> 4)),
                                         bd0 + (uint64_t)((h * kDa2BHalfBytes + k * 2048) >> 4), idesc,
                                         (uint32_t)((ch != ch0) || (k != 0)));
                    }
                    umma2_commit_mc(&empty_b[stage], 3);     // frees the stage in both CTAs
                }
                __syncwarp();
                if (++stage == kDa2Stages) { stage = 0; phase ^= 1; }
            }
            if (elect_one()) umma2_commit_mc(tmem_full, 3);
            __syncwarp();
        }
    } else {                                             // ---------------- epilogue warps 0..3
        const int q = warp & 3;
        const int k = mt * 128 + q * 32 + lane;
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
        const float inv = pow2i(-g_scale_exp(__ldg(gmax_bits)));          // undo the scale of G
        for (int c = 0; c < C; c += 32) {
            uint32_t r[32];
            tmem_ld_32x32(lane_addr + (uint32_t)c, r);
            tmem_ld_wait();
            if (k < N_a) {
                float* o = d_ahat + (size_t)k * C + c;
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                    red_add_v4(o + i, inv * __uint_as_float(r[i]), inv * __uint_as_float(r[i + 1]),
                               inv * __uint_as_float(r[i + 2]), inv * __uint_as_float(r[i + 3]));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();                          // nobody leaves while the pair may still touch its smem / TMEM
    if (warp == kDaWarpMma) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512);
    }
}

// ------------------------------------------------------------------ launch sequence
int launch_grad_cells_f16(const __half* vt, const __half* a_h, const int32_t* pos_idx,
                           int col_offset, int N_v, int N_a, int n_a_pad, int C, int P, HeadParams hp,
                           const float* sim, const float* den, const float* gsim, const float* GA,
                           __half* G, const unsigned int* gmax_bits, cudaStream_t st);

// The dv and d_ahat contractions both consume G and are independent of each other: d_ahat is
// enqueued on a side stream so that its CTAs fill the SMs the dv kernel's last wave leaves idle
// (matters for small shards: at 128 rows per rank the two kernels are ~1.3 waves each).  Plain
// event fork / join, so it is captured into CUDA graphs like any cross-stream dependency.
//
// One side context PER (device, caller stream): two host threads that drive the same device on
// their own streams (the nn.DataParallel caller of SURVEY 8b) never share an event, so one
// caller's record can not be consumed by the other's wait.  The table only grows with the number
// of distinct caller streams (bounded; past the bound the d_ahat kernel simply stays on the
// caller's stream) and the mutex guards the lookup alone.
struct SideStream {
    int dev = -1;
    cudaStream_t owner = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr, fork0 = nullptr, tdone = nullptr;
};
constexpr int kMaxSideCtx = 128;
static SideStream* side_stream(int dev, cudaStream_t owner) {
    static std::mutex mu;
    static SideStream ctx[kMaxSideCtx];
    static int used = 0;
    std::lock_guard<std::mutex> lock(mu);
    for (int i = 0; i < used; ++i)
        if (ctx[i].dev == dev && ctx[i].owner == owner) return &ctx[i];
    if (used == kMaxSideCtx) return nullptr;
    // creating a stream / events is illegal while `owner` is capturing only for the legacy
    // stream; ours are non-blocking streams and plain events, allowed under relaxed capture
    cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
    cudaThreadExchangeStreamCaptureMode(&mode);
    SideStream c;
    c.dev = dev; c.owner = owner;
    const bool ok = cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking) == cudaSuccess &&
                    cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming) == cudaSuccess &&
                    cudaEventCreateWithFlags(&c.join, cudaEventDisableTiming) == cudaSuccess &&
                    cudaEventCreateWithFlags(&c.fork0, cudaEventDisableTiming) == cudaSuccess &&
                    cudaEventCreateWithFlags(&c.tdone, cudaEventDisableTiming) == cudaSuccess;
    cudaThreadExchangeStreamCaptureMode(&mode);
    if (!ok) { (void)cudaGetLastError(); return nullptr; }
    ctx[used] = c;
    return &ctx[used++];
}

// AVT_NO_SIDE_STREAM / AVT_NO_PAIR are development switches: read ONCE per process
struct EnvSwitches { bool no_side, no_pair; int dv_groups; bool dv_bt; int dv_pairs, da_splits; bool no_cosched; };
static const EnvSwitches& env_switches() {
    static const EnvSwitches e{getenv("AVT_NO_SIDE_STREAM") != nullptr, getenv("AVT_NO_PAIR") != nullptr,
                                 getenv("AVT_DV_GROUPS") ? atoi(getenv("AVT_DV_GROUPS")) : 2,
                                 getenv("AVT_DV_BT") != nullptr,     // development: the first version's transposed-audio operand
                                 getenv("AVT_DV_PAIRS") ? atoi(getenv("AVT_DV_PAIRS")) : 0,
                                 getenv("AVT_DA_SPLITS") ? atoi(getenv("AVT_DA_SPLITS")) : 0,
                                 getenv("AVT_NO_COSCHED") != nullptr};
    return e;
}

// joins the side stream back into the caller's stream on EVERY exit path once it has forked, so an
// error return can not leave a capturing stream with an unjoined branch
struct SideJoin {
    SideStream* side; cudaStream_t st; bool forked = false;
    ~SideJoin() {
        if (!side || !forked) return;
        if (cudaEventRecord(side->join, side->stream) == cudaSuccess) cudaStreamWaitEvent(st, side->join, 0);
    }
};

int launch_bwd_f16(const float* v, const __half* vt, const __half* a_h,
                    __half* at_h, __half* g_h, const float* a_hat, const float* rnorm_v,
                    const int32_t* pos_idx, int col_offset, int N_v, int N_a, int n_a_pad, int C, int P,
                    HeadParams hp, const float* sim, const float* den, const float* gsim,
                    const float* GA, float* dv, float* d_ahat, const unsigned int* gmax_bits,
                    cudaStream_t dv_stream, cudaStream_t st) {
    (void)a_hat; (void)v;
    AVT_REQUIRE(C % 64 == 0 && C <= 512, "fp16 backward: C=%d must be a multiple of 64, <= 512", C);
    const int n_rows = N_v * P;
    int dev = 0, sms = 148;
    AVT_CHECK_CUDA(cudaGetDevice(&dev));
    AVT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));

    // Stage 2 (dv) and stage 3 (d_ahat) both consume G and are independent.  Two ways to overlap them:
    //  * caller-provided dv_stream: dv (and the audio transpose it needs) go THERE, d_ahat stays on
    //    `st`, and nothing is joined here -- the sharded step starts its reduce_scatter of d_ahat the
    //    moment stage 3 is done, under the dv kernel, and joins dv_stream when it needs dv;
    //  * otherwise an internal side stream carries the transpose and stage 3 and is joined back
    //    into `st` before this function returns.
    // With stage profiling on, everything runs on `st` so that the stage times are those of the
    // kernels alone.
    const EnvSwitches& env = env_switches();
    // the pair dv kernel reads a_hat itself as an MN-major operand; only the 1-CTA kernel (C other than
    // 256 / 512) still wants the transposed copy
    const bool dv_pair = (C == 256 || C == 512) && !env.no_pair;
    const bool b_mn = dv_pair && !env.dv_bt;
    const bool serial = prof_enabled() || env.no_side;
    // Co-scheduling.  The dv kernel alternates between a load phase (MMAs) and a store phase (its
    // epilogue holds both accumulator halves: the Jacobian needs the whole row), so alone it keeps the
    // tensor pipe 44 % and HBM 65 % busy; the d_ahat kernel is tensor-bound and reads the SAME G and
    // v_hat.  Run side by side on disjoint SMs (5/8 of the SM pairs for dv, the rest for d_ahat), both
    // sweeping the rows of G from the END (where the cell-gradient kernel has just left them in L2) and
    // in step (d_ahat's K chunks dealt round robin over its splits), each fills the other's idle
    // resource and G / v_hat come from HBM once for both: backward 637 -> 582 us at B = 1024,
    // 120 -> 105 us for a 128-row shard (profiles/r2_ab_cosched.txt).
    const bool cosched = dv_pair && !serial && !env.no_cosched;
    const bool reverse = cosched;
    SideStream* side = serial ? nullptr : side_stream(dev, st);
    const bool ext = side && dv_stream && dv_stream != st;
    cudaStream_t st_dv = ext ? dv_stream : st;
    cudaStream_t st_da = (side && !ext) ? side->stream : st;
    cudaStream_t st_tr = ext ? dv_stream : st_da;        // stream of the audio transpose
    SideJoin joiner{ext ? nullptr : side, st};
    if (!b_mn) {
        if (side) {
            AVT_CHECK_CUDA(cudaEventRecord(side->fork0, st));
            AVT_CHECK_CUDA(cudaStreamWaitEvent(st_tr, side->fork0, 0));
            joiner.forked = true;
        }
        dim3 grid(n_a_pad / 32, C / 32), block(32, 8);
        transpose_audio_kernel<<<grid, block, 0, st_tr>>>(a_h, n_a_pad, C, at_h);
        AVT_LAUNCH_CHECK();
        if (side && !ext) AVT_CHECK_CUDA(cudaEventRecord(side->tdone, side->stream));
    }
    // stage 1: cell gradients
    int rc = launch_grad_cells_f16(vt, a_h, pos_idx, col_offset, N_v, N_a, n_a_pad, C, P, hp, sim, den,
                                    gsim, GA, g_h, gmax_bits, st);
    if (rc != AVT_OK) return rc;
    if (side) {
        if (!ext && !b_mn) AVT_CHECK_CUDA(cudaStreamWaitEvent(st, side->tdone, 0));
        AVT_CHECK_CUDA(cudaEventRecord(side->fork, st));
        AVT_CHECK_CUDA(cudaStreamWaitEvent(ext ? st_dv : side->stream, side->fork, 0));
        joiner.forked = true;
    }
    // stage 2: dv
    int dv_ctas = 0;
    {
        ProfScope _ps(kProfBwdDv, st_dv);
        const bool pair = dv_pair;
        CUtensorMap map_g, map_at;
        const bool ok_b = b_mn ? make_tmap_f16_blocks(&map_at, a_h, (uint64_t)n_a_pad, (uint64_t)C, (uint64_t)C * 2, kDvBK, 2)
                               : make_tmap_f16_2d(&map_at, at_h, (uint64_t)C, (uint64_t)n_a_pad, (uint64_t)n_a_pad * 2, kDvBK,
                                                  pair ? 128 : 256);
        if (!make_tmap_f16_2d(&map_g, g_h, (uint64_t)n_rows, (uint64_t)n_a_pad, (uint64_t)n_a_pad * 2, kDvBK, kDvBM) || !ok_b) {
            set_error("fp16 backward: cuTensorMapEncodeTiled failed (dv)");
            return AVT_ERR_CUDA;
        }
        const int n_tiles = (n_rows + kDvBM - 1) / kDvBM;
        if (pair) {
            const int n_pairs = (n_tiles + 1) / 2;
            int grid_pairs = n_pairs < sms / 2 ? n_pairs : sms / 2;
            if (cosched) {
                // dv's share of the SM pairs: its store-bound epilogue costs the same per row whatever N_a is,
                // the d_ahat contraction shrinks with N_a -- 5/8 at >= 1024 audio rows, 2/3 at 512, 3/4 below
                // (profiles/r2_ab_cosched_small.txt: backward 63.6 -> 52.3 us at B = 128, 114.8 -> 101.3 at 256)
                const int slots = sms / 2;
                const int share = n_a_pad >= 1024 ? (slots * 5 + 4) / 8 : n_a_pad >= 512 ? (slots * 2 + 1) / 3
                                                                                       : (slots * 3 + 2) / 4;
                const int want = env.dv_pairs > 0 ? env.dv_pairs : share;
                if (want >= 1 && want < grid_pairs) grid_pairs = want;
            }
            dv_ctas = 2 * grid_pairs;
#define AVT_DV(G, PP, BM)                                                                                      \
    do {                                                                                                       \
        AVT_CHECK_CUDA(cudaFuncSetAttribute(bwd_dv_pair_kernel<G, PP, BM>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                            (int)kDv2SmemBytes));                                              \
        bwd_dv_pair_kernel<G, PP, BM><<<2 * grid_pairs, 64 + 128 * G, kDv2SmemBytes, st_dv>>>(                   \
            map_g, map_at, vt, rnorm_v, dv, n_rows, n_pairs, n_a_pad / kDvBK, C, P, reverse ? 1 : 0, gmax_bits);  \
    } while (0)
#define AVT_DV_P(G, BM) do { if (P == 196) AVT_DV(G, 196, BM); else AVT_DV(G, 0, BM); } while (0)
            if (env.dv_groups == 4) { if (b_mn) AVT_DV_P(4, true); else AVT_DV_P(4, false); }
            else { if (b_mn) AVT_DV_P(2, true); else AVT_DV_P(2, false); }
#undef AVT_DV_P
#undef AVT_DV
        } else {
            const int grid = n_tiles < sms ? n_tiles : sms;
            AVT_CHECK_CUDA(cudaFuncSetAttribute(bwd_dv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)kDvSmemBytes));
            bwd_dv_kernel<<<grid, kDvThreads, kDvSmemBytes, st_dv>>>(map_g, map_at, vt, rnorm_v, dv, n_rows,
                                                                 n_tiles, n_a_pad / kDvBK, C, P, gmax_bits);
        }
        AVT_LAUNCH_CHECK();
    }
    // stage 3: d_ahat (zeroed by the caller)
    {
        ProfScope _ps(kProfBwdDa, st_da);
        const int n_mt = n_a_pad / kDaBM;
        const bool pair = (n_mt % 2 == 0) && (C == 256 || C == 512) && !env.no_pair;
        const int bk = pair ? kDa2BK : kDaBK;
        CUtensorMap map_g, map_vt;
        bool ok;
        if (pair)
            ok = make_tmap_f16_blocks(&map_g, g_h, (uint64_t)n_rows, (uint64_t)n_a_pad, (uint64_t)n_a_pad * 2, bk, 2) &&
                 make_tmap_f16_blocks(&map_vt, vt, (uint64_t)n_rows, (uint64_t)C, (uint64_t)C * 2, bk, 2);
        else
            ok = make_tmap_f16_2d(&map_g, g_h, (uint64_t)n_rows, (uint64_t)n_a_pad, (uint64_t)n_a_pad * 2, 64, bk) &&
                 make_tmap_f16_2d(&map_vt, vt, (uint64_t)n_rows, (uint64_t)C, (uint64_t)C * 2, 64, bk);
        if (!ok) {
            set_error("fp16 backward: cuTensorMapEncodeTiled failed (d_ahat)");
            return AVT_ERR_CUDA;
        }
        const int n_chunks = (n_rows + bk - 1) / bk;
        int splits = sms / n_mt;                 // one split per SM: fewer splits cut the atomics but lose
                                                 // more in streaming bandwidth (measured at 128..512 rows/rank)
        if (cosched && dv_ctas > 0 && dv_ctas < sms) {      // the SMs the dv kernel leaves free
            const int want = env.da_splits > 0 ? env.da_splits : (sms - dv_ctas) / n_mt;
            if (want >= 1 && want < splits) splits = want;
        }
        if (splits < 1) splits = 1;
        if (splits > n_chunks) splits = n_chunks;
        if (pair) {
            AVT_CHECK_CUDA(cudaFuncSetAttribute(bwd_da_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)kDa2SmemBytes));
            bwd_da_pair_kernel<<<n_mt * splits, kDa2Threads, kDa2SmemBytes, st_da>>>(
                map_g, map_vt, d_ahat, N_a, n_mt, splits, n_chunks, C, gmax_bits, (cosched ? 1 : 0) | (reverse ? 2 : 0));
        } else {
            AVT_CHECK_CUDA(cudaFuncSetAttribute(bwd_da_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)kDaSmemBytes));
            bwd_da_kernel<<<n_mt * splits, kDaThreads, kDaSmemBytes, st_da>>>(map_g, map_vt, d_ahat, N_a, n_mt,
                                                                          splits, n_chunks, C, gmax_bits);
        }
        AVT_LAUNCH_CHECK();
    }
    return AVT_OK;                                   // `joiner` joins the side stream
}

}  // namespace avt

#ifdef AVT_TC_TRACE
extern "C" int avt_debug_dv_trace(unsigned long long* host_out, int n_blocks) {
    return cudaMemcpyFromSymbol(host_out, avt::g_dv_trace, sizeof(unsigned long long) * 8 * (size_t)n_blocks) == cudaSuccess ? 0 : -2;
}
#endif
