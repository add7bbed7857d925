// Evaluation kernels: the reference's per-clip glue (train_hardway_1frame.py:198-207 ==
// test.py:97-107) as ONE launch -- cv2.resize 14->224 (INTER_LINEAR), normalize_img(-h),
// pred = 1 - h, exact median-rank threshold, binarise, cal_CIOU -- plus GT rasterisation
// (utils.py:241-309), the success-curve / AUC reduction (utils.py:216-229) and mTC
// (utils.py:311-318).
//
// One block of 224 threads per map; nothing but the 14x14 map (and the GT) is read from
// HBM and the only large writes are the optional 224x224 outputs.  The k-th order statistic
// is found without sorting: every pixel value is recomputable from 12.5 KB of
// horizontally interpolated rows in shared memory, so a 56x56 sub-sample brackets the
// median, one sweep over the pixels counts / collects the few thousand candidates inside
// the bracket, and an exact radix refinement on that list returns the threshold.  A miss
// of the bracket falls back to counted bisection over the full image (still exact).
//
// Bit-exactness: pred = f(up) with f monotone non-decreasing under round-to-nearest
// (negate, subtract vmin, divide by a positive range, subtract from 1), hence
// sorted(pred)[k] == f(sorted(up)[k]) and {pred > thr or pred >= 1} is an up-set of `up`
// whose boundary is located by searching the float key space with the very same f.
#include "avt_common.cuh"

namespace avt {

constexpr int kSrc = 14;
constexpr int kDst = 224;
constexpr int kNpix = kDst * kDst;               // 50176
constexpr int kRank = kNpix / 2;                 // int(224*224/2), train_hardway_1frame.py:202
constexpr int kEvThreads = 224;                  // 56 column quads x 4 row groups
constexpr int kEvWarps = kEvThreads / 32;        // 7
constexpr int kRowsPerGroup = kDst / 4;          // 56
constexpr int kCap = 15 * kEvThreads;            // 3360 candidates (4-bit private counters)
constexpr int kSampleSide = 56;                  // every 4th pixel in x and y
constexpr int kNs = kSampleSide * kSampleSide;   // 3136 samples (14 per thread)
constexpr int kMargin = 80;                      // sample ranks either side of the median
constexpr int kBins = 64;
constexpr int kMaxBoxesSmem = 32;

struct EvalSmem {
    float m[kSrc * kSrc];                        // the 14x14 map
    float rows[kSrc][kDst];                      // horizontally interpolated rows
    uint32_t list[kCap];                         // sample keys, later candidate keys
    uint32_t cnt[kBins / 8][kEvThreads];         // private nibble counters, bank = thread
    uint32_t whist[kEvWarps][kBins];
    uint32_t red[4][kEvWarps];
    uint32_t bcast[4];
    int n_list;
    int16_t boxes[kMaxBoxesSmem][4];
};

// total order on floats as unsigned keys (-0 < +0)
__device__ __forceinline__ uint32_t f2key(float f) {
    uint32_t b = __float_as_uint(f);
    return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

// cv2 / oracle taps for one destination coordinate: f = (x + .5)/16 - .5, clamped
__device__ __forceinline__ void tap(int x, int& s0, int& s1, float& w0, float& w1) {
    int t = 2 * x - 15;                          // f = t / 32
    int s = t >> 5;
    float w = (float)(t & 31) * (1.0f / 32.0f);
    if (s < 0) { s = 0; w = 0.f; }
    if (s >= kSrc - 1) { s = kSrc - 1; w = 0.f; }
    s0 = s;
    s1 = min(s + 1, kSrc - 1);
    w1 = w;
    w0 = 1.0f - w;                               // exact: w is a multiple of 1/32
}
// a*w0 + b*w1 with every product and the sum rounded to fp32 (no FMA contraction)
__device__ __forceinline__ float lerp_rn(float a, float w0, float b, float w1) {
    return __fadd_rn(__fmul_rn(a, w0), __fmul_rn(b, w1));
}

__device__ __forceinline__ uint32_t block_sum_u(uint32_t x, uint32_t* red) {
    x = __reduce_add_sync(0xffffffffu, x);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = x;
    __syncthreads();
    uint32_t t = 0;
#pragma unroll
    for (int i = 0; i < kEvWarps; ++i) t += red[i];
    return t;
}
__device__ __forceinline__ uint32_t block_min_u(uint32_t x, uint32_t* red) {
    x = __reduce_min_sync(0xffffffffu, x);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = x;
    __syncthreads();
    uint32_t t = red[0];
#pragma unroll
    for (int i = 1; i < kEvWarps; ++i) t = min(t, red[i]);
    return t;
}
__device__ __forceinline__ uint32_t block_max_u(uint32_t x, uint32_t* red) {
    x = __reduce_max_sync(0xffffffffu, x);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = x;
    __syncthreads();
    uint32_t t = red[0];
#pragma unroll
    for (int i = 1; i < kEvWarps; ++i) t = max(t, red[i]);
    return t;
}

// r-th smallest (0-based, among the keys inside [lo, hi]) of list[0..n).  Exact: each
// round histograms the top <= 6 undecided bits into 64 bins with conflict-free private
// 4-bit counters (a thread sees <= 15 keys), then descends into the bin holding rank r.
__device__ uint32_t list_select(EvalSmem& S, int n, uint32_t r, uint32_t lo, uint32_t hi) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int round = 0; round < 8 && lo != hi; ++round) {      // <= 6 rounds cover 32 bits
        const uint32_t span = hi - lo;
        const int s = max(0, (32 - __clz(span)) - 6);        // (span >> s) < 64
#pragma unroll
        for (int w = 0; w < kBins / 8; ++w) S.cnt[w][tid] = 0u;
        for (int i = tid; i < n; i += kEvThreads) {
            uint32_t key = S.list[i];
            if (key >= lo && key <= hi) {
                uint32_t b = (key - lo) >> s;
                S.cnt[b >> 3][tid] += 1u << ((b & 7u) * 4u);
            }
        }
        uint32_t mine0 = 0, mine1 = 0;
#pragma unroll
        for (int w = 0; w < kBins / 8; ++w) {
            uint32_t word = S.cnt[w][tid];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                uint32_t tot = __reduce_add_sync(0xffffffffu, (word >> (4 * q)) & 15u);
                const int b = w * 8 + q;
                if ((b & 31) == lane) { if (b < 32) mine0 = tot; else mine1 = tot; }
            }
        }
        S.whist[warp][lane] = mine0;
        S.whist[warp][lane + 32] = mine1;
        __syncthreads();
        if (warp == 0) {
            uint32_t c0 = 0, c1 = 0;
#pragma unroll
            for (int w = 0; w < kEvWarps; ++w) { c0 += S.whist[w][lane]; c1 += S.whist[w][lane + 32]; }
            uint32_t i0 = c0, i1 = c1;                       // inclusive scans over lanes
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t a = __shfl_up_sync(0xffffffffu, i0, o);
                uint32_t b = __shfl_up_sync(0xffffffffu, i1, o);
                if (lane >= o) { i0 += a; i1 += b; }
            }
            const uint32_t tot0 = __shfl_sync(0xffffffffu, i0, 31);
            const uint32_t e0 = i0 - c0, e1 = tot0 + i1 - c1;    // exclusive prefix of each bin
            const bool h0 = (r >= e0 && r < e0 + c0), h1 = (r >= e1 && r < e1 + c1);
            if (h0) { S.bcast[0] = lane; S.bcast[1] = e0; }
            if (h1) { S.bcast[0] = lane + 32; S.bcast[1] = e1; }
            if (!__any_sync(0xffffffffu, h0 || h1) && lane == 0) { S.bcast[0] = 0; S.bcast[1] = 0; }
        }
        __syncthreads();
        const uint32_t b = S.bcast[0];
        r -= S.bcast[1];
        const uint32_t nlo = lo + (b << s);
        uint32_t nhi = nlo + ((1u << s) - 1u);
        if (nhi < nlo) nhi = 0xffffffffu;
        lo = nlo;
        hi = min(hi, nhi);
    }
    return lo;
}

// One sweep over all 50176 pixels: c_lt = #(key < blo), c_in = #(blo <= key <= bhi),
// key range, and (optionally) the in-bracket keys appended to S.list (first kCap only).
__device__ void full_pass(EvalSmem& S, uint32_t blo, uint32_t bhi, bool collect,
                          uint32_t& c_lt, uint32_t& c_in, uint32_t& kmin, uint32_t& kmax) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int cg = tid % 56, rg = tid / 56;
    uint32_t n_lt = 0, n_in = 0, mn = 0xffffffffu, mx = 0u;
    if (tid == 0) S.n_list = 0;
    __syncthreads();
    for (int yy = 0; yy < kRowsPerGroup; ++yy) {
        const int y = rg * kRowsPerGroup + yy;
        int s0, s1; float w0, w1;
        tap(y, s0, s1, w0, w1);
        const float4 a = *reinterpret_cast<const float4*>(&S.rows[s0][4 * cg]);
        const float4 b = *reinterpret_cast<const float4*>(&S.rows[s1][4 * cg]);
        const float u[4] = {lerp_rn(a.x, w0, b.x, w1), lerp_rn(a.y, w0, b.y, w1),
                            lerp_rn(a.z, w0, b.z, w1), lerp_rn(a.w, w0, b.w, w1)};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t key = f2key(u[j]);
            mn = min(mn, key); mx = max(mx, key);
            n_lt += (key < blo);
            const bool in = (key >= blo) && (key <= bhi);
            n_in += in;
            if (collect) {
                const uint32_t hit = __ballot_sync(0xffffffffu, in);
                if (hit) {
                    int base = 0;
                    const int leader = __ffs(hit) - 1;
                    if (lane == leader) base = atomicAdd(&S.n_list, __popc(hit));
                    base = __shfl_sync(0xffffffffu, base, leader);
                    if (in) {
                        const int idx = base + __popc(hit & ((1u << lane) - 1u));
                        if (idx < kCap) S.list[idx] = key;
                    }
                }
            }
        }
    }
    c_lt = block_sum_u(n_lt, S.red[0]);
    c_in = block_sum_u(n_in, S.red[1]);
    kmin = block_min_u(mn, S.red[2]);
    kmax = block_max_u(mx, S.red[3]);
}

// f of the header comment: utils.py:234-239 applied to -u, then pred = 1 - h
__device__ __forceinline__ float pred_of(float u, bool normalised, float vmin_h, float range) {
    float h = -u;
    if (normalised) h = __fdiv_rn(__fsub_rn(h, vmin_h), range);
    return __fsub_rn(1.0f, h);
}

template <bool kDenseGt>
__global__ void __launch_bounds__(kEvThreads)
eval_maps_kernel(const float* __restrict__ maps14, int N, const uint8_t* __restrict__ gt_dense,
                 const int16_t* __restrict__ boxes, const uint8_t* __restrict__ box_count,
                 int max_boxes, int gt_mode, float* __restrict__ up224, uint8_t* __restrict__ mask,
                 int32_t* __restrict__ inter_w, int32_t* __restrict__ union_w,
                 double* __restrict__ ciou) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EvalSmem& S = *reinterpret_cast<EvalSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cg = tid % 56, rg = tid / 56;

    for (int map = blockIdx.x; map < N; map += gridDim.x) {
        __syncthreads();
        if (tid < kSrc * kSrc) S.m[tid] = maps14[(size_t)map * kSrc * kSrc + tid];
        int nb = 0;
        if (!kDenseGt) {
            nb = min((int)box_count[map], min(max_boxes, kMaxBoxesSmem));
            if (tid < nb * 4)
                (&S.boxes[0][0])[tid] = boxes[((size_t)map * max_boxes) * 4 + tid];
        }
        __syncthreads();
        // horizontal pass: rows[j][x] = m[j][sx0]*(1-wx) + m[j][sx1]*wx
        {
            int s0, s1; float w0, w1;
            tap(tid, s0, s1, w0, w1);                        // tid == destination column
#pragma unroll
            for (int j = 0; j < kSrc; ++j)
                S.rows[j][tid] = lerp_rn(S.m[j * kSrc + s0], w0, S.m[j * kSrc + s1], w1);
        }
        __syncthreads();
        // ---- bracket the median from a 56x56 sub-sample
        uint32_t smn = 0xffffffffu, smx = 0u;
        for (int i = tid; i < kNs; i += kEvThreads) {
            const int sy = 4 * (i / kSampleSide) + 2, sx = 4 * (i % kSampleSide) + 1;
            int s0, s1; float w0, w1;
            tap(sy, s0, s1, w0, w1);
            const uint32_t key = f2key(lerp_rn(S.rows[s0][sx], w0, S.rows[s1][sx], w1));
            S.list[i] = key;
            smn = min(smn, key); smx = max(smx, key);
        }
        smn = block_min_u(smn, S.red[0]);
        smx = block_max_u(smx, S.red[1]);
        const uint32_t blo = list_select(S, kNs, kNs / 2 - kMargin, smn, smx);
        const uint32_t bhi = list_select(S, kNs, kNs / 2 + kMargin, smn, smx);
        // ---- count and collect
        uint32_t c_lt, c_in, kmin, kmax;
        full_pass(S, blo, bhi, true, c_lt, c_in, kmin, kmax);
        uint32_t kstar;
        if (c_lt <= (uint32_t)kRank && (uint32_t)kRank < c_lt + c_in && c_in <= (uint32_t)kCap) {
            kstar = list_select(S, (int)c_in, (uint32_t)kRank - c_lt, blo, bhi);
        } else {
            // exact fallback: counted bisection over the key range until the rest fits the list
            uint32_t lo, hi, c_lo, n_in;
            if ((uint32_t)kRank < c_lt) { lo = kmin; hi = blo - 1u; c_lo = 0; n_in = c_lt; }
            else if ((uint32_t)kRank >= c_lt + c_in) { lo = bhi + 1u; hi = kmax; c_lo = c_lt + c_in; n_in = kNpix - c_lo; }
            else { lo = blo; hi = bhi; c_lo = c_lt; n_in = c_in; }
            for (;;) {
                if (lo == hi) { kstar = lo; break; }
                uint32_t a, b, c, d;
                if (n_in <= (uint32_t)kCap) {
                    full_pass(S, lo, hi, true, a, b, c, d);
                    kstar = list_select(S, (int)n_in, (uint32_t)kRank - c_lo, lo, hi);
                    break;
                }
                const uint32_t mid = lo + (hi - lo) / 2u;
                full_pass(S, lo, mid, false, a, b, c, d);        // b = #(lo <= key <= mid)
                if ((uint32_t)kRank - c_lo < b) { hi = mid; n_in = b; }
                else { lo = mid + 1u; c_lo += b; n_in -= b; }
            }
        }
        // ---- threshold in pred space, and the first key whose pixel is switched on
        const float umin = key2f(kmin), umax = key2f(kmax);
        const float vmin_h = -umax;
        const float range = __fsub_rn(-umin, vmin_h);           // vmax - vmin, utils.py:237
        const bool normalised = !(range == 0.0f);
        const float thr = pred_of(key2f(kstar), normalised, vmin_h, range);
        if (warp == 0) {
            // smallest key K in [kmin, kmax+1] with (pred > thr or pred >= 1); 32-way search
            uint32_t lo = kmin, hi = kmax + 1u;
            while (lo < hi) {
                const uint32_t step = (hi - lo) / 32u + 1u;
                const uint64_t pk = (uint64_t)lo + (uint64_t)lane * step;
                bool on = true;
                if (pk < (uint64_t)hi) {
                    const float pr = pred_of(key2f((uint32_t)pk), normalised, vmin_h, range);
                    on = (pr > thr) || (pr >= 1.0f);
                }
                const uint32_t ball = __ballot_sync(0xffffffffu, on);
                const int j = ball ? __ffs(ball) - 1 : 32;       // 32: every probe was valid and off
                if (j == 0) { hi = lo; }
                else {
                    const uint64_t nh = (uint64_t)lo + (uint64_t)j * step;
                    const uint32_t nl = lo + (uint32_t)(j - 1) * step + 1u;
                    if (nh < (uint64_t)hi) hi = (uint32_t)nh;
                    lo = nl;
                }
            }
            if (lane == 0) S.bcast[2] = lo;
        }
        __syncthreads();
        const uint32_t k_on = S.bcast[2];
        // ---- final sweep: materialise, binarise, count
        const int gt_den = (gt_mode == AVT_GT_FLICKR) ? 2 : 1;
        uint32_t inter = 0, gsum = 0, zc = 0;
        float* up_o = up224 ? up224 + (size_t)map * kNpix : nullptr;
        uint8_t* mk_o = mask ? mask + (size_t)map * kNpix : nullptr;
        const uint8_t* gt_i = kDenseGt ? gt_dense + (size_t)map * kNpix : nullptr;
        for (int yy = 0; yy < kRowsPerGroup; ++yy) {
            const int y = rg * kRowsPerGroup + yy;
            int s0, s1; float w0, w1;
            tap(y, s0, s1, w0, w1);
            const float4 a = *reinterpret_cast<const float4*>(&S.rows[s0][4 * cg]);
            const float4 b = *reinterpret_cast<const float4*>(&S.rows[s1][4 * cg]);
            float4 u;
            u.x = lerp_rn(a.x, w0, b.x, w1); u.y = lerp_rn(a.y, w0, b.y, w1);
            u.z = lerp_rn(a.z, w0, b.z, w1); u.w = lerp_rn(a.w, w0, b.w, w1);
            const size_t o = (size_t)y * kDst + 4 * cg;
            if (up_o) __stcs(reinterpret_cast<float4*>(up_o + o), u);
            uint32_t m4 = (f2key(u.x) >= k_on ? 1u : 0u) | (f2key(u.y) >= k_on ? 0x100u : 0u) |
                          (f2key(u.z) >= k_on ? 0x10000u : 0u) | (f2key(u.w) >= k_on ? 0x1000000u : 0u);
            if (mk_o) __stcs(reinterpret_cast<unsigned int*>(mk_o + o), m4);
            uint32_t g4;
            if (kDenseGt) {
                g4 = __ldcs(reinterpret_cast<const unsigned int*>(gt_i + o));
            } else {
                g4 = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int x = 4 * cg + j;
                    int c = 0;
                    for (int q = 0; q < nb; ++q)
                        c += (x >= S.boxes[q][0] && x < S.boxes[q][2] && y >= S.boxes[q][1] && y < S.boxes[q][3]);
                    if (gt_mode == AVT_GT_VGGSS) c = min(c, 1);
                    else if (gt_mode == AVT_GT_FLICKR) c = min(c, 2);
                    else c = min(c, 255);
                    g4 |= (uint32_t)c << (8 * j);
                }
            }
            inter = __dp4a(m4, g4, inter);
            gsum = __dp4a(g4, 0x01010101u, gsum);
            zc = __dp4a(m4, __vcmpeq4(g4, 0u) & 0x01010101u, zc);
        }
        inter = block_sum_u(inter, S.red[0]);
        gsum = block_sum_u(gsum, S.red[1]);
        zc = block_sum_u(zc, S.red[2]);
        if (tid == 0) {
            const uint32_t uni = gsum + (uint32_t)gt_den * zc;
            inter_w[map] = (int32_t)inter;
            union_w[map] = (int32_t)uni;
            ciou[map] = (double)inter / (double)uni;             // 0/0 -> NaN as numpy
        }
    }
}

// utils.py:241-274 / 276-309 -- boxes -> dense weights.  One block per map.
__global__ void rasterise_gt_kernel(const int16_t* __restrict__ boxes,
                                    const uint8_t* __restrict__ box_count, int max_boxes,
                                    int gt_mode, uint8_t* __restrict__ gt_dense) {
    __shared__ int16_t sb[kMaxBoxesSmem][4];
    const int map = blockIdx.x;
    const int nb = min((int)box_count[map], min(max_boxes, kMaxBoxesSmem));
    if (threadIdx.x < nb * 4) (&sb[0][0])[threadIdx.x] = boxes[((size_t)map * max_boxes) * 4 + threadIdx.x];
    __syncthreads();
    for (int i = threadIdx.x; i < kNpix / 4; i += blockDim.x) {
        const int y = (4 * i) / kDst, x0 = (4 * i) % kDst;
        uint32_t g4 = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int x = x0 + j;
            int c = 0;
            for (int q = 0; q < nb; ++q)
                c += (x >= sb[q][0] && x < sb[q][2] && y >= sb[q][1] && y < sb[q][3]);
            if (gt_mode == AVT_GT_VGGSS) c = min(c, 1);
            else if (gt_mode == AVT_GT_FLICKR) c = min(c, 2);
            else c = min(c, 255);
            g4 |= (uint32_t)c << (8 * j);
        }
        reinterpret_cast<unsigned int*>(gt_dense + (size_t)map * kNpix)[i] = g4;
    }
}

// train_hardway_1frame.py:208-215: counts[i] = #(ciou >= 0.05*i); cIoU@0.5; trapezoid AUC
__global__ void auc_kernel(const double* __restrict__ ciou, int N, int32_t* __restrict__ counts,
                           double* __restrict__ out) {
    __shared__ int s_cnt[22];
    if (threadIdx.x < 22) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    int loc[22];
#pragma unroll
    for (int i = 0; i < 22; ++i) loc[i] = 0;
    for (int j = threadIdx.x; j < N; j += blockDim.x) {
        const double c = ciou[j];
#pragma unroll
        for (int i = 0; i < 21; ++i) loc[i] += (c >= 0.05 * (double)i);   // NaN never counts
        loc[21] += (c >= 0.5);
    }
#pragma unroll
    for (int i = 0; i < 22; ++i) {
        int t = __reduce_add_sync(0xffffffffu, loc[i]);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(&s_cnt[i], t);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double auc = 0.0;
        for (int i = 0; i < 20; ++i) {
            const double dx = 0.05 * (double)(i + 1) - 0.05 * (double)i;
            const double y0 = (double)s_cnt[i] / (double)N, y1 = (double)s_cnt[i + 1] / (double)N;
            auc += dx * (y1 + y0) / 2.0;
        }
        for (int i = 0; i < 21; ++i) counts[i] = s_cnt[i];
        out[0] = (double)s_cnt[21] / (double)N;
        out[1] = auc;
    }
}

// utils.py:311-318 mTC: block i scores cal_CIOU(mask[i], mask[i+1], 0.5)
__global__ void mtc_pairs_kernel(const uint8_t* __restrict__ masks, int32_t* __restrict__ pair_inter,
                                 int32_t* __restrict__ pair_union) {
    __shared__ uint32_t red[2][8];
    const int i = blockIdx.x;
    const unsigned int* a = reinterpret_cast<const unsigned int*>(masks + (size_t)i * kNpix);
    const unsigned int* b = reinterpret_cast<const unsigned int*>(masks + (size_t)(i + 1) * kNpix);
    uint32_t inter = 0, uni = 0;
    for (int j = threadIdx.x; j < kNpix / 4; j += blockDim.x) {
        // infer = (a >= 0.5) -> non-zero bytes of a; gt = b as stored (0/1)
        const uint32_t a4 = __vcmpne4(a[j], 0u) & 0x01010101u, b4 = b[j];
        inter = __dp4a(a4, b4, inter);
        uni = __dp4a(b4, 0x01010101u, uni);
        uni = __dp4a(a4, __vcmpeq4(b4, 0u) & 0x01010101u, uni);
    }
    inter = __reduce_add_sync(0xffffffffu, inter);
    uni = __reduce_add_sync(0xffffffffu, uni);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = inter; red[1][threadIdx.x >> 5] = uni; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t x = 0, y = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { x += red[0][w]; y += red[1][w]; }
        pair_inter[i] = (int32_t)x;
        pair_union[i] = (int32_t)y;
    }
}
__global__ void mtc_mean_kernel(const int32_t* __restrict__ pair_inter,
                                const int32_t* __restrict__ pair_union, int T, float* __restrict__ out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float acc = 0.f;                 // the reference accumulates into a float32 tensor
        for (int i = 0; i < T - 1; ++i)
            acc += (float)((double)pair_inter[i] / (double)pair_union[i]);
        *out = acc / (float)(T - 1);
    }
}

// Evaluator.cal_CIOU (utils.py:209-214) on ready-made maps: infer = (pred >= thres)
__global__ void ciou_counts_kernel(const float* __restrict__ infer, float thres,
                                   const uint8_t* __restrict__ gt_w, int gt_den,
                                   int32_t* __restrict__ inter_w, int32_t* __restrict__ union_w,
                                   double* __restrict__ ciou) {
    __shared__ uint32_t red[3][8];
    const int map = blockIdx.x;
    const float4* a = reinterpret_cast<const float4*>(infer + (size_t)map * kNpix);
    const unsigned int* g = reinterpret_cast<const unsigned int*>(gt_w + (size_t)map * kNpix);
    uint32_t inter = 0, gsum = 0, zc = 0;
    for (int j = threadIdx.x; j < kNpix / 4; j += blockDim.x) {
        const float4 x = a[j];
        const uint32_t m4 = (x.x >= thres ? 1u : 0u) | (x.y >= thres ? 0x100u : 0u) |
                            (x.z >= thres ? 0x10000u : 0u) | (x.w >= thres ? 0x1000000u : 0u);
        const uint32_t g4 = g[j];
        inter = __dp4a(m4, g4, inter);
        gsum = __dp4a(g4, 0x01010101u, gsum);
        zc = __dp4a(m4, __vcmpeq4(g4, 0u) & 0x01010101u, zc);
    }
    inter = __reduce_add_sync(0xffffffffu, inter);
    gsum = __reduce_add_sync(0xffffffffu, gsum);
    zc = __reduce_add_sync(0xffffffffu, zc);
    if ((threadIdx.x & 31) == 0) {
        red[0][threadIdx.x >> 5] = inter; red[1][threadIdx.x >> 5] = gsum; red[2][threadIdx.x >> 5] = zc;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t x = 0, y = 0, z = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { x += red[0][w]; y += red[1][w]; z += red[2][w]; }
        const uint32_t uni = y + (uint32_t)gt_den * z;
        inter_w[map] = (int32_t)x;
        union_w[map] = (int32_t)uni;
        ciou[map] = (double)x / (double)uni;
    }
}

}  // namespace avt

using namespace avt;

extern "C" int avt_eval_maps(const float* maps14, int N, const uint8_t* gt_dense,
                             const int16_t* boxes, const uint8_t* box_count, int max_boxes,
                             int gt_mode, float* up224, uint8_t* mask, int32_t* inter_w,
                             int32_t* union_w, double* ciou, void* stream) {
    AVT_REQUIRE(N >= 0, "avt_eval_maps: N=%d", N);
    if (N == 0) return AVT_OK;
    AVT_REQUIRE(maps14 && inter_w && union_w && ciou, "avt_eval_maps: null pointer");
    AVT_REQUIRE(gt_dense || (boxes && box_count && max_boxes > 0 && max_boxes <= kMaxBoxesSmem),
                "avt_eval_maps: need gt_dense, or boxes + box_count with 1 <= max_boxes <= %d",
                kMaxBoxesSmem);
    AVT_REQUIRE(gt_mode >= AVT_GT_VGGSS && gt_mode <= AVT_GT_FLICKR_FRAME, "avt_eval_maps: gt_mode=%d", gt_mode);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t smem = sizeof(EvalSmem);
    int dev = 0, sms = 148;
    AVT_CHECK_CUDA(cudaGetDevice(&dev));
    AVT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = N < sms * 6 ? N : sms * 6;           // 6 resident blocks per SM
    ProfScope _ps(kProfEval, st);
    if (gt_dense) {
        AVT_CHECK_CUDA(cudaFuncSetAttribute(eval_maps_kernel<true>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        eval_maps_kernel<true><<<grid, kEvThreads, smem, st>>>(maps14, N, gt_dense, boxes, box_count,
                                                              max_boxes, gt_mode, up224, mask, inter_w,
                                                              union_w, ciou);
    } else {
        AVT_CHECK_CUDA(cudaFuncSetAttribute(eval_maps_kernel<false>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        eval_maps_kernel<false><<<grid, kEvThreads, smem, st>>>(maps14, N, gt_dense, boxes, box_count,
                                                               max_boxes, gt_mode, up224, mask, inter_w,
                                                               union_w, ciou);
    }
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

extern "C" int avt_rasterise_gt(const int16_t* boxes, const uint8_t* box_count, int N, int max_boxes,
                                int gt_mode, uint8_t* gt_dense, void* stream) {
    AVT_REQUIRE(N >= 0, "avt_rasterise_gt: N=%d", N);
    if (N == 0) return AVT_OK;
    AVT_REQUIRE(boxes && box_count && gt_dense, "avt_rasterise_gt: null pointer");
    AVT_REQUIRE(max_boxes > 0 && max_boxes <= kMaxBoxesSmem, "avt_rasterise_gt: max_boxes=%d (1..%d)",
                max_boxes, kMaxBoxesSmem);
    AVT_REQUIRE(gt_mode >= AVT_GT_VGGSS && gt_mode <= AVT_GT_FLICKR_FRAME, "avt_rasterise_gt: gt_mode=%d", gt_mode);
    rasterise_gt_kernel<<<N, 256, 0, static_cast<cudaStream_t>(stream)>>>(boxes, box_count, max_boxes,
                                                                         gt_mode, gt_dense);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

extern "C" int avt_auc(const double* ciou, int N, int32_t* counts, double* out, void* stream) {
    AVT_REQUIRE(ciou && counts && out && N > 0, "avt_auc: null pointer or N=%d", N);
    auc_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(ciou, N, counts, out);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

extern "C" int avt_mtc(const uint8_t* masks, int T, int32_t* pair_inter, int32_t* pair_union,
                       float* out, void* stream) {
    AVT_REQUIRE(masks && pair_inter && pair_union && out, "avt_mtc: null pointer");
    AVT_REQUIRE(T >= 2, "avt_mtc: T=%d needs at least two frames", T);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    mtc_pairs_kernel<<<T - 1, 256, 0, st>>>(masks, pair_inter, pair_union);
    AVT_LAUNCH_CHECK();
    mtc_mean_kernel<<<1, 32, 0, st>>>(pair_inter, pair_union, T, out);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

extern "C" int avt_ciou_counts(const float* infer, float thres, const uint8_t* gt_w, int gt_den,
                               int N, int32_t* inter_w, int32_t* union_w, double* ciou, void* stream) {
    AVT_REQUIRE(N >= 0, "avt_ciou_counts: N=%d", N);
    if (N == 0) return AVT_OK;
    AVT_REQUIRE(infer && gt_w && inter_w && union_w && ciou, "avt_ciou_counts: null pointer");
    AVT_REQUIRE(gt_den >= 1, "avt_ciou_counts: gt_den=%d", gt_den);
    ciou_counts_kernel<<<N, 256, 0, static_cast<cudaStream_t>(stream)>>>(infer, thres, gt_w, gt_den,
                                                                        inter_w, union_w, ciou);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}
