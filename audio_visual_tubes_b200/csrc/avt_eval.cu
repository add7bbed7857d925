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
#include <atomic>
#include <mutex>

#ifndef AVT_EVAL_REGS
#define AVT_EVAL_REGS 64
#endif
#ifdef AVT_EVAL_PLAIN_ST
#define AVT_ST(p, v) (*(p) = (v))
#else
#define AVT_ST(p, v) __stcs((p), (v))
#endif

namespace avt {

constexpr int kSrc = 14;
constexpr int kDst = 224;
constexpr int kNpix = kDst * kDst;               // 50176
constexpr int kRank = kNpix / 2;                 // int(224*224/2), train_hardway_1frame.py:202
constexpr int kEvThreads = 224;                  // 56 column quads x 4 row groups
constexpr int kEvWarps = kEvThreads / 32;        // 7
constexpr int kQuads = kDst / 4;                 // 56 column quads per row
constexpr int kStrips = 28;                      // 8-row strips per column quad (2 edges + 13 x 2 halves)
constexpr int kChunk = 128;                      // list entries a warp takes at a time (32 quads = one full row step)
constexpr int kCap = 7168;                       // candidate keys kept for the exact refinement
constexpr int kSampleSide = 56;                  // every 4th pixel in x and y
constexpr int kNs = kSampleSide * kSampleSide;   // 3136 samples (14 per thread)
constexpr int kSampPerThread = kNs / kEvThreads; // 14
#ifndef AVT_EVAL_MARGIN
#define AVT_EVAL_MARGIN 24
#endif
constexpr int kMargin = AVT_EVAL_MARGIN;                      // sample ranks either side of the median
constexpr int kHist = 256;
constexpr int kMaxBoxesSmem = 32;
static_assert(kEvThreads == kDst, "rasterise_planes: one thread per destination row");

// [0] maps that took the generic (bisection) selection, [1] maps evaluated -- dev statistics
__device__ unsigned long long g_eval_stats[2];

// Work tickets: maps are claimed one at a time (their cost varies, a static stride leaves a tail).
// The ticket word belongs to the CALL (caller-owned workspace, zeroed on the stream right before
// the launch), so launches in flight, captured graphs and aborted kernels can not see each other.

// Thread -> (column quad, row group): a warp walks a 128-pixel row segment, so its global accesses
// are 512 / 128 contiguous bytes.  (An 8-quad x 4-strip patch per warp crosses the median contour
// on fewer row steps -- 71 % vs 86 % -- but measured slower: AVT_EVAL_PATCH_LANES keeps it around.)
#ifndef AVT_EVAL_PATCH_LANES
__device__ __forceinline__ int quad_col() { return threadIdx.x % kQuads; }
__device__ __forceinline__ int row_group() { return threadIdx.x / kQuads; }
#else
__device__ __forceinline__ int quad_col() { return 8 * (threadIdx.x >> 5) + (threadIdx.x & 7); }
__device__ __forceinline__ int row_group() { return (threadIdx.x >> 3) & 3; }
#endif

struct EvalSmem {
    float m[kSrc * kSrc];                        // the 14x14 map
    float rows[kSrc][kDst];                      // horizontally interpolated rows
    uint32_t list[kCap];                         // candidate keys inside the bracket
    uint32_t hist[2][kHist];                     // shared-atomic histograms (two, to save a barrier)
    uint32_t red[4][kEvWarps];
    uint32_t bcast[4];
    uint32_t geo[kStrips];                       // packed strip geometry (strip_pack)
    int n_list;
    int next_map;
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

// Four pixels at once.  The kernel is instruction-issue bound (ncu, round 1: issue slots 70 % busy,
// 19 % of all instructions were these products and sums), and sm_100 multiplies two fp32 pairs per
// instruction (FMUL2): 4 FMUL2 + 4 FADD instead of 8 FMUL + 4 FADD, every product and every sum
// still rounded on its own -- the same bits as lerp_rn.  (A packed add.rn.f32x2 would save two more
// instructions, but ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2, which is NOT the
// reference's arithmetic; the scalar __fadd_rn is never contracted.)
__device__ __forceinline__ float2 fmul2_rn(float2 a, float w) {
    float2 d;
    asm("{.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %4};\n\t"
        "mul.rn.f32x2 rd, ra, rb;\n\tmov.b64 {%0, %1}, rd;}" : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(w));
    return d;
}
__device__ __forceinline__ float4 lerp4_rn(const float4& a, float w0, const float4& b, float w1) {
    const float2 p0 = fmul2_rn(make_float2(a.x, a.y), w0), p1 = fmul2_rn(make_float2(a.z, a.w), w0);
    const float2 q0 = fmul2_rn(make_float2(b.x, b.y), w1), q1 = fmul2_rn(make_float2(b.z, b.w), w1);
    return make_float4(__fadd_rn(p0.x, q0.x), __fadd_rn(p0.y, q0.y), __fadd_rn(p1.x, q1.x), __fadd_rn(p1.y, q1.y));
}
// three-input min / max (FMNMX3)
__device__ __forceinline__ float max3(float a, float b, float c) { float m; asm("max.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(a), "f"(b), "f"(c)); return m; }
__device__ __forceinline__ float min3(float a, float b, float c) { float m; asm("min.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(a), "f"(b), "f"(c)); return m; }

// Three block reductions behind ONE pair of barriers: sum of x, min of lo, max of hi.
__device__ __forceinline__ void block_sum_min_max(EvalSmem& S, uint32_t& x, uint32_t& lo, uint32_t& hi) {
    x = __reduce_add_sync(0xffffffffu, x);
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        const int w = threadIdx.x >> 5;
        S.red[0][w] = x; S.red[1][w] = lo; S.red[2][w] = hi;
    }
    __syncthreads();
    x = 0; lo = 0xffffffffu; hi = 0u;
#pragma unroll
    for (int i = 0; i < kEvWarps; ++i) {
        x += S.red[0][i]; lo = min(lo, S.red[1][i]); hi = max(hi, S.red[2][i]);
    }
}
__device__ __forceinline__ void block_sum3(EvalSmem& S, uint32_t& a, uint32_t& b, uint32_t& c) {
    a = __reduce_add_sync(0xffffffffu, a);
    b = __reduce_add_sync(0xffffffffu, b);
    c = __reduce_add_sync(0xffffffffu, c);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        const int w = threadIdx.x >> 5;
        S.red[0][w] = a; S.red[1][w] = b; S.red[2][w] = c;
    }
    __syncthreads();
    a = b = c = 0;
#pragma unroll
    for (int i = 0; i < kEvWarps; ++i) { a += S.red[0][i]; b += S.red[1][i]; c += S.red[2][i]; }
}

// Every warp scans the same 256-bin histogram (8 bins per lane) and returns the bin that
// holds rank r (0-based) plus the number of entries in the bins before it.  r >= total
// returns bin 255.  No barrier: the caller has already synchronised on the histogram.
__device__ __forceinline__ void scan256(const uint32_t* hist, uint32_t r, int& bin, uint32_t& before) {
    const int lane = threadIdx.x & 31;
    const uint4 c0 = reinterpret_cast<const uint4*>(hist)[2 * lane];
    const uint4 c1 = reinterpret_cast<const uint4*>(hist)[2 * lane + 1];
    const uint32_t c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    uint32_t tot = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) tot += c[q];
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    const uint32_t ball = __ballot_sync(0xffffffffu, incl > r);
    const int owner = ball ? __ffs(ball) - 1 : 31;
    uint32_t e = incl - tot;
    int b = 0;
#pragma unroll
    for (int q = 0; q < 7; ++q)
        if (b == q && r >= e + c[q]) { e += c[q]; ++b; }
    bin = 8 * owner + __shfl_sync(0xffffffffu, b, owner);
    before = __shfl_sync(0xffffffffu, e, owner);
}

// the same for TWO ranks r0 <= r1 of one histogram with ONE prefix scan (the bracket needs both of its
// edges; the scan's five dependent shuffles were 15.7 % of the kernel's stall samples in round 1)
__device__ __forceinline__ void scan256_pair(const uint32_t* hist, uint32_t r0, uint32_t r1, int& bin0, int& bin1) {
    const int lane = threadIdx.x & 31;
    const uint4 c0 = reinterpret_cast<const uint4*>(hist)[2 * lane];
    const uint4 c1 = reinterpret_cast<const uint4*>(hist)[2 * lane + 1];
    const uint32_t c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    uint32_t tot = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) tot += c[q];
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    auto locate = [&](uint32_t r) -> int {
        const uint32_t ball = __ballot_sync(0xffffffffu, incl > r);
        const int owner = ball ? __ffs(ball) - 1 : 31;
        uint32_t e = incl - tot;
        int b = 0;
#pragma unroll
        for (int q = 0; q < 7; ++q)
            if (b == q && r >= e + c[q]) { e += c[q]; ++b; }
        return 8 * owner + __shfl_sync(0xffffffffu, b, owner);
    };
    bin0 = locate(r0);
    bin1 = locate(r1);
}

// r-th smallest (0-based) of the keys in list[0..n) that lie inside [lo, hi].  Exact radix
// descent, 8 bits per round, histogram by shared-memory atomics (n is a few hundred to a few
// thousand keys spread over the bins, so conflicts are rare).
__device__ uint32_t list_select(EvalSmem& S, int n, uint32_t r, uint32_t lo, uint32_t hi) {
    const int tid = threadIdx.x;
    while (lo != hi) {
        const uint32_t span = hi - lo;
        const int s = max(0, (32 - __clz(span)) - 8);            // (span >> s) < 256
        __syncthreads();                                         // previous round's readers are done
        for (int i = tid; i < kHist; i += kEvThreads) S.hist[0][i] = 0u;
        __syncthreads();
        for (int i = tid; i < n; i += kEvThreads) {
            const uint32_t key = S.list[i];
            if (key >= lo && key <= hi) atomicAdd(&S.hist[0][(key - lo) >> s], 1u);
        }
        __syncthreads();
        int b; uint32_t before;
        scan256(S.hist[0], r, b, before);
        r -= before;
        const uint32_t nlo = lo + ((uint32_t)b << s);
        uint32_t nhi = nlo + ((1u << s) - 1u);
        if (nhi < nlo) nhi = 0xffffffffu;
        lo = nlo;
        hi = min(hi, nhi);
    }
    return lo;
}

// Visit every destination row of this thread's 7 strips (4 pixels wide, 8 rows tall).
// Vertical weights repeat with period 16, so a strip between source rows s and s+1 loads its
// two float4 operands ONCE and uses compile-time weights (2r+1)/32; the second half of an
// interval is the first half mirrored (operands swapped, rows visited downwards) -- the same
// products and the same sum as tap()/lerp_rn, bit for bit.  Rows 0..7 and 216..223 clamp.
template <int kUnroll, class StripFn, class RowFn>
__device__ __forceinline__ void for_each_row(const EvalSmem& S, StripFn&& pre, RowFn&& fn) {
    const int cq = quad_col(), rg = row_group();
#pragma unroll 1
    for (int j = 0; j < kStrips / 4; ++j) {
        const int strip = 4 * j + rg;
        const bool edge = (strip == 0 || strip == kStrips - 1);
        const int s = (strip - 1) >> 1, half = (strip - 1) & 1;
        // operand rows: edges clamp (value = a*1 + b*0 for all 8 rows); second halves swap
        const int ra = edge ? (strip ? kSrc - 1 : 0) : s + half;
        const int rb = edge ? (strip ? kSrc - 1 : 1) : s + 1 - half;
        const float4 a = *reinterpret_cast<const float4*>(&S.rows[ra][4 * cq]);
        const float4 b = *reinterpret_cast<const float4*>(&S.rows[rb][4 * cq]);
        int y = edge ? (strip ? kDst - 8 : 0) : 16 * s + 8 + (half ? 15 : 0);
        const int ystep = (!edge && half) ? -1 : 1;
        const float dw = edge ? 0.0f : (1.0f / 16.0f);
        float w1 = edge ? 0.0f : (1.0f / 32.0f), w0 = 1.0f - w1;    // multiples of 1/32: exact
        pre(y, ystep);
#pragma unroll kUnroll
        for (int r = 0; r < 8; ++r) {
            const float4 u = lerp4_rn(a, w0, b, w1);
            fn(r, y, u);
            y += ystep; w1 += dw; w0 -= dw;
        }
    }
}

// Generic counting sweep in KEY space (fallback path only): c_in = #(blo <= key <= bhi),
// c_lt = #(key < blo), key range, and optionally the in-bracket keys appended to S.list.
__device__ void full_pass_keys(EvalSmem& S, uint32_t blo, uint32_t bhi, bool collect,
                               uint32_t& c_lt, uint32_t& c_in, uint32_t& kmin, uint32_t& kmax) {
    uint32_t n_lt = 0, n_in = 0, mn = 0xffffffffu, mx = 0u;
    __syncthreads();
    if (threadIdx.x == 0) S.n_list = 0;
    __syncthreads();
    for_each_row<1>(S, [](int, int) {}, [&](int, int, const float4& u) {
        const float uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t key = f2key(uu[j]);
            mn = min(mn, key); mx = max(mx, key);
            n_lt += (key < blo);
            const bool in = (key >= blo) && (key <= bhi);
            n_in += in;
            if (collect && in) {
                const int idx = atomicAdd(&S.n_list, 1);
                if (idx < kCap) S.list[idx] = key;
            }
        }
    });
    block_sum_min_max(S, n_lt, mn, mx);
    c_lt = n_lt; kmin = mn; kmax = mx;
    uint32_t z0 = 0, z1 = 0;
    block_sum3(S, n_in, z0, z1);
    c_in = n_in;
}

// Exact k-th key by counted bisection over the key range until the rest fits the list.
// Slow (several sweeps) but unconditional: NaN / Inf maps, constant maps, huge tie groups
// and the rare miss of the sampled bracket come here.
__device__ uint32_t select_generic(EvalSmem& S, uint32_t& kmin, uint32_t& kmax) {
    uint32_t a, b;
    full_pass_keys(S, 0u, 0xffffffffu, false, a, b, kmin, kmax);
    uint32_t lo = kmin, hi = kmax, c_lo = 0, n_in = kNpix;
    for (;;) {
        if (lo == hi) return lo;
        uint32_t c, d;
        if (n_in <= (uint32_t)kCap) {
            full_pass_keys(S, lo, hi, true, a, b, c, d);
            return list_select(S, (int)n_in, (uint32_t)kRank - c_lo, lo, hi);
        }
        const uint32_t mid = lo + (hi - lo) / 2u;
        full_pass_keys(S, lo, mid, false, a, b, c, d);            // b = #(lo <= key <= mid)
        if ((uint32_t)kRank - c_lo < b) { hi = mid; n_in = b; }
        else { lo = mid + 1u; c_lo += b; n_in -= b; }
    }
}

// Second-chance counting sweep for the few maps whose touching quads overflow the list: exact
// per-pixel filter, only the pixels inside [tlo, thi) are appended (a few hundred).  One more
// sweep instead of the generic bisection.  Returns this thread's #(u < tlo).
__device__ __noinline__ uint32_t count_sweep_exact(EvalSmem& S, float tlo, float thi) {
    uint32_t n_lt = 0;
    for_each_row<1>(S, [](int, int) {}, [&](int, int, const float4& u) {
        const float uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool below = uu[j] < tlo;
            n_lt += below;
            if (!below && uu[j] < thi) {
                const int idx = atomicAdd(&S.n_list, 1);
                if (idx < kCap) S.list[idx] = __float_as_uint(uu[j]);
            }
        }
    });
    return n_lt;
}

// f of the header comment: utils.py:234-239 applied to -u, then pred = 1 - h
__device__ __forceinline__ float pred_of(float u, bool normalised, float vmin_h, float range) {
    float h = -u;
    if (normalised) h = __fdiv_rn(__fsub_rn(h, vmin_h), range);
    return __fsub_rn(1.0f, h);
}

// One thread's strip: 4 pixels wide, 8 consecutive rows y0..y0+7 between source rows a and b
// with weights w1 = w1_0 + r/16 on b and 1 - w1 on a (exact multiples of 1/32), which is tap()
// written per strip.  Positions p = 4*j + rg, j = 0..6: p < 26 are the interior half-intervals
// (source interval p >> 1, half p & 1 -- the half is rg & 1, so a thread's weights never change),
// p = 26 / 27 are the clamped rows 0..7 / 216..223 (value = a*1 + b*0 on every row).
struct Strip {
    float4 a, b;
    int y0;
    float w1_0, dw;
};
// Strip geometry is the same for every map: 28 packed words in shared memory, written once per
// block.  bits 0-7 y0, bits 12-15 / 16-19 operand rows, bit 20 edge, bit 21 second half.
__device__ __forceinline__ uint32_t strip_pack(int p) {
    const bool edge = p >= kStrips - 2;
    const int s = p >> 1, half = p & 1;
    const int ra = edge ? (half ? kSrc - 1 : 0) : s;
    const int rb = edge ? (half ? kSrc - 1 : 1) : s + 1;
    const int y0 = edge ? (half ? kDst - 8 : 0) : 16 * s + 8 + 8 * half;
    return (uint32_t)y0 | ((uint32_t)ra << 12) | ((uint32_t)rb << 16) | (edge ? 0x100000u : 0u) |
           ((!edge && half) ? 0x200000u : 0u);
}
__device__ __forceinline__ int strip_y0(const EvalSmem& S, int j) {
    return (int)(S.geo[4 * j + row_group()] & 0xffu);
}
__device__ __forceinline__ Strip strip_load(const EvalSmem& S, int j) {
    const int cq = quad_col();
    const uint32_t g = S.geo[4 * j + row_group()];
    Strip t;
    t.a = *reinterpret_cast<const float4*>(&S.rows[(g >> 12) & 15u][4 * cq]);
    t.b = *reinterpret_cast<const float4*>(&S.rows[(g >> 16) & 15u][4 * cq]);
    t.y0 = (int)(g & 0xffu);
    const bool edge = (g & 0x100000u) != 0u;
    t.dw = edge ? 0.0f : (1.0f / 16.0f);
    t.w1_0 = edge ? 0.0f : ((g & 0x200000u) ? (17.0f / 32.0f) : (1.0f / 32.0f));
    return t;
}
__device__ __forceinline__ float4 strip_row(const Strip& t, int r) {
    const float w1 = fmaf((float)r, t.dw, t.w1_0), w0 = 1.0f - w1;   // multiples of 1/32: exact
    return lerp4_rn(t.a, w0, t.b, w1);
}

// Box ground truth as two bit planes in the (by then idle) candidate list: plane 0 = covered by
// at least one box, plane 1 = by at least two (utils.py:264-273 needs the union, :243-262 the
// count clipped at two annotators).  Thread y rasterises row y: 7 words per plane.
constexpr int kPlaneWords = kDst * (kDst / 32);   // 1568
__device__ __forceinline__ void rasterise_planes(EvalSmem& S, int nb) {
    const int y = threadIdx.x;                     // kEvThreads == kDst
#pragma unroll
    for (int w = 0; w < kDst / 32; ++w) {
        uint32_t c0 = 0, c1 = 0;
        for (int q = 0; q < nb; ++q) {
            if (y >= S.boxes[q][1] && y < S.boxes[q][3]) {
                const int lo = max((int)S.boxes[q][0], 32 * w) - 32 * w;
                const int hi = min((int)S.boxes[q][2], 32 * w + 32) - 32 * w;
                if (hi > lo) {
                    const uint32_t m = ((hi >= 32) ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
                    c1 |= c0 & m;
                    c0 |= m;
                }
            }
        }
        S.list[y * (kDst / 32) + w] = c0;
        S.list[kPlaneWords + y * (kDst / 32) + w] = c1;
    }
}
// the 4 GT weights of quad cq in row y from the planes: 0 / 1 (VGG-SS) or 0 / 1 / 2 (Flickr)
__device__ __forceinline__ uint32_t planes_quad(const EvalSmem& S, int gt_mode, int cq, int y) {
    const int i = y * (kDst / 32) + (cq >> 3), sh = 4 * (cq & 7);
    uint32_t g4 = (((S.list[i] >> sh) & 15u) * 0x00204081u) & 0x01010101u;
    if (gt_mode == AVT_GT_FLICKR) g4 += (((S.list[kPlaneWords + i] >> sh) & 15u) * 0x00204081u) & 0x01010101u;
    return g4;
}

__device__ __forceinline__ uint32_t boxes_quad(const EvalSmem& S, int nb, int gt_mode, int x0, int y) {
    uint32_t g4 = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int x = x0 + j;
        int c = 0;
        for (int q = 0; q < nb; ++q)
            c += (x >= S.boxes[q][0] && x < S.boxes[q][2] && y >= S.boxes[q][1] && y < S.boxes[q][3]);
        if (gt_mode == AVT_GT_VGGSS) c = min(c, 1);
        else if (gt_mode == AVT_GT_FLICKR) c = min(c, 2);
        else c = min(c, 255);
        g4 |= (uint32_t)c << (8 * j);
    }
    return g4;
}

// Final sweep: materialise, binarise, read / rasterise the GT, count.  Straight-line per strip:
// the 8 GT words of a strip are requested before any arithmetic, the up-sampled row is computed
// once and feeds the store, the compare and the counts.
// {u >= t} of four floats as 0 / 1 bytes: four SETs (all-ones / zero), three byte permutes, one AND
__device__ __forceinline__ uint32_t ge_mask4(const float4& u, float t) {
    uint32_t m0, m1, m2, m3, p, q, r;
    asm("set.ge.u32.f32 %0, %1, %2;" : "=r"(m0) : "f"(u.x), "f"(t));
    asm("set.ge.u32.f32 %0, %1, %2;" : "=r"(m1) : "f"(u.y), "f"(t));
    asm("set.ge.u32.f32 %0, %1, %2;" : "=r"(m2) : "f"(u.z), "f"(t));
    asm("set.ge.u32.f32 %0, %1, %2;" : "=r"(m3) : "f"(u.w), "f"(t));
    asm("prmt.b32 %0, %1, %2, 0x0040;" : "=r"(p) : "r"(m0), "r"(m1));
    asm("prmt.b32 %0, %1, %2, 0x0040;" : "=r"(q) : "r"(m2), "r"(m3));
    asm("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(r) : "r"(p), "r"(q));
    return r & 0x01010101u;
}

// kSmallGt: GT weights are at most 2 (VGG-SS union, Flickr clipped count), so "weight == 0" is
// three integer ops per quad instead of a byte-wise compare.
template <bool kDenseGt, bool kUp, bool kMask, bool kSmallGt>
__device__ __forceinline__ void final_sweep(const EvalSmem& S, float t_on, int nb, int gt_mode,
                                            float* __restrict__ up_o, uint8_t* __restrict__ mk_o,
                                            const uint8_t* __restrict__ gt_i,
                                            uint32_t& inter, uint32_t& gsum, uint32_t& zc) {
    const int cq = quad_col();
    uint4 g = make_uint4(0, 0, 0, 0), gn = g;
    // rows of a strip ascend, so every access of a half-strip is base + r * 224 elements: immediates
    auto request = [&](int k) -> uint4 {                        // GT words of half-strip k
        const unsigned int* q = reinterpret_cast<const unsigned int*>(
            gt_i + ((strip_y0(S, k >> 1) + (k & 1) * 4) * kDst + 4 * cq));
        uint4 d;
        d.x = __ldcs(q); d.y = __ldcs(q + kDst / 4); d.z = __ldcs(q + 2 * (kDst / 4)); d.w = __ldcs(q + 3 * (kDst / 4));
        return d;
    };
    if (kDenseGt) g = request(0);
#pragma unroll 1
    for (int k = 0; k < 2 * (kStrips / 4); ++k) {
        if (kDenseGt && k + 1 < 2 * (kStrips / 4)) gn = request(k + 1);   // one half-strip ahead
        const Strip t = strip_load(S, k >> 1);
        const int h = (k & 1) * 4;
        const int ob = (t.y0 + h) * kDst + 4 * cq;
        float* const up_p = up_o + ob;
        uint8_t* const mk_p = mk_o + ob;
        float w1 = fmaf((float)h, t.dw, t.w1_0);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const float w0 = 1.0f - w1;
            const float4 u = lerp4_rn(t.a, w0, t.b, w1);
            w1 += t.dw;
            if (kUp) AVT_ST(reinterpret_cast<float4*>(up_p + r * kDst), u);
            const uint32_t m4 = ge_mask4(u, t_on);
            if (kMask) AVT_ST(reinterpret_cast<unsigned int*>(mk_p + r * kDst), m4);
            const uint32_t gr = (r == 0) ? g.x : (r == 1) ? g.y : (r == 2) ? g.z : g.w;
            const uint32_t g4 = kDenseGt ? gr
                              : (gt_mode == AVT_GT_FLICKR_FRAME ? boxes_quad(S, nb, gt_mode, 4 * cq, t.y0 + h + r)
                                                                : planes_quad(S, gt_mode, cq, t.y0 + h + r));
            inter = __dp4a(m4, g4, inter);
            gsum = __dp4a(g4, 0x01010101u, gsum);
            if (kSmallGt) zc = __dp4a(m4, ~((g4 + 0x7f7f7f7fu) >> 7) & 0x01010101u, zc);
            else zc = __dp4a(m4, __vcmpeq4(g4, 0u) & 0x01010101u, zc);
        }
        if (kDenseGt) g = gn;
    }
}

// Key-compare variant of the final sweep for the rare maps that need it (NaN / Inf values, a
// threshold exactly at +0): same outputs, pointers tested at run time, no unrolling.
template <bool kDenseGt>
__device__ __noinline__ uint3 final_sweep_keys(const EvalSmem& S, uint32_t k_on, int nb, int gt_mode,
                                               float* __restrict__ up_o, uint8_t* __restrict__ mk_o,
                                               const uint8_t* __restrict__ gt_i) {
    const int cq = quad_col();
    uint32_t inter = 0, gsum = 0, zc = 0;
    for_each_row<1>(S, [](int, int) {}, [&](int, int y, const float4& u) {
        const int o = y * kDst + 4 * cq;
        if (up_o) __stcs(reinterpret_cast<float4*>(up_o + o), u);
        const uint32_t m4 = (f2key(u.x) >= k_on ? 1u : 0u) | (f2key(u.y) >= k_on ? 0x100u : 0u) |
                            (f2key(u.z) >= k_on ? 0x10000u : 0u) | (f2key(u.w) >= k_on ? 0x1000000u : 0u);
        if (mk_o) __stcs(reinterpret_cast<unsigned int*>(mk_o + o), m4);
        const uint32_t g4 = kDenseGt ? __ldcs(reinterpret_cast<const unsigned int*>(gt_i + o))
                                     : boxes_quad(S, nb, gt_mode, 4 * cq, y);
        inter = __dp4a(m4, g4, inter);
        gsum = __dp4a(g4, 0x01010101u, gsum);
        zc = __dp4a(m4, __vcmpeq4(g4, 0u) & 0x01010101u, zc);
    });
    return make_uint3(inter, gsum, zc);
}

template <bool kDenseGt>
__global__ void __maxnreg__(AVT_EVAL_REGS)
eval_maps_kernel(const float* __restrict__ maps14, int N, const uint8_t* __restrict__ gt_dense,
                 const int16_t* __restrict__ boxes, const uint8_t* __restrict__ box_count,
                 int max_boxes, int gt_mode, float* __restrict__ up224, uint8_t* __restrict__ mask,
                 int32_t* __restrict__ inter_w, int32_t* __restrict__ union_w,
                 double* __restrict__ ciou, unsigned int* __restrict__ ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EvalSmem& S = *reinterpret_cast<EvalSmem*>(smem_raw);
    const int tid = threadIdx.x;
    if (tid < kStrips) S.geo[tid] = strip_pack(tid);           // published by the first barrier below
    if (tid == 0) S.next_map = (int)atomicAdd(&ticket[0], 1u);
    __syncthreads();
    int map = S.next_map;

    while (map < N) {
        // ---- load the map; source range and finiteness (one barrier pair)
        uint32_t bad = 0, kmn = 0xffffffffu, kmx = 0u;
        if (tid < kSrc * kSrc) {
            const float x = maps14[(size_t)map * kSrc * kSrc + tid];
            S.m[tid] = x;
            bad = !(fabsf(x) <= 3.402823466e38f);
            kmn = kmx = f2key(x);
        }
        int nb = 0;
        if (!kDenseGt) {
            nb = min((int)box_count[map], min(max_boxes, kMaxBoxesSmem));
            if (tid < nb * 4)
                (&S.boxes[0][0])[tid] = boxes[((size_t)map * max_boxes) * 4 + tid];
        }
        for (int i = tid; i < 2 * kHist; i += kEvThreads) (&S.hist[0][0])[i] = 0u;
        if (tid == 0) { S.n_list = 0; S.next_map = (int)atomicAdd(&ticket[0], 1u); }   // claim one ahead
        block_sum_min_max(S, bad, kmn, kmx);                     // also publishes S.m, S.next_map
        const int next_map = S.next_map;
        // the GT tile is not needed before the final sweep: ask L2 for it now (392 lines of 128 B)
        if (kDenseGt && tid < kNpix / 256) {
            const uint8_t* gp = gt_dense + (size_t)map * kNpix + tid * 256;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(gp));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(gp + 128));
        }
        // ---- horizontal pass: rows[j][x] = m[j][sx0]*(1-wx) + m[j][sx1]*wx
        {
            int s0, s1; float w0, w1;
            tap(tid, s0, s1, w0, w1);                            // tid == destination column
#pragma unroll
            for (int j = 0; j < kSrc; ++j)
                S.rows[j][tid] = lerp_rn(S.m[j * kSrc + s0], w0, S.m[j * kSrc + s1], w1);
        }
        __syncthreads();

        uint32_t kstar = 0, kmin = 0, kmax = 0;
        const float smin = key2f(kmn), smax = key2f(kmx);
        const float span = smax - smin;
        bool done = false;
        if (bad == 0u && span > 0.0f && span <= 3.402823466e38f) {
            // ---- bracket the median from a 56x56 sub-sample held in registers.  Two rounds of a
            // 256-bin linear histogram; the bracket only has to CONTAIN the true rank (checked
            // below with exact counts), so its construction is free to be approximate.
            float sm[kSampPerThread];
#pragma unroll
            for (int k = 0; k < kSampPerThread; ++k) {
                const int p = ((tid + kEvThreads * k) * 1117) % kNs;      // scatter lanes over the image
                const int sy = 4 * (p / kSampleSide) + 2, sx = 4 * (p % kSampleSide) + 1;
                int s0, s1; float w0, w1;
                tap(sy, s0, s1, w0, w1);
                sm[k] = lerp_rn(S.rows[s0][sx], w0, S.rows[s1][sx], w1);
            }
            const float scale1 = 256.0f / span, inv1 = span * (1.0f / 256.0f);
#pragma unroll
            for (int k = 0; k < kSampPerThread; ++k) {
                const int b = min(kHist - 1, max(0, __float2int_rd((sm[k] - smin) * scale1)));
                atomicAdd(&S.hist[0][b], 1u);
            }
            __syncthreads();
            int bl, bh; uint32_t before;
            scan256_pair(S.hist[0], (uint32_t)(kNs / 2 - kMargin), (uint32_t)(kNs / 2 + kMargin), bl, bh);
            const float lo2 = smin + (float)bl * inv1, hi2 = smin + (float)(bh + 1) * inv1;
            const float span2 = hi2 - lo2;
            const float scale2 = 254.0f / span2, inv2 = span2 * (1.0f / 254.0f);
            // bins 1..254 tile [lo2, hi2); bin 0 = below, bin 255 = above
#pragma unroll
            for (int k = 0; k < kSampPerThread; ++k) {
                const int b = min(kHist - 1, max(0, __float2int_rd((sm[k] - lo2) * scale2) + 1));
                atomicAdd(&S.hist[1][b], 1u);
            }
            __syncthreads();
            int sl, sh;
            scan256_pair(S.hist[1], (uint32_t)(kNs / 2 - kMargin), (uint32_t)(kNs / 2 + kMargin), sl, sh);
            const float ninf = __uint_as_float(0xff800000u), pinf = __uint_as_float(0x7f800000u);
            const float tlo = (sl == 0 || !(span2 > 0.0f)) ? ninf : lo2 + (float)(sl - 1) * inv2;
            const float thi = (sh == kHist - 1 || !(span2 > 0.0f)) ? pinf : lo2 + (float)sh * inv2;
            // ---- counting sweep: quads entirely below the bracket are counted four at a time,
            // quads entirely above cost nothing, quads that touch [tlo, thi) are appended WHOLE to
            // S.list.  The append is warp-convergent: one ballot per row, positions from the ballot,
            // space taken from the shared list in 32-quad chunks per warp (one atomic per chunk),
            // chunk tails padded with +Inf, which the post-pass ignores.
            uint32_t n_lt = 0;
            float fmn = pinf, fmx = ninf;
            const int lane = tid & 31;
            uint32_t lt_mask = (1u << lane) - 1u;
            uint32_t list_sa = (uint32_t)__cvta_generic_to_shared(&S.list[0]);
            asm volatile("" : "+r"(lt_mask), "+r"(list_sa));   // keep both in registers (no rematerialisation)
            int wpos = 0, wend = 0;                              // this warp's chunk (entries), uniform
            auto pad_tail = [&]() {
                const int i = wpos + 4 * lane;
                if (i < wend && i + 4 <= kCap) *reinterpret_cast<float4*>(&S.list[i]) = make_float4(pinf, pinf, pinf, pinf);
            };
#pragma unroll 1
            for (int j = 0; j < kStrips / 4; ++j) {
                const Strip t = strip_load(S, j);
                float w1 = t.w1_0;
#pragma unroll 4
                for (int r = 0; r < 8; ++r) {
                    const float w0 = 1.0f - w1;
                    const float4 u = lerp4_rn(t.a, w0, t.b, w1);
                    w1 += t.dw;
                    const float gmx = fmaxf(max3(u.x, u.y, u.z), u.w);
                    const float gmn = fminf(min3(u.x, u.y, u.z), u.w);
                    fmx = fmaxf(fmx, gmx); fmn = fminf(fmn, gmn);
                    const bool below = gmx < tlo;
                    n_lt += below ? 4u : 0u;
                    const bool touch = !below && gmn < thi;
                    const uint32_t m = __ballot_sync(0xffffffffu, touch);
                    if (m) {
                        const int cnt = 4 * __popc(m);
                        if (wpos + cnt > wend) {                 // uniform: take the next chunk
                            pad_tail();
                            int base = 0;
                            if (lane == 0) base = atomicAdd(&S.n_list, kChunk);
                            wpos = __shfl_sync(0xffffffffu, base, 0);
                            if (wpos + kChunk > kCap) wpos = 0;  // overflow (n_list says so): scribble on chunk 0
                            wend = wpos + kChunk;
                        }
                        const uint32_t sa = list_sa + 4u * (uint32_t)(wpos + 4 * __popc(m & lt_mask));
                        if (touch)
                            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sa), "f"(u.x), "f"(u.y),
                                         "f"(u.z), "f"(u.w) : "memory");
                        wpos += cnt;
                    }
                }
            }
            pad_tail();
            kmin = f2key(fmn); kmax = f2key(fmx);
            block_sum_min_max(S, n_lt, kmin, kmax);              // its barriers also publish the list
            // full-lane post-pass over the collected quads: keys in place, exact counts.  Float
            // order == key order except between -0 and +0, so a bracket edge at zero goes generic.
            auto exact_sweep = [&](float lo_f, float hi_f) {     // one more sweep, per-pixel filter
                __syncthreads();
                if (tid == 0) S.n_list = 0;
                __syncthreads();
                n_lt = count_sweep_exact(S, lo_f, hi_f);
                uint32_t z0 = 0, z1 = 0;
                block_sum3(S, n_lt, z0, z1);
            };
            auto settle = [&](float lo_f, float hi_f) -> bool {  // keys in place, verify, select
                const int n = S.n_list;
                if (n > kCap || lo_f == 0.0f || hi_f == 0.0f) return false;
                const uint32_t klo = f2key(lo_f), khi = f2key(hi_f) - 1u;
                uint32_t below = 0, inside = 0, unused = 0;
                for (int i = tid; i < n; i += kEvThreads) {
                    const uint32_t key = f2key(__uint_as_float(S.list[i]));
                    S.list[i] = key;
                    below += (key < klo);
                    inside += (key >= klo && key <= khi);
                }
                block_sum3(S, below, inside, unused);
                const uint32_t c_lt = n_lt + below;
                if (!(c_lt <= (uint32_t)kRank && (uint32_t)kRank < c_lt + inside)) return false;
                kstar = list_select(S, n, (uint32_t)kRank - c_lt, klo, khi);
                return true;
            };
            if (S.n_list > kCap) exact_sweep(tlo, thi);          // too many touching quads
            done = settle(tlo, thi);
            if (!done) {
                // the sampled bracket missed the rank (or sat on a zero edge): retry once with a
                // 4x margin taken from the coarse histogram, exact filter; then the generic path
                int bl2, bh2;
                scan256(S.hist[0], (uint32_t)(kNs / 2 - 4 * kMargin), bl2, before);
                scan256(S.hist[0], (uint32_t)(kNs / 2 + 4 * kMargin), bh2, before);
                const float tlo2 = (bl2 == 0) ? ninf : smin + (float)bl2 * inv1;
                const float thi2 = (bh2 == kHist - 1) ? pinf : smin + (float)(bh2 + 1) * inv1;
                exact_sweep(tlo2, thi2);
                done = settle(tlo2, thi2);
            }
        }
        if (!done) kstar = select_generic(S, kmin, kmax);
        if (tid == 0) { if (!done) atomicAdd(&g_eval_stats[0], 1ull); atomicAdd(&g_eval_stats[1], 1ull); }

        // ---- threshold in pred space, and the first key whose pixel is switched on.  pred is
        // monotone non-decreasing in the key, so gallop + bisect from kstar (every thread does
        // the same few evaluations; typically on(kstar + 1) settles it).
        const float umin = key2f(kmin), umax = key2f(kmax);
        const float vmin_h = -umax;
        const float range = __fsub_rn(-umin, vmin_h);           // vmax - vmin, utils.py:237
        const bool normalised = !(range == 0.0f);
        const float thr = pred_of(key2f(kstar), normalised, vmin_h, range);
        auto on = [&](uint64_t k) -> bool {
            if (k > (uint64_t)kmax) return true;
            const float pr = pred_of(key2f((uint32_t)k), normalised, vmin_h, range);
            return (pr > thr) || (pr >= 1.0f);
        };
        uint64_t lo, hi;                                         // keys < lo are off, hi is on
        if (on(kstar)) {
            hi = kstar; lo = kmin;
            uint64_t step = 1;
            while (hi > (uint64_t)kmin) {
                const uint64_t c = (hi - kmin > step) ? hi - step : (uint64_t)kmin;
                if (on(c)) { hi = c; step <<= 1; } else { lo = c + 1; break; }
            }
            if (hi == (uint64_t)kmin) lo = hi;
        } else {
            lo = (uint64_t)kstar + 1; hi = (uint64_t)kmax + 1;
            uint64_t step = 1;
            while (lo < hi) {
                const uint64_t c = lo + step - 1;
                if (c >= hi) break;
                if (on(c)) { hi = c; break; } else { lo = c + 1; step <<= 1; }
            }
        }
        while (lo < hi) {
            const uint64_t mid = lo + (hi - lo) / 2;
            if (on(mid)) hi = mid; else lo = mid + 1;
        }
        const bool all_off = lo > 0xffffffffull;                 // only when kmax is the last key
        const uint32_t k_on = all_off ? 0xffffffffu : (uint32_t)lo;
        const float t_on = key2f(k_on);
        const bool key_cmp = (bad != 0u) || all_off || k_on == f2key(0.0f);

        // ---- final sweep: materialise, binarise, count
        const int gt_den = (gt_mode == AVT_GT_FLICKR) ? 2 : 1;
        uint32_t inter = 0, gsum = 0, zc = 0;
        float* up_o = up224 ? up224 + (size_t)map * kNpix : nullptr;
        uint8_t* mk_o = mask ? mask + (size_t)map * kNpix : nullptr;
        const uint8_t* gt_i = kDenseGt ? gt_dense + (size_t)map * kNpix : nullptr;
#define AVT_FINAL(UP, MK)                                                                                  \
    do {                                                                                                   \
        if (gt_mode != AVT_GT_FLICKR_FRAME)                                                                \
            final_sweep<kDenseGt, UP, MK, true>(S, t_on, nb, gt_mode, up_o, mk_o, gt_i, inter, gsum, zc);  \
        else                                                                                               \
            final_sweep<kDenseGt, UP, MK, false>(S, t_on, nb, gt_mode, up_o, mk_o, gt_i, inter, gsum, zc); \
    } while (0)
        if (key_cmp) {
            const uint3 c3 = final_sweep_keys<kDenseGt>(S, k_on, nb, gt_mode, up_o, mk_o, gt_i);
            inter = c3.x; gsum = c3.y; zc = c3.z;
        } else {
            if (!kDenseGt && gt_mode != AVT_GT_FLICKR_FRAME) {       // the selection is over: reuse its list
                __syncthreads();
                rasterise_planes(S, nb);
                __syncthreads();
            }
            if (up_o) { if (mk_o) AVT_FINAL(true, true); else AVT_FINAL(true, false); }
            else { if (mk_o) AVT_FINAL(false, true); else AVT_FINAL(false, false); }
        }
#undef AVT_FINAL
        block_sum3(S, inter, gsum, zc);
        if (tid == 0) {
            const uint32_t uni = gsum + (uint32_t)gt_den * zc;
            inter_w[map] = (int32_t)inter;
            union_w[map] = (int32_t)uni;
            ciou[map] = (double)inter / (double)uni;             // 0/0 -> NaN as numpy
        }
        map = next_map;
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

// train_hardway.py:163-203, the per-video metrics of the PerFrameLabels test loop: video v owns
// frames [v*F, (v+1)*F).  out[v] = { #(ciou >= 0.5)/F  (:200),  trapezoid AUC of its 21-point
// success curve (:193-199),  mTC over its F-1 consecutive mask pairs (utils.py:311-318, float32
// accumulation as the reference's torch tensor; NaN for F == 1 as torch.div(0, 0)) }.
__global__ void video_metrics_kernel(const double* __restrict__ ciou, const int32_t* __restrict__ pair_inter,
                                     const int32_t* __restrict__ pair_union, int F,
                                     double* __restrict__ out) {
    const int v = blockIdx.x, lane = threadIdx.x;
    int loc[21];
#pragma unroll
    for (int i = 0; i < 21; ++i) loc[i] = 0;
    int half = 0;
    for (int j = lane; j < F; j += 32) {
        const double c = ciou[(size_t)v * F + j];
#pragma unroll
        for (int i = 0; i < 21; ++i) loc[i] += (c >= 0.05 * (double)i);
        half += (c >= 0.5);
    }
#pragma unroll
    for (int i = 0; i < 21; ++i) loc[i] = __reduce_add_sync(0xffffffffu, loc[i]);
    half = __reduce_add_sync(0xffffffffu, half);
    if (lane == 0) {
        double auc = 0.0;
#pragma unroll
        for (int i = 0; i < 20; ++i) {
            const double dx = 0.05 * (double)(i + 1) - 0.05 * (double)i;
            auc += dx * ((double)loc[i + 1] / (double)F + (double)loc[i] / (double)F) / 2.0;
        }
        float acc = 0.f;
        for (int i = 0; i < F - 1; ++i) {
            const size_t k = (size_t)v * F + i;
            acc += (float)((double)pair_inter[k] / (double)pair_union[k]);
        }
        out[3 * v + 0] = (double)half / (double)F;
        out[3 * v + 1] = auc;
        out[3 * v + 2] = (double)(acc / (float)(F - 1));
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
                             int32_t* union_w, double* ciou, void* workspace, void* stream) {
    AVT_REQUIRE(N >= 0, "avt_eval_maps: N=%d", N);
    if (N == 0) return AVT_OK;
    AVT_REQUIRE(maps14 && inter_w && union_w && ciou, "avt_eval_maps: null pointer");
    AVT_REQUIRE(workspace != nullptr, "avt_eval_maps: workspace is null (AVT_EVAL_WORKSPACE_BYTES device bytes)");
    AVT_REQUIRE(gt_dense || (boxes && box_count && max_boxes > 0 && max_boxes <= kMaxBoxesSmem),
                "avt_eval_maps: need gt_dense, or boxes + box_count with 1 <= max_boxes <= %d",
                kMaxBoxesSmem);
    AVT_REQUIRE(gt_mode >= AVT_GT_VGGSS && gt_mode <= AVT_GT_FLICKR_FRAME, "avt_eval_maps: gt_mode=%d", gt_mode);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t smem = sizeof(EvalSmem);
    int dev = 0, sms = 148;
    AVT_CHECK_CUDA(cudaGetDevice(&dev));
    AVT_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    ProfScope _ps(kProfEval, st);
    auto launch = [&](auto kernel) -> int {
        AVT_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        AVT_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                            (int)cudaSharedmemCarveoutMaxShared));
        int per_sm = 0;                                        // persistent grid: every resident slot, once
        AVT_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kEvThreads, smem));
        if (per_sm < 1) per_sm = 1;
        const int grid = N < sms * per_sm ? N : sms * per_sm;
        unsigned int* ticket = static_cast<unsigned int*>(workspace);
        AVT_CHECK_CUDA(cudaMemsetAsync(ticket, 0, AVT_EVAL_WORKSPACE_BYTES, st));
        kernel<<<grid, kEvThreads, smem, st>>>(maps14, N, gt_dense, boxes, box_count, max_boxes, gt_mode,
                                               up224, mask, inter_w, union_w, ciou, ticket);
        return AVT_OK;
    };
    const int rc = gt_dense ? launch(eval_maps_kernel<true>) : launch(eval_maps_kernel<false>);
    if (rc != AVT_OK) return rc;
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

extern "C" int avt_eval_stats(unsigned long long* out2, int reset) {
    AVT_REQUIRE(out2, "avt_eval_stats: null pointer");
    AVT_CHECK_CUDA(cudaDeviceSynchronize());
    AVT_CHECK_CUDA(cudaMemcpyFromSymbol(out2, g_eval_stats, sizeof(unsigned long long) * 2));
    if (reset) {
        const unsigned long long z[2] = {0ull, 0ull};
        AVT_CHECK_CUDA(cudaMemcpyToSymbol(g_eval_stats, z, sizeof(z)));
    }
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

// ---- avt_mtc without caller scratch (no counterpart in the reference; added
// for callers that may not allocate: graph capture, serving loops).  The pair counts
// live in a file-scope device array of fixed size -- every GPU that loads the library holds these
// 32 KB until the process exits.  In HOST code `g_mtc_static_scratch` is only a stand-in for the
// symbol, not a device address: the address is fetched once per device, with that device current,
// by cudaGetSymbolAddress, and that pointer is what the kernels get.
static __device__ __align__(256) int32_t g_mtc_static_scratch[2 * AVT_MTC_STATIC_MAX_FRAMES];

namespace avt {
constexpr int kMaxStaticDevices = 64;
static int mtc_static_scratch(int32_t** out) {
    static std::mutex mu;
    static int32_t* addr[kMaxStaticDevices] = {};
    int dev = 0;
    AVT_CHECK_CUDA(cudaGetDevice(&dev));
    AVT_REQUIRE(dev >= 0 && dev < kMaxStaticDevices, "avt_mtc_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        AVT_CHECK_CUDA(cudaGetSymbolAddress(&p, g_mtc_static_scratch));   // this device's copy of the array
        AVT_REQUIRE(p != nullptr, "avt_mtc_static: cudaGetSymbolAddress returned a null address");
        addr[dev] = static_cast<int32_t*>(p);
    }
    *out = addr[dev];
    return AVT_OK;
}
}  // namespace avt

extern "C" int avt_mtc_static(const uint8_t* masks, int T, float* out, void* stream) {
    AVT_REQUIRE(masks && out, "avt_mtc_static: null pointer");
    AVT_REQUIRE(T >= 2, "avt_mtc_static: T=%d needs at least two frames", T);
    AVT_REQUIRE(T <= AVT_MTC_STATIC_MAX_FRAMES,
                "avt_mtc_static: T=%d exceeds AVT_MTC_STATIC_MAX_FRAMES=%d (the built-in scratch holds no more; "
                "use avt_mtc with caller scratch)", T, AVT_MTC_STATIC_MAX_FRAMES);
    int32_t* scratch = nullptr;
    int rc = mtc_static_scratch(&scratch);
    if (rc != AVT_OK) return rc;
    // Both halves are WRITTEN for every pair [0, T-1) by mtc_pairs_kernel before mtc_mean_kernel reads
    // exactly those entries, so nothing has to start at zero and what an earlier call left behind is
    // never read.  Calls that overlap on the GPU (two streams, two threads) share the array and race.
    int32_t* pair_inter = scratch;
    int32_t* pair_union = scratch + AVT_MTC_STATIC_MAX_FRAMES;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    mtc_pairs_kernel<<<T - 1, 256, 0, st>>>(masks, pair_inter, pair_union);
    AVT_LAUNCH_CHECK();
    mtc_mean_kernel<<<1, 32, 0, st>>>(pair_inter, pair_union, T, out);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

extern "C" int avt_video_metrics(const double* ciou, const uint8_t* masks, int V, int F,
                                 int32_t* pair_inter, int32_t* pair_union, double* out, void* stream) {
    AVT_REQUIRE(V >= 0 && F >= 1, "avt_video_metrics: V=%d F=%d", V, F);
    if (V == 0) return AVT_OK;
    AVT_REQUIRE(ciou && masks && pair_inter && pair_union && out, "avt_video_metrics: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if ((size_t)V * F >= 2) {      // every consecutive pair; the V-1 pairs across videos are not read
        mtc_pairs_kernel<<<V * F - 1, 256, 0, st>>>(masks, pair_inter, pair_union);
        AVT_LAUNCH_CHECK();
    }
    video_metrics_kernel<<<V, 32, 0, st>>>(ciou, pair_inter, pair_union, F, out);
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
