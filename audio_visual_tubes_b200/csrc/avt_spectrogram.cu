// The audio front-end that feeds the path, datasets/dataloader.py:86-96:
//     samples tiled to 10 s, cut, clipped to [-1, 1]
//     scipy.signal.spectrogram(x, fs, nperseg=512, noverlap=1)   (tukey(0.25) periodic window,
//         detrend='constant', one-sided 'density' PSD: 257 bins, hop 511)
//     log(S + 1e-7), ToTensor + Normalize(mean 0, std 12)  -> [1, 257, T], later .float()
// One block = kSegs consecutive segments of one clip.  Two real segments ride one complex
// 512-point transform (re = segment 2p, im = segment 2p + 1); a transform belongs to 64 threads
// (two warps) that hold 8 points each in registers and run it as 8 x 8 x 8: three radix-8
// butterflies with two conflict-free exchanges through shared memory (re / im planes of doubles,
// padded), named barriers per warp pair instead of block-wide ones.  All arithmetic in fp64 like
// scipy; every frequency row gets kSegs consecutive floats.
// Algorithmic HBM traffic: 4 B read per sample, 4 B written per bin.  In fp64 the kernel is bound by
// the L1/shared data path (84 % busy) and the fp64 pipe (52 %), not by HBM (profiles/).
#include "avt_common.cuh"

namespace avt {

namespace {

constexpr int kSeg = 512;            // nperseg = nfft
constexpr int kHop = 511;            // nperseg - noverlap
constexpr int kBins = 257;
constexpr int kPlane = 644;          // 512 doubles + padding (layout 3 ends at 637; planes 4 banks apart)
constexpr double kRh = 0.70710678118654752440;

struct cd { double x, y; };
__device__ __forceinline__ cd cadd(cd a, cd b) { return {a.x + b.x, a.y + b.y}; }
__device__ __forceinline__ cd csub(cd a, cd b) { return {a.x - b.x, a.y - b.y}; }
__device__ __forceinline__ cd cmul(cd a, cd b) { return {a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x}; }
__device__ __forceinline__ cd mul_mi(cd a) { return {a.y, -a.x}; }                 // * (-i)

// 8-point DFT (forward, W8 = exp(-2 pi i / 8)) in place; a[b] ends up holding Y[bitrev3(b)]
__device__ __forceinline__ void dft8(cd (&a)[8]) {
    cd u0 = cadd(a[0], a[4]), u4 = csub(a[0], a[4]);
    cd u1 = cadd(a[1], a[5]), d1 = csub(a[1], a[5]);
    cd u2 = cadd(a[2], a[6]), u6 = mul_mi(csub(a[2], a[6]));
    cd u3 = cadd(a[3], a[7]), d3 = csub(a[3], a[7]);
    cd u5 = {kRh * (d1.x + d1.y), kRh * (d1.y - d1.x)};                             // * W8^1
    cd u7 = {kRh * (d3.y - d3.x), -kRh * (d3.x + d3.y)};                            // * W8^3
    cd v0 = cadd(u0, u2), v2 = csub(u0, u2);
    cd v1 = cadd(u1, u3), v3 = mul_mi(csub(u1, u3));
    cd v4 = cadd(u4, u6), v6 = csub(u4, u6);
    cd v5 = cadd(u5, u7), v7 = mul_mi(csub(u5, u7));
    a[0] = cadd(v0, v1); a[1] = csub(v0, v1);      // Y0, Y4
    a[2] = cadd(v2, v3); a[3] = csub(v2, v3);      // Y2, Y6
    a[4] = cadd(v4, v5); a[5] = csub(v4, v5);      // Y1, Y5
    a[6] = cadd(v6, v7); a[7] = csub(v6, v7);      // Y3, Y7
}
__device__ __forceinline__ constexpr int rev3(int q) { return ((q & 1) << 2) | (q & 2) | (q >> 2); }

// a[rev3(q)] *= w^q for q = 1..7 (w^2, w^4 by squaring: at most three roundings deep)
__device__ __forceinline__ void twiddle8(cd (&a)[8], cd w) {
    const cd w2 = cmul(w, w), w3 = cmul(w2, w), w4 = cmul(w2, w2);
    const cd w5 = cmul(w4, w), w6 = cmul(w4, w2), w7 = cmul(w4, w3);
    a[rev3(1)] = cmul(a[rev3(1)], w);  a[rev3(2)] = cmul(a[rev3(2)], w2);
    a[rev3(3)] = cmul(a[rev3(3)], w3); a[rev3(4)] = cmul(a[rev3(4)], w4);
    a[rev3(5)] = cmul(a[rev3(5)], w5); a[rev3(6)] = cmul(a[rev3(6)], w6);
    a[rev3(7)] = cmul(a[rev3(7)], w7);
}

// exp(-2 pi i j / 512), j = 0..63, correctly rounded (generated with 50-digit arithmetic); pass 1 uses
// entry u, pass 2 entry 8 r (W64^r)
__device__ const double2 kW512[64] = {
    {1.0, 0.0}, {0.9999247018391445, -0.012271538285719925},
    {0.9996988186962042, -0.024541228522912288}, {0.9993223845883495, -0.03680722294135883},
    {0.9987954562051724, -0.049067674327418015}, {0.9981181129001492, -0.06132073630220858},
    {0.9972904566786902, -0.07356456359966743}, {0.996312612182778, -0.0857973123444399},
    {0.9951847266721969, -0.0980171403295606}, {0.9939069700023561, -0.11022220729388306},
    {0.99247953459871, -0.1224106751992162}, {0.99090263542778, -0.1345807085071262},
    {0.989176509964781, -0.14673047445536175}, {0.9873014181578584, -0.15885814333386145},
    {0.9852776423889412, -0.17096188876030122}, {0.9831054874312163, -0.18303988795514095},
    {0.9807852804032304, -0.19509032201612828}, {0.9783173707196277, -0.20711137619221856},
    {0.9757021300385286, -0.2191012401568698}, {0.9729399522055602, -0.2310581082806711},
    {0.970031253194544, -0.2429801799032639}, {0.9669764710448521, -0.25486565960451457},
    {0.9637760657954398, -0.26671275747489837}, {0.9604305194155658, -0.2785196893850531},
    {0.9569403357322088, -0.2902846772544624}, {0.9533060403541939, -0.3020059493192281},
    {0.9495281805930367, -0.31368174039889146}, {0.9456073253805213, -0.3253102921622629},
    {0.9415440651830208, -0.33688985339222005}, {0.937339011912575, -0.34841868024943456},
    {0.9329927988347388, -0.35989503653498817}, {0.9285060804732156, -0.37131719395183754},
    {0.9238795325112867, -0.3826834323650898}, {0.9191138516900578, -0.3939920400610481},
    {0.9142097557035307, -0.40524131400498986}, {0.9091679830905224, -0.4164295600976372},
    {0.9039892931234433, -0.4275550934302821}, {0.8986744656939538, -0.43861623853852766},
    {0.8932243011955153, -0.4496113296546066}, {0.8876396204028539, -0.46053871095824},
    {0.881921264348355, -0.47139673682599764}, {0.8760700941954066, -0.4821837720791228},
    {0.8700869911087115, -0.49289819222978404}, {0.8639728561215867, -0.5035383837257176},
    {0.8577286100002721, -0.5141027441932218}, {0.8513551931052652, -0.524589682678469},
    {0.8448535652497071, -0.5349976198870973}, {0.8382247055548381, -0.5453249884220465},
    {0.8314696123025452, -0.5555702330196022}, {0.8245893027850253, -0.5657318107836132},
    {0.8175848131515837, -0.5758081914178453}, {0.8104571982525948, -0.5857978574564389},
    {0.8032075314806449, -0.5956993044924334}, {0.7958369046088836, -0.6055110414043255},
    {0.7883464276266062, -0.6152315905806268}, {0.7807372285720945, -0.6248594881423863},
    {0.773010453362737, -0.6343932841636455}, {0.765167265622459, -0.6438315428897915},
    {0.7572088465064846, -0.6531728429537768}, {0.7491363945234594, -0.6624157775901718},
    {0.7409511253549591, -0.6715589548470184}, {0.7326542716724128, -0.680600997795453},
    {0.7242470829514669, -0.6895405447370669}, {0.7157308252838187, -0.6983762494089728},
};

__device__ __forceinline__ void pair_barrier(int id) {      // the two warps of one transform
    asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory");
}

// log(x) * g for normal positive x from a per-block table: x = 2^e m, m in [1, 2); the top 7
// mantissa bits pick c = centre of m's 1/128 cell, table = (rc ~ 1/c, -g log(rc)); r = m rc - 1
// (|r| < 2^-7.9) and log m = log1p(r) - log(rc) with log1p to r^7 (remainder < 1e-18); absolute
// error ~2e-16, far below the fp32 rounding of the result.  Coefficients sit in the constant bank
// so the DFMAs take them as operands.
constexpr int kLogCells = 128;
__constant__ double kLogC[6] = {1.0 / 7.0, -1.0 / 6.0, 1.0 / 5.0, -1.0 / 4.0, 1.0 / 3.0, -0.5};
__device__ __forceinline__ double fast_log_scaled(double x, const double2* __restrict__ tab, double g,
                                                  double g_ln2) {
    const int hi = __double2hiint(x), lo = __double2loint(x);
    const int e = (hi >> 20) - 1023;
    const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, lo);
    const double2 c = tab[(hi >> 13) & (kLogCells - 1)];
    const double r = fma(m, c.x, -1.0);
    double p = fma(r, kLogC[0], kLogC[1]);
    p = fma(p, r, kLogC[2]);
    p = fma(p, r, kLogC[3]);
    p = fma(p, r, kLogC[4]);
    p = fma(p, r, kLogC[5]);
    p = fma(p * r, r, r);
    return fma((double)e, g_ln2, fma(p, g, c.y));
}

// one sample as sf.read would return it: fp32 as is, 16-bit PCM divided by 32768 (exact in fp32)
__device__ __forceinline__ float sample(const float* p) { return __ldg(p); }
__device__ __forceinline__ float sample(const int16_t* p) { return (float)__ldg(p) * (1.0f / 32768.0f); }

template <int kSegs, typename TS>
__global__ void __launch_bounds__(kSegs * 32)
spectrogram_kernel(const TS* __restrict__ samples, const int32_t* __restrict__ lengths,
                   long long stride, int n_keep, const double* __restrict__ window, double scale,
                   double log_eps, double inv_std, float* __restrict__ out, int T) {
    constexpr int kFft = kSegs / 2;
    __shared__ double sre[kFft][kPlane], sim[kFft][kPlane];
    __shared__ double ssum[kSegs][2];
    __shared__ double2 slog[kLogCells];
    const int clip = blockIdx.y;
    const int t0 = blockIdx.x * kSegs;
    const int tid = threadIdx.x, lane = tid & 31;
    const int p = tid >> 6, u = tid & 63;                 // transform, thread within it
    const TS* x = samples + (size_t)clip * stride;
    const int L = lengths ? lengths[clip] : n_keep;       // shorter clips are tiled (dataloader.py:88-90)
    double* re = sre[p];
    double* im = sim[p];
    if (tid < kLogCells) {
        const double rc = 1.0 / (1.0 + (tid + 0.5) * (1.0 / kLogCells));
        slog[tid] = make_double2(rc, -log(rc) * inv_std);
    }

    // ---- load: thread u holds points u + 64 m of segments 2p (re) and 2p + 1 (im), clipped
    // (dataloader.py:92-93); per-segment mean in fp64 (detrend='constant')
    cd a[8];
    {
        double s0 = 0.0, s1 = 0.0;
        const int ta = t0 + 2 * p, tb = ta + 1;
        if (L >= n_keep) {                                // block-uniform: a full 10 s clip, no tiling
            const TS* xa = x + (size_t)ta * kHop + u;
            float fa[8], fb[8];
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                fa[m] = ta < T ? sample(xa + 64 * m) : 0.f;
                fb[m] = tb < T ? sample(xa + kHop + 64 * m) : 0.f;
            }
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                a[m].x = (double)fminf(fmaxf(fa[m], -1.0f), 1.0f);
                a[m].y = (double)fminf(fmaxf(fb[m], -1.0f), 1.0f);
                s0 += a[m].x; s1 += a[m].y;
            }
        } else {
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                const int j = u + 64 * m;
                const float fa = ta < T ? sample(x + (ta * kHop + j) % L) : 0.f;
                const float fb = tb < T ? sample(x + (tb * kHop + j) % L) : 0.f;
                a[m].x = (double)fminf(fmaxf(fa, -1.0f), 1.0f);
                a[m].y = (double)fminf(fmaxf(fb, -1.0f), 1.0f);
            }
#pragma unroll
            for (int m = 0; m < 8; ++m) { s0 += a[m].x; s1 += a[m].y; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, o);
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        }
        const int half = (tid >> 5) & 1;
        if (lane == 0) { ssum[2 * p][half] = s0; ssum[2 * p + 1][half] = s1; }
        pair_barrier(1 + p);
        const double m0 = (ssum[2 * p][0] + ssum[2 * p][1]) * (1.0 / kSeg);
        const double m1 = (ssum[2 * p + 1][0] + ssum[2 * p + 1][1]) * (1.0 / kSeg);
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            const double w = __ldg(window + u + 64 * m);
            a[m].x = (a[m].x - m0) * w;
            a[m].y = (a[m].y - m1) * w;
        }
    }
    // ---- pass 1: n = u + 64 m -> 8 sub-transforms q of length 64 over u, twiddle W512^(u q)
    dft8(a);
    {
        const double2 t = __ldg(&kW512[u]);
        const cd w = {t.x, t.y};
        twiddle8(a, w);
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {                         // layout 1: [q][u], rows padded to 72
        re[72 * q + u] = a[rev3(q)].x;
        im[72 * q + u] = a[rev3(q)].y;
    }
    pair_barrier(1 + p);
    // ---- pass 2: thread (q, r) = (u / 8, u % 8) takes points r + 8 m of sub-transform q
    {
        const int q = u >> 3, r = u & 7;
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            a[m].x = re[72 * q + r + 8 * m];
            a[m].y = im[72 * q + r + 8 * m];
        }
        dft8(a);
        const double2 t = __ldg(&kW512[8 * r]);           // W64^r
        const cd w = {t.x, t.y};
        twiddle8(a, w);
        pair_barrier(1 + p);                              // everyone has read layout 1
#pragma unroll
        for (int q2 = 0; q2 < 8; ++q2) {                  // layout 2: row (8 q + q2) of 8, padded to 9
            re[9 * (8 * q + q2) + r] = a[rev3(q2)].x;
            im[9 * (8 * q + q2) + r] = a[rev3(q2)].y;
        }
    }
    pair_barrier(1 + p);
    // ---- pass 3: thread (q, q2) = (u / 8, u % 8) transforms its row -> X[q + 8 q2 + 64 r3]
    {
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            a[m].x = re[9 * u + m];
            a[m].y = im[9 * u + m];
        }
        dft8(a);
        pair_barrier(1 + p);                              // everyone has read layout 2
        const int k0 = (u >> 3) + 8 * (u & 7);
#pragma unroll
        for (int r3 = 0; r3 < 8; ++r3) {                  // layout 3: natural order, 10 slots per 8
            const int k = k0 + 64 * r3;
            re[k + 2 * (k >> 3)] = a[rev3(r3)].x;
            im[k + 2 * (k >> 3)] = a[rev3(r3)].y;
        }
    }
    __syncthreads();
    // ---- split the packed pairs, one-sided density PSD, log, normalise.  Thread = (transform pp,
    // bin k0 + rows i) finishes both segments of its transform: kSegs consecutive t per bin row
    {
        constexpr int kRows = kSegs * 32 / kFft;          // bins per sweep = 64
        const int pp = tid % kFft, t = t0 + 2 * pp;
        const double* pre = sre[pp];
        const double* pim = sim[pp];
        const double g_ln2 = 0.69314718055994530942 * inv_std;
        float* o = out + ((size_t)clip * kBins) * T + t;
        if (t < T) {
            const bool both = t + 1 < T;
#pragma unroll 1
            for (int k = tid / kFft; k < kBins; k += kRows) {
                const int kn = (kSeg - k) & (kSeg - 1);
                const int ik = k + 2 * (k >> 3), in = kn + 2 * (kn >> 3);
                const double zkx = pre[ik], zky = pim[ik], znx = pre[in], zny = pim[in];
                // even: X = (Z[k] + conj(Z[N-k])) / 2;  odd: X = (Z[k] - conj(Z[N-k])) / (2i)
                const double ar = zkx + znx, ai = zky - zny, br = zky + zny, bi = zkx - znx;
                const double sc = (k != 0 && k != kSeg / 2) ? 0.5 * scale : 0.25 * scale;
                const double pa = fma(fma(ar, ar, ai * ai), sc, log_eps);
                const double pb = fma(fma(br, br, bi * bi), sc, log_eps);
                float* ok = o + (size_t)k * T;
                ok[0] = (float)fast_log_scaled(pa, slog, inv_std, g_ln2);
                if (both) ok[1] = (float)fast_log_scaled(pb, slog, inv_std, g_ln2);
            }
        }
    }
}

}  // namespace
}  // namespace avt

using namespace avt;

extern "C" int avt_spectrogram_frames(int n_keep) {
    return n_keep < kSeg ? 0 : (n_keep - (kSeg - kHop)) / kHop;
}

namespace {
template <typename TS>
int log_spectrogram(const char* who, const TS* samples, const int32_t* lengths, int B, long long stride,
                    int n_keep, const double* window, double scale, double log_eps, double std,
                    float* out, int T, void* stream) {
    AVT_REQUIRE(B >= 0, "%s: B=%d", who, B);
    if (B == 0) return AVT_OK;
    AVT_REQUIRE(samples && window && out, "%s: null pointer", who);
    AVT_REQUIRE(n_keep >= kSeg && stride >= 1, "%s: n_keep=%d needs at least %d samples", who, n_keep, kSeg);
    AVT_REQUIRE(T == avt_spectrogram_frames(n_keep), "%s: T=%d but %d samples give %d frames", who,
                T, n_keep, avt_spectrogram_frames(n_keep));
    AVT_REQUIRE(std != 0.0, "%s: std is 0", who);
    // 8 segments per block: 4 segments halve the row runs (0.095 ms vs 0.078 ms for 128 clips at
    // 16 kHz), 16 (78 KB of opt-in shared memory, 512 threads) measured slower too: 0.086 ms
    constexpr int kSegs = 8;
    spectrogram_kernel<kSegs, TS><<<dim3((T + kSegs - 1) / kSegs, B), kSegs * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        samples, lengths, stride, n_keep, window, scale, log_eps, 1.0 / std, out, T);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}
}  // namespace

extern "C" int avt_log_spectrogram(const float* samples, const int32_t* lengths, int B,
                                   long long stride, int n_keep, const double* window, double scale,
                                   double log_eps, double std, float* out, int T, void* stream) {
    return log_spectrogram("avt_log_spectrogram", samples, lengths, B, stride, n_keep, window, scale,
                           log_eps, std, out, T, stream);
}

extern "C" int avt_log_spectrogram_pcm16(const int16_t* samples, const int32_t* lengths, int B,
                                         long long stride, int n_keep, const double* window, double scale,
                                         double log_eps, double std, float* out, int T, void* stream) {
    return log_spectrogram("avt_log_spectrogram_pcm16", samples, lengths, B, stride, n_keep, window, scale,
                           log_eps, std, out, T, stream);
}
