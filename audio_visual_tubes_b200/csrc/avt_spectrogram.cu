// The audio front-end that feeds the path, datasets/dataloader.py:86-96:
//     samples tiled to 10 s, cut, clipped to [-1, 1]
//     scipy.signal.spectrogram(x, fs, nperseg=512, noverlap=1)   (tukey(0.25) periodic window,
//         detrend='constant', one-sided 'density' PSD: 257 bins, hop 511)
//     log(S + 1e-7), ToTensor + Normalize(mean 0, std 12)  -> [1, 257, T], later .float()
// One block = 8 consecutive segments of one clip: the 8 real 512-point transforms run as 4
// complex radix-2 FFTs in shared memory (two real sequences per complex transform), all in
// fp64 like scipy, and every frequency row gets 8 consecutive floats (one 32-byte sector).
// HBM-bound: 4 B read per sample, 4 B written per bin.
#include "avt_common.cuh"

namespace avt {

namespace {

constexpr int kSeg = 512;            // nperseg = nfft
constexpr int kHop = 511;            // nperseg - noverlap
constexpr int kBins = 257;
constexpr int kSegPerBlock = 8;
constexpr int kSpecThreads = 256;

__global__ void __launch_bounds__(kSpecThreads)
spectrogram_kernel(const float* __restrict__ samples, const int32_t* __restrict__ lengths,
                   long long stride, int n_keep, const double* __restrict__ window, double scale,
                   double log_eps, double std, float* __restrict__ out, int T) {
    __shared__ double2 z[kSegPerBlock / 2][kSeg];      // 32 KB: FFT p holds segments 2p (re), 2p+1 (im)
    __shared__ double2 tw[kSeg / 2];                   // exp(-2 pi i j / 512)
    const int clip = blockIdx.y;
    const int t0 = blockIdx.x * kSegPerBlock;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* x = samples + (size_t)clip * stride;
    const int L = lengths ? lengths[clip] : n_keep;    // shorter clips are tiled (dataloader.py:88-90)

    {
        double s, c;
        sincospi(-(double)tid / 256.0, &s, &c);
        tw[tid] = make_double2(c, s);
    }
    // warp w loads segment w: 16 samples per lane, clipped (dataloader.py:92-93); mean in fp64
    {
        const int seg = warp, t = t0 + seg;
        double v[16];
        double sum = 0.0;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int n = t * kHop + lane + 32 * i;
            float f = 0.f;
            if (t < T) {
                const int src = (L >= n_keep) ? n : n % L;
                f = fminf(fmaxf(__ldg(x + src), -1.0f), 1.0f);
            }
            v[i] = (double)f;
            sum += v[i];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const double mean = sum / (double)kSeg;        // detrend='constant'
        double* comp = reinterpret_cast<double*>(&z[seg >> 1][0]) + (seg & 1);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int j = lane + 32 * i;
            comp[2 * j] = (v[i] - mean) * window[j];
        }
    }
    __syncthreads();
    // decimation-in-frequency radix-2, in place; X[k] ends at index bitrev9(k)
#pragma unroll 1
    for (int h = kSeg / 2, shift = 0; h >= 1; h >>= 1, ++shift) {
#pragma unroll
        for (int p = 0; p < kSegPerBlock / 2; ++p) {
            const int b = tid;                          // 256 butterflies per transform
            const int j = b & (h - 1);
            const int i0 = ((b - j) << 1) + j, i1 = i0 + h;
            const double2 a = z[p][i0], c = z[p][i1];
            const double2 w = tw[j << shift];
            const double dr = a.x - c.x, di = a.y - c.y;
            z[p][i0] = make_double2(a.x + c.x, a.y + c.y);
            z[p][i1] = make_double2(dr * w.x - di * w.y, dr * w.y + di * w.x);
        }
        __syncthreads();
    }
    // split the packed pairs, one-sided density PSD, log, normalise; 8 consecutive t per bin
    for (int idx = tid; idx < kBins * kSegPerBlock; idx += kSpecThreads) {
        const int k = idx >> 3, s = idx & 7, t = t0 + s;
        if (t >= T) continue;
        const int p = s >> 1;
        const double2 zk = z[p][__brev((unsigned)k) >> 23];
        const double2 zn = z[p][__brev((unsigned)((kSeg - k) & (kSeg - 1))) >> 23];
        double re, im;
        if ((s & 1) == 0) {      // X1 = (Z[k] + conj(Z[N-k])) / 2
            re = 0.5 * (zk.x + zn.x); im = 0.5 * (zk.y - zn.y);
        } else {                 // X2 = (Z[k] - conj(Z[N-k])) / (2i)
            re = 0.5 * (zk.y + zn.y); im = -0.5 * (zk.x - zn.x);
        }
        double psd = (re * re + im * im) * scale;
        if (k != 0 && k != kSeg / 2) psd *= 2.0;
        out[((size_t)clip * kBins + k) * T + t] = (float)(log(psd + log_eps) / std);
    }
}

}  // namespace
}  // namespace avt

using namespace avt;

extern "C" int avt_spectrogram_frames(int n_keep) {
    return n_keep < kSeg ? 0 : (n_keep - (kSeg - kHop)) / kHop;
}

extern "C" int avt_log_spectrogram(const float* samples, const int32_t* lengths, int B,
                                   long long stride, int n_keep, const double* window, double scale,
                                   double log_eps, double std, float* out, int T, void* stream) {
    AVT_REQUIRE(B >= 0, "avt_log_spectrogram: B=%d", B);
    if (B == 0) return AVT_OK;
    AVT_REQUIRE(samples && window && out, "avt_log_spectrogram: null pointer");
    AVT_REQUIRE(n_keep >= kSeg && stride >= 1, "avt_log_spectrogram: n_keep=%d needs at least %d samples",
                n_keep, kSeg);
    AVT_REQUIRE(T == avt_spectrogram_frames(n_keep), "avt_log_spectrogram: T=%d but %d samples give %d frames",
                T, n_keep, avt_spectrogram_frames(n_keep));
    AVT_REQUIRE(std != 0.0, "avt_log_spectrogram: std is 0");
    dim3 grid((T + kSegPerBlock - 1) / kSegPerBlock, B);
    spectrogram_kernel<<<grid, kSpecThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        samples, lengths, stride, n_keep, window, scale, log_eps, std, out, T);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}
