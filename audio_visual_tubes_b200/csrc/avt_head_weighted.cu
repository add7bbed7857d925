// Backward of weighted_A (model.py:148-152):
//     norm_pos   = F.normalize(Pos, dim=(2,3))            Pos = sigmoid((A - eps)/tau)
//     weighted_A = (img * norm_pos).mean(dim=1)           img = the L2-normalised visual map
// i.e. weighted_A[n,p] = msum[n,p] * Pos[n,p] / max(||Pos[n,:]||, 1e-12) with msum = mean_c v_hat.
// train_hardway.py:138-142 back-propagates an MSE and the PropagationLoss through it, so an
// upstream gradient gW [N_v,P] has two destinations:
//   (1) the positive map:  dA[n,p] += dPos * Pos (1 - Pos) / tau, dPos the normalise Jacobian
//       applied to gW * msum  -> handed to the row-gradient kernel through its dA input;
//   (2) every channel of v_hat alike:  dv_hat[n,c,p] += u[n,p], u = gW * norm_pos / C, which
//       after the visual normalise Jacobian is
//           dv[n,c,p] += rnorm * u * (1 - v_hat[n,c,p] * C * msum[n,p]).
// (1) is one block per row over P values; (2) is one streaming pass over v and dv (HBM-bound:
// 12 bytes per element), run after the main dv kernels.
#include "avt_common.cuh"

namespace avt {

namespace {

constexpr int kWThreads = 256;

__global__ void __launch_bounds__(kWThreads)
weighted_row_grad_kernel(const float* __restrict__ A, const float* __restrict__ msum,
                         const float* __restrict__ gW, const float* __restrict__ dA_in, int P, int C,
                         HeadParams hp, float* __restrict__ dA_out, float* __restrict__ u_out) {
    __shared__ float s_red[kWThreads / 32][2];
    const int n = blockIdx.x;
    const float itau = 1.0f / hp.tau;
    const int p = threadIdx.x;                 // P <= 256: one value per thread
    const bool on = p < P;
    const size_t o = (size_t)n * P + (on ? p : 0);
    const float a = on ? A[o] : 0.f;
    const float pos = on ? 1.0f / (1.0f + expf(-(a - hp.eps) * itau)) : 0.f;
    const float g = on ? gW[o] : 0.f;
    const float m = on ? msum[o] : 0.f;
    const float gnp = g * m;                   // gradient wrt norm_pos
    float spp = warp_sum(pos * pos), sgp = warp_sum(gnp * pos);
    if ((threadIdx.x & 31) == 0) { s_red[threadIdx.x >> 5][0] = spp; s_red[threadIdx.x >> 5][1] = sgp; }
    __syncthreads();
    spp = sgp = 0.f;
#pragma unroll
    for (int i = 0; i < kWThreads / 32; ++i) { spp += s_red[i][0]; sgp += s_red[i][1]; }
    if (!on) return;
    const float nrm = sqrtf(spp);
    float dpos;
    if (nrm > kNormEps) {                      // y = x/|x|: dx = (g - y (g.y)) / |x|
        const float r = 1.0f / nrm;
        dpos = (gnp - pos * r * (sgp * r)) * r;
    } else {                                   // clamped denominator is a constant
        dpos = gnp / kNormEps;
    }
    const float extra = dpos * pos * (1.0f - pos) * itau;
    dA_out[o] = (dA_in ? dA_in[o] : 0.f) + extra;
    u_out[o] = g * pos / fmaxf(nrm, kNormEps) / (float)C;
}

// dv[n,c,p] += rnorm * u * (1 - v*rnorm * C*msum); 4 consecutive p per thread, all c of a
// 64-channel slab per block row so the [n,p] operands are loaded once per slab
__global__ void __launch_bounds__(256)
weighted_dv_kernel(const float* __restrict__ v, const float* __restrict__ rnorm_v,
                   const float* __restrict__ msum, const float* __restrict__ u, int C, int P,
                   int c_per_block, bool vec, float* __restrict__ dv) {
    const int n = blockIdx.y;
    const int c_begin = blockIdx.x * c_per_block;
    const int P4 = vec ? P >> 2 : 0;           // rows are 16-byte aligned only when P % 4 == 0
    for (int q = threadIdx.x; q < P4; q += blockDim.x) {
        const size_t np = (size_t)n * P + 4 * q;
        const float4 rn = *reinterpret_cast<const float4*>(rnorm_v + np);
        const float4 ms = *reinterpret_cast<const float4*>(msum + np);
        const float4 uu = *reinterpret_cast<const float4*>(u + np);
        const float4 k1 = make_float4(rn.x * uu.x, rn.y * uu.y, rn.z * uu.z, rn.w * uu.w);
        const float fc = (float)C;
        const float4 k2 = make_float4(rn.x * fc * ms.x, rn.y * fc * ms.y, rn.z * fc * ms.z, rn.w * fc * ms.w);
        const size_t base = ((size_t)n * C + c_begin) * P + 4 * q;
#pragma unroll 4
        for (int c = 0; c < c_per_block; ++c) {
            const float4 x = __ldcs(reinterpret_cast<const float4*>(v + base + (size_t)c * P));
            float4 d = *reinterpret_cast<float4*>(dv + base + (size_t)c * P);
            d.x = fmaf(k1.x, 1.0f - x.x * k2.x, d.x);
            d.y = fmaf(k1.y, 1.0f - x.y * k2.y, d.y);
            d.z = fmaf(k1.z, 1.0f - x.z * k2.z, d.z);
            d.w = fmaf(k1.w, 1.0f - x.w * k2.w, d.w);
            *reinterpret_cast<float4*>(dv + base + (size_t)c * P) = d;
        }
    }
    // scalar path when P is not a multiple of 4 (fp32 path only)
    for (int p = 4 * P4 + threadIdx.x; p < P; p += blockDim.x) {
        const size_t np = (size_t)n * P + p;
        const float k1 = rnorm_v[np] * u[np], k2 = rnorm_v[np] * (float)C * msum[np];
        for (int c = 0; c < c_per_block; ++c) {
            const size_t i = ((size_t)n * C + c_begin + c) * P + p;
            dv[i] = fmaf(k1, 1.0f - v[i] * k2, dv[i]);
        }
    }
}

}  // namespace

int launch_weighted_row_grad(const float* A, const float* msum, const float* gW, const float* dA_in,
                             int N_v, int P, int C, HeadParams hp, float* dA_out, float* u_out,
                             cudaStream_t st) {
    weighted_row_grad_kernel<<<N_v, kWThreads, 0, st>>>(A, msum, gW, dA_in, P, C, hp, dA_out, u_out);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

int launch_weighted_dv(const float* v, const float* rnorm_v, const float* msum, const float* u,
                       int N_v, int C, int P, float* dv, cudaStream_t st) {
    const int c_per_block = 64;                 // C % 64 == 0 is checked at the API
    dim3 grid(C / c_per_block, N_v);
    const bool vec = P % 4 == 0;
    weighted_dv_kernel<<<grid, vec ? 64 : 256, 0, st>>>(v, rnorm_v, msum, u, C, P, c_per_block, vec, dv);
    AVT_LAUNCH_CHECK();
    return AVT_OK;
}

}  // namespace avt
