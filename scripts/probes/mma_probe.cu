// Micro-probe (dev tool): cycles per tcgen05.mma kind::f16 at M = 128 (cta_group::1) for several N,
// back-to-back on whatever is in shared memory, plus the SM clock seen while doing so.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_probe mma_probe.cu && ./mma_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../audio_visual_tubes_b200/csrc/avt_tc.cuh"
using namespace avt::tc;

__global__ void __launch_bounds__(128 + 512, 1) probe(int N, int iters, long long* out, int ld_warps, int mufu) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t holder;
    for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(base)[i] = 0x3c003c00u;  // 1.0h
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (threadIdx.x < 32) tmem_alloc(&holder, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = holder;
    if (threadIdx.x == 0) {
        const uint32_t idesc = make_idesc_f16(128, N, 0, 0);
        const uint32_t a = smem_u32(base), b = smem_u32(base + 16384);
        uint64_t g0 = global_ns();
        long long c0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_f16_ss(tm + (it & 1) * 256, make_smem_desc(a + k * 32, 16, 1024), make_smem_desc(b + k * 32, 16, 1024),
                            idesc, (uint32_t)(k != 0));
        }
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        long long c1 = clock64();
        uint64_t g1 = global_ns();
        if (blockIdx.x == 0) { out[0] = c1 - c0; out[1] = (long long)(g1 - g0); }
    }
    else if (threadIdx.x >= 128 && threadIdx.x < 128 + 32 * ld_warps) {
        // "epilogue" warps: stream tcgen05.ld over the accumulator region while the MMAs run
        const int w = (threadIdx.x >> 5) & 3;
        float acc = 0.f;
        volatile int* flag = reinterpret_cast<volatile int*>(&bar);
        (void)flag;
        for (int it = 0; it < iters / 8; ++it) {
            for (int c = 0; c < 192; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(tm + ((uint32_t)(w * 32) << 16) + 256 * (it & 1) + c, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float x = __uint_as_float(r[i]);
                    if (mufu) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(x) : "f"(y + 1.0f)); }
                    acc += x;
                }
            }
        }
        if (acc == 12345.f) out[2] = 1;
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
    long long* d; cudaMalloc(&d, 32);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    const int iters = 20000;
    for (int grid : {148}) {
        for (int ldw : {0, 4, 8, 16}) for (int mufu : {0, 1}) {
          for (int N : {208, 256}) {
            long long h[2];
            probe<<<grid, 128 + 512, 100 * 1024>>>(N, iters, d, ldw, mufu);
            cudaDeviceSynchronize();
            probe<<<grid, 128 + 512, 100 * 1024>>>(N, iters, d, ldw, mufu);
            cudaError_t e = cudaDeviceSynchronize();
            cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
            printf("ld warps %2d mufu %d N %3d: %.1f clk per MMA (K=16), formula %d, SM clock %.0f MHz  %s\n", ldw, mufu, N,
                   (double)h[0] / (iters * 4.0), N / 2, (double)h[0] / (double)h[1] * 1e3, cudaGetErrorString(e));
          }
        }
    }
    return 0;
}
