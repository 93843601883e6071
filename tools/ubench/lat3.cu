// tools/ubench/lat3.cu -- cost of fetching the 5 emission values per lattice step from shared memory
//   Y=0: constants   Y=1: fp64 rows, LDS.64 gathers   Y=2: fp32 rows + cvt.f64.f32   Y=3: fp32 rows + integer widening
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../deepspeech.pytorch_b200/csrc/ctc_device.cuh"
using namespace ctcb200;
#define N 2048
__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }
constexpr int K = 8;
__device__ __forceinline__ double widen(float x) {  // exact for positive normal floats
    const unsigned b = __float_as_uint(x);
    return __hiloint2double((int)((b >> 3) + 0x38000000u), (int)(b << 29));
}
template <int Y>
__global__ void k_step(double *out, long long *cyc, int C) {
    __shared__ double yd[8 * 32];
    __shared__ float yf[8 * 32];
    __shared__ unsigned yp[8 * 64];
    __shared__ double slot[64];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) yp[i] = (i & 32) ? 0x20000000u * (i & 7) : 0x3fd00000u + (i % 29) * 4096;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) { yd[i] = 0.3 + 1e-3 * (i % 29); yf[i] = 0.3f + 1e-3f * (i % 29); }
    __syncthreads();
    double a[K];
    for (int i = 0; i < K; ++i) a[i] = 1.0 + threadIdx.x * 1e-3 + i;
    double skip[K / 2];
    int idx[K / 2];
    for (int h = 0; h < K / 2; ++h) { skip[h] = (threadIdx.x + h) & 1 ? 1.0 : 0.0; idx[h] = 1 + ((threadIdx.x * 7 + h * 13 + (threadIdx.x >> 2)) % (C - 1)); }
    double pr0 = 0.5;
    const int P = (Y == 1) ? 30 : 32;
    YRow<K> y;
    y.yb = 0.3; for (int h = 0; h < K / 2; ++h) y.yl[h] = 0.31;
    long long t0 = clk();
    for (int n = 0; n < N; n += 8) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            YRow<K> yn = y;
            const int rn = (r + 1) & 7;
            if (Y == 1) { yn.yb = yd[rn * P]; for (int h = 0; h < K / 2; ++h) yn.yl[h] = yd[rn * P + idx[h]]; }
            if (Y == 2) { yn.yb = (double)yf[rn * P]; for (int h = 0; h < K / 2; ++h) yn.yl[h] = (double)yf[rn * P + idx[h]]; }
            if (Y >= 4) { const unsigned *row = yp + rn * 64; yn.yb = __hiloint2double(row[0], row[32]); for (int h = 0; h < K / 2; ++h) yn.yl[h] = __hiloint2double(row[idx[h]], row[32 + idx[h]]); }
            if (Y == 3) { yn.yb = widen(yf[rn * P]); for (int h = 0; h < K / 2; ++h) yn.yl[h] = widen(yf[rn * P + idx[h]]); }
            const double p1 = pr0;
            const double top = alpha_top<K>(a, y, skip);
            pr0 = __shfl_up_sync(kFull, top, 1);
            if (Y == 5) { if ((threadIdx.x & 31) == 31) slot[(threadIdx.x >> 5) * 8 + r] = top; }
            alpha_rest<K>(a, top, p1, y, skip);
            if (Y == 5) { if ((threadIdx.x & 31) == 0) pr0 = slot[((threadIdx.x >> 5) ^ 1) * 8 + r]; }
            y = yn;
        }
        // keep values finite: cheap rescale
#pragma unroll
        for (int i = 0; i < K; ++i) a[i] = a[i] * 0.25 + 1.0;
    }
    long long t1 = clk();
    double s = 0;
    for (int i = 0; i < K; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double *out; long long *cyc, h;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
#define RUN(name, call) call; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); { cudaError_t e = cudaGetLastError(); printf("%-52s %8.2f clk/step %s\n", name, (double)h / N, e ? cudaGetErrorString(e) : ""); }
    RUN("Y0 constants, 4 warps", (k_step<0><<<1, 128>>>(out, cyc, 29)));
    RUN("Y1 fp64 rows LDS.64 gather, 4 warps", (k_step<1><<<1, 128>>>(out, cyc, 29)));
    RUN("Y2 fp32 rows + cvt, 4 warps", (k_step<2><<<1, 128>>>(out, cyc, 29)));
    RUN("Y3 fp32 rows + int widen, 4 warps", (k_step<3><<<1, 128>>>(out, cyc, 29)));
    RUN("Y4 planar hi/lo 10 LDS.32, 4 warps", (k_step<4><<<1, 128>>>(out, cyc, 29)));
    RUN("Y5 planar + slot sts/lds, 4 warps", (k_step<5><<<1, 128>>>(out, cyc, 29)));
    RUN("Y4 planar hi/lo 10 LDS.32, 1 warp", (k_step<4><<<1, 32>>>(out, cyc, 29)));
    RUN("Y1 fp64 rows LDS.64 gather, 8 warps", (k_step<1><<<1, 256>>>(out, cyc, 29)));
    RUN("Y2 fp32 rows + cvt, 8 warps", (k_step<2><<<1, 256>>>(out, cyc, 29)));
    RUN("Y3 fp32 rows + int widen, 8 warps", (k_step<3><<<1, 256>>>(out, cyc, 29)));
    RUN("Y0 constants, 8 warps", (k_step<0><<<1, 256>>>(out, cyc, 29)));
    return 0;
}
