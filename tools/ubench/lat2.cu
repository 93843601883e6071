// tools/ubench/lat2.cu -- which ingredient of the real lattice step costs what (1 warp per SMSP, 4 warps/SM)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../deepspeech.pytorch_b200/csrc/ctc_device.cuh"
using namespace ctcb200;
#define N 2048
__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }
constexpr int K = 8;
// MODE 0: shuffle m only, no exponent logic.  1: + E shuffle + inflow_scale.  2: + renorm every 8.  3: + group_rebase
template <int MODE>
__global__ void k_step(double *out, long long *cyc, double yv, int e0) {
    Cells<K> c;
    for (int i = 0; i < K; ++i) c.a[i] = 1.0 + threadIdx.x * 1e-3 + i;
    c.E = e0;
    double skip[K / 2];
    for (int h = 0; h < K / 2; ++h) skip[h] = (threadIdx.x + h) & 1 ? 1.0 : 0.0;
    YRow<K> y;
    y.yb = yv;
    for (int h = 0; h < K / 2; ++h) y.yl[h] = yv + 1e-9 * h;
    double pr0 = 0.5;
    int En = e0;
    long long t0 = clk();
    for (int n = 0; n < N; n += 8) {
        if (MODE >= 3) group_rebase<K>(c, En);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            double p1;
            if (MODE >= 1) p1 = pr0 * inflow_scale<K>(c, En); else p1 = pr0;
            const double top = alpha_top<K>(c.a, y, skip);
            double sh_m = 0; int sh_E = 0;
            const bool last = (r == 7) && MODE >= 2;
            if (!last) { sh_m = __shfl_up_sync(kFull, top, 1); if (MODE >= 1) sh_E = __shfl_up_sync(kFull, c.E, 1); }
            alpha_rest<K>(c.a, top, p1, y, skip);
            if (last) { renorm<K>(c); sh_m = __shfl_up_sync(kFull, c.a[K - 1], 1); sh_E = __shfl_up_sync(kFull, c.E, 1); }
            pr0 = sh_m; if (MODE >= 1) En = sh_E;
        }
    }
    long long t1 = clk();
    double s = 0;
    for (int i = 0; i < K; ++i) s += c.a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + c.E;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double *out; long long *cyc, h;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
#define RUN(name, call) call; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); { cudaError_t e = cudaGetLastError(); printf("%-44s %8.2f clk/step %s\n", name, (double)h / N, e ? cudaGetErrorString(e) : ""); }
    RUN("mode0 m-shuffle only, 1 warp", (k_step<0><<<1, 32>>>(out, cyc, 0.97, 0)));
    RUN("mode0 m-shuffle only, 4 warps", (k_step<0><<<1, 128>>>(out, cyc, 0.97, 0)));
    RUN("mode1 + E shuffle + scale, 4 warps", (k_step<1><<<1, 128>>>(out, cyc, 0.97, 0)));
    RUN("mode2 + renorm/8, 4 warps", (k_step<2><<<1, 128>>>(out, cyc, 0.97, 0)));
    RUN("mode3 + group_rebase, 4 warps", (k_step<3><<<1, 128>>>(out, cyc, 0.97, 0)));
    RUN("mode3, 5 warps", (k_step<3><<<1, 160>>>(out, cyc, 0.97, 0)));
    RUN("mode3, 8 warps", (k_step<3><<<1, 256>>>(out, cyc, 0.97, 0)));
    return 0;
}
