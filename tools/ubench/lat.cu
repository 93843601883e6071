// tools/ubench/lat.cu -- B200 micro-benchmarks behind the lattice-kernel design decisions (fp64 pipe, shuffle,
// shared-memory hand-off).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat lat.cu ; ./lat
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N 4096
__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }

template <int CH>
__global__ void k_dfma(double *out, long long *cyc, double m, double a) {
    double x[CH];
    for (int i = 0; i < CH; ++i) x[i] = threadIdx.x * 1e-9 + i;
    long long t0 = clk();
    for (int n = 0; n < N; ++n) {
#pragma unroll
        for (int i = 0; i < CH; ++i) x[i] = fma(x[i], m, a);
    }
    long long t1 = clk();
    double s = 0;
    for (int i = 0; i < CH; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int CH>
__global__ void k_dadd(double *out, long long *cyc, double a) {
    double x[CH];
    for (int i = 0; i < CH; ++i) x[i] = threadIdx.x * 1e-9 + i;
    long long t0 = clk();
    for (int n = 0; n < N; ++n) {
#pragma unroll
        for (int i = 0; i < CH; ++i) x[i] = x[i] + a;
    }
    long long t1 = clk();
    double s = 0;
    for (int i = 0; i < CH; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_shfl32(int *out, long long *cyc) {
    int x = threadIdx.x;
    long long t0 = clk();
    for (int n = 0; n < N; ++n) x = __shfl_up_sync(0xffffffffu, x, 1) + 1;
    long long t1 = clk();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_shfl64(double *out, long long *cyc) {
    double x = threadIdx.x;
    long long t0 = clk();
    for (int n = 0; n < N; ++n) x = __shfl_up_sync(0xffffffffu, x, 1) + 1.0;
    long long t1 = clk();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_lds64(double *out, long long *cyc) {
    __shared__ double s[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s[i] = (double)((i * 7 + 1) & 1023);
    __syncthreads();
    int idx = threadIdx.x;
    long long t0 = clk();
    for (int n = 0; n < N; ++n) idx = (int)s[idx] ;
    long long t1 = clk();
    out[threadIdx.x] = idx;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_bar(int *out, long long *cyc, int nthr) {
    long long t0 = clk();
    for (int n = 0; n < N; ++n) asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory");
    long long t1 = clk();
    if (threadIdx.x == 0) *cyc = t1 - t0;
    out[threadIdx.x] = 0;
}
// ping-pong through shared memory flags between warp 0 and warp 1 (volatile), round trip
__global__ void k_pingpong(int *out, long long *cyc) {
    __shared__ volatile int f0, f1;
    if (threadIdx.x == 0) { f0 = 0; f1 = 0; }
    __syncthreads();
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long t0 = clk();
    if (w == 0) {
        for (int n = 1; n <= N; ++n) {
            if (lane == 0) { f0 = n; while (f1 < n) {} }
            __syncwarp();
        }
    } else if (w == 1) {
        for (int n = 1; n <= N; ++n) {
            if (lane == 0) { while (f0 < n) {} f1 = n; }
            __syncwarp();
        }
    }
    long long t1 = clk();
    if (threadIdx.x == 0) *cyc = t1 - t0;
    out[threadIdx.x] = 0;
}
// a lattice-like step: shfl -> dfma -> dmul chain plus 7 independent cells (K=8), no sync
__global__ void k_cell(double *out, long long *cyc, double y, double f) {
    double a[8];
    for (int i = 0; i < 8; ++i) a[i] = 1.0 + threadIdx.x * 1e-3 + i;
    long long t0 = clk();
    for (int n = 0; n < N; ++n) {
        double p1 = __shfl_up_sync(0xffffffffu, a[7], 1);
#pragma unroll
        for (int k = 7; k >= 0; --k) {
            double prev1 = k ? a[k - 1] : p1 * f;
            if (k & 1) {
                double prev2 = k >= 2 ? a[k - 2] : p1 * f;
                a[k] = fma(f, prev2, a[k] + prev1) * y;
            } else a[k] = (a[k] + prev1) * y;
        }
    }
    long long t1 = clk();
    double s = 0;
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double *out; long long *cyc, h;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
#define RUN(name, call) call; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); { cudaError_t e = cudaGetLastError(); printf("%-44s %8.2f clk/iter %s\n", name, (double)h / N, e ? cudaGetErrorString(e) : ""); }
    RUN("dfma chain x1 (latency), 1 warp", (k_dfma<1><<<1, 32>>>(out, cyc, 1.0000001, 1e-9)));
    RUN("dfma x8 indep, 1 warp (/8 = issue)", (k_dfma<8><<<1, 32>>>(out, cyc, 1.0000001, 1e-9)));
    RUN("dfma x8 indep, 4 warps/SM", (k_dfma<8><<<1, 128>>>(out, cyc, 1.0000001, 1e-9)));
    RUN("dfma x8 indep, 8 warps/SM", (k_dfma<8><<<1, 256>>>(out, cyc, 1.0000001, 1e-9)));
    RUN("dfma x8 indep, 16 warps/SM", (k_dfma<8><<<1, 512>>>(out, cyc, 1.0000001, 1e-9)));
    RUN("dadd chain x1 (latency), 1 warp", (k_dadd<1><<<1, 32>>>(out, cyc, 1e-9)));
    RUN("dadd x8 indep, 1 warp", (k_dadd<8><<<1, 32>>>(out, cyc, 1e-9)));
    RUN("shfl32 + iadd chain", (k_shfl32<<<1, 32>>>((int *)out, cyc)));
    RUN("shfl64 + dadd chain", (k_shfl64<<<1, 32>>>(out, cyc)));
    RUN("lds64 -> cvt -> addr chain", (k_lds64<<<1, 32>>>(out, cyc)));
    RUN("named bar.sync, 4 warps", (k_bar<<<1, 128>>>((int *)out, cyc, 128)));
    RUN("named bar.sync, 8 warps", (k_bar<<<1, 256>>>((int *)out, cyc, 256)));
    RUN("smem flag ping-pong round trip (2 warps)", (k_pingpong<<<1, 64>>>((int *)out, cyc)));
    RUN("lattice step K=8, 1 warp", (k_cell<<<1, 32>>>(out, cyc, 0.97, 1.0)));
    RUN("lattice step K=8, 4 warps/SM", (k_cell<<<1, 128>>>(out, cyc, 0.97, 1.0)));
    RUN("lattice step K=8, 8 warps/SM", (k_cell<<<1, 256>>>(out, cyc, 0.97, 1.0)));
    RUN("lattice step K=8, 16 warps/SM", (k_cell<<<1, 512>>>(out, cyc, 0.97, 1.0)));
    return 0;
}
