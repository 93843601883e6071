// PCIe probe for the host-buffer entry (ctc_b200_compute_host): DMA copies (contiguous and the strided 2-D form the
// utterance slices use) against kernels that read / write PINNED host memory directly.  Build on the box:
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pcie_probe pcie_probe.cu && ./pcie_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void copy_k(const float4 *__restrict__ src, float4 *__restrict__ dst, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = __ldg(src + i);
}

// one warp per (utterance-in-slice, frame): rows of C floats, `B*C` floats apart in the host tensor [T][B][C],
// gathered into / scattered from a compact [nb][T][C] device buffer (what a length-bucketed slice needs when the
// batch is not sorted by length)
__global__ void gather_rows_k(const float *__restrict__ host, float *__restrict__ dev, const int *ids, int nb, int T,
                              int B, int C, int to_host) {
    const int lane = threadIdx.x & 31;
    const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long r = w; r < (long long)nb * T; r += nw) {
        const int u = (int)(r / T), t = (int)(r - (long long)u * T);
        const float *h = host + ((size_t)t * B + ids[u]) * C;
        float *d = dev + r * C;
        if (lane < C) {
            if (to_host) const_cast<float *>(h)[lane] = d[lane];
            else d[lane] = __ldg(h + lane);
        }
    }
}

int main() {
    const int T = 1500, B = 64, C = 29, NCH = 4;
    const size_t n = (size_t)T * B * C, bytes = n * 4;
    float *h_in, *h_out, *d_a, *d_b;
    CK(cudaHostAlloc(&h_in, bytes, cudaHostAllocDefault));
    CK(cudaHostAlloc(&h_out, bytes, cudaHostAllocDefault));
    CK(cudaMalloc(&d_a, bytes));
    CK(cudaMalloc(&d_b, bytes));
    for (size_t i = 0; i < n; ++i) h_in[i] = (float)(i & 1023);
    cudaStream_t s0, s1;
    CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    auto timeit = [&](const char *name, auto fn) {
        float best = 1e9f;
        for (int it = 0; it < 8; ++it) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0, s0));
            fn();
            CK(cudaEventRecord(e1, s0));
            CK(cudaEventSynchronize(e1));
            CK(cudaDeviceSynchronize());
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (it >= 2 && ms < best) best = ms;
        }
        printf("%-46s %8.1f us  %6.1f GB/s\n", name, best * 1e3, bytes / (best * 1e-3) / 1e9);
    };
    const size_t pitch = (size_t)B * C * 4, width = pitch / NCH;
    timeit("H2D contiguous", [&] { CK(cudaMemcpyAsync(d_a, h_in, bytes, cudaMemcpyHostToDevice, s0)); });
    timeit("D2H contiguous", [&] { CK(cudaMemcpyAsync(h_out, d_a, bytes, cudaMemcpyDeviceToHost, s0)); });
    timeit("H2D 4 x 2-D slices (1856 B rows)", [&] {
        for (int k = 0; k < NCH; ++k)
            CK(cudaMemcpy2DAsync((char *)d_a + k * width * T, width, (char *)h_in + k * width, pitch, width, T,
                                 cudaMemcpyHostToDevice, s0));
    });
    timeit("D2H 4 x 2-D slices (1856 B rows)", [&] {
        for (int k = 0; k < NCH; ++k)
            CK(cudaMemcpy2DAsync((char *)h_out + k * width, pitch, (char *)d_a + k * width * T, width, width, T,
                                 cudaMemcpyDeviceToHost, s0));
    });
    timeit("H2D 4 x contiguous time chunks", [&] {
        for (int k = 0; k < NCH; ++k)
            CK(cudaMemcpyAsync((char *)d_a + k * (bytes / NCH), (char *)h_in + k * (bytes / NCH), bytes / NCH,
                               cudaMemcpyHostToDevice, s0));
    });
    for (int grid : {148, 592, 2368}) {
        char nm[96];
        snprintf(nm, sizeof nm, "kernel reads pinned host -> HBM, grid %d x 256", grid);
        timeit(nm, [&] { copy_k<<<grid, 256, 0, s0>>>((const float4 *)h_in, (float4 *)d_b, n / 4); });
        snprintf(nm, sizeof nm, "kernel writes HBM -> pinned host, grid %d x 256", grid);
        timeit(nm, [&] { copy_k<<<grid, 256, 0, s0>>>((const float4 *)d_a, (float4 *)h_out, n / 4); });
    }
    // duplex: DMA both ways at once, and kernel both ways at once
    cudaEvent_t j;
    CK(cudaEventCreateWithFlags(&j, cudaEventDisableTiming));
    timeit("duplex DMA: H2D (s0) + D2H (s1)", [&] {
        CK(cudaMemcpyAsync(h_out, d_b, bytes, cudaMemcpyDeviceToHost, s1));
        CK(cudaMemcpyAsync(d_a, h_in, bytes, cudaMemcpyHostToDevice, s0));
        CK(cudaEventRecord(j, s1));
        CK(cudaStreamWaitEvent(s0, j, 0));
    });
    timeit("duplex kernels: read host (s0) + write host (s1)", [&] {
        copy_k<<<592, 256, 0, s1>>>((const float4 *)d_b, (float4 *)h_out, n / 4);
        copy_k<<<592, 256, 0, s0>>>((const float4 *)h_in, (float4 *)d_a, n / 4);
        CK(cudaEventRecord(j, s1));
        CK(cudaStreamWaitEvent(s0, j, 0));
    });
    {
        // 16 of the 64 utterances (every 4th), all T frames: 16 * 1500 * 116 B = 2.78 MB per direction
        const int nb = 16;
        int hid[nb], *did;
        for (int i = 0; i < nb; ++i) hid[i] = 4 * i + 1;
        CK(cudaMalloc(&did, sizeof(hid)));
        CK(cudaMemcpy(did, hid, sizeof(hid), cudaMemcpyHostToDevice));
        const double gb = (double)nb * T * C * 4;
        for (int grid : {148, 592, 1184}) {
            float best[2] = {1e9f, 1e9f};
            for (int dir = 0; dir < 2; ++dir)
                for (int it = 0; it < 6; ++it) {
                    CK(cudaDeviceSynchronize());
                    CK(cudaEventRecord(e0, s0));
                    gather_rows_k<<<grid, 256, 0, s0>>>(dir ? h_out : h_in, d_b, did, nb, T, B, C, dir);
                    CK(cudaEventRecord(e1, s0));
                    CK(cudaEventSynchronize(e1));
                    float ms;
                    CK(cudaEventElapsedTime(&ms, e0, e1));
                    if (it >= 2 && ms < best[dir]) best[dir] = ms;
                }
            printf("kernel row gather host->HBM (116 B rows), grid %4d  %8.1f us %6.1f GB/s | scatter HBM->host %8.1f us %6.1f GB/s\n",
                   grid, best[0] * 1e3, gb / (best[0] * 1e-3) / 1e9, best[1] * 1e3, gb / (best[1] * 1e-3) / 1e9);
        }
        // the same 16 utterances as 16 DMA 2-D copies of 116-byte rows
        float best = 1e9f;
        for (int it = 0; it < 5; ++it) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0, s0));
            for (int i = 0; i < nb; ++i)
                CK(cudaMemcpy2DAsync((char *)d_b + (size_t)i * T * C * 4, C * 4, (char *)h_in + (size_t)hid[i] * C * 4, pitch,
                                     C * 4, T, cudaMemcpyHostToDevice, s0));
            CK(cudaEventRecord(e1, s0));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (it >= 1 && ms < best) best = ms;
        }
        printf("DMA 16 x 2-D copies of 116 B rows              %8.1f us %6.1f GB/s\n", best * 1e3, gb / (best * 1e-3) / 1e9);
    }
    size_t bad = 0;
    for (size_t i = 0; i < n; ++i) bad += (h_out[i] != h_in[i]);
    printf("mismatches %zu\n", bad);
    return 0;
}
