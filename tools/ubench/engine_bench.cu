// tools/ubench/engine_bench.cu -- standalone timing harness for the three kernels (no torch): synthetic headline-like
// batch, CUDA-event time per kernel, optional in-kernel clock breakdown (-DCTC_PROFILE).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo [-DCTC_PROFILE] -o engine_bench engine_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <random>
#include <algorithm>
#include "../../deepspeech.pytorch_b200/csrc/ctc_kernels.cuh"
using namespace ctcb200;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
#ifndef KK
#define KK 8
#endif
constexpr int kK = KK;
int main(int argc, char **argv) {
    int B = argc > 1 ? atoi(argv[1]) : 64, T = argc > 2 ? atoi(argv[2]) : 1500, C = argc > 3 ? atoi(argv[3]) : 29;
    int R = 8;
    size_t eng_pad = argc > 4 ? (size_t)atoi(argv[4]) * 1024 : 0;  // force engine smem >= this (occupancy experiment)
    double rlo = 0.15, rhi = 0.30;
    int Lcap = 450;
    std::mt19937 rng(1234);
    std::uniform_real_distribution<double> U(0, 1);
    std::vector<int> al(B), ll(B), lab;
    for (int b = 0; b < B; ++b) { al[b] = std::max(1, (int)std::lround(T * (0.35 + 0.65 * U(rng)))); }
    al[0] = T;
    int Lmax = 0;
    for (int b = 0; b < B; ++b) { ll[b] = std::min(Lcap, std::max(1, (int)std::lround(al[b] * (rlo + (rhi - rlo) * U(rng))))); Lmax = std::max(Lmax, ll[b]); }
    for (int b = 0; b < B; ++b) for (int i = 0; i < ll[b]; ++i) lab.push_back(1 + rng() % (C - 1));
    std::vector<float> acts((size_t)T * B * C);
    std::normal_distribution<float> N(0, 1);
    for (auto &v : acts) v = N(rng);
    CtcParams p{};
    p.T = T; p.B = B; p.C = C; p.Cp = 32; p.blank = 0; p.bpos = 0; p.big = 0; p.zero_infinity = 1; p.grad_scale = 1.f; p.Lmax = Lmax;
    p.nwl = (2 * Lmax + 1 + 32 * kK - 1) / (32 * kK); p.R = R; p.nblk = (T + R - 1) / R; p.score_only = 0;
    const int NT = p.nwl * 32;
    float *d_acts, *d_grads, *d_costs; int *d_lab, *d_ll, *d_al;
    CK(cudaMalloc(&d_acts, acts.size() * 4)); CK(cudaMalloc(&d_grads, acts.size() * 4)); CK(cudaMalloc(&d_costs, B * 4));
    CK(cudaMalloc(&d_lab, lab.size() * 4)); CK(cudaMalloc(&d_ll, B * 4)); CK(cudaMalloc(&d_al, B * 4));
    CK(cudaMemcpy(d_acts, acts.data(), acts.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_lab, lab.data(), lab.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_ll, ll.data(), B * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_al, al.data(), B * 4, cudaMemcpyHostToDevice));
    p.acts = d_acts; p.grads = d_grads; p.costs = d_costs; p.labels = d_lab; p.label_lens = d_ll; p.act_lens = d_al;
    CK(cudaMalloc(&p.y, sizeof(unsigned) * (size_t)T * B * 2 * p.Cp));
    size_t ck = sizeof(int) * (size_t)B * p.nblk * NT * kK, ckE = sizeof(int) * (size_t)B * p.nblk * NT;
    CK(cudaMalloc(&p.ckA, ck)); CK(cudaMalloc(&p.ckB, ck)); CK(cudaMalloc(&p.ckAE, ckE)); CK(cudaMalloc(&p.ckBE, ckE));
    CK(cudaMalloc(&p.Pm, 8 * B)); CK(cudaMalloc(&p.Pe, 4 * B)); CK(cudaMalloc(&p.feas, 4 * B)); CK(cudaMalloc(&p.laboff, 4 * B));
    CK(cudaMalloc(&p.tlab, 16 * (size_t)B * NT));
#ifdef CTC_PROFILE
    CK(cudaMalloc(&p.dbg, 8 * 2 * B * 16 * 8 + 8192)); CK(cudaMemset(p.dbg, 0, 8 * 2 * B * 16 * 8 + 8192));
#endif
    p.nqmax = (2 * Lmax + 1 + 223) / 224;
    p.gq_lo = 1; p.gLmax = Lmax;
    size_t s_prep = prep_smem(C, Lmax), s_eng = engine_smem(p.Cp, p.nwl).total, s_grad = grad_smem(C, p.Cp, Lmax, p.nqmax, kK, false).total;
    { int resident = (int)(227 * 1024 / s_grad); if (resident < 1) resident = 1; p.nsplit = (2 * 148 * resident + B - 1) / B; if (p.nsplit > p.nblk) p.nsplit = p.nblk; if (argc > 5) p.nsplit = atoi(argv[5]); }
    CK(cudaFuncSetAttribute(ctc_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    CK(cudaFuncSetAttribute(ctc_engine_kernel<kK, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    CK(cudaFuncSetAttribute(ctc_engine2_kernel<kK, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    const int engv = getenv("ENGINE") ? atoi(getenv("ENGINE")) : 2;
    printf("engine form %d\n", engv);
#if KK == 8
    CK(cudaFuncSetAttribute(ctc_grad_kernel<kK, 32, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
#endif
    if (s_eng < eng_pad) s_eng = eng_pad;
    printf("nqmax %d nsplit %d\n", p.nqmax, p.nsplit);
    printf("B=%d T=%d C=%d Lmax=%d nwl=%d R=%d smem prep %zu engine %zu grad %zu\n", B, T, C, Lmax, p.nwl, R, s_prep, s_eng, s_grad);
    cudaEvent_t e[4];
    for (auto &x : e) cudaEventCreate(&x);
    const int nrb = (int)(((long long)T * B + kPrepRowsCta - 1) / kPrepRowsCta);
    float best[3] = {1e9, 1e9, 1e9};
    for (int it = 0; it < 6; ++it) {
        cudaEventRecord(e[0]);
        ctc_prep_kernel<<<nrb + B, kPrepRows, s_prep>>>(p);
        cudaEventRecord(e[1]);
        if (engv == 1) ctc_engine_kernel<kK, 32><<<2 * B, (p.nwl + 1) * 32, s_eng>>>(p);
        else ctc_engine2_kernel<kK, 32><<<2 * B, (p.nwl + 1) * 32, s_eng>>>(p);
        cudaEventRecord(e[2]);
        if (getenv("SCORE_ONLY")) { CtcParams q = p; q.score_only = 1; cudaEventRecord(e[1]); ctc_engine_kernel<kK, 32><<<B, (p.nwl + 1) * 32, s_eng>>>(q); cudaEventRecord(e[2]); }
#if KK == 8
        if (getenv("CLASSES") && p.nqmax >= 2) {  // small warp class on its own stream, full class on the default one
            static cudaStream_t s2 = nullptr; static cudaEvent_t ev, ev2;
            if (!s2) { cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking); cudaEventCreate(&ev); cudaEventCreate(&ev2); }
            CtcParams ps = p, pf = p;
            ps.nqmax = p.nqmax / 2; ps.gLmax = grad_class_lmax(ps.nqmax, Lmax); pf.gq_lo = ps.nqmax + 1;
            const size_t sm = grad_smem(C, p.Cp, ps.gLmax, ps.nqmax, kK, false).total;
            int res = (int)(227 * 1024 / sm); if (res > 8) res = 8;
            ps.nsplit = std::min(p.nblk / 2, (7 * 148 * res + B - 1) / B);
            if (getenv("NSPLIT_SMALL")) ps.nsplit = atoi(getenv("NSPLIT_SMALL"));
            cudaEventRecord(ev, 0); cudaStreamWaitEvent(s2, ev, 0);
            ctc_grad_kernel<kK, 32, false><<<dim3(ps.nsplit, B), ps.nqmax * 32, sm, s2>>>(ps);
            ctc_grad_kernel<kK, 32, false><<<dim3(pf.nsplit, B), pf.nqmax * 32, s_grad>>>(pf);
            cudaEventRecord(ev2, s2); cudaStreamWaitEvent(0, ev2, 0);
        } else
        ctc_grad_kernel<kK, 32, false><<<dim3(p.nsplit, B), p.nqmax * 32, s_grad>>>(p);
#endif
        cudaEventRecord(e[3]);
        CK(cudaDeviceSynchronize());
        for (int k = 0; k < 3; ++k) { float ms; cudaEventElapsedTime(&ms, e[k], e[k + 1]); best[k] = std::min(best[k], ms); }
    }
    printf("prep %.1f us  engine %.1f us  grad %.1f us   (best of 6)\n", best[0] * 1e3, best[1] * 1e3, best[2] * 1e3);
    std::vector<float> costs(B);
    CK(cudaMemcpy(costs.data(), d_costs, B * 4, cudaMemcpyDeviceToHost));
    double s = 0; for (float c : costs) s += c;
    printf("sum costs %.4f  cost[0] %.4f (T=%d L=%d)\n", s, costs[0], al[0], ll[0]);
#ifdef CTC_PROFILE
    std::vector<long long> dbg(2 * B * 16 * 8 + 1024);
    CK(cudaMemcpy(dbg.data(), p.dbg, dbg.size() * 8, cudaMemcpyDeviceToHost));
    for (int cta = 0; cta < 2; ++cta)
        for (int w = 0; w < p.nwl + 1; ++w) {
            long long *d = &dbg[(cta * 16 + w) * 8];
            printf("cta %d warp %d: total %lld  work %lld  barrier %lld  groups %lld  loop %lld steps %lld renorm %lld settle %lld\n", cta, w, d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7]);
        }
#endif
#ifdef CTC_PROFILE
    for (int w = 0; w < p.nqmax; ++w) {
        long long *d = &dbg[(32 + w) * 8];
        printf("grad cta(0,0) warp %d: total %lld setup %lld ywait %lld alpha %lld beta %lld reduce %lld items %lld bar1 %lld bar2 %lld\n", w, d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7] / 1000000, d[7] % 1000000);
    }
#endif
    return 0;
}
