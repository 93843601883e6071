// csrc/ctc_device.cuh -- device-side building blocks of the B200 CTC lattice kernels (sm_100a only).
//
// Numeric representation (validated on the CPU by tests/numeric_model.py against oracle/):
//   alpha / beta live in the LINEAR domain.  Each lattice thread owns K adjacent states of the blank-extended
//   label sequence l'  as fp64 mantissas plus ONE int exponent E: value = m * 2^E.
//   Every kRenorm time steps a thread rescales its mantissas by an exact power of two so that its largest one
//   lies in [1,2).  Inflow from the neighbouring thread is multiplied by 2^(E_nbr - E_me).
//   B200 issues fp64 at half the fp32 rate (64 lanes/SM/clk), so a lattice cell costs 2-3 DADD/DFMA/DMUL and
//   no transcendental; fp32 state cannot meet "grad abs <= 1e-5" (SURVEY.md 7.3).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ctcb200 {

constexpr int kRenorm = 8;      // time steps between renormalisations
constexpr int kLift = 500;      // max exponent gap tolerated towards the inflow neighbour
// R (template parameter of both kernels): time steps per checkpoint block == emission rows per smem tile;
// a multiple of kRenorm so that checkpoints fall on freshly renormalised states.
constexpr int kYStages = 3;     // emission tile ring depth (engine kernel)
constexpr int kRawStages = 3;   // raw-logit tile ring depth (engine kernel)
constexpr int kMaxSmallC = 128; // alphabet bound of the class-gather path
constexpr int kMaxLatticeWarps = 8;
constexpr unsigned kFull = 0xffffffffu;


// ---------------------------------------------------------------- exponent helpers
// 2^d as a double for d in [-1022, 1023]; smaller d gives 0 (flush), larger d clamps.
__device__ __forceinline__ double pow2i(int d) {
    int dd = min(d, 1023);
    int hi = (dd + 1023) << 20;
    hi = (dd < -1022) ? 0 : hi;
    return __hiloint2double(hi, 0);
}
// biased exponent field of a non-negative double (0 => zero/denormal)
__device__ __forceinline__ int expo_field(double x) { return __double2hiint(x) >> 20; }

// ---------------------------------------------------------------- barriers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------- async copies
__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
// TMA: 2-D tiled tensor load global -> shared, completion on an mbarrier (SASS: UTMALDG)
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const void *tensor_map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(smem_dst)), "l"(tensor_map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}

// ---------------------------------------------------------------- softmax of emission rows (one warp)
// raw: shared fp32 [nrows][raw_stride] logits; y: shared fp64 [nrows][y_stride] <- softmax probabilities.
// 8 lanes per row, 4 rows per pass: float loads are conflict-free when raw_stride % 32 == 8.
// y is clamped to >= 2^-126 so that a feasible alignment never gets probability exactly 0 from fp32 underflow.
__device__ __forceinline__ void softmax_rows_warp(const float *raw, int raw_stride, double *y, int y_stride, int nrows,
                                                  int C, int lane) {
    const int sub = lane & 7, rloc = lane >> 3;
    for (int r0 = 0; r0 < nrows; r0 += 4) {
        const int r = min(r0 + rloc, nrows - 1);  // duplicate the last row instead of diverging
        const float *x = raw + r * raw_stride;
        double *yr = y + r * y_stride;
        float m = -INFINITY;
        for (int c = sub; c < C; c += 8) m = fmaxf(m, x[c]);
        m = fmaxf(m, __shfl_xor_sync(kFull, m, 1));
        m = fmaxf(m, __shfl_xor_sync(kFull, m, 2));
        m = fmaxf(m, __shfl_xor_sync(kFull, m, 4));
        float sum = 0.f;
        for (int c = sub; c < C; c += 8) sum += expf(x[c] - m);
        sum += __shfl_xor_sync(kFull, sum, 1);
        sum += __shfl_xor_sync(kFull, sum, 2);
        sum += __shfl_xor_sync(kFull, sum, 4);
        const float inv = 1.0f / sum;
        if (r0 + rloc < nrows)
            for (int c = sub; c < C; c += 8) yr[c] = (double)fmaxf(expf(x[c] - m) * inv, 1.17549435e-38f);
    }
}

// ---------------------------------------------------------------- renormalisation
// Rescale the K mantissas of this thread and settle its exponent.  FWD: inflow comes from lane-1 (alpha),
// otherwise from lane+1 (beta).  carryE/has_carry: exponent of the neighbouring WARP's boundary thread.
//   live  thread : largest mantissa -> [1,2)
//   dead  thread : (all zero/denormal) flush, adopt the exponent of the nearest live thread on the inflow side
//   lift         : never let the inflow neighbour sit more than 2^kLift above (prefix-max scan), so that the
//                  boundary factor 2^(E_nbr-E_me) cannot overflow between two renormalisations
// Afterwards f = 2^(E_nbr - E_me) for the in-warp neighbour (the cross-warp lane recomputes it every step).
template <int K, bool FWD>
__device__ __forceinline__ void renorm(double (&a)[K], int &E, double &f, int lane, int carryE, bool has_carry) {
    int hm = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) hm = max(hm, __double2hiint(a[k]));
    const int ex = hm >> 20;
    const bool live = ex != 0;
    const int shift = live ? 1023 - ex : 0;
    const int E1 = E - shift;
    const unsigned lm = __ballot_sync(kFull, live);
    const unsigned side = FWD ? (lm & ((1u << lane) - 1u)) : (lane == 31 ? 0u : (lm & ~((2u << lane) - 1u)));
    const int src = side ? (FWD ? 31 - __clz(side) : __ffs(side) - 1) : lane;
    const int Es = __shfl_sync(kFull, E1, src);
    const bool sourced = live || side != 0u || has_carry;
    const int E2 = live ? E1 : (side ? Es : (has_carry ? carryE : E1));
    int E3 = sourced ? E2 : -(1 << 28);  // threads with nothing flowing in must not lift anybody
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = FWD ? __shfl_up_sync(kFull, E3, d) : __shfl_down_sync(kFull, E3, d);
        const bool ok = FWD ? (lane >= d) : (lane + d < 32);
        if (ok) E3 = max(E3, o - d * kLift);
    }
    if (has_carry) E3 = max(E3, carryE - (FWD ? lane + 1 : 32 - lane) * kLift);
    const int Enew = sourced ? E3 : E2;
    const double sc = live ? pow2i(shift - (Enew - E2)) : 0.0;
#pragma unroll
    for (int k = 0; k < K; ++k) a[k] *= sc;
    E = Enew;
    const int En = FWD ? __shfl_up_sync(kFull, Enew, 1) : __shfl_down_sync(kFull, Enew, 1);
    const bool edge = FWD ? (lane == 0) : (lane == 31);
    f = edge ? 1.0 : pow2i(En - Enew);
}

// ---------------------------------------------------------------- lattice cells
// alpha_t(s) = y_t(l'_s) * (alpha_{t-1}(s) + alpha_{t-1}(s-1) + [skip] alpha_{t-1}(s-2))      
// Thread base state is even, so k even = blank, k odd = label.  p1 = alpha_{t-1}(base-1) in this thread's scale.
template <int K>
__device__ __forceinline__ void alpha_cells(double (&a)[K], double p1, const double *__restrict__ yrow, int blank_idx,
                                            const int (&idx)[K / 2], const double (&skip)[K / 2]) {
    const double yb = yrow[blank_idx];
#pragma unroll
    for (int k = K - 1; k >= 0; --k) {
        const double prev1 = (k >= 1) ? a[k - 1] : p1;
        if (k & 1) {
            const double prev2 = (k >= 2) ? a[k - 2] : p1;
            a[k] = fma(skip[k >> 1], prev2, a[k] + prev1) * yrow[idx[k >> 1]];
        } else {
            a[k] = (a[k] + prev1) * yb;
        }
    }
}

// betabar_t(s) = beta_{t+1}(s) + beta_{t+1}(s+1) + [skip] beta_{t+1}(s+2);  beta_t(s) = y_t(l'_s) * betabar_t(s)
// n1/n2 = beta_{t+1}(base+K), beta_{t+1}(base+K+1) in this thread's scale.  bb receives betabar_t.
template <int K>
__device__ __forceinline__ void beta_cells(double (&a)[K], double (&bb)[K], double n1, double n2,
                                           const double *__restrict__ yrow, int blank_idx, const int (&idx)[K / 2],
                                           const double (&skip)[K / 2]) {
    const double yb = yrow[blank_idx];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const double nx1 = (k + 1 < K) ? a[k + 1] : n1;
        if (k & 1) {
            const double nx2 = (k + 2 < K) ? a[k + 2] : n2;
            bb[k] = fma(skip[k >> 1], nx2, a[k] + nx1);
            a[k] = bb[k] * yrow[idx[k >> 1]];
        } else {
            bb[k] = a[k] + nx1;
            a[k] = bb[k] * yb;
        }
    }
}

// Per-thread constants of the blank-extended sequence: emission index and skip masks of the K/2 label states.
// slab: shared int [L] labels of this utterance.  Invalid (padding) label states point at `pad_idx`.
template <int K>
__device__ __forceinline__ void thread_labels(const int *slab, int L, int thread, int C, int blank,
                                              int (&idx)[K / 2], double (&skipA)[K / 2], double (&skipB)[K / 2],
                                              int (&lab_i)[K / 2]) {
#pragma unroll
    for (int h = 0; h < K / 2; ++h) {
        const int i = (thread * K) / 2 + h;  // label position of state thread*K + 2h+1
        const bool valid = i < L;
        int c = valid ? slab[i] : blank;
        c = min(max(c, 0), C - 1);
        idx[h] = c;
        lab_i[h] = valid ? i : -1;
        skipA[h] = (valid && (i == 0 || min(max(slab[i - 1], 0), C - 1) != c)) ? 1.0 : 0.0;
        skipB[h] = (valid && i + 1 < L && min(max(slab[i + 1], 0), C - 1) != c) ? 1.0 : 0.0;
    }
}

}  // namespace ctcb200
