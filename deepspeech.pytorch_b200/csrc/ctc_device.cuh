// csrc/ctc_device.cuh -- device-side building blocks of the B200 CTC lattice kernels (sm_100a only).
//
// Numeric representation (modelled on the CPU by tests/numeric_model.py):
//   alpha / beta live in the LINEAR domain.  Each lattice thread owns K adjacent states of the blank-extended
//   label sequence l'  as fp64 mantissas plus ONE int exponent E: value = m * 2^E.
//   * every kGroup (= 8) time steps a thread rescales its own mantissas by an exact power of two so that the
//     largest lies in [1,2) (renorm, thread-local); a thread whose mantissas are all zero is DEAD (E = kDeadE).
//     Emissions are clamped to >= 2^-126, so nothing can underflow within one group.
//   * right after the renorm the warp SETTLES its exponents (settle): dead threads adopt the exponent of the nearest
//     live thread on their inflow side, threads sitting more than 2^kLift below their inflow neighbour re-base onto
//     it.  Exponents are then fixed for the whole next group, so a step moves mantissas only and scales the value
//     flowing in from the neighbouring thread by the per-group constant f = 2^(E_nbr - E_me).
//   Measured on B200 (tools/ubench/lat*.cu): DFMA latency 8 clk, 16 lanes/clk/SMSP; SHFL 27 clk; bar.sync (4 warps)
//   39-52 clk; an E-shuffle plus rescale in every step costs +32 clk/step, which is why exponents are per group.
//   A K=8 lattice step is ~21 fp64 + ~30 other instructions, ~90-100 clk for the one warp an SMSP hosts, so
//   neighbour values are exchanged one step AHEAD of their use and the CTA synchronises once per group.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ctcb200 {

constexpr int kGroup = 8;            // time steps per group == renormalisation period == emission rows per tile
constexpr int kLift = 250;           // max exponent gap tolerated towards the inflow neighbour
constexpr int kDeadE = -(1 << 30);   // exponent of a thread that holds only zeros
constexpr int kMaxSmallC = 128;      // alphabet bound of the class-row (gather) emission path
constexpr int kMaxLatticeWarps = 8;  // K=8: label length <= 1023
constexpr int kRing = 16;            // boundary-slot ring between adjacent lattice warps (2 groups)
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

// High word of a non-negative double, rounded to nearest: the full exponent range with 21 significant bits.  Used for
// values that only seed / enter products of the gradient pass (parked rows, checkpoints): relative error <= 2^-22.
__device__ __forceinline__ int hi_round(double x) { return __double2hiint(x) + (int)((unsigned)__double2loint(x) >> 31); }

// ---------------------------------------------------------------- barriers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
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

// ---------------------------------------------------------------- TMA bulk copy (SASS: UBLKCP)
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completion counted on an mbarrier.
__device__ __forceinline__ void tma_bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ---------------------------------------------------------------- lattice state of one thread
template <int K>
struct Cells {
    double a[K];
    int E;
};

// Thread-local renormalisation (end of every group): largest mantissa -> [1,2); all-zero thread -> dead.
template <int K>
__device__ __forceinline__ void renorm(Cells<K> &c) {
    int h[K];
#pragma unroll
    for (int k = 0; k < K; ++k) h[k] = __double2hiint(c.a[k]);
#pragma unroll
    for (int w = 1; w < K; w <<= 1)
#pragma unroll
        for (int k = 0; k + w < K; k += 2 * w) h[k] = max(h[k], h[k + w]);
    const int ex = h[0] >> 20;
    const bool live = ex != 0;
    const double sc = live ? pow2i(1023 - ex) : 0.0;
#pragma unroll
    for (int k = 0; k < K; ++k) c.a[k] *= sc;
    c.E = live ? c.E - (1023 - ex) : kDeadE;
}

// Exponent settlement between two groups (after renorm), one warp, all lanes.  Exponents are then FIXED for the
// whole next group, so the per-step path carries mantissas only and scales the inflow by a per-group constant f.
//   * a dead thread adopts the exponent of the nearest live thread on its inflow side (ballot + one indexed
//     shuffle; beyond the warp: E_edge, the settled exponent of the neighbouring warp's boundary thread, which ran
//     this group one super-step earlier) -- so whatever flows into it during the group arrives in a valid scale;
//   * a thread whose inflow neighbour sits more than 2^kLift above it re-bases onto the neighbour's exponent (its
//     own content is negligible next to what is about to flow in); repeated while any lane re-based (rare).
// UP = inflow comes from lane-1 (alpha); otherwise from lane+1 (beta).  Returns f = 2^(E_nbr - E_me).
template <int K, bool UP>
__device__ __forceinline__ double settle(Cells<K> &c, int E_edge, int lane) {
    int E = c.E;
    const bool live = E != kDeadE;
    const unsigned lm = __ballot_sync(kFull, live);
    const unsigned side = UP ? (lm & ((1u << lane) - 1u)) : (lane == 31 ? 0u : (lm & ~((2u << lane) - 1u)));
    const int src = side ? (UP ? 31 - __clz(side) : __ffs(side) - 1) : lane;
    const int Es = __shfl_sync(kFull, E, src);
    if (!live) E = side ? Es : E_edge;
    int En;
    bool lifted;
    do {
        En = UP ? __shfl_up_sync(kFull, E, 1) : __shfl_down_sync(kFull, E, 1);
        if (UP ? (lane == 0) : (lane == 31)) En = E_edge;
        lifted = false;
#ifndef CTC_EXP_NOLIFT
        if (En != kDeadE && En - E > kLift) {  // E == kDeadE with a live neighbour cannot happen after adoption
            const double s = pow2i(E - En);
#pragma unroll
            for (int k = 0; k < K; ++k) c.a[k] *= s;
            E = En;
            lifted = true;
        }
#endif
    } while (__any_sync(kFull, lifted));
    c.E = E;
    return (En == kDeadE) ? 0.0 : pow2i(En - E);
}

// ---------------------------------------------------------------- lattice cells
// Emission values one thread needs for one time step, loaded from a shared-memory row one step AHEAD of use.
template <int K>
struct YRow {
    double yb;         // y_t(blank)
    double yl[K / 2];  // y_t(label of state 2h+1)
};
// Emission rows are stored PLANAR: CP high words then CP low words of the fp64 values (CP = 32 for C <= 32), so a
// per-lane gather is two bank-conflict-free single-wavefront LDS.32 that land directly in a register pair.
// (Measured, tools/ubench/lat3.cu: interleaved fp64 rows need 2-4 wavefronts per LDS.64 gather and make the step
// LSU-bound, 78 vs 43 clk; fp32 rows + conversion cost either the FP64 pipe (cvt) or ~20 extra integer/move
// instructions per step.)  Row r of a tile starts at word r * 2 * CP: all offsets fold into immediates.
// CTC_Y_PLANES = 1: only the high-word plane exists -- the emission probabilities rounded to 21 significant bits
// (relative 2^-22): half the emission bytes, tile copies and gathers.  tools/y_rounding_study.py: on a T = 1500,
// L = 400 utterance the cost moves by 1.5e-8 relative and the per-class occupancies by <= 6.6e-7.
#ifndef CTC_Y_PLANES
#define CTC_Y_PLANES 1
#endif
constexpr int kYP = CTC_Y_PLANES;
template <int K, int CP>
__device__ __forceinline__ void load_y(YRow<K> &y, const unsigned *tile, int r, int blank_idx, const int (&idx)[K / 2]) {
#ifdef CTC_EXP_NOY
    y.yb = 0.9 + 1e-9 * blank_idx;
#pragma unroll
    for (int h = 0; h < K / 2; ++h) y.yl[h] = 0.9 + 1e-9 * idx[h];
    (void)tile; (void)r;
#else
    const unsigned *row = tile + r * kYP * CP;
    y.yb = __hiloint2double((int)row[blank_idx], kYP == 2 ? (int)row[CP + blank_idx] : 0);
#pragma unroll
    for (int h = 0; h < K / 2; ++h)
        y.yl[h] = __hiloint2double((int)row[idx[h]], kYP == 2 ? (int)row[CP + idx[h]] : 0);
#endif
}
// exact fp32 -> planar fp64 words (positive normal floats only: y is clamped to >= 2^-126) and back
// (one plane: the high word is rounded to nearest, ties to EVEN -- round-half-up is biased by 1/16 ulp, which adds up
// over the frames of a confidently predicted utterance: 600 frames at y ~ 0.995 put 1e-5 relative on a cost of 2.8;
// a carry into the exponent field is the right answer)
__device__ __forceinline__ unsigned y_hi_word(float x) {
    const unsigned b = __float_as_uint(x);
    return ((b + (kYP == 2 ? 0u : 3u + ((b >> 3) & 1u))) >> 3) + 0x38000000u;
}
__device__ __forceinline__ unsigned y_lo_word(float x) { return __float_as_uint(x) << 29; }
__device__ __forceinline__ float y_from_words(unsigned hi, unsigned lo) {
    return __uint_as_float(((hi - 0x38000000u) << 3) | (kYP == 2 ? (lo >> 29) : 0u));
}
// the emission probability of class cc in row r of a tile at pitch CP, as fp32
__device__ __forceinline__ float y_of_tile(const unsigned *tile, int r, int CP, int cc) {
    const unsigned *row = tile + r * kYP * CP;
    return y_from_words(row[cc], kYP == 2 ? row[CP + cc] : 0u);
}

// alpha_t(s) = y_t(l'_s) * (alpha_{t-1}(s) + alpha_{t-1}(s-1) + [skip] alpha_{t-1}(s-2))      
// Thread base state is even, so k even = blank, k odd = label.  The step is split so that the value the next
// thread needs (new a[K-1], which does not depend on this step's inflow when K >= 4) is produced first and can be
// shuffled while the remaining cells are computed; the inflow p1 = alpha_{t-1}(base-1) is consumed last.
template <int K>
__device__ __forceinline__ double alpha_top(const double (&a)[K], const YRow<K> &y, const double (&skip)[K / 2]) {
    static_assert(K >= 4, "K >= 4");
    return fma(skip[K / 2 - 1], a[K - 3], a[K - 1] + a[K - 2]) * y.yl[K / 2 - 1];
}
template <int K>
__device__ __forceinline__ void alpha_rest(double (&a)[K], double top, double p1, const YRow<K> &y,
                                           const double (&skip)[K / 2]) {
#pragma unroll
    for (int k = K - 2; k >= 0; --k) {
        const double prev1 = (k >= 1) ? a[k - 1] : p1;
        if (k & 1) {
            const double prev2 = (k >= 2) ? a[k - 2] : p1;
            a[k] = fma(skip[k >> 1], prev2, a[k] + prev1) * y.yl[k >> 1];
        } else {
            a[k] = (a[k] + prev1) * y.yb;
        }
    }
    a[K - 1] = top;
}

// betabar_t(s) = beta_{t+1}(s) + beta_{t+1}(s+1) + [skip] beta_{t+1}(s+2);  beta_t(s) = y_t(l'_s) * betabar_t(s)
// Bottom two cells first (what the previous thread needs); n1/n2 = beta_{t+1}(base+K), beta_{t+1}(base+K+1) in
// this thread's scale are consumed last.  bb receives betabar_t.
template <int K>
__device__ __forceinline__ void beta_bottom(const double (&a)[K], const YRow<K> &y, const double (&skip)[K / 2],
                                            double &b0, double &b1, double &n0, double &n1v) {
    static_assert(K >= 4, "K >= 4");
    b0 = a[0] + a[1];
    b1 = fma(skip[0], a[3], a[1] + a[2]);
    n0 = b0 * y.yb;
    n1v = b1 * y.yl[0];
}
template <int K>
__device__ __forceinline__ void beta_rest(double (&a)[K], double (&bb)[K], double b0, double b1, double n0, double n1v,
                                          double n1, double n2, const YRow<K> &y, const double (&skip)[K / 2]) {
#pragma unroll
    for (int k = 2; k < K; ++k) {
        const double nx1 = (k + 1 < K) ? a[k + 1] : n1;
        if (k & 1) {
            const double nx2 = (k + 2 < K) ? a[k + 2] : n2;
            bb[k] = fma(skip[k >> 1], nx2, a[k] + nx1);
            a[k] = bb[k] * y.yl[k >> 1];
        } else {
            bb[k] = a[k] + nx1;
            a[k] = bb[k] * y.yb;
        }
    }
    bb[0] = b0;
    bb[1] = b1;
    a[0] = n0;
    a[1] = n1v;
}

}  // namespace ctcb200
