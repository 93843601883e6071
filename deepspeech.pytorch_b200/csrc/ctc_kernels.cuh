// csrc/ctc_kernels.cuh -- the two kernels of the B200 CTC loss (sm_100a).
//
//   ctc_engine_kernel : grid 2B (score-only: B).  CTA 2b runs the alpha recursion of utterance b forwards in time,
//                       CTA 2b+1 the beta recursion backwards, concurrently on different SMs.  Lattice warps hold
//                       K states per thread in registers; a dedicated producer warp streams logit rows into a
//                       shared-memory ring (cp.async / TMA) and turns them into fp64 softmax rows, handing tiles
//                       over through mbarriers.  Every R steps the state is checkpointed to the workspace.
//                       The alpha CTA finishes with P = p(l|x) -> costs[b].
//   ctc_grad_kernel   : grid (ceil(T/R), B).  One CTA per (utterance, block of R time steps): recomputes
//                       the alpha rows of the block from its checkpoint into shared memory, sweeps beta backwards
//                       from the beta checkpoint, forms posteriors alpha*betabar/P, reduces them per class in
//                       shared memory and writes grad = scale * (softmax - occupancy) once, one coalesced row
//                       segment per (t,b).  Rows t >= act_lens[b] are written as zeros by the same kernel.
//
// Replaces the arithmetic behind `criterion(logits, targets, output_sizes.cpu(), target_sizes)`
// (/root/reference/train.py:600; eval :352,:457), i.e. warp-ctc's softmax + compute_alpha + compute_betas_and_grad
// stages [external], with a different algorithm: linear-domain fp64 block-exponent recursion, bidirectional
// engines, checkpoint + recompute instead of an S x T lattice in global memory (which at B=64,T=1500 would be
// 256 MB of traffic against 22 MB of algorithmic bytes).
#pragma once
#include "ctc_device.cuh"

namespace ctcb200 {

struct CtcParams {
    const float *acts;
    float *grads;
    float *costs;
    const int *labels;      // device, flat
    const int *label_lens;  // device [B]
    const int *act_lens;    // device [B]
    int T, B, C, blank, zero_infinity;
    float grad_scale;
    int nwl;   // lattice warps per CTA
    int nblk;  // ceil(T / R)
    int Lmax;  // max label length (sizes the shared label array)
    int score_only;
    // workspace
    double *ckA, *ckB;  // [B][nblk][nwl*32*K] mantissas
    int *ckAE, *ckBE;   // [B][nblk][nwl*32]   exponents
    double *Pm;         // [B] mantissa of p(l|x) in [1,2)
    int *Pe;            // [B] exponent of p(l|x)
    int *feas;          // [B] 1 = feasible
    int *laboff;        // [B] offset of utterance b in `labels`
    int *lpos;          // [B][Lmax] position of label i in class-sorted order
    int *coff;          // [B][C+1]  class segment offsets in that order
};

struct Slot {
    double m0, m1;
    int E;
    int pad;
};

__host__ __device__ inline int raw_stride_of(int C) { return ((C + 31) / 32) * 32 + 8; }
__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

// ------------------------------------------------------------------------------------------------ engine
struct EngineSmem {
    size_t ytile, raw, slab, cnt, slots, bars, misc, total;
};
__host__ __device__ inline EngineSmem engine_smem(int C, int Lmax, int R) {
    EngineSmem s;
    size_t o = 0;
    s.ytile = o; o = align16(o + sizeof(double) * kYStages * R * C);
    s.raw = o;   o = align16(o + sizeof(float) * kRawStages * R * raw_stride_of(C));
    s.slab = o;  o = align16(o + sizeof(int) * (Lmax > 0 ? Lmax : 1));
    s.cnt = o;   o = align16(o + sizeof(int) * C);
    s.slots = o; o = align16(o + sizeof(Slot) * 2 * kMaxLatticeWarps);
    s.bars = o;  o = align16(o + sizeof(uint64_t) * 2 * kYStages);
    s.misc = o;  o = align16(o + 64);
    s.total = o;
    return s;
}

// Class-sorted positions of the labels of one utterance (one warp): lpos[i] = rank of label i when labels are
// stably sorted by class; coff[c] = first position of class c.  Lets the grad kernel reduce posteriors per class
// over contiguous shared-memory segments, deterministically.
__device__ __forceinline__ void sort_labels_warp(const int *slab, int L, int C, int *cnt, int *lpos_g, int *coff_g,
                                                 int lane) {
    for (int c = lane; c < C; c += 32) cnt[c] = 0;
    __syncwarp();
    for (int i = lane; i < L; i += 32) atomicAdd(&cnt[min(max(slab[i], 0), C - 1)], 1);
    __syncwarp();
    if (lane == 0) {
        int run = 0;
        for (int c = 0; c < C; ++c) {
            const int n = cnt[c];
            cnt[c] = run;
            coff_g[c] = run;
            run += n;
        }
        coff_g[C] = run;
    }
    __syncwarp();
    for (int base = 0; base < L; base += 32) {
        const int i = base + lane;
        const bool act = i < L;
        const int c = act ? min(max(slab[i], 0), C - 1) : (-1 - lane);
        const unsigned m = __match_any_sync(kFull, c);
        const int rank = __popc(m & ((1u << lane) - 1u));
        const int start = act ? cnt[c] : 0;
        __syncwarp();
        if (act) lpos_g[i] = start + rank;
        if (act && rank == 0) cnt[c] = start + __popc(m);
        __syncwarp();
    }
}

template <int K, int R>
__global__ void __launch_bounds__((kMaxLatticeWarps + 1) * 32) ctc_engine_kernel(const CtcParams p) {
    static_assert(R % kRenorm == 0 && K % 2 == 0, "bad tile shape");
    extern __shared__ __align__(16) unsigned char smem[];
    const int b = p.score_only ? blockIdx.x : (blockIdx.x >> 1);
    const bool fwd = p.score_only ? true : !(blockIdx.x & 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwl = p.nwl, C = p.C;
    const int NT = nwl * 32;
    const EngineSmem lay = engine_smem(C, p.Lmax, R);
    double *ytile = reinterpret_cast<double *>(smem + lay.ytile);
    float *raw = reinterpret_cast<float *>(smem + lay.raw);
    int *slab = reinterpret_cast<int *>(smem + lay.slab);
    int *cnt = reinterpret_cast<int *>(smem + lay.cnt);
    Slot *slots = reinterpret_cast<Slot *>(smem + lay.slots);
    uint64_t *y_full = reinterpret_cast<uint64_t *>(smem + lay.bars);
    uint64_t *y_empty = y_full + kYStages;
    double *pm = reinterpret_cast<double *>(smem + lay.misc);  // [2]
    int *pe = reinterpret_cast<int *>(smem + lay.misc + 16);   // [2]
    int *s_off = reinterpret_cast<int *>(smem + lay.misc + 32);

    const int Tb = min(max(p.act_lens[b], 0), p.T);
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int S = 2 * L + 1;
    const int nwa = min(nwl, (S + 32 * K - 1) / (32 * K));  // lattice warps this utterance needs
    const int RS = raw_stride_of(C);

    // ---- label offset (prefix sum of label_lens) and labels -> shared
    if (tid == 0) {
        *s_off = 0;
        for (int i = 0; i < kYStages; ++i) {
            mbar_init(&y_full[i], 1);
            mbar_init(&y_empty[i], nwa);
        }
        fence_mbar_init();
    }
    __syncthreads();
    {
        int part = 0;
        for (int i = tid; i < b; i += blockDim.x) part += max(p.label_lens[i], 0);
#pragma unroll
        for (int d = 16; d; d >>= 1) part += __shfl_xor_sync(kFull, part, d);
        if (lane == 0 && part) atomicAdd(s_off, part);
    }
    __syncthreads();
    const int off = *s_off;
    for (int i = tid; i < L; i += blockDim.x) slab[i] = p.labels[off + i];
    if (tid < 2 * kMaxLatticeWarps) {
        slots[tid].m0 = 0.0;
        slots[tid].m1 = 0.0;
        slots[tid].E = 0;
    }
    __syncthreads();

    if (Tb == 0) {  // nothing to recurse over
        if (fwd && tid == 0) {
            const bool ok = (L == 0);
            p.feas[b] = ok;
            p.laboff[b] = off;
            p.Pm[b] = 1.0;
            p.Pe[b] = 0;
            p.costs[b] = ok ? 0.f : (p.zero_infinity ? 0.f : INFINITY);
        }
        return;
    }

    const int ntiles = (Tb + R - 1) / R;

    if (warp == nwl) {
        // =========================================================== producer warp
        if (fwd) {
            if (!p.score_only)
                sort_labels_warp(slab, L, C, cnt, p.lpos + (size_t)b * p.Lmax, p.coff + (size_t)b * (C + 1), lane);
            if (lane == 0) p.laboff[b] = off;
        }
        const float *abase = p.acts + (size_t)b * C;
        const size_t rowstride = (size_t)p.B * C;
        auto issue = [&](int tile) {
            if (tile < ntiles) {
                float *dst = raw + (tile % kRawStages) * R * RS;
                const int n0 = tile * R;
                const int nr = min(R, Tb - n0);
                for (int r = 0; r < nr; ++r) {
                    const int t = fwd ? (n0 + r) : (Tb - 1 - n0 - r);
                    const float *src = abase + (size_t)t * rowstride;
                    for (int c = lane; c < C; c += 32) cp_async4(dst + r * RS + c, src + c);
                }
            }
            cp_async_commit();
        };
        issue(0);
        issue(1);
        for (int tile = 0; tile < ntiles; ++tile) {
            issue(tile + 2);
            cp_async_wait<2>();
            __syncwarp();
            const int ys = tile % kYStages;
            mbar_wait(&y_empty[ys], ((tile / kYStages) & 1) ^ 1);
            const int nr = min(R, Tb - tile * R);
            softmax_rows_warp(raw + (tile % kRawStages) * R * RS, RS, ytile + ys * R * C, C, nr, C,
                              lane);
            __syncwarp();
            if (lane == 0) mbar_arrive(&y_full[ys]);
        }
        cp_async_wait<0>();
        return;
    }
    if (warp >= nwa) return;  // lattice warps beyond this utterance's S

    // =============================================================== lattice warps
    const int thread = warp * 32 + lane;  // lattice thread index; owns states [thread*K, thread*K+K)
    int idx[K / 2], lab_i[K / 2];
    double skipA[K / 2], skipB[K / 2];
    thread_labels<K>(slab, L, thread, C, p.blank, idx, skipA, skipB, lab_i);
    (void)lab_i;

    double a[K];
#pragma unroll
    for (int k = 0; k < K; ++k) a[k] = 0.0;
    int E = 0;
    double f = 1.0;
    const int nbar = nwa * 32;
    double *ck = (fwd ? p.ckA : p.ckB) + (size_t)b * p.nblk * NT * K;
    int *ckE = (fwd ? p.ckAE : p.ckBE) + (size_t)b * p.nblk * NT;

    if (!fwd) {  // virtual beta_T: 1 at state S-1
        const int s1 = S - 1;
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (thread * K + k == s1) a[k] = 1.0;
        // publish the initial boundary state: step 0 of the warp below reads it (S-1 may sit alone in a warp)
        if (lane == 0 && warp > 0) {
            slots[warp].m0 = a[0];
            slots[warp].m1 = a[1];
            slots[warp].E = 0;
        }
        if (nwa > 1) named_bar_sync(1, nbar);
    }

    for (int n = 0; n < Tb; ++n) {
        const int r = n % R, tile = n / R, ys = tile % kYStages;
        if (r == 0) mbar_wait(&y_full[ys], (tile / kYStages) & 1);
        const double *yrow = ytile + (ys * R + r) * C;
        const Slot *rd = slots + (n & 1) * kMaxLatticeWarps;
        Slot *wr = slots + ((n + 1) & 1) * kMaxLatticeWarps;
        if (fwd) {
            double p1 = __shfl_up_sync(kFull, a[K - 1], 1) * f;
            int carryE = 0;
            if (lane == 0) {
                if (warp == 0) {
                    p1 = (n == 0) ? 1.0 : 0.0;
                } else {
                    const Slot s = rd[warp - 1];
                    carryE = s.E;
                    p1 = s.m0 * pow2i(s.E - E);
                }
            }
            alpha_cells<K>(a, p1, yrow, p.blank, idx, skipA);
            if (((n + 1) % kRenorm) == 0) {
                carryE = __shfl_sync(kFull, carryE, 0);
                renorm<K, true>(a, E, f, lane, carryE, warp > 0);
            }
            if (lane == 31 && warp + 1 < nwa) {
                wr[warp].m0 = a[K - 1];
                wr[warp].E = E;
            }
            const int t1 = n + 1;
            if ((t1 % R) == 0 && t1 < Tb && !p.score_only) {
                double *dst = ck + ((size_t)(t1 / R) * NT + thread) * K;
#pragma unroll
                for (int k = 0; k < K; k += 2) *reinterpret_cast<double2 *>(dst + k) = make_double2(a[k], a[k + 1]);
                ckE[(size_t)(t1 / R) * NT + thread] = E;
            }
        } else {
            const int t = Tb - 1 - n;
            double n1 = __shfl_down_sync(kFull, a[0], 1) * f;
            double n2 = __shfl_down_sync(kFull, a[1], 1) * f;
            int carryE = 0;
            const bool has_up = warp + 1 < nwa;
            if (lane == 31) {
                if (!has_up) {
                    n1 = 0.0;
                    n2 = 0.0;
                } else {
                    const Slot s = rd[warp + 1];
                    carryE = s.E;
                    const double fs = pow2i(s.E - E);
                    n1 = s.m0 * fs;
                    n2 = s.m1 * fs;
                }
            }
            double bb[K];
            beta_cells<K>(a, bb, n1, n2, yrow, p.blank, idx, skipB);
            if ((t % kRenorm) == 0) {
                carryE = __shfl_sync(kFull, carryE, 31);
                renorm<K, false>(a, E, f, lane, carryE, has_up);
            }
            if (lane == 0 && warp > 0) {
                wr[warp].m0 = a[0];
                wr[warp].m1 = a[1];
                wr[warp].E = E;
            }
            if ((t % R) == 0 && t > 0) {  // beta_t for the block that ends at row t-1
                const int kb = t / R - 1;
                double *dst = ck + ((size_t)kb * NT + thread) * K;
#pragma unroll
                for (int k = 0; k < K; k += 2) *reinterpret_cast<double2 *>(dst + k) = make_double2(a[k], a[k + 1]);
                ckE[(size_t)kb * NT + thread] = E;
            }
        }
        if (r == R - 1 || n == Tb - 1) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&y_empty[ys]);
        }
        if (nwa > 1) named_bar_sync(1, nbar);
    }

    if (fwd) {
        // ---- P = alpha_{T-1}(S-1) + alpha_{T-1}(S-2)
        const int s1 = S - 1, s2 = S - 2;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (thread * K + k == s1) { pm[0] = a[k]; pe[0] = E; }
            if (thread * K + k == s2) { pm[1] = a[k]; pe[1] = E; }
        }
        if (nwa > 1) named_bar_sync(1, nbar); else __syncwarp();
        if (thread == 0) {
            double m = pm[0];
            int e = pe[0];
            if (S > 1) {
                const int e2 = pe[1];
                const int em = max(e, e2);
                m = pm[0] * pow2i(e - em) + pm[1] * pow2i(e2 - em);
                e = em;
            }
            const int ex = expo_field(m);
            const bool ok = ex != 0 && ex != 2047;
            if (ok) {
                m *= pow2i(1023 - ex);
                e += ex - 1023;
            }
            p.feas[b] = ok;
            p.Pm[b] = ok ? m : 1.0;
            p.Pe[b] = ok ? e : 0;
            const double cost = -(log(m) + (double)e * 0.69314718055994530942);
            p.costs[b] = ok ? (float)cost : (p.zero_infinity ? 0.f : INFINITY);
        }
    }
}


// ------------------------------------------------------------------------------------------------ grad
struct GradSmem {
    size_t ytile, raw, slab, arow, Ea, psort, bpart, coff, slots, total;
};
__host__ __device__ inline GradSmem grad_smem(int C, int Lmax, int NT, int K, int R) {
    GradSmem s;
    size_t o = 0;
    s.ytile = o; o = align16(o + sizeof(double) * R * C);
    s.raw = o;   o = align16(o + sizeof(float) * R * raw_stride_of(C));
    s.slab = o;  o = align16(o + sizeof(int) * (Lmax > 0 ? Lmax : 1));
    s.arow = o;  o = align16(o + sizeof(double) * R * NT * K);
    s.Ea = o;    o = align16(o + sizeof(int) * R * NT);
    s.psort = o; o = align16(o + sizeof(float) * R * (Lmax > 0 ? Lmax : 1));
    s.bpart = o; o = align16(o + sizeof(float) * R * NT);
    s.coff = o;  o = align16(o + sizeof(int) * (C + 1));
    s.slots = o; o = align16(o + sizeof(Slot) * 2 * kMaxLatticeWarps);
    s.total = o;
    return s;
}

template <int K, int R>
__global__ void __launch_bounds__(kMaxLatticeWarps * 32) ctc_grad_kernel(const CtcParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int blk = blockIdx.x, b = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwl = p.nwl, C = p.C, NT = nwl * 32;
    const int Tb = min(max(p.act_lens[b], 0), p.T);
    const int t0 = blk * R;
    const int rows_out = min(R, p.T - t0);  // rows this CTA must write
    float *gbase = p.grads + ((size_t)t0 * p.B + b) * C;
    const size_t rowstride = (size_t)p.B * C;

    if (t0 >= Tb || !p.feas[b]) {  // padding frames / infeasible utterance: zeros, written once
        for (int r = warp; r < rows_out; r += nwl)
            for (int c = lane; c < C; c += 32) gbase[(size_t)r * rowstride + c] = 0.f;
        return;
    }
    const int nrow = min(R, Tb - t0);  // valid rows
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int S = 2 * L + 1;
    const int nwa = min(nwl, (S + 32 * K - 1) / (32 * K));
    const int RS = raw_stride_of(C);
    const GradSmem lay = grad_smem(C, p.Lmax, NT, K, R);
    double *ytile = reinterpret_cast<double *>(smem + lay.ytile);
    float *raw = reinterpret_cast<float *>(smem + lay.raw);
    int *slab = reinterpret_cast<int *>(smem + lay.slab);
    double *arow = reinterpret_cast<double *>(smem + lay.arow);
    int *Ea = reinterpret_cast<int *>(smem + lay.Ea);
    float *psort = reinterpret_cast<float *>(smem + lay.psort);
    float *bpart = reinterpret_cast<float *>(smem + lay.bpart);
    int *coff = reinterpret_cast<int *>(smem + lay.coff);
    Slot *slots = reinterpret_cast<Slot *>(smem + lay.slots);
    const int Lp = p.Lmax > 0 ? p.Lmax : 1;

    // ---- stage 1: logits rows (cp.async), labels, class offsets -> shared
    {
        const float *abase = p.acts + ((size_t)t0 * p.B + b) * C;
        for (int r = warp; r < nrow; r += nwl)
            for (int c = lane; c < C; c += 32) cp_async4(raw + r * RS + c, abase + (size_t)r * rowstride + c);
        cp_async_commit();
        const int off = p.laboff[b];
        for (int i = tid; i < L; i += blockDim.x) slab[i] = p.labels[off + i];
        const int *cg = p.coff + (size_t)b * (C + 1);
        for (int c = tid; c <= C; c += blockDim.x) coff[c] = cg[c];
        if (tid < 2 * kMaxLatticeWarps) {
            slots[tid].m0 = 0.0;
            slots[tid].m1 = 0.0;
            slots[tid].E = 0;
        }
        cp_async_wait<0>();
    }
    __syncthreads();
    // ---- stage 2: softmax rows -> fp64 emission tile
    for (int r0 = warp * 4; r0 < nrow; r0 += nwl * 4)
        softmax_rows_warp(raw + r0 * RS, RS, ytile + r0 * C, C, min(4, nrow - r0), C, lane);
    __syncthreads();

    const int thread = warp * 32 + lane;
    const int nbar = nwa * 32;
    if (warp < nwa) {
        int idx[K / 2], lab_i[K / 2], lpos[K / 2];
        double skipA[K / 2], skipB[K / 2];
        thread_labels<K>(slab, L, thread, C, p.blank, idx, skipA, skipB, lab_i);
        {
            const int *lg = p.lpos + (size_t)b * p.Lmax;
#pragma unroll
            for (int h = 0; h < K / 2; ++h) lpos[h] = lab_i[h] >= 0 ? lg[lab_i[h]] : -1;
        }
        double a[K];
        int E = 0;
        double f = 1.0;
        // ---- stage 3: recompute alpha rows t0 .. t0+nrow-1 from the checkpoint "after row t0-1"
        if (blk == 0) {
#pragma unroll
            for (int k = 0; k < K; ++k) a[k] = 0.0;
        } else {
            const double *src = p.ckA + (((size_t)b * p.nblk + blk) * NT + thread) * K;
#pragma unroll
            for (int k = 0; k < K; k += 2) {
                const double2 v = *reinterpret_cast<const double2 *>(src + k);
                a[k] = v.x;
                a[k + 1] = v.y;
            }
            E = p.ckAE[((size_t)b * p.nblk + blk) * NT + thread];
            const int En = __shfl_up_sync(kFull, E, 1);
            f = lane == 0 ? 1.0 : pow2i(En - E);
            if (lane == 31 && warp + 1 < nwa) {
                slots[warp].m0 = a[K - 1];
                slots[warp].E = E;
            }
        }
        if (nwa > 1) named_bar_sync(1, nbar);
        for (int r = 0; r < nrow; ++r) {
            const int t = t0 + r;
            const Slot *rd = slots + (r & 1) * kMaxLatticeWarps;
            Slot *wr = slots + ((r + 1) & 1) * kMaxLatticeWarps;
            double p1 = __shfl_up_sync(kFull, a[K - 1], 1) * f;
            int carryE = 0;
            if (lane == 0) {
                if (warp == 0) {
                    p1 = (t == 0) ? 1.0 : 0.0;
                } else {
                    const Slot s = rd[warp - 1];
                    carryE = s.E;
                    p1 = s.m0 * pow2i(s.E - E);
                }
            }
            alpha_cells<K>(a, p1, ytile + r * C, p.blank, idx, skipA);
            if (((t + 1) % kRenorm) == 0) {
                carryE = __shfl_sync(kFull, carryE, 0);
                renorm<K, true>(a, E, f, lane, carryE, warp > 0);
            }
            if (lane == 31 && warp + 1 < nwa) {
                wr[warp].m0 = a[K - 1];
                wr[warp].E = E;
            }
            double *ar = arow + (size_t)r * NT * K;
#pragma unroll
            for (int h = 0; h < K / 2; ++h)
                *reinterpret_cast<double2 *>(ar + ((size_t)h * NT + thread) * 2) = make_double2(a[2 * h], a[2 * h + 1]);
            Ea[r * NT + thread] = E;
            if (nwa > 1) named_bar_sync(1, nbar);
        }
        // ---- stage 4: beta sweep backwards over the block, posteriors
        const double invP = 1.0 / p.Pm[b];
        const int eP = p.Pe[b];
        E = 0;
        f = 1.0;
        if (t0 + nrow == Tb) {  // virtual beta_T: 1 at state S-1
#pragma unroll
            for (int k = 0; k < K; ++k) a[k] = (thread * K + k == S - 1) ? 1.0 : 0.0;
        } else {
            const double *src = p.ckB + (((size_t)b * p.nblk + blk) * NT + thread) * K;
#pragma unroll
            for (int k = 0; k < K; k += 2) {
                const double2 v = *reinterpret_cast<const double2 *>(src + k);
                a[k] = v.x;
                a[k + 1] = v.y;
            }
            E = p.ckBE[((size_t)b * p.nblk + blk) * NT + thread];
            const int En = __shfl_down_sync(kFull, E, 1);
            f = lane == 31 ? 1.0 : pow2i(En - E);
        }
        if (lane == 0 && warp > 0) {
            slots[warp].m0 = a[0];
            slots[warp].m1 = a[1];
            slots[warp].E = E;
        }
        if (nwa > 1) named_bar_sync(1, nbar);
        for (int q = 0; q < nrow; ++q) {
            const int r = nrow - 1 - q, t = t0 + r;
            const Slot *rd = slots + (q & 1) * kMaxLatticeWarps;
            Slot *wr = slots + ((q + 1) & 1) * kMaxLatticeWarps;
            double n1 = __shfl_down_sync(kFull, a[0], 1) * f;
            double n2 = __shfl_down_sync(kFull, a[1], 1) * f;
            int carryE = 0;
            const bool has_up = warp + 1 < nwa;
            if (lane == 31) {
                if (!has_up) {
                    n1 = 0.0;
                    n2 = 0.0;
                } else {
                    const Slot s = rd[warp + 1];
                    carryE = s.E;
                    const double fs = pow2i(s.E - E);
                    n1 = s.m0 * fs;
                    n2 = s.m1 * fs;
                }
            }
            double bb[K];
            beta_cells<K>(a, bb, n1, n2, ytile + r * C, p.blank, idx, skipB);
            // posterior(t,s) = alpha_t(s) * betabar_t(s) / P
            const double g = pow2i(min(Ea[r * NT + thread] + E - eP, 1000)) * invP;
            const double *ar = arow + (size_t)r * NT * K;
            double bsum = 0.0;
#pragma unroll
            for (int h = 0; h < K / 2; ++h) {
                const double2 av = *reinterpret_cast<const double2 *>(ar + ((size_t)h * NT + thread) * 2);
                bsum = fma(av.x, bb[2 * h], bsum);
                const double pl = av.y * bb[2 * h + 1] * g;
                if (lpos[h] >= 0) psort[r * Lp + lpos[h]] = (float)pl;
            }
            bpart[r * NT + thread] = (float)(bsum * g);
            if ((t % kRenorm) == 0) {
                carryE = __shfl_sync(kFull, carryE, 31);
                renorm<K, false>(a, E, f, lane, carryE, has_up);
            }
            if (lane == 0 && warp > 0) {
                wr[warp].m0 = a[0];
                wr[warp].m1 = a[1];
                wr[warp].E = E;
            }
            if (nwa > 1) named_bar_sync(1, nbar);
        }
    }
    __syncthreads();
    // ---- stage 5: per-class occupancy, gradient rows written once
    const float scale = p.grad_scale;
    const int nthr_a = nwa * 32;
    for (int r = warp; r < rows_out; r += nwl) {
        float *grow = gbase + (size_t)r * rowstride;
        if (r >= nrow) {
            for (int c = lane; c < C; c += 32) grow[c] = 0.f;
            continue;
        }
        float ob = 0.f;
        for (int j = lane; j < nthr_a; j += 32) ob += bpart[r * NT + j];
#pragma unroll
        for (int d = 16; d; d >>= 1) ob += __shfl_xor_sync(kFull, ob, d);
        for (int c = lane; c < C; c += 32) {
            float occ = 0.f;
            if (c == p.blank) {
                occ = ob;
            } else {
                const int q1 = coff[c + 1];
                for (int q = coff[c]; q < q1; ++q) occ += psort[r * Lp + q];
            }
            grow[c] = scale * ((float)ytile[r * C + c] - occ);
        }
    }
}

}  // namespace ctcb200
