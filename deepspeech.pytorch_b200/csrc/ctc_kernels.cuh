// csrc/ctc_kernels.cuh -- the kernels of the B200 CTC loss (sm_100a).
//
//   ctc_prep_kernel   : grid ceil(T*B/128) + B (+1).  Row blocks: fused softmax over C per (t,b) -- coalesced float4
//                       loads of 128 consecutive rows into shared memory, one thread per row for max / sum, coalesced
//                       stores of the emission rows y[b,t,:] as fp64 HIGH words (kYP = 1; kYP = 2: + a plane of low words;
//                       utterance-major, so the 8 rows of a group are ONE TMA bulk copy); padding frames get their
//                       zero gradient rows here.  Utterance blocks: label offsets and the per-thread label table.
//                       Extra block of a sliced launch: utterances ranked by length (sort_by_length).
//   ctc_engine_kernel : grid 2 * nb (score-only: nb) per slice.  CTA 2i runs the alpha recursion of utterance i
//                       forwards in time, CTA 2i+1 the beta recursion backwards, concurrently on different SMs.
//                       Lattice warps hold K = 8 states per thread in registers and form a SYSTOLIC pipeline over
//                       groups of 8 time steps: in super-step i warp q works on group i-q, boundary values cross warps
//                       through a shared-memory ring one group late and the CTA synchronises once per group.  A
//                       producer warp TMA-prefetches the emission tiles of the coming groups into a ring.  After every
//                       group the renormalised, settled state is checkpointed.  The alpha CTA ends with P = p(l|x).
//   ctc_grad_kernel   : grid (nsplit, nb) per slice, 4 CTAs/SM.  CTA (z, i) takes time groups z, z+nsplit, ... of
//                       utterance i: warp q recomputes alpha over its 224 states (+16-state halos) from the
//                       checkpoint into shared memory, sweeps beta backwards from the beta checkpoint forming
//                       posteriors alpha*betabar/P and adding them, as 2^-30 fixed point, into [row][class]
//                       occupancy tables in shared memory (integer reductions: order-independent, deterministic);
//                       then the CTA writes grad = scale * (softmax - occupancy) once, as coalesced rows.
//   large alphabets (C > 128): ctc_prep_utt_kernel + ctc_prep_big_kernel (streaming softmax that also writes the dense
//                       part of the gradient; emission rows gathered by label POSITION) -> the same engine / grad
//                       kernels at pitch 256, the grad kernel subtracting the occupancies in place.
//
// Replaces the arithmetic behind `criterion(logits, targets, output_sizes.cpu(), target_sizes)`
// (/root/reference/train.py:600; eval :352,:457), i.e. warp-ctc's softmax + compute_alpha + compute_betas_and_grad
// stages [external], with a different algorithm: linear-domain fp64 block-exponent recursion, bidirectional
// engines, checkpoint + recompute instead of an S x T lattice in global memory.
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
    int T, B, C, Cp, blank, zero_infinity;
    float grad_scale;
    int nwl;   // lattice warps per CTA (from Lmax)
    int R;     // time steps per checkpoint block (8 or 16)
    int nblk;  // ceil(T / R)
    int Lmax;  // max label length
    int score_only;
    // workspace
    unsigned *y;        // [B][T][kYP*Cp] softmax rows: Cp high words of the fp64 values (kYP = 2: then Cp low words; exact widening
                        // of the fp32 softmax, clamped to >= 2^-126)
    int *ckA, *ckB;     // [B][nblk][nwl][K/4][32] int4: checkpoint mantissas as rounded high words (hi_round)
    int *ckAE, *ckBE;   // [B][nblk][nwl*32]   exponents
    double *Pm2;        // [B][2] cluster engine: mantissas of alpha_{T-1}(S-1), alpha_{T-1}(S-2)
    int *Pe2;           // [B][2] ... and their exponents
    double *Pm;         // [B] mantissa of p(l|x) in [1,2)
    int *Pe;            // [B] exponent of p(l|x)
    int *feas;          // [B] 1 = feasible
    int *laboff;        // [B] offset of utterance b in `labels`
    uint4 *tlab;        // [B][nwl*32] per lattice thread (K = 8): packed label classes, skip/valid bits, lpos
    int bpos;           // index of the blank inside an emission row (class rows: blank; position rows: 0)
    int big;            // 1: large-alphabet path (emission rows indexed by label POSITION, see ctc_prep_big_kernel)
    int2 *ctab;         // big: [B][Lmax] per class-sorted label slot j: (class, run length of the class if j is its
                        // first slot, else 0) -- the gradient kernel subtracts one merged occupancy per class
    int *path;          // optional [B][T]: argmax class of every valid frame (best path), by-product of the softmax pass
    int *dec_ids, *dec_off, *dec_len;  // with path: collapsed best path [B][T], frame offsets [B][T] (nullable), lengths [B]
    int *bord;          // [B] utterances ordered by decreasing act_len (length bucketing, see sort_by_length)
    int slice0;         // engine / grad kernels: this launch covers bord[slice0 .. slice0 + gridDim) (bord == 0: b itself)
    int slice_nb;       // utterances of this launch (large-alphabet row kernel: grid = T * slice_nb)
    int nsplit;         // grad kernel: time blocks are dealt round-robin to this many CTAs per utterance
    int nqmax;          // grad kernel: warps per CTA of THIS launch (= its shared-memory layout)
    int gq_lo;          // grad kernel: this launch takes the utterances that need gq_lo .. nqmax lattice warps (the
                        // host launches one grid per warp class, so that every resident warp has work)
    int gLmax;          // grad kernel: longest label sequence of this launch's class (shared-memory layout)
    int nan_to_zero;    // 1: a NaN logit is read as 0 (the reference's train.py:595-598 guard, folded into the softmax pass)
    int *nan_count;     // optional: += number of NaN logits seen in valid frames (nan_to_zero only)
    int phase;          // second engine form: 0 = whole recursion; 1 / 2 = its first / second part (engine_phase_range)
    double *est;        // [B][2 directions][12][NT]: per-thread state handed from part 1 to part 2
    int prep_dbg;       // timing experiments (CTC_B200_PREP_DBG): 1 = row blocks return at once, 2 = table blocks do
    int y_ready;        // 1: the emission rows y were written by fc_logits_kernel; the row blocks only zero padding grads
#ifdef CTC_PROFILE
    long long *dbg;     // [ctas][16 warps][8] clock counters (tools/ubench/engine_bench.cu only)
#endif
};

#ifdef CTC_PROFILE
__device__ __forceinline__ long long prof_clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }
#define PROF(x) x
#else
#define PROF(x)
#endif

struct __align__(8) Slot {
    double m0, m1;
    int E;
    int pad;
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

// Posterior rows of the gradient kernel (one per time step of a group): [0, Lmax) label posteriors at their
// class-sorted positions, then nq * 32 per-thread partial sums of the blank posteriors.  The pitch is odd so
// that the 8 rows of a group start in 8 different banks.
constexpr int kCoreThreads = 28; // engine threads per gradient-kernel warp (224 states) + 2 halo threads each side
__host__ __device__ inline int grad_nq(int L) { return (2 * L + 1 + 8 * kCoreThreads - 1) / (8 * kCoreThreads); }
__host__ __device__ inline int grad_row_pitch(int Lmax, int nqmax) {
    const int n = (Lmax > 0 ? Lmax : 1) + nqmax * 32;
#ifdef CTC_PITCH_EXTRA  // tools/ubench experiment: other odd pitches.  Measured on the headline shape (engine_bench,
                        // gradient kernel alone): +0 98.4 us, +2 / +4 / +8 / +16 / +24 / +32 floats 105.3-106.5 us --
                        // a step at 55.5 KB of shared memory per CTA, not a bank-conflict effect
    return (n | 1) + 2 * (CTC_PITCH_EXTRA);
#endif
    return n | 1;  // any ODD pitch puts the 8 rows of a group in 8 different residues mod 8, hence (with units 8 banks
                   // apart) the reduction's (unit, row) lanes in 32 different banks; the smallest one saves up to 31
                   // floats per row -- at B=64,T=1000,C=5000 (Lmax 183) exactly what lets a 4th CTA fit on the SM
}
// Longest label sequence nq gradient warps can hold, and where the blank block of an utterance's posterior row starts:
// behind the longest label sequence of the utterance's own warp class (the same in every launch that can take it).
__host__ __device__ inline int grad_class_lmax(int nq, int Lmax) {
    const int l = 4 * kCoreThreads * nq - 1;
    return l < Lmax ? l : Lmax;
}
__host__ __device__ inline int grad_blank_off(int L, int Lmax) {
    const int l = grad_class_lmax(grad_nq(L), Lmax);
    return l > 0 ? l : 1;
}
// ================================================================================================ prep
constexpr int kPrepRows = 128;    // THREADS of a prep CTA (the name dates from the round-1 one-thread-per-row softmax)
#ifndef CTC_PREP_ROWS_CTA
#define CTC_PREP_ROWS_CTA 32
#endif
constexpr int kPrepRowsCta = CTC_PREP_ROWS_CTA;  // softmax rows per row block: 4x the warps in flight of 128-row blocks

__host__ __device__ inline size_t prep_smem(int C, int Lmax) {
    const size_t rows = 0;  // (the row blocks no longer stage anything in shared memory)
    const size_t utt = 2 * sizeof(int) * (size_t)(Lmax > 0 ? Lmax : 1);  // sorted positions + staged labels
    return (rows > utt ? rows : utt) + sizeof(int) * (C + 2) + 64;  // >= 4 * B bytes (sort block): see make_plan
}

// Utterance block of the prep kernels (128 threads): label offset and the per-lattice-thread label table (K = 8:
// thread e owns labels 4e .. 4e+3); large alphabets also rank the labels by class (the gradient kernel's
// class-sorted posterior row).  The labels are staged in shared memory first, so the block is three rounds of global
// latency (lengths, labels, results) rather than one per label.  lps, labs: shared int [Lmax] each; cnt: shared
// int [C + 2].
__device__ __forceinline__ void prep_utterance(const CtcParams &p, int b, int *lps, int *labs, int *cnt) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int C = p.C;
    int *s_off = cnt + C + 1;
    if (tid == 0) *s_off = 0;
    __syncthreads();
    int part = 0;
    for (int i = tid; i < b; i += blockDim.x) part += max(p.label_lens[i], 0);
#pragma unroll
    for (int d = 16; d; d >>= 1) part += __shfl_xor_sync(kFull, part, d);
    if (lane == 0 && part) atomicAdd(s_off, part);
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    __syncthreads();
    const int off = *s_off;
    if (tid == 0) p.laboff[b] = off;
    for (int i = tid; i < L; i += blockDim.x) labs[i] = min(max(p.labels[off + i], 0), C - 1);
    __syncthreads();
    const int *lab = labs;
    if (p.big) {
        // large alphabets: a counting sort would walk C counters; rank the L <= 255 labels against each other instead
        // (stable: ties by position).  ctab[slot] = (class, run length at the class's first slot, else 0).
        for (int i = tid; i < L; i += blockDim.x) {
            const int c = lab[i];
            int less = 0, eq_before = 0, eq = 0;
            for (int j = 0; j < L; ++j) {
                const int cj = lab[j];
                less += cj < c;
                eq += cj == c;
                eq_before += (cj == c) & (j < i);
            }
            lps[i] = less + eq_before;
            p.ctab[(size_t)b * p.Lmax + less + eq_before] = make_int2(c, eq_before == 0 ? eq : 0);
        }
    }
    __syncthreads();
    const int NT = p.nwl * 32;
    for (int e = tid; e < NT; e += blockDim.x) {
        unsigned cls = 0, bits = 0, lp[4] = {0, 0, 0, 0};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            const int i = 4 * e + h;
            const bool valid = i < L;
            const int c = valid ? lab[i] : min(max(p.blank, 0), C - 1);
            // emission-row index: the class itself, or (large alphabets) the label's own position 1 + i
            cls |= (unsigned)(p.big ? (valid ? 1 + i : 0) : c) << (8 * h);
            if (valid) {
                bits |= 1u << (8 + h);
                if (i == 0 || lab[i - 1] != c) bits |= 1u << h;            // skip into s
                if (i + 1 < L && lab[i + 1] != c) bits |= 1u << (4 + h);  // skip out of s
                if (p.big) lp[h] = (unsigned)lps[i];  // class-sorted slot of the label's posterior
            }
        }
        p.tlab[(size_t)b * NT + e] = make_uint4(cls, bits, lp[0] | (lp[1] << 16), lp[2] | (lp[3] << 16));
    }
}

// Length bucketing on the device (one CTA; lengths may live only on the device, so the host never sees them):
// bord[rank] = utterance, ranked by decreasing number of frames (ties by index).  Rank by counting, O(B^2 / 128) per
// thread with the lengths cached in shared memory -- a few microseconds for B <= 1024, hidden behind the softmax
// blocks of the same launch.  The host then launches the recursion and gradient kernels per SLICE of ranks on side
// streams: slices of short utterances finish their recursion early and their gradient work fills the SMs while
// the long ones are still stepping through time.
constexpr int kMaxSortB = 2048;
__device__ __forceinline__ void sort_by_length(const CtcParams &p, int *keys) {
    const int tid = threadIdx.x, B = p.B;
    for (int i = tid; i < B; i += blockDim.x) keys[i] = min(max(p.act_lens[i], 0), p.T);
    __syncthreads();
    for (int i = tid; i < B; i += blockDim.x) {
        const int ki = keys[i];
        int rank = 0;
        for (int j = 0; j < B; ++j) {
            const int kj = keys[j];
            rank += (kj > ki || (kj == ki && j < i)) ? 1 : 0;
        }
        p.bord[rank] = i;
    }
}

// NaN guard of the training loop (/root/reference/train.py:595-598: `logits[torch.isnan(logits)] = 0` ahead of the
// criterion): with nan_to_zero set every read of a logit goes through this, so all passes see the same value.
__device__ __forceinline__ float nan_fix(float v, bool ntz) { return (ntz && v != v) ? 0.f : v; }
// softmax value clamped to >= 2^-126; a row whose sum is NaN (`bad`) stays NaN instead of collapsing onto the clamp
constexpr unsigned kNanHi = 0x7ff80000u;  // high word of a quiet fp64 NaN (planar emission rows)
__device__ __forceinline__ float clamp_y(float v, bool bad) {
    return bad ? __int_as_float(0x7fc00000) : fmaxf(v, 1.17549435e-38f);
}

__global__ void __launch_bounds__(kPrepRows) ctc_prep_kernel(const CtcParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int C = p.C, Cp = p.Cp;
    const long long nrows_total = (long long)p.T * p.B;
    const int nrb = (int)((nrows_total + kPrepRowsCta - 1) / kPrepRowsCta);
    const size_t rows_bytes = 0;
    const size_t utt_bytes = 2 * sizeof(int) * (size_t)(p.Lmax > 0 ? p.Lmax : 1);
    int *cnt = reinterpret_cast<int *>(smem + (rows_bytes > utt_bytes ? rows_bytes : utt_bytes));

    // The utterance / sort blocks are long dependent chains of small loads: they take the LOWEST block indices so that
    // they are scheduled first and run under the softmax blocks instead of forming the tail of the launch.
    const int nextra = (int)gridDim.x - nrb;  // B utterance blocks (+1 sort block when the launch is sliced)
    if (p.prep_dbg && ((int)blockIdx.x < nextra) == (p.prep_dbg == 2)) return;
    if ((int)blockIdx.x < nextra) {
        if ((int)blockIdx.x == p.B)
            sort_by_length(p, reinterpret_cast<int *>(smem));
        else
            prep_utterance(p, blockIdx.x, reinterpret_cast<int *>(smem),
                           reinterpret_cast<int *>(smem) + (p.Lmax > 0 ? p.Lmax : 1), cnt);
        return;
    }
    // ---------------------------------------------------- row block: softmax of 128 consecutive (t,b) rows
    float *xs = reinterpret_cast<float *>(smem);
    const long long row0 = (long long)((int)blockIdx.x - nextra) * kPrepRowsCta;
    const int nrows = (int)min((long long)kPrepRowsCta, nrows_total - row0);
    if (p.y_ready) {
        // the emission rows came with the logits (fc_logits_kernel's epilogue): the logits are not read at all; what
        // is left of this block's work is the zero gradient of its padding frames
        if (p.grads)
            for (int i = tid; i < nrows * C; i += kPrepRows) {
                const long long row = row0 + i / C;
                const int t = (int)(row / p.B), b = (int)(row - (long long)t * p.B);
                if (t >= min(max(p.act_lens[b], 0), p.T)) p.grads[row0 * C + i] = 0.f;
            }
        return;
    }
    // Rows straight from global memory, no shared-memory stage (round 1 staged the 128 rows, synchronised the CTA and
    // then made 8 dependent passes per warp: 18.6 us for 11 MB in / 25 MB out on the headline, and this kernel is on
    // the critical path of every step).  Cp / 4 lanes per row (8 for C <= 32), 4 adjacent classes per lane; the loads
    // of TWO passes are issued before the first is consumed; max and sum by shuffles inside the lane group, every exp
    // evaluated once, the emission row written as two 16-byte stores per lane.
    (void)xs;
    const int lpr = Cp >> 2, rpw = 32 / lpr;  // lanes per row, rows per warp pass
    const int j = lane & (lpr - 1);
    constexpr int U = kPrepRowsCta / 16;  // passes whose loads are in flight together (C <= 32: the whole block)
    const int stride = (kPrepRows / 32) * rpw;
    for (int r0 = warp * rpw; r0 < nrows; r0 += U * stride) {
        float xv[U][4];
        int tt[U], bb[U], al[U];
        bool vv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int r = r0 + u * stride + lane / lpr;
            vv[u] = r < nrows && (r0 + u * stride) < nrows;
            const long long row = row0 + (vv[u] ? r : 0);
            if (nrows_total < (1ll << 31)) {  // 32-bit division (the 64-bit one is ~100 instructions per lane)
                tt[u] = (int)((unsigned)row / (unsigned)p.B);
                bb[u] = (int)((unsigned)row - (unsigned)tt[u] * (unsigned)p.B);
            } else {
                tt[u] = (int)(row / p.B);
                bb[u] = (int)(row - (long long)tt[u] * p.B);
            }
            al[u] = vv[u] ? __ldg(p.act_lens + bb[u]) : 0;
            const float *src = p.acts + row * C + 4 * j;
#pragma unroll
            for (int i = 0; i < 4; ++i) xv[u][i] = (vv[u] && 4 * j + i < C) ? __ldg(src + i) : -INFINITY;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if ((r0 + u * stride) >= nrows) break;  // warp-uniform
            const bool valid = vv[u];
            const int t = tt[u], b = bb[u];
            const long long row = (long long)t * p.B + b;
            const bool active = valid && t < min(max(al[u], 0), p.T);
            float x[4] = {xv[u][0], xv[u][1], xv[u][2], xv[u][3]};
            if (p.nan_to_zero) {
                int nn = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (x[i] != x[i]) { x[i] = 0.f; ++nn; }
                if (nn && active && p.nan_count) atomicAdd(p.nan_count, nn);
            }
            float m = fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3]));
            for (int d = 1; d < lpr; d <<= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, d));
            if (!valid) m = 0.f;
            if (p.path != nullptr) {
                // best path: first class attaining the row maximum, from the values already in
                // registers -- the decoder needs no second read of the logits
                int cand = 0x7fffffff;
#pragma unroll
                for (int i = 3; i >= 0; --i)
                    if (x[i] == m) cand = 4 * j + i;
                for (int d = 1; d < lpr; d <<= 1) cand = min(cand, __shfl_xor_sync(kFull, cand, d));
                if (active && j == 0) p.path[(size_t)b * p.T + t] = cand;
            }
            float e[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) e[i] = (4 * j + i < C) ? expf(x[i] - m) : 0.f;
            float sum = (e[0] + e[1]) + (e[2] + e[3]);
            for (int d = 1; d < lpr; d <<= 1) sum += __shfl_xor_sync(kFull, sum, d);
            const float inv = 1.0f / sum;
            if (active) {
                // y clamped to >= 2^-126: a feasible alignment never gets probability exactly 0 from fp32 underflow
                unsigned *ydst = p.y + ((size_t)b * p.T + t) * kYP * Cp;  // utterance-major: a tile of rows is one bulk copy
                uint4 hi, lo;
                float v = fmaxf(e[0] * inv, 1.17549435e-38f); hi.x = y_hi_word(v); lo.x = y_lo_word(v);
                v = fmaxf(e[1] * inv, 1.17549435e-38f); hi.y = y_hi_word(v); lo.y = y_lo_word(v);
                v = fmaxf(e[2] * inv, 1.17549435e-38f); hi.z = y_hi_word(v); lo.z = y_lo_word(v);
                v = fmaxf(e[3] * inv, 1.17549435e-38f); hi.w = y_hi_word(v); lo.w = y_lo_word(v);
                if (sum != sum) {
                    // a NaN or +Inf logit (not scrubbed by nan_to_zero) makes the row sum NaN: the whole row becomes NaN
                    // and propagates through the recursions into cost and gradient of this utterance, as it would in
                    // the reference's criterion (whose caller then reports it, train.py:604-611) -- never a finite
                    // wrong value (fmaxf would have returned the clamp)
                    hi = make_uint4(kNanHi, kNanHi, kNanHi, kNanHi);
                    lo = make_uint4(0u, 0u, 0u, 0u);
                }
                *reinterpret_cast<uint4 *>(ydst + 4 * j) = hi;
                if (kYP == 2) *reinterpret_cast<uint4 *>(ydst + Cp + 4 * j) = lo;
            } else if (valid && p.grads) {
                // padding frame (t >= act_lens[b]): never read by the lattice kernels; its gradient row is zero
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (4 * j + i < C) p.grads[row * C + 4 * j + i] = 0.f;
            }
        }
    }
}

// Decode one entry of the prep kernels' label table (4 labels = 8 lattice states per entry).  K = 8: the whole
// entry; K = 4: labels 2*half .. 2*half+1 of it.
template <int K>
__device__ __forceinline__ void decode_tlab(const uint4 tl, int (&idx)[K / 2], double (&skipA)[K / 2],
                                            double (&skipB)[K / 2], int (&lpos)[K / 2], bool want_lpos, int half = 0) {
    static_assert(K == 8 || K == 4, "label table entries hold 4 labels");
#pragma unroll
    for (int hh = 0; hh < K / 2; ++hh) {
        const int h = (K == 8) ? hh : 2 * half + hh;
        idx[hh] = (tl.x >> (8 * h)) & 0xff;
        skipA[hh] = ((tl.y >> h) & 1) ? 1.0 : 0.0;
        skipB[hh] = ((tl.y >> (4 + h)) & 1) ? 1.0 : 0.0;
        const unsigned lw = (h < 2) ? tl.z : tl.w;
        lpos[hh] = (want_lpos && ((tl.y >> (8 + h)) & 1)) ? (int)((lw >> (16 * (h & 1))) & 0xffff) : -1;
    }
}

// Large alphabets index their emission rows by label POSITION (0: blank, 1 + i: label i); the table's 8-bit field
// cannot hold positions beyond 255, and the position follows from the lattice thread index anyway (thread e owns
// labels 4e .. 4e+3), so it is recomputed here.
template <int K>
__device__ __forceinline__ void position_idx(const uint4 tl, int e, int (&idx)[K / 2]) {
#pragma unroll
    for (int h = 0; h < K / 2; ++h) idx[h] = ((tl.y >> (8 + h)) & 1) ? 1 + 4 * e + h : 0;
}

// ================================================================================================ engine
#ifndef CTC_LOOKAHEAD
#define CTC_LOOKAHEAD 4
#endif
constexpr int kLookAhead = CTC_LOOKAHEAD;  // emission tiles in flight ahead of the first lattice warp
// ... except at pitch 512 (large alphabets with 255 < L <= 511): a tile is 32 KB there, so one tile ahead
__host__ __device__ constexpr int look_ahead(int Cp) { return Cp >= 512 ? 1 : kLookAhead; }

struct EngineSmem {
    size_t ytile, bars, slots, misc, xlink, total;
    int nstage;
};
__host__ __device__ inline EngineSmem engine_smem(int Cp, int nwl) {
    EngineSmem s;
    s.nstage = nwl + look_ahead(Cp) + 1;
    size_t o = 0;
    s.ytile = o; o = align16(o + sizeof(unsigned) * s.nstage * kGroup * kYP * Cp);
    s.bars = o;  o = align16(o + sizeof(uint64_t) * s.nstage);
    s.slots = o; o = align16(o + sizeof(Slot) * kMaxLatticeWarps * kRing);
    s.misc = o;  o = align16(o + 64);
    // cluster form only: receive ring of the cross-CTA link (2 rings of 320 B) + 8 mbarriers
    s.xlink = o; o = align16(o + 2 * 320 + 8 * sizeof(uint64_t));
    s.total = o;
    return s;
}


// One group (<= 8 time steps) of the alpha recursion for one lattice warp.  Exponents are fixed during a group, so a
// step moves mantissas only: the inflow (pr0, from the thread below) is scaled by the per-group constant f.
// Neighbour values are exchanged one step AHEAD of use (shuffle in-warp, shared-memory slot across warps) and are
// consumed at the END of the next step; emission values are loaded one step ahead.  After the last step the thread
// renormalises, the warp settles its exponents (-> new f) and only then sends the boundary mantissa.
#ifdef CTC_PROFILE
struct ProfAcc { long long steps, ren, settle, send; };
#define PROF_ARG , ProfAcc &pa
#define PROF_PASS , pa
#else
#define PROF_ARG
#define PROF_PASS
#endif
template <int K, int CP, bool FULL>
__device__ __forceinline__ void alpha_group(Cells<K> &c, double &pr0, double &f, const unsigned *yt, int nr,
                                            int t0, int blank, const int (&idx)[K / 2], const double (&skipA)[K / 2],
                                            int lane, bool has_lower, bool has_upper, Slot *slot_mine,
                                            const Slot *slot_in PROF_ARG) {
    YRow<K> y;
    load_y<K, CP>(y, yt, 0, blank, idx);
    PROF(long long q0 = prof_clk();)
#pragma unroll
    for (int r = 0; r < kGroup; ++r) {
        if (FULL || r < nr) {
            const bool last = FULL ? (r == kGroup - 1) : (r == nr - 1);
            const int t = t0 + r;
            YRow<K> yn;
            load_y<K, CP>(yn, yt, FULL ? min(r + 1, kGroup - 1) : min(r + 1, nr - 1), blank, idx);
            const double p1 = pr0 * f;
            const double top = alpha_top<K>(c.a, y, skipA);
            double sh = 0.0;
            if (!last) {
                sh = __shfl_up_sync(kFull, top, 1);
#ifndef CTC_EXP_NOSLOT
                if (lane == 31 && has_upper) slot_mine[t & (kRing - 1)].m0 = top;
#endif
            }
            alpha_rest<K>(c.a, top, p1, y, skipA);
            if (!last) {
                pr0 = sh;
#ifndef CTC_EXP_NOSLOT
                if (lane == 0) pr0 = has_lower ? slot_in[t & (kRing - 1)].m0 : 0.0;
#endif
            }
            y = yn;
        }
    }
    const int tl = (t0 + nr - 1) & (kRing - 1);
    PROF(long long q1 = prof_clk();)
    renorm<K>(c);
    PROF(long long q2 = prof_clk();)
    f = settle<K, true>(c, has_lower ? slot_in[tl].E : kDeadE, lane);
    PROF(long long q3 = prof_clk(); pa.steps += q1 - q0; pa.ren += q2 - q1; pa.settle += q3 - q2;)
    pr0 = __shfl_up_sync(kFull, c.a[K - 1], 1);
    if (lane == 31 && has_upper) {
        slot_mine[tl].m0 = c.a[K - 1];
        slot_mine[tl].E = c.E;
    }
    if (lane == 0) pr0 = has_lower ? slot_in[tl].m0 : 0.0;
}

// Mirror image for beta: rows are visited t0+nr-1 down to t0, inflow (two values) comes from the thread above.
template <int K, int CP, bool FULL>
__device__ __forceinline__ void beta_group(Cells<K> &c, double &pr0, double &pr1, double &f, const unsigned *yt,
                                           int nr, int t0, int blank, const int (&idx)[K / 2],
                                           const double (&skipB)[K / 2], int lane, bool has_lower, bool has_upper,
                                           Slot *slot_mine, const Slot *slot_in) {
    YRow<K> y;
    load_y<K, CP>(y, yt, nr - 1, blank, idx);
#pragma unroll
    for (int rr = 0; rr < kGroup; ++rr) {
        const int r = (FULL ? kGroup : nr) - 1 - rr;
        if (FULL || r >= 0) {
            const bool last = (r == 0);
            const int t = t0 + r;
            YRow<K> yn;
            load_y<K, CP>(yn, yt, max(r - 1, 0), blank, idx);
            const double n1 = pr0 * f, n2 = pr1 * f;
            double b0, b1, s0, s1v;
            beta_bottom<K>(c.a, y, skipB, b0, b1, s0, s1v);
            double sh0 = 0.0, sh1 = 0.0;
            if (!last) {
                sh0 = __shfl_down_sync(kFull, s0, 1);
                sh1 = __shfl_down_sync(kFull, s1v, 1);
                if (lane == 0 && has_lower) {
                    slot_mine[t & (kRing - 1)].m0 = s0;
                    slot_mine[t & (kRing - 1)].m1 = s1v;
                }
            }
            double bb[K];
            beta_rest<K>(c.a, bb, b0, b1, s0, s1v, n1, n2, y, skipB);
            if (!last) {
                pr0 = sh0;
                pr1 = sh1;
                if (lane == 31) {
                    pr0 = has_upper ? slot_in[t & (kRing - 1)].m0 : 0.0;
                    pr1 = has_upper ? slot_in[t & (kRing - 1)].m1 : 0.0;
                }
            }
            y = yn;
        }
    }
    const int tl = t0 & (kRing - 1);
    renorm<K>(c);
    f = settle<K, false>(c, has_upper ? slot_in[tl].E : kDeadE, lane);
    pr0 = __shfl_down_sync(kFull, c.a[0], 1);
    pr1 = __shfl_down_sync(kFull, c.a[1], 1);
    if (lane == 0 && has_lower) {
        slot_mine[tl].m0 = c.a[0];
        slot_mine[tl].m1 = c.a[1];
        slot_mine[tl].E = c.E;
    }
    if (lane == 31) {
        pr0 = has_upper ? slot_in[tl].m0 : 0.0;
        pr1 = has_upper ? slot_in[tl].m1 : 0.0;
    }
}

template <int K, int CP>
__global__ void __launch_bounds__((kMaxLatticeWarps + 1) * 32) ctc_engine_kernel(const CtcParams p) {
    static_assert(K % 2 == 0 && K >= 4, "K must be even and >= 4");
    extern __shared__ __align__(16) unsigned char smem[];
    const int bi = p.slice0 + (p.score_only ? blockIdx.x : (blockIdx.x >> 1));
    const int b = p.bord ? p.bord[bi] : bi;
    const bool fwd = p.score_only ? true : !(blockIdx.x & 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwl = p.nwl, C = p.C;
    constexpr int Cp = CP;
    const int NT = nwl * 32;
    const EngineSmem lay = engine_smem(Cp, nwl);
    const int NS = lay.nstage;
    unsigned *ytile = reinterpret_cast<unsigned *>(smem + lay.ytile);
    uint64_t *y_full = reinterpret_cast<uint64_t *>(smem + lay.bars);
    Slot *slots = reinterpret_cast<Slot *>(smem + lay.slots);
    double *pm = reinterpret_cast<double *>(smem + lay.misc);  // [2]
    int *pe = reinterpret_cast<int *>(smem + lay.misc + 16);   // [2]
    int *ck = (fwd ? p.ckA : p.ckB) + (size_t)b * p.nblk * NT * K;
    int *ckE = (fwd ? p.ckAE : p.ckBE) + (size_t)b * p.nblk * NT;

    const int Tb = min(max(p.act_lens[b], 0), p.T);
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int S = 2 * L + 1;
    const int nwa = min(nwl, (S + 32 * K - 1) / (32 * K));  // lattice warps this utterance needs

    if (Tb == 0) {  // nothing to recurse over
        if (fwd && tid == 0) {
            const bool ok = (L == 0);
            p.feas[b] = ok;
            p.Pm[b] = 1.0;
            p.Pe[b] = 0;
            p.costs[b] = ok ? 0.f : (p.zero_infinity ? 0.f : INFINITY);
        }
        return;
    }
    if (warp >= nwa && warp != nwl) return;  // lattice warps beyond this utterance's S

    const int ngroups = (Tb + kGroup - 1) / kGroup;
    const int nsuper = ngroups + nwa - 1;
    const int nbar = (nwa + 1) * 32;

    if (warp == nwl) {
        // =========================================================== producer warp: TMA emission-tile prefetch
        // y is utterance-major, so the <= 8 rows of a group are contiguous: ONE bulk copy per tile.
        if (lane == 0) {
            for (int i = 0; i < NS; ++i) mbar_init(&y_full[i], 1);
            fence_mbar_init();
        }
        __syncwarp();
        const unsigned *ybase = p.y + (size_t)b * p.T * kYP * Cp;
        const uint32_t rowbytes = (uint32_t)(kYP * Cp * sizeof(unsigned));
        int ist = 0;  // ring slot of the next tile to issue (issue() is called with j = 0, 1, 2, ...)
        auto issue = [&](int j) {
            const int st = ist;
            ist = (ist + 1 == NS) ? 0 : ist + 1;
            if (j < ngroups && lane == 0) {
                const int G = fwd ? j : (ngroups - 1 - j);
                const int t0 = G * kGroup;
                const int nr = min(kGroup, Tb - t0);
                mbar_arrive_expect_tx(&y_full[st], rowbytes * nr);
                tma_bulk_g2s(ytile + (size_t)(st * kGroup) * kYP * Cp, ybase + (size_t)t0 * kYP * Cp, rowbytes * nr,
                             &y_full[st]);
            }
        };
        int wst = 0, wph = 0;  // ring slot / mbarrier phase of the next tile to wait for
        auto wait_tile = [&]() {
            mbar_wait(&y_full[wst], wph);
            if (++wst == NS) { wst = 0; wph ^= 1; }
        };
        PROF(long long tp0 = prof_clk(); long long t_work = 0; long long t_bar = 0;)
        for (int j = 0; j <= kLookAhead; ++j) issue(j);
        wait_tile();
        named_bar_sync(1, nbar);
        for (int i = 0; i < nsuper; ++i) {
            PROF(long long ta = prof_clk();)
            fence_proxy_async();
            issue(i + kLookAhead + 1);
            if (i + 1 < ngroups) wait_tile();
            PROF(long long tb = prof_clk();)
            named_bar_sync(1, nbar);
            PROF(long long tc = prof_clk(); t_work += tb - ta; t_bar += tc - tb;)
        }
        PROF(if (lane == 0 && blockIdx.x < 2) { long long *d = p.dbg + (blockIdx.x * 16 + warp) * 8; d[0] = prof_clk() - tp0; d[1] = t_work; d[2] = t_bar; d[3] = nsuper; })
        return;
    }

    // =============================================================== lattice warps
    const int thread = warp * 32 + lane;  // lattice thread index; owns states [thread*K, thread*K+K)
    int idx[K / 2], lab_i[K / 2];
    double skipA[K / 2], skipB[K / 2];
    decode_tlab<K>(p.tlab[(size_t)b * NT + (K == 8 ? thread : (thread >> 1))], idx, skipA, skipB, lab_i, false,
                   thread & 1);
    if (K == 8 && p.big) position_idx<K>(p.tlab[(size_t)b * NT + thread], thread, idx);
    (void)lab_i;
    (void)C;

    // initial state: everything in scale 2^0 (f = 1); the first settlement happens after the first group
    Cells<K> c;
#pragma unroll
    for (int k = 0; k < K; ++k) c.a[k] = 0.0;
    c.E = 0;
    double f = 1.0, pr0 = 0.0, pr1 = 0.0;
    const int q = fwd ? warp : (nwa - 1 - warp);  // position in the systolic pipeline
    const bool has_lower = warp > 0, has_upper = warp + 1 < nwa;
    Slot *slot_mine = slots + warp * kRing;
    const Slot *slot_in = slots + (fwd ? warp - 1 : warp + 1) * kRing;

    if (fwd) {
        if (thread == 0) pr0 = 1.0;  // virtual alpha_{-1}(-1) = 1
    } else {
        const int s1 = S - 1;  // virtual beta_T: 1 at state S-1
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (thread * K + k == s1) c.a[k] = 1.0;
        pr0 = __shfl_down_sync(kFull, c.a[0], 1);
        pr1 = __shfl_down_sync(kFull, c.a[1], 1);
        if (lane == 31) {
            pr0 = ((thread + 1) * K == s1) ? 1.0 : 0.0;  // S-1 is state 0 of the next thread
            pr1 = 0.0;
        }
    }
    named_bar_sync(1, nbar);
    PROF(long long tp0 = prof_clk(); long long t_work = 0; long long t_bar = 0; long long ngr = 0; long long t_loop = 0; ProfAcc pa; pa.steps = 0; pa.ren = 0; pa.settle = 0; pa.send = 0;)

    // Systolic schedule: this warp runs group j in super-step q + j and idles (barrier only) before and after.  All
    // per-group indices advance incrementally -- no multiplies, divisions or 64-bit address rebuilds per group.
    for (int i = 0; i < q; ++i) named_bar_sync(1, nbar);
    int st = 0;                                       // ring slot of this warp's next emission tile
    const unsigned *yt = ytile;
    int t0 = fwd ? 0 : (ngroups - 1) * kGroup;        // first row of the current group
    const int dt = fwd ? kGroup : -kGroup;
    const int ntail = Tb - (ngroups - 1) * kGroup;    // rows of the ragged last group (1..8)
    // checkpoints (settled exponents): alpha after row t0+7 -> block G+1 ; beta_{t0} -> block G-1, i.e. every group
    // but this direction's last one.  They only seed the gradient kernel's 8-step recomputation, so the mantissas are
    // stored as rounded high words (hi_round: full exponent range, 21 bits): half the bytes of the largest stream of
    // the whole pass.  Global layout per (block, lattice warp): [K/4][32 lanes] int4 -- every store (and every load of
    // the gradient kernel) is one fully coalesced 512-byte access.
    const int cb0 = fwd ? 1 : ngroups - 2;
    int4 *ckp = reinterpret_cast<int4 *>(ck + ((ptrdiff_t)cb0 * NT + warp * 32) * K) + lane;
    int *ckEp = ckE + (ptrdiff_t)cb0 * NT + thread;
    const ptrdiff_t ckstep = (fwd ? 1 : -1) * (ptrdiff_t)NT * (K / 4), ckEstep = (fwd ? 1 : -1) * (ptrdiff_t)NT;
    const bool want_ck = !p.score_only;
    for (int j = 0; j < ngroups; ++j) {
        PROF(long long ta = prof_clk(); ++ngr;)
        const bool tail = fwd ? (j == ngroups - 1) : (j == 0);
        PROF(long long tl0 = prof_clk();)
        if (fwd) {
            if (!tail || ntail == kGroup)
                alpha_group<K, CP, true>(c, pr0, f, yt, kGroup, t0, p.bpos, idx, skipA, lane, has_lower, has_upper,
                                         slot_mine, slot_in PROF_PASS);
            else
                alpha_group<K, CP, false>(c, pr0, f, yt, ntail, t0, p.bpos, idx, skipA, lane, has_lower, has_upper,
                                          slot_mine, slot_in PROF_PASS);
        } else {
            if (!tail || ntail == kGroup)
                beta_group<K, CP, true>(c, pr0, pr1, f, yt, kGroup, t0, p.bpos, idx, skipB, lane, has_lower,
                                        has_upper, slot_mine, slot_in);
            else
                beta_group<K, CP, false>(c, pr0, pr1, f, yt, ntail, t0, p.bpos, idx, skipB, lane, has_lower,
                                         has_upper, slot_mine, slot_in);
        }
        PROF(t_loop += prof_clk() - tl0;)
        if (j < ngroups - 1) {
            if (want_ck) {
#pragma unroll
                for (int k = 0; k < K; k += 4)
                    ckp[(k / 4) * 32] = make_int4(hi_round(c.a[k]), hi_round(c.a[k + 1]), hi_round(c.a[k + 2]),
                                                  hi_round(c.a[k + 3]));
                *ckEp = c.E;
            }
            ckp += ckstep;
            ckEp += ckEstep;
        } else if (fwd) {
            // ---- P = alpha_{T-1}(S-1) + alpha_{T-1}(S-2): the owning threads publish (mantissa, exponent)
            const int s1 = S - 1, s2 = S - 2;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (thread * K + k == s1) { pm[0] = c.a[k]; pe[0] = c.E; }
                if (thread * K + k == s2) { pm[1] = c.a[k]; pe[1] = c.E; }
            }
        }
        t0 += dt;
        yt += kGroup * kYP * Cp;
        if (++st == NS) { st = 0; yt = ytile; }
        PROF(long long tb = prof_clk();)
        named_bar_sync(1, nbar);
        PROF(long long tc = prof_clk(); t_work += tb - ta; t_bar += tc - tb;)
    }
    for (int i = q + ngroups; i < nsuper; ++i) named_bar_sync(1, nbar);
    PROF(if (lane == 0 && blockIdx.x < 2) { long long *d = p.dbg + (blockIdx.x * 16 + warp) * 8; d[0] = prof_clk() - tp0; d[1] = t_work; d[2] = t_bar; d[3] = ngr; d[4] = t_loop; d[5] = pa.steps; d[6] = pa.ren; d[7] = pa.settle; })

    if (fwd && thread == 0) {
        double m = pm[0];
        int e = pe[0];
        if (S > 1) {
            const int e2 = pe[1];
            const int em = max(e, e2);
            m = pm[0] * pow2i(e - em) + pm[1] * pow2i(e2 - em);
            e = em;
        }
        const int ex = expo_field(m);
        const bool isnan_p = m != m;  // a non-finite emission row reached the lattice: cost and gradient are NaN
        const bool ok = ex != 0 && ex != 2047 && e > kDeadE / 2;
        if (ok) {
            m *= pow2i(1023 - ex);
            e += ex - 1023;
        }
        p.feas[b] = ok || isnan_p;
        p.Pm[b] = (ok || isnan_p) ? m : 1.0;
        p.Pe[b] = ok ? e : 0;
        const double cost = -(log(m) + (double)e * 0.69314718055994530942);
        p.costs[b] = isnan_p ? __int_as_float(0x7fc00000) : (ok ? (float)cost : (p.zero_infinity ? 0.f : INFINITY));
    }
}

// ================================================================================================ engine, second form
// Same recursion, same arithmetic and the same shared-memory / checkpoint protocol as ctc_engine_kernel above, with
// the per-GROUP overhead cut (ncu on the first form, headline shape: of ~1650 clk per group of 8 steps only ~780 were
// the steps; ~340 went into a prologue that re-derived direction, ring slots and addresses, ~260 into renormalise +
// settle + send, ~220 into checkpoint + loop bookkeeping):
//   * one instantiation per DIRECTION (no direction tests, no re-derived block indices inside the loop);
//   * boundary slots laid out planar and indexed by group parity + step, so every slot address of a group is one base
//     register plus an immediate;
//   * threads without an inflow neighbour (state 0's thread for alpha, the last warp's lane 31 for beta) carry
//     f = 0 instead of having their inflow zeroed every step; alpha starts from a virtual alpha_{-1}(0) = 1;
//   * renormalise + settle in ONE pass with 3 dependent warp collectives on the fast path (ballot -> indexed shuffle
//     -> vote) instead of 5: the exponent a dead thread adopts and the exponent of its inflow neighbour are the same
//     number (the nearest live thread on the inflow side), so one shuffle serves both; the boundary mantissa is
//     shuffled while the exponents settle; the lift loop is entered only when some lane needs it;
//   * the first emission row of the NEXT group is fetched before the group barrier (the producer runs one more tile
//     ahead), so a group no longer starts with an exposed shared-memory latency.
struct SlotRing {
    double m0[kRing];
    double m1[kRing];
    int E[kRing];
};
static_assert(sizeof(SlotRing) <= sizeof(Slot) * kRing, "the v2 rings live in the v1 slot area");
constexpr int kRingM1 = (int)offsetof(SlotRing, m1), kRingE = (int)offsetof(SlotRing, E);

// The slot rings are addressed through 32-bit shared-window addresses that the compiler cannot see through (it
// otherwise re-derives them from %tid / the CTA's shared window in EVERY step: measured, ~18 instructions per access)
// and accessed with predicated instructions, so the one lane per warp that talks to the neighbouring warp costs the
// other 31 no branch.
__device__ __forceinline__ uint32_t opaque_u32(uint32_t x) {
    asm volatile("" : "+r"(x));
    return x;
}
__device__ __forceinline__ void sts_f64_if(uint32_t addr, double v, bool pred) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p st.shared.f64 [%0], %1;\n\t}" ::"r"(addr), "d"(v),
                 "r"((int)pred));
}
__device__ __forceinline__ void sts_b32_if(uint32_t addr, int v, bool pred) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p st.shared.b32 [%0], %1;\n\t}" ::"r"(addr), "r"(v),
                 "r"((int)pred));
}
// v = pred ? *addr : v
__device__ __forceinline__ void lds_f64_if(double &v, uint32_t addr, bool pred) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p ld.shared.f64 %0, [%1];\n\t}" : "+d"(v) : "r"(addr),
                 "r"((int)pred));
}
__device__ __forceinline__ int lds_b32(uint32_t addr) {
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// ---- thread-block cluster / distributed shared memory (long label sequences: the lattice warps of one utterance and
// direction are split over the two CTAs of a cluster, see ctc_engine2c_kernel)
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
// shared-window address of the SAME variable in CTA `rank` of the cluster (a shared::cluster address)
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void stc_f64_if(uint32_t caddr, double v, bool pred) {  // store into a peer CTA's smem
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p st.shared::cluster.f64 [%0], %1;\n\t}" ::"r"(caddr),
                 "d"(v), "r"((int)pred));
}
__device__ __forceinline__ void stc_b32_if(uint32_t caddr, int v, bool pred) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p st.shared::cluster.b32 [%0], %1;\n\t}" ::"r"(caddr),
                 "r"(v), "r"((int)pred));
}
// arrive on an mbarrier of a peer CTA; release at cluster scope: this thread's earlier DSMEM stores are visible to
// whoever acquires the completed phase
__device__ __forceinline__ void mbar_arrive_remote_if(uint32_t caddr, bool pred) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %1, 0;\n\t"
                 "@p mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n\t}" ::"r"(caddr), "r"((int)pred)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t addr, uint32_t parity) {  // local mbarrier, cluster acquire
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!done);
}

// Link between the two CTAs of a cluster (CL instantiations of engine_lattice only).  The boundary values that cross
// from the last lattice warp of one CTA to the first of the other travel as DSMEM stores into a ring that lives in
// the RECEIVER's shared memory (stores are fire-and-forget; a remote load would put ~200 clk on the receiver), four
// groups deep (two SlotRings back to back), guarded by one "full" mbarrier per group slot in the receiver (the sender
// arrives remotely, release.cluster, once a group's 8 step values + boundary mantissa + exponent are written) and one
// "empty" mbarrier per slot in the sender (the receiver arrives remotely when it is done with the slot).
constexpr int kXSlots = 4;
struct EngLink {
    int rank;           // this CTA's rank in the cluster
    int gw0;            // global index of this CTA's lattice warp 0
    int nwa_l;          // lattice warps of this CTA that this utterance needs
    uint32_t xring;     // local shared address: SlotRing[2], the receive ring of this CTA
    uint32_t xbars;     // local shared address: uint64 full[kXSlots], then empty[kXSlots]
    double *Pm2;        // global [2]: mantissas of alpha_{T-1}(S-1), alpha_{T-1}(S-2)
    int *Pe2;           // global [2]: their exponents
};

// thread-local renormalisation + exponent settlement (see renorm / settle in ctc_device.cuh for the rules).
// bnd0 / bnd1: the mantissas the inflow-side neighbour needs next (alpha: a[K-1]; beta: a[0], a[1]); returned in
// pr0 / pr1 as received from the in-warp neighbour (the caller overrides the warp-edge lane from the slot ring).
template <int K, bool UP>
__device__ __forceinline__ double renorm_settle(Cells<K> &c, int E_edge, int lane, double &pr0, double &pr1) {
    int h[K];
#pragma unroll
    for (int k = 0; k < K; ++k) h[k] = __double2hiint(c.a[k]);
#pragma unroll
    for (int w = 1; w < K; w <<= 1)
#pragma unroll
        for (int k = 0; k + w < K; k += 2 * w) h[k] = max(h[k], h[k + w]);
    const int ex = h[0] >> 20;
    const bool live = ex != 0;
    const unsigned lm = __ballot_sync(kFull, live);
    const double sc = live ? pow2i(1023 - ex) : 0.0;
#pragma unroll
    for (int k = 0; k < K; ++k) c.a[k] *= sc;
    int E = live ? c.E - (1023 - ex) : kDeadE;
    if (UP) {
        pr0 = __shfl_up_sync(kFull, c.a[K - 1], 1);
    } else {
        pr0 = __shfl_down_sync(kFull, c.a[0], 1);
        pr1 = __shfl_down_sync(kFull, c.a[1], 1);
    }
    // nearest live thread strictly on the inflow side: its exponent is what a dead thread adopts AND what the inflow
    // neighbour of this thread ends up with (if that neighbour is dead, it adopts exactly this one)
    const unsigned side = UP ? (lm & ((1u << lane) - 1u)) : (lane == 31 ? 0u : (lm & ~((2u << lane) - 1u)));
    const int src = side ? (UP ? 31 - __clz(side) : __ffs(side) - 1) : lane;
    int En = __shfl_sync(kFull, E, src);
    if (!side) En = E_edge;
    if (!live) E = En;
    bool lifted = En != kDeadE && En - E > kLift;
    if (__any_sync(kFull, lifted)) {
        // rare: some thread sits more than 2^kLift below its inflow neighbour -- the general loop of settle()
        do {
            En = UP ? __shfl_up_sync(kFull, E, 1) : __shfl_down_sync(kFull, E, 1);
            if (UP ? (lane == 0) : (lane == 31)) En = E_edge;
            lifted = false;
            if (En != kDeadE && En - E > kLift) {
                const double s = pow2i(E - En);
#pragma unroll
                for (int k = 0; k < K; ++k) c.a[k] *= s;
                E = En;
                lifted = true;
            }
        } while (__any_sync(kFull, lifted));
        if (UP) {
            pr0 = __shfl_up_sync(kFull, c.a[K - 1], 1);
        } else {
            pr0 = __shfl_down_sync(kFull, c.a[0], 1);
            pr1 = __shfl_down_sync(kFull, c.a[1], 1);
        }
    }
    c.E = E;
    return (En == kDeadE) ? 0.0 : pow2i(En - E);
}

// 8 (FULL) or nr < 8 steps of the alpha recursion; y holds the emission values of the group's first row on entry.
// hb = slot-ring base of this group (0 or 8); send: lane 31 of a warp with an upper neighbour; recv: lane 0 of a warp
// with a lower neighbour.
template <int K, int CP, bool FULL, bool CL>
__device__ __forceinline__ void alpha_steps(Cells<K> &c, double &pr0, const double f, YRow<K> y, const unsigned *yt,
                                            int nr, int blank, const int (&idx)[K / 2], const double (&skipA)[K / 2],
                                            bool send, bool recv, uint32_t mine_m0, uint32_t in_m0, bool sendx,
                                            uint32_t peer_m0) {
#pragma unroll
    for (int r = 0; r < kGroup; ++r) {
        if (FULL || r < nr) {
            const bool last = FULL ? (r == kGroup - 1) : (r == nr - 1);
            YRow<K> yn;
            load_y<K, CP>(yn, yt, FULL ? min(r + 1, kGroup - 1) : min(r + 1, nr - 1), blank, idx);
            const double p1 = pr0 * f;
            const double top = alpha_top<K>(c.a, y, skipA);
            double sh = 0.0;
            if (!last) {
                sh = __shfl_up_sync(kFull, top, 1);
                sts_f64_if(mine_m0 + 8 * r, top, send);
                if (CL) stc_f64_if(peer_m0 + 8 * r, top, sendx);
            }
            alpha_rest<K>(c.a, top, p1, y, skipA);
            if (!last) {
                lds_f64_if(sh, in_m0 + 8 * r, recv);
                pr0 = sh;
            }
            y = yn;
        }
    }
}

template <int K, int CP, bool FULL, bool CL>
__device__ __forceinline__ void beta_steps(Cells<K> &c, double &pr0, double &pr1, const double f, YRow<K> y,
                                           const unsigned *yt, int nr, int blank, const int (&idx)[K / 2],
                                           const double (&skipB)[K / 2], bool send, bool recv, uint32_t mine_m0,
                                           uint32_t in_m0, bool sendx, uint32_t peer_m0) {
#pragma unroll
    for (int rr = 0; rr < kGroup; ++rr) {
        const int r = (FULL ? kGroup : nr) - 1 - rr;
        if (FULL || r >= 0) {
            const bool last = (r == 0);
            YRow<K> yn;
            load_y<K, CP>(yn, yt, max(r - 1, 0), blank, idx);
            const double n1 = pr0 * f, n2 = pr1 * f;
            double b0, b1, s0, s1v;
            beta_bottom<K>(c.a, y, skipB, b0, b1, s0, s1v);
            double sh0 = 0.0, sh1 = 0.0;
            if (!last) {
                sh0 = __shfl_down_sync(kFull, s0, 1);
                sh1 = __shfl_down_sync(kFull, s1v, 1);
                sts_f64_if(mine_m0 + 8 * r, s0, send);
                sts_f64_if(mine_m0 + kRingM1 + 8 * r, s1v, send);
                if (CL) {
                    stc_f64_if(peer_m0 + 8 * r, s0, sendx);
                    stc_f64_if(peer_m0 + kRingM1 + 8 * r, s1v, sendx);
                }
            }
            double bb[K];
            beta_rest<K>(c.a, bb, b0, b1, s0, s1v, n1, n2, y, skipB);
            if (!last) {
                lds_f64_if(sh0, in_m0 + 8 * r, recv);
                lds_f64_if(sh1, in_m0 + kRingM1 + 8 * r, recv);
                pr0 = sh0;
                pr1 = sh1;
            }
            y = yn;
        }
    }
}

// Two-part recursion.  A time block's gradient needs the alpha checkpoint at its start and the beta checkpoint at its
// end; both exist once each direction has passed the block, i.e. at max(x, 1 - x) of the recursion for a block at
// relative position x.  So the engines can be run in two launches -- groups [0, cut) of each direction's processing
// order, then [cut, n) -- and the gradient kernel launched in between for the blocks that are already complete (the
// middle half for cut = 3/4 n) runs under the second part instead of queueing up behind the whole recursion.  The
// systolic schedule restarts cleanly: every warp re-enters at its own super-step and all a thread carries across is
// its 8 mantissas, exponent, inflow scale and inflow values (kept in full fp64, so the two-part recursion is
// bit-identical to the one-part one).  Utterances with fewer than 8 groups do everything in part 1.
__host__ __device__ inline int engine_phase_cut(int ngroups) { return ngroups >= 8 ? (3 * ngroups + 3) / 4 : ngroups; }
__host__ __device__ inline void engine_phase_range(int phase, int ngroups, int &jb, int &je) {
    const int cut = engine_phase_cut(ngroups);
    jb = phase == 2 ? cut : 0;
    je = phase == 1 ? cut : ngroups;
}
constexpr int kEngStateFields = 12;  // a[8], pr0, pr1, f, E

// the lattice warps of one direction.  CL: the utterance's lattice warps are split over the two CTAs of a cluster
// (`warp` is then the CTA-local index, lk.gw0 + warp the global one); nwa = lattice warps the utterance needs in all.
template <int K, int CP, bool FWD, bool CL>
__device__ __forceinline__ void engine_lattice(const CtcParams &p, int b, int warp, int lane, int nwa, int Tb, int S,
                                               unsigned *ytile, int NS, SlotRing *rings, double *pm, int *pe,
                                               const EngLink &lk, int jb = 0, int je = -1) {
    constexpr int Cp = CP;
    const int NT = p.nwl * 32;
    const int gw = CL ? lk.gw0 + warp : warp;      // global lattice warp
    const int nwa_l = CL ? lk.nwa_l : nwa;         // lattice warps of this CTA
    const int nbar = (nwa_l + 1) * 32;
    const int thread = gw * 32 + lane;  // lattice thread index; owns states [thread*K, thread*K+K)
    const int ngroups = (Tb + kGroup - 1) / kGroup;
    if (je < 0) je = ngroups;  // groups [jb, je) of this direction's processing order
    const int nsuper = (je - jb) + nwa_l - 1;
    int idx[K / 2], lab_i[K / 2];
    double skipA[K / 2], skipB[K / 2];
    {
        const uint4 tl = p.tlab[(size_t)b * NT + thread];
        decode_tlab<K>(tl, idx, skipA, skipB, lab_i, false, 0);
        if (p.big) position_idx<K>(tl, thread, idx);
    }
    (void)lab_i;
    const int blank = p.bpos;

    Cells<K> c;
#pragma unroll
    for (int k = 0; k < K; ++k) c.a[k] = 0.0;
    c.E = 0;
    double pr0 = 0.0, pr1 = 0.0;
    const int q = FWD ? warp : (nwa_l - 1 - warp);  // position in this CTA's systolic pipeline
    const bool has_lower = gw > 0, has_upper = gw + 1 < nwa;
    const bool has_in = FWD ? has_lower : has_upper, has_out = FWD ? has_upper : has_lower;
    // does the inflow / outflow neighbour live in the peer CTA?
    const bool in_x = CL && has_in && (FWD ? warp == 0 : warp == nwa_l - 1);
    const bool out_x = CL && has_out && (FWD ? warp == nwa_l - 1 : warp == 0);
    // the warp-edge lane that talks to the neighbouring warp through the slot rings
    // (kept opaque: the compiler otherwise re-derives them from %tid inside the loop)
    const bool send_lane = FWD ? (lane == 31) : (lane == 0), recv_lane = FWD ? (lane == 0) : (lane == 31);
    const bool send = opaque_u32(send_lane && has_out && !out_x) != 0;   // into this CTA's own ring
    const bool sendx = CL && opaque_u32(send_lane && out_x) != 0;        // into the peer CTA's receive ring
    const bool recv = opaque_u32(recv_lane && has_in) != 0;
    const bool edge = recv_lane;
    // shared-window addresses of this warp's ring and of the ring it reads (the inflow-side neighbour's, or this
    // CTA's receive ring when that neighbour is in the peer CTA); only lanes / warps with such a neighbour use it
    const uint32_t mine = opaque_u32(smem_u32(rings + warp));
    const uint32_t in_l = opaque_u32(in_x ? lk.xring : smem_u32(rings + (FWD ? warp - 1 : warp + 1)));
    const uint32_t peer_ring = CL ? opaque_u32(mapa_u32(lk.xring, lk.rank ^ 1)) : 0u;
    // everything starts in scale 2^0; a thread without an inflow neighbour has f = 0 for good
    double f = (FWD ? (thread == 0) : (edge && !has_upper)) ? 0.0 : 1.0;
    if (FWD) {
        if (thread == 0) c.a[0] = 1.0;  // virtual alpha_{-1}: all mass in state 0
    } else {
        const int s1 = S - 1;  // virtual beta_T: 1 at state S-1
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (thread * K + k == s1) c.a[k] = 1.0;
        pr0 = __shfl_down_sync(kFull, c.a[0], 1);
        pr1 = __shfl_down_sync(kFull, c.a[1], 1);
        if (lane == 31) {
            pr0 = ((thread + 1) * K == s1) ? 1.0 : 0.0;  // S-1 is state 0 of the next thread
            pr1 = 0.0;
        }
    }
    double *est = nullptr;
    if (!CL && p.est) est = p.est + ((size_t)b * 2 + (FWD ? 0 : 1)) * kEngStateFields * NT + thread;
    if (!CL && jb > 0) {  // second part: continue from the state part 1 left
#pragma unroll
        for (int k = 0; k < K; ++k) c.a[k] = est[(size_t)k * NT];
        pr0 = est[(size_t)8 * NT];
        pr1 = est[(size_t)9 * NT];
        f = est[(size_t)10 * NT];
        c.E = (int)est[(size_t)11 * NT];
    }
    named_bar_sync(1, nbar);
    for (int i = 0; i < q; ++i) named_bar_sync(1, nbar);

    int st = 0;  // ring slot of this warp's current emission tile
    const unsigned *yt = ytile;
    const int ntail = Tb - (ngroups - 1) * kGroup;  // rows of the ragged last group (1..8)
    // checkpoints: alpha after row t0+7 -> block G+1 ; beta_{t0} -> block G-1 (see ctc_engine_kernel)
    int *ck = (FWD ? p.ckA : p.ckB) + (size_t)b * p.nblk * NT * K;
    int *ckE = (FWD ? p.ckAE : p.ckBE) + (size_t)b * p.nblk * NT;
    const int cb0 = FWD ? 1 + jb : ngroups - 2 - jb;
    int4 *ckp = reinterpret_cast<int4 *>(ck + ((ptrdiff_t)cb0 * NT + gw * 32) * K) + lane;
    int *ckEp = ckE + (ptrdiff_t)cb0 * NT + thread;
    const ptrdiff_t ckstep = (FWD ? 1 : -1) * (ptrdiff_t)NT * (K / 4), ckEstep = (FWD ? 1 : -1) * (ptrdiff_t)NT;
    const bool want_ck = !p.score_only;

    // emission values of the first row this direction visits
    YRow<K> y;
    {
        const int G0 = FWD ? jb : ngroups - 1 - jb;                 // first group processed here
        const int nr0 = (G0 == ngroups - 1) ? ntail : kGroup;       // its rows
        load_y<K, CP>(y, yt, FWD ? 0 : nr0 - 1, blank, idx);
    }
    for (int j = jb; j < je; ++j) {
        const bool tail = FWD ? (j == ngroups - 1) : (j == 0);
        const int nr = tail ? ntail : kGroup;
        // slots of this group's rows: the j-th group this direction processes uses half (j & 1) of a ring; the
        // cross-CTA ring is two rings back to back, group j in ring (j >> 1) & 1
        const int hb = (j & 1) * kGroup;
        const uint32_t xo = (uint32_t)((j >> 1) & 1) * (uint32_t)sizeof(SlotRing);
        const uint32_t mg = mine + 8 * hb;
        const uint32_t ig = in_l + (in_x ? xo : 0u) + 8 * hb;
        const uint32_t pg = CL ? peer_ring + xo + 8 * hb : 0u;
        // (CL: that the peer's values for group j have landed in the receive ring, and that the peer is done with the
        // slot this group's values go to, was established by this CTA's producer warp before the group barrier that
        // started this super-step -- the lattice warps themselves never touch a cluster-scope operation)
        if (FWD) {
            if (nr == kGroup)
                alpha_steps<K, CP, true, CL>(c, pr0, f, y, yt, kGroup, blank, idx, skipA, send, recv, mg, ig, sendx, pg);
            else
                alpha_steps<K, CP, false, CL>(c, pr0, f, y, yt, nr, blank, idx, skipA, send, recv, mg, ig, sendx, pg);
        } else {
            if (nr == kGroup)
                beta_steps<K, CP, true, CL>(c, pr0, pr1, f, y, yt, kGroup, blank, idx, skipB, send, recv, mg, ig, sendx,
                                            pg);
            else
                beta_steps<K, CP, false, CL>(c, pr0, pr1, f, y, yt, nr, blank, idx, skipB, send, recv, mg, ig, sendx,
                                             pg);
        }
        // ---- group boundary: slot index of the row the boundary values belong to (alpha: last row; beta: first)
        const int tl = hb + (FWD ? nr - 1 : 0);
        // settled exponent of the neighbouring warp's boundary thread (that warp ran this group a super-step ago);
        // every lane may need it: a lane whose inflow side holds no live thread inside the warp adopts it
        const uint32_t in_g = in_l + (in_x ? xo : 0u);
        const int E_edge = has_in ? lds_b32(in_g + kRingE + 4 * tl) : kDeadE;
        f = renorm_settle<K, FWD>(c, E_edge, lane, pr0, pr1);
        if (FWD) {
            sts_f64_if(mine + 8 * tl, c.a[K - 1], send);
            lds_f64_if(pr0, in_g + 8 * tl, recv);
        } else {
            sts_f64_if(mine + 8 * tl, c.a[0], send);
            sts_f64_if(mine + kRingM1 + 8 * tl, c.a[1], send);
            lds_f64_if(pr0, in_g + 8 * tl, recv);
            lds_f64_if(pr1, in_g + kRingM1 + 8 * tl, recv);
        }
        sts_b32_if(mine + kRingE + 4 * tl, c.E, send);
        if (CL) {
            const uint32_t pr = peer_ring + xo;
            if (FWD) {
                stc_f64_if(pr + 8 * tl, c.a[K - 1], sendx);
            } else {
                stc_f64_if(pr + 8 * tl, c.a[0], sendx);
                stc_f64_if(pr + kRingM1 + 8 * tl, c.a[1], sendx);
            }
            stc_b32_if(pr + kRingE + 4 * tl, c.E, sendx);
        }
        if (j < ngroups - 1) {
            if (want_ck) {
#pragma unroll
                for (int k = 0; k < K; k += 4)
                    ckp[(k / 4) * 32] = make_int4(hi_round(c.a[k]), hi_round(c.a[k + 1]), hi_round(c.a[k + 2]),
                                                  hi_round(c.a[k + 3]));
                *ckEp = c.E;
            }
            ckp += ckstep;
            ckEp += ckEstep;
            // next tile: its first row is fetched here, before the barrier (the producer guarantees tile j + 1 of
            // this warp has landed: it waits one tile further ahead than the lattice needs)
            yt += kGroup * kYP * Cp;
            if (++st == NS) { st = 0; yt = ytile; }
#ifndef CTC_V2_NOPREFETCH
            if (j + 1 < je) load_y<K, CP>(y, yt, FWD ? 0 : kGroup - 1, blank, idx);  // (the next part loads its own)
#endif
        } else if (FWD) {
            // ---- P = alpha_{T-1}(S-1) + alpha_{T-1}(S-2): the owning threads publish (mantissa, exponent)
            const int s1 = S - 1, s2 = S - 2;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (thread * K + k == s1) {
                    if (CL) { lk.Pm2[0] = c.a[k]; lk.Pe2[0] = c.E; } else { pm[0] = c.a[k]; pe[0] = c.E; }
                }
                if (thread * K + k == s2) {
                    if (CL) { lk.Pm2[1] = c.a[k]; lk.Pe2[1] = c.E; } else { pm[1] = c.a[k]; pe[1] = c.E; }
                }
            }
        }
        named_bar_sync(1, nbar);
#ifdef CTC_V2_NOPREFETCH
        if (j + 1 < je) load_y<K, CP>(y, yt, FWD ? 0 : kGroup - 1, blank, idx);
#endif
    }
    for (int i = q + (je - jb); i < nsuper; ++i) named_bar_sync(1, nbar);
    if (!CL && je < ngroups) {  // first part: hand the state to the second
#pragma unroll
        for (int k = 0; k < K; ++k) est[(size_t)k * NT] = c.a[k];
        est[(size_t)8 * NT] = pr0;
        est[(size_t)9 * NT] = pr1;
        est[(size_t)10 * NT] = f;
        est[(size_t)11 * NT] = (double)c.E;
    }
}

// cost and p(l|x) of one utterance from the published (mantissa, exponent) pairs of alpha_{T-1}(S-1), alpha_{T-1}(S-2)
__device__ __forceinline__ void engine_finish(const CtcParams &p, int b, int S, double m0, int e0, double m1, int e1) {
    double m = m0;
    int e = e0;
    if (S > 1) {
        const int em = max(e0, e1);
        m = m0 * pow2i(e0 - em) + m1 * pow2i(e1 - em);
        e = em;
    }
    const int ex = expo_field(m);
    const bool isnan_p = m != m;  // a non-finite emission row reached the lattice: cost and gradient are NaN
    const bool ok = ex != 0 && ex != 2047 && e > kDeadE / 2;
    if (ok) {
        m *= pow2i(1023 - ex);
        e += ex - 1023;
    }
    p.feas[b] = ok || isnan_p;
    p.Pm[b] = (ok || isnan_p) ? m : 1.0;
    p.Pe[b] = ok ? e : 0;
    const double cost = -(log(m) + (double)e * 0.69314718055994530942);
    p.costs[b] = isnan_p ? __int_as_float(0x7fc00000) : (ok ? (float)cost : (p.zero_infinity ? 0.f : INFINITY));
}

// emission-tile producer warp of the second engine form (TMA bulk copies into a ring of NS tiles, kLookAhead + 1
// tiles in flight; the tile the lattice may touch is waited for one super-step EARLIER than in ctc_engine_kernel:
// lattice warps fetch the first row of their next tile before the group barrier).  nwa_l: lattice warps of this CTA.
// CL (cluster form): the producer warp also performs every cluster-scope operation of the cross-CTA link, so that the
// lattice warps only issue plain DSMEM stores / local loads.  xrecv: this CTA's first pipeline warp reads from the
// peer (before the barrier that starts the super-step in which that warp runs group j the producer has seen the peer's
// "full" arrival for group j, acquire.cluster; after the barrier that ends it, it returns the slot with a remote
// "empty" arrival).  xsend: this CTA's last pipeline warp writes to the peer (before the barrier that starts group j
// there, j >= 4, the producer has seen the peer's "empty" arrival for group j - 4; after the barrier that ends it, it
// publishes the group with a remote "full" arrival, release.cluster -- the CTA barrier orders the lattice warp's
// DSMEM stores before it, and release is cumulative).
template <int CP, bool CL>
__device__ __forceinline__ void engine_producer(const CtcParams &p, int b, bool fwd, int lane, int nwa_l, int Tb,
                                                unsigned *ytile, uint64_t *y_full, int NS, const EngLink &lk,
                                                bool xrecv, bool xsend, int jb = 0, int je = -1) {
    constexpr int Cp = CP;
    const int ngroups_all = (Tb + kGroup - 1) / kGroup;
    if (je < 0) je = ngroups_all;
    const int ngroups = je - jb;  // tiles of this launch: tile j is group jb + j of the direction's processing order
    const int nsuper = ngroups + nwa_l - 1;
    const int nbar = (nwa_l + 1) * 32;
    const uint32_t full_l = CL ? lk.xbars : 0u, empty_l = CL ? lk.xbars + 8 * kXSlots : 0u;
    const uint32_t full_p = CL ? mapa_u32(full_l, lk.rank ^ 1) : 0u, empty_p = CL ? mapa_u32(empty_l, lk.rank ^ 1) : 0u;
    // group the link-side lattice warp runs in super-step i: the first pipeline warp (receiver) runs group i, the
    // last one (sender) group i - (nwa_l - 1)
    auto link_before = [&](int inext) {  // ahead of the barrier that STARTS super-step inext
        if (!CL) return;
        if (xrecv && inext < ngroups)
            mbar_wait_cluster(full_l + 8 * (inext & (kXSlots - 1)), (uint32_t)(inext / kXSlots) & 1u);
        const int js = inext - (nwa_l - 1);
        if (xsend && js >= kXSlots && js < ngroups)
            mbar_wait_cluster(empty_l + 8 * (js & (kXSlots - 1)), (uint32_t)(js / kXSlots - 1) & 1u);
    };
    auto link_after = [&](int idone) {  // behind the barrier that ENDS super-step idone
        if (!CL) return;
        if (xrecv && idone < ngroups) mbar_arrive_remote_if(empty_p + 8 * (idone & (kXSlots - 1)), lane == 0);
        const int js = idone - (nwa_l - 1);
        if (xsend && js >= 0 && js < ngroups) mbar_arrive_remote_if(full_p + 8 * (js & (kXSlots - 1)), lane == 0);
    };
    if (lane == 0) {
        for (int i = 0; i < NS; ++i) mbar_init(&y_full[i], 1);
        fence_mbar_init();
    }
    __syncwarp();
    const unsigned *ybase = p.y + (size_t)b * p.T * kYP * Cp;
    const uint32_t rowbytes = (uint32_t)(kYP * Cp * sizeof(unsigned));
    int ist = 0;
    auto issue = [&](int j) {
        const int st = ist;
        ist = (ist + 1 == NS) ? 0 : ist + 1;
        if (j < ngroups && lane == 0) {
            const int G = fwd ? (jb + j) : (ngroups_all - 1 - (jb + j));
            const int t0 = G * kGroup;
            const int nr = min(kGroup, Tb - t0);
            mbar_arrive_expect_tx(&y_full[st], rowbytes * nr);
            tma_bulk_g2s(ytile + (size_t)(st * kGroup) * kYP * Cp, ybase + (size_t)t0 * kYP * Cp, rowbytes * nr,
                         &y_full[st]);
        }
    };
    int wst = 0, wph = 0, nwaited = 0;
    auto wait_tile = [&]() {
        if (nwaited < ngroups) {
            mbar_wait(&y_full[wst], wph);
            if (++wst == NS) { wst = 0; wph ^= 1; }
            ++nwaited;
        }
    };
    constexpr int LA = look_ahead(CP);
    for (int j = 0; j <= LA; ++j) issue(j);
    wait_tile();  // tile 0
    wait_tile();  // tile 1: its first row is read during super-step 0
    link_before(0);
    named_bar_sync(1, nbar);
    for (int i = 0; i < nsuper; ++i) {
        fence_proxy_async();
        issue(i + LA + 1);
        wait_tile();  // tile i + 2
        link_before(i + 1);
        named_bar_sync(1, nbar);
        link_after(i);
    }
}

template <int K, int CP>
__global__ void __launch_bounds__((kMaxLatticeWarps + 1) * 32) ctc_engine2_kernel(const CtcParams p) {
    static_assert(K == 8, "label-table entries hold 4 labels = 8 states");
    extern __shared__ __align__(16) unsigned char smem[];
    const int bi = p.slice0 + (p.score_only ? blockIdx.x : (blockIdx.x >> 1));
    const int b = p.bord ? p.bord[bi] : bi;
    const bool fwd = p.score_only ? true : !(blockIdx.x & 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwl = p.nwl;
    const EngineSmem lay = engine_smem(CP, nwl);
    const int NS = lay.nstage;
    unsigned *ytile = reinterpret_cast<unsigned *>(smem + lay.ytile);
    uint64_t *y_full = reinterpret_cast<uint64_t *>(smem + lay.bars);
    SlotRing *rings = reinterpret_cast<SlotRing *>(smem + lay.slots);
    double *pm = reinterpret_cast<double *>(smem + lay.misc);  // [2]
    int *pe = reinterpret_cast<int *>(smem + lay.misc + 16);   // [2]

    const int Tb = min(max(p.act_lens[b], 0), p.T);
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int S = 2 * L + 1;
    const int nwa = min(nwl, (S + 32 * K - 1) / (32 * K));  // lattice warps this utterance needs

    if (Tb == 0) {  // nothing to recurse over
        if (fwd && tid == 0 && p.phase != 2) {
            const bool ok = (L == 0);
            p.feas[b] = ok;
            p.Pm[b] = 1.0;
            p.Pe[b] = 0;
            p.costs[b] = ok ? 0.f : (p.zero_infinity ? 0.f : INFINITY);
        }
        return;
    }
    if (warp >= nwa && warp != nwl) return;  // lattice warps beyond this utterance's S
    const int ngroups = (Tb + kGroup - 1) / kGroup;
    int jb, je;
    engine_phase_range(p.phase, ngroups, jb, je);
    if (jb >= je) return;  // second part of an utterance that did everything in the first
    if (warp == nwl) {
        const EngLink nolink = {0, 0, nwa, 0u, 0u, nullptr, nullptr};
        engine_producer<CP, false>(p, b, fwd, lane, nwa, Tb, ytile, y_full, NS, nolink, false, false, jb, je);
        return;
    }
    const EngLink nolink = {0, 0, nwa, 0u, 0u, nullptr, nullptr};
    if (fwd)
        engine_lattice<K, CP, true, false>(p, b, warp, lane, nwa, Tb, S, ytile, NS, rings, pm, pe, nolink, jb, je);
    else
        engine_lattice<K, CP, false, false>(p, b, warp, lane, nwa, Tb, S, ytile, NS, rings, pm, pe, nolink, jb, je);
    if (fwd && tid == 0 && je == ngroups) engine_finish(p, b, S, pm[0], pe[0], pm[1], pe[1]);
}

// Long label sequences (more than 4 lattice warps, L > 511): the lattice warps of one (utterance, direction) are split
// over the TWO CTAs of a thread-block cluster, so that every lattice warp has a scheduler (SMSP) and an FP64 pipe to
// itself (5 warps in one CTA put two of them on one SMSP: +21 % per step on the first engine form).  Each CTA runs
// its own systolic pipeline over its W = ceil(nwl / 2) warps with its own producer warp, emission-tile ring and
// group barrier; the one boundary that crosses the CTAs goes through distributed shared memory (EngLink): DSMEM
// stores into the receiver's ring + remote mbarrier arrivals, no cluster-wide barrier inside the loop.  The systolic
// skew means the receiver consumes a group a whole super-step after it was sent, so neither the ~200 clk DSMEM
// latency nor the mbarrier wake-up sits on the critical path.
template <int K, int CP>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__((kMaxLatticeWarps / 2 + 1) * 32)
    ctc_engine2c_kernel(const CtcParams p) {
    static_assert(K == 8, "label-table entries hold 4 labels = 8 states");
    extern __shared__ __align__(16) unsigned char smem[];
    const int rank = (int)cluster_ctarank();
    const int ud = (int)(blockIdx.x >> 1);  // (utterance, direction) of this cluster
    const int bi = p.slice0 + (p.score_only ? ud : (ud >> 1));
    const int b = p.bord ? p.bord[bi] : bi;
    const bool fwd = p.score_only ? true : !(ud & 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwl = p.nwl, W = (nwl + 1) >> 1;  // lattice warps per CTA
    const EngineSmem lay = engine_smem(CP, W);
    const int NS = lay.nstage;
    unsigned *ytile = reinterpret_cast<unsigned *>(smem + lay.ytile);
    uint64_t *y_full = reinterpret_cast<uint64_t *>(smem + lay.bars);
    SlotRing *rings = reinterpret_cast<SlotRing *>(smem + lay.slots);
    unsigned char *xr = smem + lay.xlink;  // SlotRing[2] receive ring, then full[4] / empty[4] mbarriers

    const int Tb = min(max(p.act_lens[b], 0), p.T);
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int S = 2 * L + 1;
    const int nwa = min(nwl, (S + 32 * K - 1) / (32 * K));  // lattice warps this utterance needs (both CTAs)
    const int nwa_l = min(max(nwa - rank * W, 0), W);       // ... of which this CTA runs
    if (Tb == 0) {  // nothing to recurse over (the whole cluster leaves: no cluster barrier has been touched yet)
        if (fwd && tid == 0 && rank == 0) {
            const bool ok = (L == 0);
            p.feas[b] = ok;
            p.Pm[b] = 1.0;
            p.Pe[b] = 0;
            p.costs[b] = ok ? 0.f : (p.zero_infinity ? 0.f : INFINITY);
        }
        return;
    }
    EngLink lk;
    lk.rank = rank;
    lk.gw0 = rank * W;
    lk.nwa_l = nwa_l;
    lk.xring = smem_u32(xr);
    lk.xbars = smem_u32(xr + 2 * sizeof(SlotRing));
    lk.Pm2 = p.Pm2 + 2 * (size_t)b;
    lk.Pe2 = p.Pe2 + 2 * (size_t)b;
    if (tid == 0) {
        uint64_t *xb = reinterpret_cast<uint64_t *>(xr + 2 * sizeof(SlotRing));
        for (int i = 0; i < 2 * kXSlots; ++i) mbar_init(&xb[i], 1);
        fence_mbar_init();
    }
    cluster_sync_all();  // both CTAs' link barriers exist before either sends
    if (warp < nwa_l) {
        if (fwd)
            engine_lattice<K, CP, true, true>(p, b, warp, lane, nwa, Tb, S, ytile, NS, rings, nullptr, nullptr, lk);
        else
            engine_lattice<K, CP, false, true>(p, b, warp, lane, nwa, Tb, S, ytile, NS, rings, nullptr, nullptr, lk);
    } else if (warp == W && nwa_l > 0) {
        // the link exists iff the utterance needs warps in both CTAs; alpha flows rank 0 -> 1, beta rank 1 -> 0
        const bool linked = nwa > W;
        const bool xrecv = linked && (fwd ? rank == 1 : rank == 0), xsend = linked && (fwd ? rank == 0 : rank == 1);
        engine_producer<CP, true>(p, b, fwd, lane, nwa_l, Tb, ytile, y_full, NS, lk, xrecv, xsend);
    }
    // no CTA leaves while its peer may still store into its shared memory; also publishes Pm2 / Pe2 (global memory,
    // written by whichever CTA owns states S-1 and S-2) to rank 0
    cluster_sync_all();
    if (fwd && rank == 0 && tid == 0) engine_finish(p, b, S, lk.Pm2[0], lk.Pe2[0], lk.Pm2[1], lk.Pe2[1]);
}

// ================================================================================================ grad
// One CTA per (utterance, every nsplit-th group of 8 time steps).  Warp q owns the 224 lattice states
// [224q, 224q+224) (28 core threads) and recomputes a HALO of 16 states (2 threads) on either side, which is exactly
// how far information travels in 8 steps (2 states per step): the warps of a CTA never exchange boundary values and
// synchronise only around the per-class reduction.  The checkpoints are renormalised and exponent-settled, so inside
// a group every thread works with ONE alpha exponent and ONE beta exponent.  The next group's checkpoints
// (registers) and emission tile (TMA, double buffer) are prefetched during the current one.
constexpr int kMaxGradWarps = 10;  // ceil((2*1023+1) / 224)
#ifndef CTC_OCC_COPIES
#define CTC_OCC_COPIES 4
#endif
// occupancy tables per CTA (lane % copies picks one: fewer same-address collisions).  Gradient kernel alone, B=64 /
// B=1024: 1 copy 84.1 / 909 us, 2: 86.2 / 948, 4: 82.0 / 885, 8: 82.0 / 913
constexpr int kOccCopies = CTC_OCC_COPIES;

struct GradSmem {
    size_t ytile, bar, psort, arow, total;
    int pitch;
};
__host__ __device__ inline GradSmem grad_smem(int C, int Cp, int Lmax, int nq, int K, bool big) {
    GradSmem s;
    s.pitch = grad_row_pitch(Lmax, nq);
    size_t o = 0;
    s.ytile = o; o = align16(o + sizeof(unsigned) * 2 * kGroup * kYP * Cp);
    s.bar = o;   o = align16(o + 2 * sizeof(uint64_t));
    {
        const size_t rows = sizeof(float) * kGroup * s.pitch, occ = sizeof(unsigned) * kOccCopies * (kGroup * Cp + 8);
        s.psort = o; o = align16(o + (big ? rows : occ));  // small alphabets: occupancy tables instead of posterior rows
    }
    s.arow = o;  o = align16(o + sizeof(int) * (size_t)nq * kGroup * 32 * K);
    s.total = o;
    return s;
}

// Rows parked in shared memory are only ever multiplied into posteriors (the recursions continue from registers in
// full fp64), so they are stored as the HIGH WORD of the double, rounded to nearest: the complete exponent range and
// 21 significant bits (relative error <= 2^-22 on a posterior <= 1, i.e. < 2.4e-7 on a gradient entry) at half the
// shared memory and half the load/store wavefronts -- which is what lets a third CTA fit on the SM.
// Thread-private slots, [K/4][32 lanes] int4: conflict-free 16-byte accesses.
template <int K>
__device__ __forceinline__ void park_row(int *row, int lane, const double (&v)[K]) {
#pragma unroll
    for (int h = 0; h < K / 4; ++h) {
        int4 w;
        w.x = hi_round(v[4 * h + 0]);
        w.y = hi_round(v[4 * h + 1]);
        w.z = hi_round(v[4 * h + 2]);
        w.w = hi_round(v[4 * h + 3]);
        *reinterpret_cast<int4 *>(row + (h * 32 + lane) * 4) = w;
    }
}
template <int K>
__device__ __forceinline__ void fetch_row(const int *row, int lane, double (&v)[K]) {
#pragma unroll
    for (int h = 0; h < K / 4; ++h) {
        const int4 w = *reinterpret_cast<const int4 *>(row + (h * 32 + lane) * 4);
        v[4 * h + 0] = __hiloint2double(w.x, 0);
        v[4 * h + 1] = __hiloint2double(w.y, 0);
        v[4 * h + 2] = __hiloint2double(w.z, 0);
        v[4 * h + 3] = __hiloint2double(w.w, 0);
    }
}

// alpha rows of one group recomputed into shared memory
template <int K, int CP, bool FULL>
__device__ __forceinline__ void grad_alpha_rows(Cells<K> &c, double pr0, double f, const unsigned *yt, int nrow,
                                                int blank, const int (&idx)[K / 2], const double (&skipA)[K / 2],
                                                int lane, int *arow) {
    YRow<K> y;
    load_y<K, CP>(y, yt, 0, blank, idx);
#pragma unroll
    for (int r = 0; r < kGroup; ++r) {
        if (FULL || r < nrow) {
            YRow<K> yn;
            load_y<K, CP>(yn, yt, FULL ? min(r + 1, kGroup - 1) : min(r + 1, nrow - 1), blank, idx);
            const double p1 = pr0 * f;
            const double top = alpha_top<K>(c.a, y, skipA);
            pr0 = __shfl_up_sync(kFull, top, 1);
            if (lane == 0) pr0 = 0.0;
            alpha_rest<K>(c.a, top, p1, y, skipA);
            park_row<K>(arow + r * 32 * K, lane, c.a);
            y = yn;
        }
    }
}

// beta sweep over one group with posteriors: label-state posteriors -> psort (class-sorted slots), blank-state
// posteriors summed per thread -> bpart
template <int K, int CP, bool FULL, bool OCC>
__device__ __forceinline__ void grad_beta_rows(Cells<K> &c, double pr0, double pr1, double f, double g,
                                               const unsigned *yt, int nrow, int blank, const int (&idx)[K / 2],
                                               const double (&skipB)[K / 2], const int (&lpos)[K / 2], bool core,
                                               int lane, const int *arow, float *psort, int Lp, float *bpart_w) {
    YRow<K> y;
    load_y<K, CP>(y, yt, nrow - 1, blank, idx);
#pragma unroll
    for (int qq = 0; qq < kGroup; ++qq) {
        const int r = (FULL ? kGroup : nrow) - 1 - qq;
        if (FULL || r >= 0) {
            YRow<K> yn;
            load_y<K, CP>(yn, yt, max(r - 1, 0), blank, idx);
            double av[K];
            fetch_row<K>(arow + r * 32 * K, lane, av);
            const double n1 = pr0 * f, n2 = pr1 * f;
            double b0, b1, s0, s1v;
            beta_bottom<K>(c.a, y, skipB, b0, b1, s0, s1v);
            pr0 = __shfl_down_sync(kFull, s0, 1);
            pr1 = __shfl_down_sync(kFull, s1v, 1);
            if (lane == 31) { pr0 = 0.0; pr1 = 0.0; }
            double bb[K];
            beta_rest<K>(c.a, bb, b0, b1, s0, s1v, n1, n2, y, skipB);
            double bsum = 0.0;
            if (OCC) {
                // small alphabets (CTC_GRAD_OCC): occupancies accumulate as 2^-30 fixed point in [row][class] tables,
                // integer shared-memory reductions (order-independent, hence deterministic); g carries the 2^30;
                // round-to-nearest through the low word of (x + 2^52): one DFMA, no conversion instruction
                unsigned *occ = reinterpret_cast<unsigned *>(psort) + r * CP + (lane % kOccCopies) * (kGroup * CP + 8);
#pragma unroll
                for (int h = 0; h < K / 2; ++h) {
                    bsum = fma(av[2 * h], bb[2 * h], bsum);
                    const double pl = fma(av[2 * h + 1] * bb[2 * h + 1], g, 4503599627370496.0);
                    // (timing experiment: with these reductions compiled out the kernel is 9-11 % faster -- 85.0 ->
                    // 75.8 us at B=64, 1096 -> 995 at B=1024 -- so that is all a cheaper scheme could bring)
                    if (lpos[h] >= 0) atomicAdd(occ + idx[h], (unsigned)__double2loint(pl));
                }
                const unsigned bv =
                    __reduce_add_sync(kFull, core ? (unsigned)__double2loint(fma(bsum, g, 4503599627370496.0)) : 0u);
                if (lane == 0) atomicAdd(occ + blank, bv);
            } else {
#pragma unroll
                for (int h = 0; h < K / 2; ++h) {
                    bsum = fma(av[2 * h], bb[2 * h], bsum);
                    const double pl = av[2 * h + 1] * bb[2 * h + 1] * g;
                    if (lpos[h] >= 0) psort[r * Lp + lpos[h]] = (float)pl;
                }
                bpart_w[r * Lp + lane] = core ? (float)(bsum * g) : 0.f;
            }
            y = yn;
        }
    }
}

// 128 registers x 128 threads and ~56 KB of shared memory per CTA on the headline shape: 4 CTAs (16 warps) per SM.
// The kernel is latency-bound per warp, so residency is what buys throughput (measured: 2 CTAs/SM 172 us, 3: 123 us,
// 4: 103 us); an interleaved alpha/beta "meet in the middle" variant needed 168 registers and lost to this one.
#ifndef CTC_GRAD_MAXREG
#define CTC_GRAD_MAXREG 128
#endif
#define CTC_GRAD_ATTR __maxnreg__(CTC_GRAD_MAXREG)
template <int K, int CP, bool BIG>
__global__ void CTC_GRAD_ATTR ctc_grad_kernel(const CtcParams p) {
    static_assert(K == 8, "the label table and the 224-state chunking are laid out for K = 8");
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int R = kGroup, Cp = CP;
    const int z = blockIdx.x, b = p.bord ? p.bord[p.slice0 + blockIdx.y] : p.slice0 + (int)blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, q = tid >> 5;
    const int C = p.C, NT = p.nwl * 32;
    const int Tb = min(max(p.act_lens[b], 0), p.T);
    const int nblk_b = (Tb + R - 1) / R;
    if (z >= nblk_b) return;  // padding frames were zero-filled by the prep kernel
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int S = 2 * L + 1;
    const int nq = (S + K * kCoreThreads - 1) / (K * kCoreThreads);  // lattice warps of this utterance
    if (nq < p.gq_lo || nq > p.nqmax) return;  // another warp class's launch takes it
    const size_t rowstride = (size_t)p.B * C;
    if (!p.feas[b]) {  // infeasible utterance: zero gradient
        for (int blk = z; blk < nblk_b; blk += p.nsplit)
            for (int r = q; r < min(R, Tb - blk * R); r += blockDim.x / 32)
                for (int c = lane; c < C; c += 32) p.grads[((size_t)(blk * R + r) * p.B + b) * C + c] = 0.f;
        return;
    }
    const GradSmem lay = grad_smem(C, Cp, p.gLmax, p.nqmax, K, BIG);
    unsigned *ytile = reinterpret_cast<unsigned *>(smem + lay.ytile);
    uint64_t *ybar = reinterpret_cast<uint64_t *>(smem + lay.bar);
    float *psort = reinterpret_cast<float *>(smem + lay.psort);
    int *arow = reinterpret_cast<int *>(smem + lay.arow) + (size_t)q * R * 32 * K;
    const int Lp = lay.pitch;                    // pitch of a posterior row
    const int Lb = grad_blank_off(L, p.Lmax);    // where the blank partial sums start inside a row
    float *bpart = psort + Lb;                   // [row * Lp + warp * 32 + lane]
    const uint32_t rowbytes = (uint32_t)(kYP * Cp * sizeof(unsigned));
    const unsigned *ybase = p.y + (size_t)b * p.T * kYP * Cp;

    if (tid == 0) {
        mbar_init(&ybar[0], 1);
        mbar_init(&ybar[1], 1);
        fence_mbar_init();
        const int nr0 = min(R, Tb - z * R);
        mbar_arrive_expect_tx(&ybar[0], rowbytes * nr0);
        tma_bulk_g2s(ytile, ybase + (size_t)(z * R) * kYP * Cp, rowbytes * nr0, &ybar[0]);
    }
    constexpr bool kOcc = !BIG;  // small alphabets: occupancies accumulate in fixed-point [row][class] tables
    if (kOcc)
        for (int i = tid; i < kOccCopies * (R * CP + 8); i += blockDim.x) reinterpret_cast<unsigned *>(psort)[i] = 0u;
    __syncthreads();  // mbarrier init + cleared occupancy tables visible
    if (q >= nq) return;  // warps beyond this utterance's states
    const int nbar = nq * 32;

    // ---- per-thread constants: engine thread e = 28q - 2 + lane owns states [8e, 8e+8)
    const int e = kCoreThreads * q - 2 + lane;
    // the engine runs (and checkpoints) only the warps this utterance needs: threads beyond them hold no state
    const int ne = min(p.nwl, (S + 32 * K - 1) / (32 * K)) * 32;
    const bool valid_e = e >= 0 && e < ne;
    const bool core = valid_e && lane >= 2 && lane < 2 + kCoreThreads;
    int idx[K / 2], lpos[K / 2];
    double skipA[K / 2], skipB[K / 2];
    {
        uint4 tl = make_uint4(0x01010101u * (unsigned)p.bpos, 0u, 0u, 0u);
        if (valid_e) tl = p.tlab[(size_t)b * NT + e];
        decode_tlab<K>(tl, idx, skipA, skipB, lpos, core);
        if (BIG && valid_e) position_idx<K>(tl, e, idx);
    }
    const double invP = 1.0 / p.Pm[b];
    const int eP = p.Pe[b];
    const int s_last = S - 1;

    // checkpoint registers of the NEXT group to process (prefetched one group ahead)
    int4 nA[K / 4], nB[K / 4];
    int nEa = 0, nEb = 0;
    auto fetch = [&](int blk) {
        const size_t cb = ((size_t)b * p.nblk + blk) * NT;
        const bool hasA = valid_e && blk > 0, hasB = valid_e && (blk + 1) * R < Tb;
        const size_t base = (cb + (e & ~31)) * K;
#pragma unroll
        for (int h = 0; h < K / 4; ++h) {
            nA[h] = make_int4(0, 0, 0, 0);
            nB[h] = make_int4(0, 0, 0, 0);
            if (hasA) nA[h] = *reinterpret_cast<const int4 *>(p.ckA + base + (h * 32 + (e & 31)) * 4);
            if (hasB) nB[h] = *reinterpret_cast<const int4 *>(p.ckB + base + (h * 32 + (e & 31)) * 4);
        }
        nEa = hasA ? p.ckAE[cb + e] : 0;
        nEb = hasB ? p.ckBE[cb + e] : 0;
    };
    fetch(z);

    PROF(long long g0 = prof_clk(); long long ga = 0; long long gb = 0; long long gw = 0; long long gr = 0; long long gs = 0; long long gb1 = 0; long long gb2 = 0;)
    int it = 0;
    for (int blk = z; blk < nblk_b; blk += p.nsplit, ++it) {
        PROF(long long h0 = prof_clk();)
        const int cur = it & 1;
        const int t0 = blk * R;
        const int nrow = min(R, Tb - t0);
        const unsigned *yt = ytile + (size_t)cur * R * kYP * Cp;
        // ---- current checkpoints out of the prefetch registers, then prefetch the next group
        Cells<K> c;
        double bsave[K];
#pragma unroll
        for (int h = 0; h < K / 4; ++h) {
            c.a[4 * h + 0] = __hiloint2double(nA[h].x, 0); bsave[4 * h + 0] = __hiloint2double(nB[h].x, 0);
            c.a[4 * h + 1] = __hiloint2double(nA[h].y, 0); bsave[4 * h + 1] = __hiloint2double(nB[h].y, 0);
            c.a[4 * h + 2] = __hiloint2double(nA[h].z, 0); bsave[4 * h + 2] = __hiloint2double(nB[h].z, 0);
            c.a[4 * h + 3] = __hiloint2double(nA[h].w, 0); bsave[4 * h + 3] = __hiloint2double(nB[h].w, 0);
        }
        c.E = nEa;
        const int Eb0 = nEb;
        const int nblk_next = blk + p.nsplit;
        if (nblk_next < nblk_b) {
            if (tid == 0) {
                const int nrn = min(R, Tb - nblk_next * R);
                fence_proxy_async();
                mbar_arrive_expect_tx(&ybar[cur ^ 1], rowbytes * nrn);
                tma_bulk_g2s(ytile + (size_t)(cur ^ 1) * R * kYP * Cp, ybase + (size_t)(nblk_next * R) * kYP * Cp,
                             rowbytes * nrn, &ybar[cur ^ 1]);
            }
            fetch(nblk_next);
        }
        // ---- alpha state "after row t0-1" and beta state "after row t0+nrow" with their inflows
        double pa, fa, pb0, pb1, fb;
        {
            const int En = __shfl_up_sync(kFull, c.E, 1);
            pa = __shfl_up_sync(kFull, c.a[K - 1], 1);
            fa = (En == kDeadE) ? 0.0 : pow2i(En - c.E);
            if (lane == 0) pa = 0.0;
            if (blk == 0) {  // everything in scale 2^0; virtual alpha_{-1}(-1) = 1 flows into state 0
                fa = 1.0;
                pa = (e == 0) ? 1.0 : 0.0;
            }
        }
        Cells<K> cbeta;
        if (t0 + nrow == Tb) {  // virtual beta_T: 1 at state S-1, everything in scale 2^0
            cbeta.E = 0;
#pragma unroll
            for (int k = 0; k < K; ++k) cbeta.a[k] = (valid_e && e * K + k == s_last) ? 1.0 : 0.0;
            fb = 1.0;
        } else {
#pragma unroll
            for (int k = 0; k < K; ++k) cbeta.a[k] = bsave[k];
            cbeta.E = Eb0;
            const int En = __shfl_down_sync(kFull, cbeta.E, 1);
            fb = (En == kDeadE) ? 0.0 : pow2i(En - cbeta.E);
        }
        pb0 = __shfl_down_sync(kFull, cbeta.a[0], 1);
        pb1 = __shfl_down_sync(kFull, cbeta.a[1], 1);
        if (lane == 31) { pb0 = 0.0; pb1 = 0.0; }
        // posterior(t,s) = alpha_t(s) * betabar_t(s) / P ; both exponents are constant over the group
        // (kOcc: times 2^30, the fixed-point scale of the occupancy tables)
        const double g = pow2i(min(max(c.E, -(1 << 28)) + max(cbeta.E, -(1 << 28)) - eP, 960) + (kOcc ? 30 : 0)) * invP;
        float *bpart_w = bpart + q * 32;
        PROF(long long h1 = prof_clk();)
        mbar_wait(&ybar[cur], (it >> 1) & 1);
        PROF(long long h2 = prof_clk(); long long h3 = h2;)
        if (nrow == R) {
            grad_alpha_rows<K, CP, true>(c, pa, fa, yt, nrow, p.bpos, idx, skipA, lane, arow);
            PROF(h3 = prof_clk();)
            grad_beta_rows<K, CP, true, kOcc>(cbeta, pb0, pb1, fb, g, yt, nrow, p.bpos, idx, skipB, lpos, core, lane, arow,
                                        psort, Lp, bpart_w);
        } else {
            grad_alpha_rows<K, CP, false>(c, pa, fa, yt, nrow, p.bpos, idx, skipA, lane, arow);
            PROF(h3 = prof_clk();)
            grad_beta_rows<K, CP, false, kOcc>(cbeta, pb0, pb1, fb, g, yt, nrow, p.bpos, idx, skipB, lpos, core, lane, arow,
                                         psort, Lp, bpart_w);
        }
        PROF(long long h4 = prof_clk();)
        named_bar_sync(1, nbar);
        PROF(float probe = *reinterpret_cast<volatile float *>(bpart + lane); long long h4b = prof_clk() + (probe > 1e30f ? 1 : 0); gb1 += h4b - h4;)
        // ---- per-class occupancy, gradient rows written once
        if (kOcc) {
            // the occupancy tables are complete: gradient rows leave as coalesced stores, the tables are cleared for
            // the next group.  A NaN likelihood (NaN / Inf logits) makes every gradient of the utterance NaN.
            const float scale = p.grad_scale;
            unsigned *occ = reinterpret_cast<unsigned *>(psort);
            const float bad = (invP == invP) ? 0.f : __int_as_float(0x7fc00000);
            for (int i = tid; i < R * CP; i += nbar) {
                const int r = i / CP, cc = i % CP;
                unsigned v = 0;
#pragma unroll
                for (int k = 0; k < kOccCopies; ++k) {
                    v += occ[i + k * (R * CP + 8)];
                    occ[i + k * (R * CP + 8)] = 0u;
                }
                if (cc < C && r < nrow)
                    p.grads[((size_t)(t0 + r) * p.B + b) * C + cc] =
                        scale * (y_of_tile(yt, r, Cp, cc) -
                                 ((float)v * (1.f / 1073741824.f) + bad));
            }
        } else {
            // large alphabets: the dense rows scale * softmax are already in place (ctc_prep_big_kernel); subtract the
            // occupancy of each class on the lattice.  One reduction (RED.ADD, no return value, so no DRAM round trip
            // on this thread) per (row, distinct class): a single addend per address, hence deterministic and equal
            // to g - scale * occupancy bit for bit.  Repeated classes sit in adjacent class-sorted slots and are
            // summed in slot order first.
            const float nscale = -p.grad_scale;
            const int2 *ct = p.ctab + (size_t)b * p.Lmax;
            for (int r = q; r < nrow; r += nq) {
                float *grow = p.grads + ((size_t)(t0 + r) * p.B + b) * C;
                const float *prow = psort + r * Lp;
                float ob = 0.f;
                for (int j = 0; j < nq; ++j) ob += bpart[r * Lp + j * 32 + lane];
#pragma unroll
                for (int d = 16; d; d >>= 1) ob += __shfl_xor_sync(kFull, ob, d);
                if (lane == 0) atomicAdd(grow + p.blank, nscale * ob);
                for (int j = lane; j < L; j += 32) {
                    const int2 e = __ldg(ct + j);
                    if (e.y > 0) {
                        float sum = prow[j];
                        for (int k = 1; k < e.y; ++k) sum += prow[j + k];
                        atomicAdd(grow + e.x, nscale * sum);
                    }
                }
            }
        }
        PROF(long long h5 = prof_clk();)
        named_bar_sync(1, nbar);  // occupancy tables / posterior rows / emission buffer are reused by the next group
        PROF(long long h6 = prof_clk(); gs += h1 - h0; gw += h2 - h1; ga += h3 - h2; gb += h4 - h3; gr += h5 - h4b; gb2 += h6 - h5;)
    }
    PROF(if (lane == 0 && blockIdx.x == 0 && blockIdx.y == 0) { long long *d = p.dbg + (32 + q) * 8; d[0] = prof_clk() - g0; d[1] = gs; d[2] = gw; d[3] = ga; d[4] = gb; d[5] = gr; d[6] = it; d[7] = gb1 * 1000000 + gb2; })
    (void)rowstride;
}


// ================================================================================================ large alphabets
// C > 128 (BPE-style vocabularies): the softmax is the HBM-bound part (the row is C floats, the lattice touches only
// 1 + L of them), so every logit is read ONCE and every gradient element is written ONCE:
//   ctc_prep_big_kernel  one CTA per (t,b) row: float4 streaming loads into a shared-memory copy of the row, max and
//                        sum by warp shuffles + a shared-memory step, the dense part of the gradient
//                        scale * softmax written straight from that copy (streaming float4 stores), and the 1 + L
//                        emissions the lattice needs gathered into a compact POSITION-indexed row (0: blank, 1+i:
//                        label i), so the lattice kernels run unchanged with pitch 256 (L <= 255).
//   ctc_grad_kernel<BIG> subtracts the <= 1 + L occupancies of a row in place with one fire-and-forget reduction
//                        per distinct class (a sparse read-modify-write of 32-byte sectors done by the L2,
//                        ~ (1+L) * 64 B per row against 8 * C).
// DRAM traffic per valid frame: 4C read + 4C written + ~64 (1+L) + the position rows; per padding frame 4C written.
constexpr int kBigThreads = 256;
constexpr int kBigStageMaxC = 16384;  // rows up to 64 KB are staged in shared memory (>= 3 CTAs per SM); longer rows
                                      // are re-read (L2) by the later passes

__device__ __forceinline__ float block_reduce(float v, float *red, bool is_max) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        const float o = __shfl_xor_sync(kFull, v, d);
        v = is_max ? fmaxf(v, o) : v + o;
    }
    __syncthreads();  // red may still be read by the previous reduction
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = red[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = is_max ? fmaxf(r, red[w]) : r + red[w];
    return r;
}

__host__ __device__ inline size_t prep_big_smem(int C, int Lmax) {
    const size_t utt = 2 * sizeof(int) * (size_t)(Lmax > 0 ? Lmax : 1) + sizeof(int) * (C + 2) + 64;
    return utt > 256 ? utt : 256;
}
__host__ __device__ inline size_t prep_big_row_smem(int C) {
    return 256 + (C <= kBigStageMaxC ? align16(sizeof(float) * (size_t)C) : 0);
}

// label tables only (one CTA per utterance); the large-alphabet path runs it ahead of the row kernel
__global__ void __launch_bounds__(kPrepRows) ctc_prep_utt_kernel(const CtcParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    int *lps = reinterpret_cast<int *>(smem);
    if ((int)blockIdx.x >= p.B) {  // extra CTA of a sliced launch: length bucketing
        sort_by_length(p, lps);
        return;
    }
    const int Lp = p.Lmax > 0 ? p.Lmax : 1;
    prep_utterance(p, blockIdx.x, lps, lps + Lp, lps + 2 * Lp);
}

__device__ __forceinline__ float4 nan_fix4(float4 v, bool ntz, int &nn) {
    if (ntz) {
        nn += (v.x != v.x) + (v.y != v.y) + (v.z != v.z) + (v.w != v.w);
        v.x = nan_fix(v.x, true); v.y = nan_fix(v.y, true); v.z = nan_fix(v.z, true); v.w = nan_fix(v.w, true);
    }
    return v;
}

// Register-resident form of the row kernel for 16-byte-aligned rows of C <= 1024 * NV classes (C % 4 == 0): thread i
// keeps float4 i, i + 256, ... of the row in registers across the three passes, so the row is read from DRAM once;
// the exponentials are also dropped into shared memory (one store per float4) for the 1 + L gathered emissions.  Same arithmetic and summation order
// as the float4 form of the generic kernel below.
template <int CP, int NV, bool NTZ>
__global__ void __launch_bounds__(kBigThreads) ctc_prep_big_reg_kernel(const CtcParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    float *red = reinterpret_cast<float *>(smem);
    float4 *s4 = reinterpret_cast<float4 *>(smem + 256);  // exponentials of the row, for the emission gather
    const int tid = threadIdx.x;
    const int C = p.C;
    const int t = (int)(blockIdx.x / (unsigned)p.slice_nb), bi = (int)(blockIdx.x - (unsigned)t * (unsigned)p.slice_nb);
    const int b = p.bord ? p.bord[p.slice0 + bi] : p.slice0 + bi;
    const long long row = (long long)t * p.B + b;
    const float4 *x4 = reinterpret_cast<const float4 *>(p.acts + row * C);
    float4 *g4 = p.grads ? reinterpret_cast<float4 *>(p.grads + row * C) : nullptr;
    const int n4 = C >> 2;
    if (t >= min(max(p.act_lens[b], 0), p.T)) {  // padding frame: zero gradient row, nothing else
        if (g4) {
#pragma unroll
            for (int k = 0; k < NV; ++k)
                if (tid + k * kBigThreads < n4) __stcs(g4 + tid + k * kBigThreads, make_float4(0.f, 0.f, 0.f, 0.f));
        }
        return;
    }
    float4 v[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const int i = tid + k * kBigThreads;
        v[k] = (i < n4) ? __ldcs(x4 + i) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
    int nn = 0;
    float m = -INFINITY;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        if (NTZ) v[k] = nan_fix4(v[k], true, nn);
        m = fmaxf(fmaxf(m, fmaxf(v[k].x, v[k].y)), fmaxf(v[k].z, v[k].w));
    }
    if (NTZ && nn && p.nan_count) atomicAdd(p.nan_count, nn);
    m = block_reduce(m, red, true);
    if (p.path != nullptr) {  // best path by-product: first class attaining the row maximum
        int *s_arg = reinterpret_cast<int *>(red) + 40;
        if (tid == 0) *s_arg = 0x7fffffff;
        __syncthreads();
        int cand = 0x7fffffff;
#pragma unroll
        for (int k = NV - 1; k >= 0; --k) {
            const int c0 = 4 * (tid + k * kBigThreads);
            if (v[k].w == m) cand = c0 + 3;
            if (v[k].z == m) cand = c0 + 2;
            if (v[k].y == m) cand = c0 + 1;
            if (v[k].x == m) cand = c0;
        }
        if (cand != 0x7fffffff) atomicMin(s_arg, cand);
        __syncthreads();
        if (tid == 0) p.path[(size_t)b * p.T + t] = *s_arg;
    }
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        if (tid + k * kBigThreads < n4) {
            v[k].x = expf(v[k].x - m); v[k].y = expf(v[k].y - m); v[k].z = expf(v[k].z - m); v[k].w = expf(v[k].w - m);
            sum += (v[k].x + v[k].y) + (v[k].z + v[k].w);
            s4[tid + k * kBigThreads] = v[k];
        }
    }
    sum = block_reduce(sum, red, false);  // its barriers also publish the exponentials to the whole CTA
    const float inv = 1.0f / sum;
    const bool bad = sum != sum;  // NaN / +Inf logit in the row: everything derived from it is NaN (see ctc_prep_kernel)
    if (g4) {
        const float scale = p.grad_scale;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = tid + k * kBigThreads;
            if (i < n4) {
                float4 o;
                o.x = scale * clamp_y(v[k].x * inv, bad);
                o.y = scale * clamp_y(v[k].y * inv, bad);
                o.z = scale * clamp_y(v[k].z * inv, bad);
                o.w = scale * clamp_y(v[k].w * inv, bad);
                __stcs(g4 + i, o);
            }
        }
    }
    // compact emission row, planar fp64 words.  (Re-reading the 1 + L logits from global memory instead cost 240 MB
    // of DRAM reads per step at C = 5000: the streaming loads above do not keep the row in L2.)
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int off = p.laboff[b];  // written by ctc_prep_utt_kernel, launched before this kernel
    const float *srow = reinterpret_cast<const float *>(s4);
    unsigned *ydst = p.y + ((size_t)b * p.T + t) * kYP * CP;
    for (int j = tid; j < CP; j += kBigThreads) {
        float y = 1.17549435e-38f;
        if (j <= L) {
            const int c = (j == 0) ? p.blank : min(max(p.labels[off + j - 1], 0), C - 1);
            y = fmaxf(srow[c] * inv, 1.17549435e-38f);
        }
        ydst[j] = bad ? kNanHi : y_hi_word(y);
        if (kYP == 2) ydst[CP + j] = bad ? 0u : y_lo_word(y);
    }
}

// grid: T * slice_nb CTAs; CTA i handles frame t = i / slice_nb of utterance bord[slice0 + i % slice_nb]
template <int CP>
__global__ void __launch_bounds__(kBigThreads) ctc_prep_big_kernel(const CtcParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x;
    const int C = p.C;
    float *red = reinterpret_cast<float *>(smem);
    float *srow = reinterpret_cast<float *>(smem + 256);
    const bool staged = C <= kBigStageMaxC;
    const int t = (int)(blockIdx.x / (unsigned)p.slice_nb), bi = (int)(blockIdx.x - (unsigned)t * (unsigned)p.slice_nb);
    const int b = p.bord ? p.bord[p.slice0 + bi] : p.slice0 + bi;
    const long long row = (long long)t * p.B + b;
    const float *x = p.acts + row * C;
    float *g = p.grads ? p.grads + row * C : nullptr;
    const bool gvec = g && ((reinterpret_cast<uintptr_t>(g) & 15) == 0);
    if (t >= min(max(p.act_lens[b], 0), p.T)) {  // padding frame: zero gradient row, nothing else
        if (g) {
            const int n4 = gvec ? (C >> 2) : 0;
            for (int i = tid; i < n4; i += kBigThreads) __stcs(reinterpret_cast<float4 *>(g) + i, make_float4(0.f, 0.f, 0.f, 0.f));
            for (int c = (n4 << 2) + tid; c < C; c += kBigThreads) g[c] = 0.f;
        }
        return;
    }
    const bool vec = ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    const int n4 = vec ? (C >> 2) : 0;
    const bool ntz = p.nan_to_zero != 0;
    const float4 *x4 = reinterpret_cast<const float4 *>(x);
    float4 *s4 = reinterpret_cast<float4 *>(srow);
    // ---- pass 1 (the only DRAM read of the row): NaN guard, row maximum, shared-memory copy
    float m = -INFINITY;
    int nn = 0;
#pragma unroll 4
    for (int i = tid; i < n4; i += kBigThreads) {
        const float4 v = nan_fix4(__ldcs(x4 + i), ntz, nn);
        if (staged) s4[i] = v;
        m = fmaxf(fmaxf(m, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    }
    for (int c = (n4 << 2) + tid; c < C; c += kBigThreads) {
        float v = __ldcs(x + c);
        nn += (ntz && v != v);
        v = nan_fix(v, ntz);
        if (staged) srow[c] = v;
        m = fmaxf(m, v);
    }
    if (nn && p.nan_count) atomicAdd(p.nan_count, nn);
    m = block_reduce(m, red, true);
    if (p.path != nullptr) {  // best path by-product: first class attaining the row maximum
        int *s_arg = reinterpret_cast<int *>(red) + 40;
        if (tid == 0) *s_arg = 0x7fffffff;
        __syncthreads();
        int cand = 0x7fffffff;
        for (int c = tid; c < C; c += kBigThreads)
            if ((staged ? srow[c] : nan_fix(__ldg(x + c), ntz)) == m) { cand = c; break; }
        if (cand != 0x7fffffff) atomicMin(s_arg, cand);
        __syncthreads();
        if (tid == 0) p.path[(size_t)b * p.T + t] = *s_arg;
    }
    // ---- pass 2 (shared memory, or L2 for rows too long to stage): exponentials in place, row sum
    float sum = 0.f;
    for (int i = tid; i < n4; i += kBigThreads) {
        int dummy = 0;
        const float4 v = staged ? s4[i] : nan_fix4(__ldg(x4 + i), ntz, dummy);
        float4 e;
        e.x = expf(v.x - m); e.y = expf(v.y - m); e.z = expf(v.z - m); e.w = expf(v.w - m);
        if (staged) s4[i] = e;
        sum += (e.x + e.y) + (e.z + e.w);
    }
    for (int c = (n4 << 2) + tid; c < C; c += kBigThreads) {
        const float e = expf((staged ? srow[c] : nan_fix(__ldg(x + c), ntz)) - m);
        if (staged) srow[c] = e;
        sum += e;
    }
    sum = block_reduce(sum, red, false);  // its barriers also publish the exponentials to the whole CTA
    const float inv = 1.0f / sum;
    const bool bad = sum != sum;  // NaN / +Inf logit in the row: everything derived from it is NaN (see ctc_prep_kernel)
    // ---- pass 3: the dense part of the gradient, scale * softmax, written once (the occupancies of the <= 1 + L
    // classes on the lattice are subtracted by the gradient kernel)
    if (g) {
        const float scale = p.grad_scale;
        const int g4 = (vec && gvec) ? n4 : 0;
        for (int i = tid; i < g4; i += kBigThreads) {
            float4 e;
            if (staged) {
                e = s4[i];
            } else {
                int dummy = 0;
                const float4 v = nan_fix4(__ldg(x4 + i), ntz, dummy);
                e.x = expf(v.x - m); e.y = expf(v.y - m); e.z = expf(v.z - m); e.w = expf(v.w - m);
            }
            float4 o;
            o.x = scale * clamp_y(e.x * inv, bad);
            o.y = scale * clamp_y(e.y * inv, bad);
            o.z = scale * clamp_y(e.z * inv, bad);
            o.w = scale * clamp_y(e.w * inv, bad);
            __stcs(reinterpret_cast<float4 *>(g) + i, o);
        }
        for (int c = (g4 << 2) + tid; c < C; c += kBigThreads) {
            const float e = staged ? srow[c] : expf(nan_fix(__ldg(x + c), ntz) - m);
            g[c] = scale * clamp_y(e * inv, bad);
        }
    }
    // ---- compact emission row, planar fp64 words
    const int L = min(max(p.label_lens[b], 0), p.Lmax);
    const int *lab = p.labels;
    const int off = p.laboff[b];  // written by ctc_prep_utt_kernel, launched before this kernel
    unsigned *ydst = p.y + ((size_t)b * p.T + t) * kYP * CP;
    for (int j = tid; j < CP; j += kBigThreads) {
        float v = 1.17549435e-38f;
        if (j <= L) {
            const int c = (j == 0) ? p.blank : min(max(lab[off + j - 1], 0), C - 1);
            const float e = staged ? srow[c] : expf(nan_fix(__ldg(x + c), ntz) - m);
            v = fmaxf(e * inv, 1.17549435e-38f);
        }
        ydst[j] = bad ? kNanHi : y_hi_word(v);
        if (kYP == 2) ydst[CP + j] = bad ? 0u : y_lo_word(v);
    }
}

// ================================================================================================ summed loss
// loss = sum_b costs[b], the one-element value the criterion returns (train.py:600).  One warp, fixed order, fp64
// accumulation: deterministic and independent of how the batch was sliced.  Runs beside the gradient kernels.
//
// Multi-GPU (replaces reduce_tensor, /root/reference/data/utils.py:40-44): with a communicator the SAME kernel also
// performs the path's one exchange, the all-reduce of the summed loss, over NVLink peer memory -- no collective
// launch follows.  One-shot protocol: every rank owns 2 x kCommMaxWorld slots of 8 bytes; call number e (a counter in
// the rank's own buffer, so CUDA-graph replays advance it too) stores (e << 32 | float bits of the local sum) with ONE
// 64-bit store into slot [e & 1][rank] of every peer, then polls its own slots [e & 1][0 .. world) until all carry
// epoch e and adds the values in rank order (every rank gets the same bits).  Value and epoch travel in one word,
// so no fence orders them; the two parities make reuse safe (a rank cannot enter call e+2 before every peer has
// entered call e+1, i.e. finished reading call e).  A late peer is waited for; one that has not arrived after the
// communicator's wall-clock timeout turns the result into NaN and sets the communicator's sticky flag.
constexpr int kCommMaxWorld = 32;
struct CommArgs {
    unsigned long long *local;          // this rank's slots [2][kCommMaxWorld], then the call counter
    unsigned long long *const *peers;   // device array [world]: slot base of every rank (peers[rank] == local)
    int rank, world;
    unsigned long long timeout_ns;      // wall-clock bound on the wait for the peers (globaltimer)
    int *timed_out;                     // host-mapped sticky flag: set when a peer never arrived (see ctc_api.cu)
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(32) ctc_sum_costs_kernel(const float *costs, int B, float *loss_sum,
                                                           float *loss_all, CommArgs ca) {
    const int lane = threadIdx.x;
    double acc = 0.0;
    for (int base = 0; base < B; base += 32) {
        const int i = base + lane;
        double v = i < B ? (double)costs[i] : 0.0;
#pragma unroll
        for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(kFull, v, d);
        acc += v;
    }
    if (lane == 0 && loss_sum) *loss_sum = (float)acc;  // this rank's own sum
    if (!loss_all) return;
    if (ca.world <= 1) {
        if (lane == 0) *loss_all = (float)acc;
        return;
    }
    unsigned *ctr = reinterpret_cast<unsigned *>(ca.local + 2 * kCommMaxWorld);
    unsigned e = 0;
    if (lane == 0) {
        e = *ctr + 1;
        *ctr = e;
    }
    e = __shfl_sync(kFull, e, 0);
    const int par = (int)(e & 1u) * kCommMaxWorld;
    const unsigned long long word = ((unsigned long long)e << 32) | (unsigned long long)__float_as_uint((float)acc);
    if (lane < ca.world) {
        volatile unsigned long long *dst = ca.peers[lane] + par + ca.rank;
        *dst = word;
    }
    __threadfence_system();
    float mine = 0.f;
    bool ok = true;
    if (lane < ca.world) {
        volatile unsigned long long *src = ca.local + par + lane;
        unsigned long long w = *src;
        // A peer that is late (checkpoint save, validation, a data-loader stall) is simply waited for, like a
        // collective would; only after timeout_ns of WALL-CLOCK time (default 120 s, ctc_b200_comm_set_timeout) is
        // the exchange given up: the result is NaN and the communicator's sticky flag makes every later call
        // return an error status instead of running the protocol out of step.
        if ((unsigned)(w >> 32) != e) {
            const unsigned long long t0 = global_ns();
            for (unsigned spins = 0;; ++spins) {
                w = *src;
                if ((unsigned)(w >> 32) == e) break;
                if ((spins & 255u) == 255u && global_ns() - t0 > ca.timeout_ns) break;
            }
        }
        ok = (unsigned)(w >> 32) == e;
        mine = __uint_as_float((unsigned)w);
    }
    ok = __all_sync(kFull, ok);
    double tot = 0.0;
    for (int j = 0; j < ca.world; ++j) tot += (double)__shfl_sync(kFull, mine, j);
    if (lane == 0) {
        *loss_all = ok ? (float)tot : __int_as_float(0x7fc00000);
        if (!ok && ca.timed_out) {
            *reinterpret_cast<volatile int *>(ca.timed_out) = 1;
            __threadfence_system();
        }
    }
}

// Gradient rows times a DEVICE scalar, in place -- the autograd backward of the drop-in module (the reference
// binding returns `cached_grads * grad_output`, train.py:619-620 feeding 1/B through `loss / B`).  The scalar is
// read on the device, so nothing waits on the host; when it is exactly 1 (a caller that folded the 1/B into the
// criterion: size_average / grad_scale) every CTA leaves at once and the pass costs a launch, not 2 x [T,B,C] bytes.
__global__ void __launch_bounds__(256) ctc_scale_grads_kernel(float *g, size_t n, const float *scale) {
    const float s = __ldg(scale);
    if (s == 1.0f) return;
    const size_t n4 = (reinterpret_cast<uintptr_t>(g) & 15) == 0 ? n >> 2 : 0;
    float4 *g4 = reinterpret_cast<float4 *>(g);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 v = g4[i];
        v.x *= s; v.y *= s; v.z *= s; v.w *= s;
        g4[i] = v;
    }
    for (size_t i = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) g[i] *= s;
}

// ================================================================================================ greedy decoding
// SURVEY 8(f) rank 1 -- replaces GreedyDecoder.decode (/root/reference/decoder.py:182-197: argmax over classes per
// frame) and process_string (:165-180: drop blanks and frames equal to the previous frame, keep the frame index of
// every surviving character).  Integer work, bit-exact against oracle/greedy_oracle.py.

// Stand-alone best path (eval call sites: no loss wanted).  8 lanes per row, classes strided over the lanes; ties go
// to the lowest class index (what torch.max returns).  Element (t,b,c) of the input sits at t*st + b*sb + c, so both
// the loss layout [T,B,C] and the reference decoder's batch-first [B,T,C] are read in place.
__global__ void __launch_bounds__(128) ctc_argmax_kernel(const float *acts, long long st, long long sb,
                                                         const int *act_lens, int T, int B, int C, int *path) {
    const int lane = threadIdx.x & 31, j = lane & 7;
    const long long row = (long long)blockIdx.x * 16 + (threadIdx.x >> 3);
    const bool valid = row < (long long)T * B;
    const int t = valid ? (int)(row / B) : 0, b = valid ? (int)(row - (long long)t * B) : 0;
    const bool active = valid && t < (act_lens ? min(max(act_lens[b], 0), T) : T);
    const float *x = acts + (long long)t * st + (long long)b * sb;
    float best = -INFINITY;
    int arg = 0x7fffffff;
    if (active)
        for (int c = j; c < C; c += 8) {
            const float v = __ldg(x + c);
            if (v > best || arg == 0x7fffffff) { best = v; arg = c; }  // strict >: the first maximum wins
        }
#pragma unroll
    for (int d = 1; d < 8; d <<= 1) {
        const float ob = __shfl_xor_sync(kFull, best, d);
        const int oa = __shfl_xor_sync(kFull, arg, d);
        if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
    }
    if (active && j == 0) path[(size_t)b * T + t] = arg;
}

// Collapse of the best path, one CTA of 128 threads per utterance: keep frame t iff path[t] != blank and (t == 0 or
// path[t] != path[t-1]).  Every thread owns a contiguous segment of frames: count its survivors, block-wide exclusive
// scan of the counts (shuffle scan per warp + 4 partial sums), then write ids / offsets in order.  Two short rounds
// of independent loads instead of a serial walk over T.
constexpr int kCollapseThreads = 128;
__global__ void __launch_bounds__(kCollapseThreads) ctc_collapse_kernel(const int *path, const int *act_lens, int T,
                                                                        int B, int blank, int *ids, int *offsets,
                                                                        int *out_lens) {
    __shared__ int wsum[kCollapseThreads / 32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int Tb = act_lens ? min(max(act_lens[b], 0), T) : T;
    const int *pb = path + (size_t)b * T;
    const int seg = (Tb + kCollapseThreads - 1) / kCollapseThreads;
    const int t0 = min(tid * seg, Tb), t1 = min(t0 + seg, Tb);
    int cnt = 0;
    {
        int prev = t0 > 0 ? pb[t0 - 1] : -1;
        for (int t = t0; t < t1; ++t) {
            const int a = pb[t];
            cnt += (a != blank && a != prev) ? 1 : 0;
            prev = a;
        }
    }
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int pos = incl - cnt;
    for (int w = 0; w < warp; ++w) pos += wsum[w];
    {
        int prev = t0 > 0 ? pb[t0 - 1] : -1;
        for (int t = t0; t < t1; ++t) {
            const int a = pb[t];
            if (a != blank && a != prev) {
                ids[(size_t)b * T + pos] = a;
                if (offsets) offsets[(size_t)b * T + pos] = t;
                ++pos;
            }
            prev = a;
        }
    }
    if (tid == kCollapseThreads - 1) out_lens[b] = pos;
}

}  // namespace ctcb200
