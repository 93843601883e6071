// csrc/ctc_metrics.cuh -- batched edit distances of the greedy transcripts against the targets.
//
// Replaces, per utterance, what the reference computes in Python on every training and eval batch
// (/root/reference/train.py:369-385, 577-591 through data/utils.py:47-57 get_cer_wer and decoder.py:44-73):
//   cer      = Levenshtein distance of the two character strings with the spaces removed      (decoder.py:64-73)
//   wer      = Levenshtein distance of the two WORD sequences (split on runs of spaces)         (decoder.py:44-62)
//   wer_ref  = number of reference words or 1, cer_ref = number of reference non-space chars or 1 (utils.py:50-51)
// The reference's strings are 1:1 images of the label ids (decoder.py:165-180 maps every id through int_to_char), so
// the same numbers are obtained on the ids; integer work, bit-exact against oracle/wer_oracle.py.
//
// One CTA per utterance.  Both sequences are staged in shared memory; two warps compact the non-space symbols and
// mark the word starts with ballot prefix sums; the distances are anti-diagonal wavefronts of the classic DP (three
// rotating diagonals in shared memory, one CTA barrier per diagonal).  Words are compared exactly (length, then
// symbols), never by hash.
#pragma once
#include "ctc_device.cuh"

namespace ctcb200 {

constexpr int kMetricThreads = 256;

__host__ __device__ inline size_t metrics_smem(int max_hyp, int max_ref) {
    const size_t nh = (size_t)(max_hyp > 0 ? max_hyp : 1), nr = (size_t)(max_ref > 0 ? max_ref : 1);
    // raw + compacted symbols, word starts + lengths, three diagonals indexed by the hypothesis position
    return sizeof(int) * (2 * nh + 2 * nr + 2 * (nh / 2 + 1) + 2 * (nr / 2 + 1) + 3 * (nh + 1) + 8);
}

// one warp: out[0..n') = the non-space symbols of x[0..n), ws[0..w) = positions where a word starts; returns (n', w)
__device__ __forceinline__ int2 compact_and_words(const int *x, int n, int space, int *out, int *ws, int lane) {
    int nc = 0, nw = 0;
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        const bool in = i < n;
        const int v = in ? x[i] : space;
        const bool ch = in && v != space;
        const bool start = ch && (i == 0 || x[i - 1] == space);
        const unsigned mc = __ballot_sync(kFull, ch), ms = __ballot_sync(kFull, start);
        const unsigned lt = (1u << lane) - 1u;
        if (ch) out[nc + __popc(mc & lt)] = v;
        if (start) ws[nw + __popc(ms & lt)] = i;
        nc += __popc(mc);
        nw += __popc(ms);
    }
    return make_int2(nc, nw);
}

// D[n][m] of the unit-cost edit distance, eq(i, j) comparing element i of the first with element j of the second
// sequence.  All threads of the CTA call it; diag = 3 * (n + 1) ints of shared memory.
template <typename Eq>
__device__ __forceinline__ int lev_wavefront(int n, int m, Eq eq, int *diag) {
    int *d2 = diag, *d1 = diag + (n + 1), *d0 = diag + 2 * (n + 1);  // d-2, d-1, d (indexed by i)
    for (int d = 0; d <= n + m; ++d) {
        const int lo = max(0, d - m), hi = min(n, d);
        for (int i = lo + (int)threadIdx.x; i <= hi; i += blockDim.x) {
            const int j = d - i;
            int v;
            if (i == 0) v = j;
            else if (j == 0) v = i;
            else v = min(min(d1[i - 1], d1[i]) + 1, d2[i - 1] + (eq(i - 1, j - 1) ? 0 : 1));
            d0[i] = v;
        }
        __syncthreads();
        int *t = d2; d2 = d1; d1 = d0; d0 = t;
    }
    return d1[n];  // the last diagonal written
}

// counts[b] = (wer, cer, wer_ref, cer_ref).  hyp: [B][hyp_pitch] ids + hyp_lens; ref: flat labels + ref_lens.
__global__ void __launch_bounds__(kMetricThreads) ctc_error_counts_kernel(const int *hyp, int hyp_pitch,
                                                                          const int *hyp_lens, const int *ref,
                                                                          const int *ref_lens, int B, int max_hyp,
                                                                          int max_ref, int space, int4 *counts) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nhm = max_hyp > 0 ? max_hyp : 1, nrm = max_ref > 0 ? max_ref : 1;
    int *H = reinterpret_cast<int *>(smem), *R = H + nhm, *Hc = R + nrm, *Rc = Hc + nhm;
    int *wsH = Rc + nrm, *wlH = wsH + (nhm / 2 + 1), *wsR = wlH + (nhm / 2 + 1), *wlR = wsR + (nrm / 2 + 1);
    int *diag = wlR + (nrm / 2 + 1);
    int *meta = diag + 3 * (nhm + 1);  // [0..3]: nc_h, nw_h, nc_r, nw_r; [4]: label offset
    const int nh = min(max(hyp_lens[b], 0), max_hyp), nr = min(max(ref_lens[b], 0), max_ref);
    if (tid == 0) meta[4] = 0;
    __syncthreads();
    int part = 0;
    for (int i = tid; i < b; i += blockDim.x) part += max(ref_lens[i], 0);
#pragma unroll
    for (int d = 16; d; d >>= 1) part += __shfl_xor_sync(kFull, part, d);
    if (lane == 0 && part) atomicAdd(&meta[4], part);
    __syncthreads();
    const int off = meta[4];
    for (int i = tid; i < nh; i += blockDim.x) H[i] = hyp[(size_t)b * hyp_pitch + i];
    for (int i = tid; i < nr; i += blockDim.x) R[i] = ref[off + i];
    __syncthreads();
    if (warp == 0) {
        const int2 r = compact_and_words(H, nh, space, Hc, wsH, lane);
        if (lane == 0) { meta[0] = r.x; meta[1] = r.y; }
    } else if (warp == 1) {
        const int2 r = compact_and_words(R, nr, space, Rc, wsR, lane);
        if (lane == 0) { meta[2] = r.x; meta[3] = r.y; }
    }
    __syncthreads();
    const int nch = meta[0], nwh = meta[1], ncr = meta[2], nwr = meta[3];
    for (int k = tid; k < nwh; k += blockDim.x) {
        int e = wsH[k];
        while (e < nh && H[e] != space) ++e;
        wlH[k] = e - wsH[k];
    }
    for (int k = tid; k < nwr; k += blockDim.x) {
        int e = wsR[k];
        while (e < nr && R[e] != space) ++e;
        wlR[k] = e - wsR[k];
    }
    __syncthreads();
    const int cer = lev_wavefront(nch, ncr, [&](int i, int j) { return Hc[i] == Rc[j]; }, diag);
    __syncthreads();
    const int wer = lev_wavefront(nwh, nwr, [&](int i, int j) {
        const int len = wlH[i];
        if (len != wlR[j]) return false;
        const int *a = H + wsH[i], *c = R + wsR[j];
        for (int k = 0; k < len; ++k)
            if (a[k] != c[k]) return false;
        return true;
    }, diag);
    if (tid == 0) counts[b] = make_int4(wer, cer, nwr > 0 ? nwr : 1, ncr > 0 ? ncr : 1);
}

}  // namespace ctcb200
