// csrc/fc_kernels.cuh -- the logits producer of the CTC path as ONE sm_100a kernel.
//
// Replaces, for inference / evaluation, the reference's final layer
//     fully_connected = nn.Sequential(nn.BatchNorm1d(H), nn.Linear(H, C, bias=False));  SequenceWise(fully_connected)
// (/root/reference/model.py:335-341, applied at model.py:373 to x [T,B,H] viewed as [T*B, H], model.py:28-49) and the
// softmax the CTC path applies to its output (the criterion's own, and model.py:377):
//     logits[r, c] = sum_h W[c,h] * (s[h] * x[r,h] + t[h]),   s = gamma / sqrt(var + eps),  t = beta - mean * s
// BatchNorm is folded into the weights once per call by fc_fold_kernel (W' = W * s, bias = W t), so the hot kernel is a
// [T*B, H] x [H, C] contraction -- the one dense contraction of the path, hence the one place tensor cores are right:
//   * TMA (cp.async.bulk.tensor.2d, 128-byte swizzle; SASS UTMALDG) streams 128-row x 32-column tiles of x, read from
//     HBM exactly once, and the matching W' tiles (L2-resident) into a 3-4 stage shared-memory ring;
//   * tcgen05.mma kind::tf32 (SASS UTCHMMA), M = 128, accumulators in TMEM, issued by one thread.  The reference runs
//     this layer in fp32, and one TF32 pass is 2e-3 away from that on K = 800 (measured), so the product is SPLIT:
//         x W'^T  ~=  x [W_hi | W_lo]^T  +  x_lo W_hi^T          (x_hi is what the tensor core sees of x: it ignores
//     the low 13 mantissa bits; x_lo = x - x_hi is formed by four converter warps from the staged tile; W_hi / W_lo
//     are prepared by the fold kernel).  One N = 2*Cp MMA + one N = Cp MMA per 8 columns; error ~1e-6, like fp32;
//   * epilogue warps read the accumulators with tcgen05.ld (SASS LDTM), one logits row per thread: + bias, the fp32
//     logits [T,B,C] the caller still wants, and -- optionally -- the row's softmax as the planar emission row
//     y[b][t][2*Cp] of the CTC workspace, so the lattice kernels can start without ctc_prep_kernel's softmax pass
//     re-reading the logits.
// HBM-bound: 4*H bytes per frame in (x), 4*C + 8*Cp out.
#pragma once
#include <cuda.h>
#include "ctc_device.cuh"

namespace ctcb200 {

// 1: whole rows of the NEXT tile are bulk-prefetched into L2 by the producer lane (see the producer).  Measured on
// B200 and rejected: 0.126 ms against 0.106 ms without (T=1500, B=64, H=800) -- kept as a switch for re-measurement.
#ifndef CTC_FC_L2_PREFETCH
#define CTC_FC_L2_PREFETCH 0
#endif
constexpr int kFcBM = 256;      // rows (frames) per tile: two UMMA M = 128 halves that share one W' tile
constexpr int kFcHM = 128;      // UMMA M
constexpr int kFcBK = 32;       // fp32 columns per k-block = one 128-byte swizzle row
constexpr int kFcUmmaK = 8;     // tf32 elements per tcgen05.mma
constexpr int kFcThreads = 352; // warp 0: TMA producer, warp 1: TMEM owner + MMA issuer of tile half 0, 2-5: converters,
                                // 6-9: epilogue, warp 10: MMA issuer of tile half 1

template <int N>
struct FcCfg {
    static_assert(N == 32 || N == 64, "classes are padded to 32 or 64");
    static constexpr int stages = N == 32 ? 4 : 3;          // TMA ring: x tile + W' tiles of one k-block per stage
    static constexpr int lo_bufs = 2;                       // x_lo tiles (converter -> MMA), outside the TMA ring (3 buffers
                                                            // + 3 ring stages measured: the ring runs dry, 0.101 vs 0.098 ms)
    static constexpr int nacc = N == 32 ? 2 : 1;            // accumulators per half tile, k-blocks dealt round-robin
    static constexpr int a_bytes = kFcBM * kFcBK * 4;       // 32 KB: x tile (also the size of an x_lo tile)
    static constexpr int h_bytes = kFcHM * kFcBK * 4;       // one M = 128 half of it
    static constexpr int w_bytes = 2 * N * kFcBK * 4;       // [W_hi rows ; W_lo rows] of this k-block
    static constexpr int stage_bytes = a_bytes + w_bytes;
    static constexpr int acc_cols = 2 * N;                  // one accumulator: cols [0,N) main + x_lo term, [N,2N) W_lo term
    static constexpr int tmem_cols = 2 * 2 * nacc * acc_cols;  // two tile buffers x two halves: 512 columns for both N
    static constexpr size_t smem = 1024 /*alignment slack*/ + (size_t)stages * stage_bytes +
                                   (size_t)lo_bufs * a_bytes + 256 /*barriers*/;
};
// Why 256-row tiles: the W' tiles are re-fetched from L2 for every x tile, and with 128-row tiles that was 8 KB of W'
// per 16 KB of x -- 460 MB through the L2-to-SM fabric for 307 MB of x, which is what bounded the first version
// (0.106 ms whether x came from HBM or sat in L2; ablations with conversion and MMAs switched off: 0.094 ms).
// Why several accumulators: the tensor core adds into its fp32 accumulator with truncation, so one accumulator
// taking all 2 * H/8 additions of a row drifts by ~1e-5 at H = 800, |logit| ~ 3 (measured against the reference in
// fp64).  Dealing the k-blocks round-robin to nacc accumulators and adding those once, rounded, in the epilogue
// divides both the number of truncating additions and the magnitude they act on.

// folded weights: W2 [2N][Kp] fp32 (rows [0,N): tf32-rounded W' = W*s, rows [N,2N): W' - that; rows >= C and columns
// >= H are zero), then bias [N].  Kp = H rounded up to a multiple of 4 (TMA row pitch).
__host__ __device__ inline size_t fc_ws_bytes(int H, int N) {
    const size_t Kp = (size_t)((H + 3) & ~3);
    return align16(sizeof(float) * 2 * N * Kp) + sizeof(float) * N;
}

// ---------------------------------------------------------------- BatchNorm batch statistics (training-mode forward)
// mean[h], biased var[h] over `rows` rows of x [rows, H] -- what nn.BatchNorm1d normalises with in training mode
// (model.py:336).  Two launches: partial sums per row block (coalesced, fp64 accumulation per thread), then combine.
constexpr int kFcStatRows = 256;
__global__ void __launch_bounds__(256) fc_bn_partial_kernel(const float *__restrict__ x, long long rows, int H,
                                                            double *__restrict__ part /*[blocks][2][H]*/) {
    const long long r0 = (long long)blockIdx.x * kFcStatRows;
    const int nr = (int)min((long long)kFcStatRows, rows - r0);
    for (int h = threadIdx.x; h < H; h += blockDim.x) {
        double s = 0.0, q = 0.0;
        const float *p = x + r0 * H + h;
        for (int r = 0; r < nr; ++r) {
            const double v = (double)__ldg(p + (size_t)r * H);
            s += v;
            q = fma(v, v, q);
        }
        part[((size_t)blockIdx.x * 2 + 0) * H + h] = s;
        part[((size_t)blockIdx.x * 2 + 1) * H + h] = q;
    }
}
__global__ void __launch_bounds__(256) fc_bn_combine_kernel(const double *__restrict__ part, int nblocks, long long rows,
                                                            int H, float *__restrict__ mean, float *__restrict__ var) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    double s = 0.0, q = 0.0;
    for (int b = 0; b < nblocks; ++b) {  // fixed order: deterministic
        s += part[((size_t)b * 2 + 0) * H + h];
        q += part[((size_t)b * 2 + 1) * H + h];
    }
    const double m = s / (double)rows;
    mean[h] = (float)m;
    var[h] = (float)fmax(q / (double)rows - m * m, 0.0);
}

// ---------------------------------------------------------------- BatchNorm folded into the weights
// One CTA per class row c < N: W'[c,h] = W[c,h] * s[h] split into tf32-rounded hi + remainder lo; bias[c] = sum_h
// W[c,h] * t[h] (fp64, fixed order).  gamma / beta / mean / var may be NULL (identity / 0 / 0 / 1).
__global__ void __launch_bounds__(256) fc_fold_kernel(const float *__restrict__ W, const float *__restrict__ gamma,
                                                      const float *__restrict__ beta, const float *__restrict__ mean,
                                                      const float *__restrict__ var, float eps, int H, int C, int N,
                                                      float *__restrict__ W2, float *__restrict__ bias) {
    __shared__ double red[256];
    const int c = blockIdx.x, Kp = (H + 3) & ~3;
    double acc = 0.0;
    for (int h = threadIdx.x; h < Kp; h += blockDim.x) {
        float hi = 0.f, lo = 0.f;
        if (c < C && h < H) {
            const bool bn = var != nullptr || mean != nullptr || gamma != nullptr || beta != nullptr;
            // the same fp32 arithmetic order as torch's batch_norm: (x - mean) * (gamma * rsqrt(var + eps)) + beta
            const float inv = bn ? 1.0f / sqrtf((var ? var[h] : 1.0f) + eps) : 1.0f;
            const float s = (gamma ? gamma[h] : 1.0f) * inv;
            const double t = (double)(beta ? beta[h] : 0.0f) - (double)(mean ? mean[h] : 0.0f) * (double)s;
            const float w = W[(size_t)c * H + h];
            const float ws = w * s;
            hi = __uint_as_float((__float_as_uint(ws) + 0x1000u) & 0xffffe000u);  // tf32, round to nearest
            lo = ws - hi;
            acc = fma((double)w, t, acc);
        }
        W2[(size_t)c * Kp + h] = hi;
        W2[(size_t)(N + c) * Kp + h] = lo;
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 128; d; d >>= 1) {
        if ((int)threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) bias[c] = (float)red[0];
}

// ---------------------------------------------------------------- tcgen05 / TMA building blocks (inline PTX)
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}
// hint: bring `bytes` (multiple of 16) contiguous bytes at the 16-byte-aligned `gptr` into L2
__device__ __forceinline__ void l2_prefetch_bulk(const void *gptr, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gptr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// whole warp: allocate `cols` TMEM columns (power of two >= 32), base address written to *slot (shared memory)
__device__ __forceinline__ void tmem_alloc(uint32_t *slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// the mbarrier receives one arrival when every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, tf32 inputs, fp32 accumulate
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread `lane` of the warp receives its lane's (= row's) 32 values
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor of a K-major tile whose rows are ONE 128-byte swizzle span (32 fp32) and whose
// 8-row groups lie 1024 bytes apart (what a 128B-swizzled TMA box of 32 columns produces; tile base 1024-aligned):
//   [0,14) start address >> 4, [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1), [32,46) stride byte
//   offset >> 4 (1024 >> 4), [46,48) descriptor version 1 (sm_100), [61,64) layout 2 = SWIZZLE_128B.
// Stepping 8 tf32 (32 bytes) along K inside the span adds 32 >> 4 to the start-address field.
__device__ __forceinline__ uint64_t fc_smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3ffffu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
           (2ull << 61);
}
// Instruction descriptor, kind::tf32: [4,6) D format 1 = F32, [7,10) A format 2 = TF32, [10,13) B format 2 = TF32,
// bit 15 / 16 A / B major 0 = K, [17,23) N >> 3, [24,29) M >> 4.
__host__ __device__ constexpr uint32_t fc_idesc(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

struct FcParams {
    const float *x;        // [rows][H] (also behind map_x): whole rows are prefetched into L2 one tile ahead
    long long rows;        // T * B
    int rows_per_cta;      // CTA c owns rows [c * rows_per_cta, (c+1) * rows_per_cta): equal HBM bytes per SM
    int T, B, H, C;
    const float *bias;     // [N] (folded BatchNorm shift through the weights)
    float *logits;         // [T,B,C] fp32 or nullptr
    unsigned *y;           // CTC emission rows [B][T][2*N] (planar fp64 words, see ctc_prep_kernel) or nullptr
    const int *act_lens;   // with y: rows t >= act_lens[b] are skipped (nullptr: every row is written)
    int *path;             // optional [B][T]: argmax class of every frame (greedy best path by-product)
    long long *trace;      // diagnostic (CTC_FC_TRACE): CTA 0 records clock64 per role and k-block, [kFcTraceK][8]
};
constexpr int kFcTraceK = 96;
__device__ __forceinline__ long long fc_clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }
#define FC_TRACE(slot, g) do { if (p.trace && blockIdx.x == 0 && (g) < (unsigned)kFcTraceK) p.trace[(g) * 8 + (slot)] = fc_clk(); } while (0)

// 32 x 32 transpose across a warp: on entry lane r holds v[c] = M[r][c], on exit lane c holds v[r] = M[r][c].
// Five butterfly stages of 16 shuffles, register indices static.
__device__ __forceinline__ void warp_transpose32(float (&v)[32], int lane) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const bool up = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            if ((i & s) == 0) {
                const float give = up ? v[i] : v[i + s];
                const float got = __shfl_xor_sync(kFull, give, s);
                if (up) v[i] = got; else v[i + s] = got;
            }
        }
    }
}

// ---------------------------------------------------------------- the logits kernel
// Persistent: one CTA per SM; CTA c owns the rows [c * rows_per_cta, (c + 1) * rows_per_cta) and walks them in tiles
// of 256 (two UMMA M = 128 halves).  Its last tile is usually partial: it is loaded with 32-row boxes (a second tensor map), so every SM moves
// the same number of bytes (whole 128-row tiles dealt round-robin left 10 of 148 SMs with a 6th tile: +18 %).  The
// UMMA still runs on 128 rows; rows that were not loaded hold stale data and are never written out.
// Pipelines (all mbarriers, no CTA-wide barrier after set-up):
//   full[s]   TMA -> converters + MMA : ring stage s holds the x tile and the W' tiles of a k-block     (1 arrival + tx)
//   empty[s]  MMA -> TMA              : the MMAs reading stage s have completed (tcgen05.commit x 2 issuers)  (2)
//   conv[j]   converters -> MMA       : x_lo buffer j is written (and fenced to the async proxy)              (128)
//   lofree[j] MMA -> converters       : the MMAs reading x_lo buffer j have completed (tcgen05.commit x 2)    (2)
//   tfull[b]  MMA -> epilogue         : tile buffer b holds a finished tile (tcgen05.commit x 2)              (2)
//   tempty[b] epilogue -> MMA         : tile buffer b has been read out                                       (128)
template <int N>
__global__ void __launch_bounds__(kFcThreads, 1)
    fc_logits_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_x32,
                     const __grid_constant__ CUtensorMap map_w, const FcParams p) {
    using Cfg = FcCfg<N>;
    constexpr int S = Cfg::stages, LB = Cfg::lo_bufs, NA = Cfg::nacc;
    extern __shared__ unsigned char fc_smem_raw[];
    // 128-byte swizzle: tile bases must be 1024-byte aligned
    // (as an offset from the shared array, so that the compiler keeps the shared address space: through a uintptr_t
    // round trip every access below became a generic LD / ST)
    unsigned char *smem = fc_smem_raw + ((1024u - (smem_u32(fc_smem_raw) & 1023u)) & 1023u);
    unsigned char *lo_base = smem + (size_t)S * Cfg::stage_bytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(lo_base + (size_t)LB * Cfg::a_bytes);
    uint64_t *full = bars, *empty = bars + S, *conv = bars + 2 * S, *lofree = conv + LB, *tfull = lofree + LB,
             *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nkb = (p.H + kFcBK - 1) / kFcBK;
    const long long r0 = (long long)blockIdx.x * p.rows_per_cta;
    const long long r1 = min(p.rows, r0 + p.rows_per_cta);

    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 2);   // one tcgen05.commit from each of the two MMA issuers
        }
        for (int j = 0; j < LB; ++j) {
            mbar_init(&conv[j], 128);
            mbar_init(&lofree[j], 2);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull[b], 2);
            mbar_init(&tempty[b], 128);
        }
        fence_mbar_init();
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_x32);
        tma_prefetch_desc(&map_w);
    }
    if (warp == 1) tmem_alloc(tmem_slot, Cfg::tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (p.trace && tid == 0) p.trace[kFcTraceK * 8 + 2 * blockIdx.x] = (long long)global_ns();

    if (warp == 0) {
        // ============================================================ TMA producer (one lane)
        if (lane == 0) {
            // A TMA box takes 128 bytes of each of 128 rows that lie 4*H bytes apart, and the next 128 bytes of a row
            // follow a whole k-block later: left alone, HBM sees scattered 128-byte reads (measured: 39 % of its
            // peak).  So whole ROWS of the next tile (4*H contiguous bytes each) are prefetched into L2 while the
            // current tile is consumed -- DRAM reads become sequential and the boxes hit L2.  The prefetches are
            // spread over the k-blocks of the current tile, so at most one tile per SM (~60 MB in all) waits in L2.
            const uint32_t row_bytes = (uint32_t)p.H * 4u;
            const bool pf = (row_bytes & 15u) == 0 && CTC_FC_L2_PREFETCH;
            if (pf)
                for (long long r = r0; r < min(r1, r0 + kFcBM); ++r) l2_prefetch_bulk(p.x + r * p.H, row_bytes);
            int s = 0, ph = 0;
            unsigned gp = 0;
            const int koff = (int)((blockIdx.x * 7u) % (unsigned)nkb);
            for (long long row = r0; row < r1; row += kFcBM) {
                const int nrows = (int)min((long long)kFcBM, r1 - row);
                const int nbox = (nrows + 31) >> 5;  // 32-row boxes of a partial tile
                const int nnext = (int)max(0ll, min((long long)kFcBM, r1 - (row + kFcBM)));  // rows of the next tile
                for (int kbi = 0; kbi < nkb; ++kbi) {
                    // every CTA walks the k-blocks from its own starting point (the sum over k does not care): all
                    // 148 CTAs fetching the SAME W' tile from the same L2 lines at the same moment was the limiter of
                    // the first version (ablation: 94 us with conversion and MMAs switched off)
                    const int kb = (kbi + koff) % nkb;
                    if (pf)
                        for (int i = kbi * nnext / nkb; i < (kbi + 1) * nnext / nkb; ++i)
                            l2_prefetch_bulk(p.x + (row + kFcBM + i) * p.H, row_bytes);
                    mbar_wait(&empty[s], ph ^ 1);  // passes at once the first time round the ring
                    FC_TRACE(0, gp);  // producer: slot free, about to issue
                    ++gp;
                    unsigned char *st = smem + (size_t)s * Cfg::stage_bytes;
                    if (nrows == kFcBM) {
                        mbar_arrive_expect_tx(&full[s], Cfg::a_bytes + Cfg::w_bytes);
                        tma_load_2d(st, &map_x, kb * kFcBK, (int)row, &full[s]);  // x tile [256][32]
                    } else {
                        mbar_arrive_expect_tx(&full[s], nbox * 32 * kFcBK * 4 + Cfg::w_bytes);
                        for (int j = 0; j < nbox; ++j)  // 32 rows = 4 swizzle atoms of 1024 bytes
                            tma_load_2d(st + j * 32 * kFcBK * 4, &map_x32, kb * kFcBK, (int)row + 32 * j, &full[s]);
                    }
                    tma_load_2d(st + Cfg::a_bytes, &map_w, kb * kFcBK, 0, &full[s]);  // W' tile [2N][32]
                    if (++s == S) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1 || warp == 10) {
        // ============================================================ MMA issuers (one lane each): warp 1 issues the
        // MMAs of the tile's first M = 128 half, warp 10 those of the second.  One thread issuing all 16 MMAs of a
        // k-block was the bottleneck of the kernel (clock64 trace, CTC_FC_TRACE: ~70 clk per issue, 1165 clk per
        // k-block against 670 for the converters); the halves have their own accumulators, so they need no ordering.
        const int h = warp == 1 ? 0 : 1;
        if (lane == 0) {
            constexpr uint32_t idesc2 = fc_idesc(kFcHM, 2 * N), idesc1 = fc_idesc(kFcHM, N);
            int s = 0, ph = 0, it = 0;
            unsigned g = 0;  // k-blocks issued so far (x_lo buffer = g % LB)
            for (long long row = r0; row < r1; row += kFcBM, ++it) {
                const int buf = it & 1;
                mbar_wait(&tempty[buf], ((it >> 1) & 1) ^ 1);  // the epilogue has drained this tile buffer
                tc_fence_after();
                for (int kb = 0; kb < nkb; ++kb, ++g) {
                    const int j = (int)(g % LB);
                    mbar_wait(&full[s], ph);
                    if (h == 0) FC_TRACE(4, g);  // MMA: tile landed
                    mbar_wait(&conv[j], (g / LB) & 1);
                    if (h == 0) FC_TRACE(5, g);  // MMA: x_lo ready
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)s * Cfg::stage_bytes);
                    const uint32_t sl = smem_u32(lo_base + (size_t)j * Cfg::a_bytes);
                    const uint64_t dw = fc_smem_desc(sa + Cfg::a_bytes);
                    {  // this issuer's M = 128 half of the tile (the halves share the W' tile)
                        const uint32_t d = tmem_base + (uint32_t)(((buf * 2 + h) * NA + kb % NA) * Cfg::acc_cols);
                        const uint64_t da = fc_smem_desc(sa + h * Cfg::h_bytes), dl = fc_smem_desc(sl + h * Cfg::h_bytes);
#pragma unroll
                        for (int k = 0; k < kFcBK / kFcUmmaK; ++k) {
                            const uint64_t o = (uint64_t)((k * kFcUmmaK * 4) >> 4);
                            // the first k-block of each accumulator overwrites, everything else accumulates
                            tc_mma_tf32(d, da + o, dw + o, idesc2, (kb >= NA || k > 0) ? 1u : 0u);  // x [W_hi | W_lo]^T
                            tc_mma_tf32(d, dl + o, dw + o, idesc1, 1u);                              // x_lo W_hi^T
                        }
                    }
                    tc_commit(&empty[s]);    // ring stage s may be refilled once these MMAs have read it
                    tc_commit(&lofree[j]);   // ... and x_lo buffer j rewritten
                    if (h == 0) FC_TRACE(6, g);  // MMA: issued + committed
                    if (++s == S) { s = 0; ph ^= 1; }
                }
                tc_commit(&tfull[buf]);
            }
        }
    } else if (warp < 6) {
        // ============================================================ converters: x_lo = x - (x with 13 low bits cleared)
        const int ct = tid - 64;  // 0..127
        int s = 0, ph = 0;
        unsigned g = 0;
        for (long long row = r0; row < r1; row += kFcBM) {
            for (int kb = 0; kb < nkb; ++kb, ++g) {
                const int j = (int)(g % LB);
                mbar_wait(&lofree[j], ((g / LB) & 1) ^ 1);  // passes at once for the first LB k-blocks
                if (ct == 0) FC_TRACE(1, g);  // converter: x_lo buffer free
                mbar_wait(&full[s], ph);
                if (ct == 0) FC_TRACE(2, g);  // converter: tile landed
                const float4 *a = reinterpret_cast<const float4 *>(smem + (size_t)s * Cfg::stage_bytes);
                float4 *lo = reinterpret_cast<float4 *>(lo_base + (size_t)j * Cfg::a_bytes);
                // elementwise, so the swizzled placement of the tile is irrelevant: same position in both tiles
                // all loads first (the two tiles may alias as far as the compiler knows: written naively, every load
                // waited for the previous store -- 8 dependent shared-memory round trips per k-block, which made this
                // loop the critical chain of the whole kernel: ncu, profiles/r02_fc_ncu_summary.md)
                constexpr int NV = Cfg::a_bytes / 16 / 128;
                float4 v[NV];
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] = a[ct + 128 * i];
#pragma unroll
                for (int i = 0; i < NV; ++i) {
                    float4 l;
                    l.x = v[i].x - __uint_as_float(__float_as_uint(v[i].x) & 0xffffe000u);
                    l.y = v[i].y - __uint_as_float(__float_as_uint(v[i].y) & 0xffffe000u);
                    l.z = v[i].z - __uint_as_float(__float_as_uint(v[i].z) & 0xffffe000u);
                    l.w = v[i].w - __uint_as_float(__float_as_uint(v[i].w) & 0xffffe000u);
                    lo[ct + 128 * i] = l;
                }
                fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
                mbar_arrive(&conv[j]);
                if (ct == 0) FC_TRACE(3, g);  // converter: done
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
    } else {
        // ============================================================ epilogue: one logits row per thread
        const int q = warp & 3;  // a warp may only touch TMEM lanes [32 * (warp % 4), +32)
        const int na = min(NA, nkb);  // accumulators that received at least one k-block
        int it = 0;
        for (long long row = r0; row < r1; row += kFcBM, ++it) {
            const int buf = it & 1;
            mbar_wait(&tfull[buf], (it >> 1) & 1);
            tc_fence_after();
          for (int h = 0; h < 2; ++h) {
            __syncwarp();  // lanes leave the previous half's body at different points; tcgen05.ld is warp-aligned
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((buf * 2 + h) * NA * Cfg::acc_cols);
            float v[N];
#pragma unroll
            for (int c0 = 0; c0 < N; c0 += 32) {
                float acc[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = 0.f;
                for (int a = 0; a < na; ++a) {
                    float hi[32], lo[32];
                    tmem_ld_32x32(taddr + a * Cfg::acc_cols + c0, hi);
                    tmem_ld_32x32(taddr + a * Cfg::acc_cols + N + c0, lo);
#pragma unroll
                    for (int i = 0; i < 32; ++i) acc[i] += hi[i] + lo[i];
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) v[c0 + i] = acc[i] + __ldg(p.bias + c0 + i);
            }
            if (h == 1) {
                tc_fence_before();
                mbar_arrive(&tempty[buf]);  // accumulators are in registers: the MMA warp may reuse the buffer
            }
            const long long rw = row + h * kFcHM + q * 32;  // first row of this warp's 32
            const long long r = rw + lane;
            if (p.logits) {
                // The warp's 32 rows x C logits are one contiguous block of memory.  One thread storing its row as C
                // scalars makes every store instruction touch 32 different lines (7680 L1 wavefronts per tile on the
                // load/store pipe the converters share: a 7 us stall at every tile boundary in the clock64 trace);
                // transposed through shuffles, lane c holds column c and every store instruction writes ONE row.
#pragma unroll
                for (int c0 = 0; c0 < N; c0 += 32) {
                    float w[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) w[i] = v[c0 + i];
                    warp_transpose32(w, lane);
                    const int c = c0 + lane;
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c < p.C && rw + i < r1) p.logits[(rw + i) * p.C + c] = w[i];
                }
            }
            if (r >= r1) continue;  // not this CTA's row (or past the end): never written
            if (p.y == nullptr && p.path == nullptr) continue;
            const int t = (int)(r / p.B), b = (int)(r - (long long)t * p.B);
            if (p.act_lens && t >= min(max(p.act_lens[b], 0), p.T)) continue;  // padding frame: never read
            float m = -INFINITY;
            int arg = 0;
#pragma unroll
            for (int c = 0; c < N; ++c)
                if (c < p.C && v[c] > m) { m = v[c]; arg = c; }  // strict >: the first maximum wins (torch.max)
            if (p.path) p.path[(size_t)b * p.T + t] = arg;
            if (p.y == nullptr) continue;
            // the softmax row exactly as ctc_prep_kernel forms it: expf, one division, clamp to >= 2^-126, planar words
            float sum = 0.f;
#pragma unroll
            for (int c = 0; c < N; ++c) {
                v[c] = (c < p.C) ? expf(v[c] - m) : 0.f;
                sum += v[c];
            }
            const float inv = 1.0f / sum;
            const bool bad = sum != sum;
            uint4 *yd = reinterpret_cast<uint4 *>(p.y + ((size_t)b * p.T + t) * kYP * N);
#pragma unroll
            for (int c = 0; c < N; c += 4) {
                uint4 hw, l;
                float e = fmaxf(v[c] * inv, 1.17549435e-38f);     hw.x = y_hi_word(e); l.x = y_lo_word(e);
                e = fmaxf(v[c + 1] * inv, 1.17549435e-38f);       hw.y = y_hi_word(e); l.y = y_lo_word(e);
                e = fmaxf(v[c + 2] * inv, 1.17549435e-38f);       hw.z = y_hi_word(e); l.z = y_lo_word(e);
                e = fmaxf(v[c + 3] * inv, 1.17549435e-38f);       hw.w = y_hi_word(e); l.w = y_lo_word(e);
                if (bad) { hw = make_uint4(0x7ff80000u, 0x7ff80000u, 0x7ff80000u, 0x7ff80000u); l = make_uint4(0, 0, 0, 0); }
                yd[c / 4] = hw;
                if (kYP == 2) yd[N / 4 + c / 4] = l;
            }
          }
        }
    }
    // ---- teardown: everybody is done with TMEM before its owner frees it
    tc_fence_before();
    __syncthreads();
    if (p.trace && tid == 0) p.trace[kFcTraceK * 8 + 2 * blockIdx.x + 1] = (long long)global_ns();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::tmem_cols);
    }
}


// ---------------------------------------------------------------- TS form: A operands from TMEM
// D[tmem] (+)= A[tmem] * B[smem]^T : A is M x 8 tf32 on lanes 0..M-1, 8 consecutive columns from `tmem_a`
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns: thread `lane` of the warp writes its lane's 32 values
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
        "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
        "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
// TMEM (512 columns): [0, 256) accumulators -- tile buffer b, half h at (b*2 + h) * 2N (2N <= 64 ... 128: see smem
// and column budget below); [256, 512) A operands -- k-block buffer j: x of half 0 / 1 at +0 / +32, x_lo at +64 / +96.
template <int N>
struct FcTsCfg {
    static_assert(N == 32, "the TS form is built for classes padded to 32 (2 * 2 * 2N accumulator columns + 256)");
    static constexpr int stages = 5;                        // TMA ring: x tile (32 KB) + W' tiles (8 KB) per stage
    static constexpr int a_bytes = kFcBM * kFcBK * 4;
    static constexpr int h_bytes = kFcHM * kFcBK * 4;
    static constexpr int w_bytes = 2 * N * kFcBK * 4;
    static constexpr int stage_bytes = a_bytes + w_bytes;
    static constexpr int acc_cols = 2 * N;
    static constexpr int tmem_cols = 512;
    __host__ __device__ static constexpr int a_col(int j, int h) { return 256 + j * 128 + h * 32; }
    static constexpr size_t smem = 1024 + (size_t)stages * stage_bytes + 256;
};

// The same kernel with the A operands in TMEM (see the converter role).  One accumulator per tile half (the TMEM
// columns the second one used now hold the A operands).
template <int N>
__global__ void __launch_bounds__(kFcThreads, 1)
    fc_logits_ts_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_x32,
                     const __grid_constant__ CUtensorMap map_w, const FcParams p) {
    using Cfg = FcTsCfg<N>;
    constexpr int S = Cfg::stages, LB = 2, NA = 1;
    extern __shared__ unsigned char fc_smem_raw[];
    // 128-byte swizzle: tile bases must be 1024-byte aligned
    // (as an offset from the shared array, so that the compiler keeps the shared address space: through a uintptr_t
    // round trip every access below became a generic LD / ST)
    unsigned char *smem = fc_smem_raw + ((1024u - (smem_u32(fc_smem_raw) & 1023u)) & 1023u);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)S * Cfg::stage_bytes);
    uint64_t *full = bars, *empty = bars + S, *conv = bars + 2 * S, *lofree = conv + LB, *tfull = lofree + LB,
             *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nkb = (p.H + kFcBK - 1) / kFcBK;
    const long long r0 = (long long)blockIdx.x * p.rows_per_cta;
    const long long r1 = min(p.rows, r0 + p.rows_per_cta);

    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 130);  // 128 converter threads have read the x tile + the two issuers' MMAs the W' tile
        }
        for (int j = 0; j < LB; ++j) {
            mbar_init(&conv[j], 128);
            mbar_init(&lofree[j], 2);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull[b], 2);
            mbar_init(&tempty[b], 128);
        }
        fence_mbar_init();
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_x32);
        tma_prefetch_desc(&map_w);
    }
    if (warp == 1) tmem_alloc(tmem_slot, Cfg::tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (p.trace && tid == 0) p.trace[kFcTraceK * 8 + 2 * blockIdx.x] = (long long)global_ns();

    if (warp == 0) {
        // ============================================================ TMA producer (one lane)
        if (lane == 0) {
            // A TMA box takes 128 bytes of each of 128 rows that lie 4*H bytes apart, and the next 128 bytes of a row
            // follow a whole k-block later: left alone, HBM sees scattered 128-byte reads (measured: 39 % of its
            // peak).  So whole ROWS of the next tile (4*H contiguous bytes each) are prefetched into L2 while the
            // current tile is consumed -- DRAM reads become sequential and the boxes hit L2.  The prefetches are
            // spread over the k-blocks of the current tile, so at most one tile per SM (~60 MB in all) waits in L2.
            const uint32_t row_bytes = (uint32_t)p.H * 4u;
            const bool pf = (row_bytes & 15u) == 0 && CTC_FC_L2_PREFETCH;
            if (pf)
                for (long long r = r0; r < min(r1, r0 + kFcBM); ++r) l2_prefetch_bulk(p.x + r * p.H, row_bytes);
            int s = 0, ph = 0;
            unsigned gp = 0;
            const int koff = (int)((blockIdx.x * 7u) % (unsigned)nkb);
            for (long long row = r0; row < r1; row += kFcBM) {
                const int nrows = (int)min((long long)kFcBM, r1 - row);
                const int nbox = (nrows + 31) >> 5;  // 32-row boxes of a partial tile
                const int nnext = (int)max(0ll, min((long long)kFcBM, r1 - (row + kFcBM)));  // rows of the next tile
                for (int kbi = 0; kbi < nkb; ++kbi) {
                    // every CTA walks the k-blocks from its own starting point (the sum over k does not care): all
                    // 148 CTAs fetching the SAME W' tile from the same L2 lines at the same moment was the limiter of
                    // the first version (ablation: 94 us with conversion and MMAs switched off)
                    const int kb = (kbi + koff) % nkb;
                    if (pf)
                        for (int i = kbi * nnext / nkb; i < (kbi + 1) * nnext / nkb; ++i)
                            l2_prefetch_bulk(p.x + (row + kFcBM + i) * p.H, row_bytes);
                    mbar_wait(&empty[s], ph ^ 1);  // passes at once the first time round the ring
                    FC_TRACE(0, gp);  // producer: slot free, about to issue
                    ++gp;
                    unsigned char *st = smem + (size_t)s * Cfg::stage_bytes;
                    if (nrows == kFcBM) {
                        mbar_arrive_expect_tx(&full[s], Cfg::a_bytes + Cfg::w_bytes);
                        tma_load_2d(st, &map_x, kb * kFcBK, (int)row, &full[s]);  // x tile [256][32]
                    } else {
                        mbar_arrive_expect_tx(&full[s], nbox * 32 * kFcBK * 4 + Cfg::w_bytes);
                        for (int j = 0; j < nbox; ++j)  // 32 rows = 4 swizzle atoms of 1024 bytes
                            tma_load_2d(st + j * 32 * kFcBK * 4, &map_x32, kb * kFcBK, (int)row + 32 * j, &full[s]);
                    }
                    tma_load_2d(st + Cfg::a_bytes, &map_w, kb * kFcBK, 0, &full[s]);  // W' tile [2N][32]
                    if (++s == S) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1 || warp == 10) {
        // ============================================================ MMA issuers (one lane each): warp 1 issues the
        // MMAs of the tile's first M = 128 half, warp 10 those of the second.  One thread issuing all 16 MMAs of a
        // k-block was the bottleneck of the kernel (clock64 trace, CTC_FC_TRACE: ~70 clk per issue, 1165 clk per
        // k-block against 670 for the converters); the halves have their own accumulators, so they need no ordering.
        const int h = warp == 1 ? 0 : 1;
        if (lane == 0) {
            constexpr uint32_t idesc2 = fc_idesc(kFcHM, 2 * N), idesc1 = fc_idesc(kFcHM, N);
            int s = 0, ph = 0, it = 0;
            unsigned g = 0;  // k-blocks issued so far (x_lo buffer = g % LB)
            for (long long row = r0; row < r1; row += kFcBM, ++it) {
                const int buf = it & 1;
                mbar_wait(&tempty[buf], ((it >> 1) & 1) ^ 1);  // the epilogue has drained this tile buffer
                tc_fence_after();
                for (int kb = 0; kb < nkb; ++kb, ++g) {
                    const int j = (int)(g % LB);
                    mbar_wait(&full[s], ph);
                    if (h == 0) FC_TRACE(4, g);  // MMA: tile landed
                    mbar_wait(&conv[j], (g / LB) & 1);
                    if (h == 0) FC_TRACE(5, g);  // MMA: x_lo ready
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)s * Cfg::stage_bytes);
                    const uint64_t dw = fc_smem_desc(sa + Cfg::a_bytes);
                    // A operands in TMEM: x of this half at Cfg::a_col(j, h), x_lo 64 columns further
                    const uint32_t ax = tmem_base + (uint32_t)Cfg::a_col(j, h);
                    {  // this issuer's M = 128 half of the tile (the halves share the W' tile)
                        const uint32_t d = tmem_base + (uint32_t)(((buf * 2 + h) * NA + kb % NA) * Cfg::acc_cols);
#pragma unroll
                        for (int k = 0; k < kFcBK / kFcUmmaK; ++k) {
                            const uint64_t o = (uint64_t)((k * kFcUmmaK * 4) >> 4);
                            // the first k-block of each accumulator overwrites, everything else accumulates
                            tc_mma_tf32_ts(d, ax + 8 * k, dw + o, idesc2, (kb >= NA || k > 0) ? 1u : 0u);  // x [W_hi | W_lo]^T
                            tc_mma_tf32_ts(d, ax + 64 + 8 * k, dw + o, idesc1, 1u);                       // x_lo W_hi^T
                        }
                    }
                    tc_commit(&empty[s]);    // ring stage s may be refilled once these MMAs have read it
                    tc_commit(&lofree[j]);   // ... and x_lo buffer j rewritten
                    if (h == 0) FC_TRACE(6, g);  // MMA: issued + committed
                    if (++s == S) { s = 0; ph ^= 1; }
                }
                tc_commit(&tfull[buf]);
            }
        }
    } else if (warp < 6) {
        // ============================================================ converters: one tile row per thread and half.  The
        // thread reads its row's 32 fp32 of the k-block from the swizzled tile (16-byte chunk c of row r sits at chunk
        // position c ^ (r & 7)) and writes x and x_lo = x - (x with 13 low bits cleared) into TMEM with tcgen05.st --
        // a warp may only touch TMEM lanes [32 * (warp % 4), +32), and A operands put tile row m on lane m, so warp
        // quarter q handles rows q*32 + lane of BOTH halves.  The tensor core then reads A from TMEM: no x_lo tile in
        // shared memory, no A-operand reads from shared memory (half the shared-memory traffic of the SS form).
        const int qc = warp & 3;
        const int rt = qc * 32 + lane;  // row inside a half
        int s = 0, ph = 0;
        unsigned g = 0;
        for (long long row = r0; row < r1; row += kFcBM) {
            for (int kb = 0; kb < nkb; ++kb, ++g) {
                const int j = (int)(g % LB);
                mbar_wait(&lofree[j], ((g / LB) & 1) ^ 1);  // passes at once for the first LB k-blocks
                if (tid == 64) FC_TRACE(1, g);  // converter: TMEM A buffer free
                mbar_wait(&full[s], ph);
                if (tid == 64) FC_TRACE(2, g);  // converter: tile landed
                tc_fence_after();
                const unsigned char *tile = smem + (size_t)s * Cfg::stage_bytes;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int r = h * kFcHM + rt;
                    const unsigned char *rp = tile + (size_t)r * 128;
                    uint32_t xv[32], lv[32];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const uint4 w = *reinterpret_cast<const uint4 *>(rp + ((c ^ (r & 7)) << 4));
                        xv[4 * c + 0] = w.x; xv[4 * c + 1] = w.y; xv[4 * c + 2] = w.z; xv[4 * c + 3] = w.w;
                    }
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        lv[i] = __float_as_uint(__uint_as_float(xv[i]) - __uint_as_float(xv[i] & 0xffffe000u));
                    const uint32_t ta = tmem_base + ((uint32_t)(qc * 32) << 16) + (uint32_t)Cfg::a_col(j, h);
                    tmem_st_32x32(ta, xv);
                    tmem_st_32x32(ta + 64, lv);
                }
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc_fence_before();
                mbar_arrive(&conv[j]);   // A operands of this k-block are in TMEM
                mbar_arrive(&empty[s]);  // ... and this thread is done with the x tile of ring stage s
                if (tid == 64) FC_TRACE(3, g);  // converter: done
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
    } else {
        // ============================================================ epilogue: one logits row per thread
        const int q = warp & 3;  // a warp may only touch TMEM lanes [32 * (warp % 4), +32)
        const int na = min(NA, nkb);  // accumulators that received at least one k-block
        int it = 0;
        for (long long row = r0; row < r1; row += kFcBM, ++it) {
            const int buf = it & 1;
            mbar_wait(&tfull[buf], (it >> 1) & 1);
            tc_fence_after();
          for (int h = 0; h < 2; ++h) {
            __syncwarp();  // lanes leave the previous half's body at different points; tcgen05.ld is warp-aligned
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((buf * 2 + h) * NA * Cfg::acc_cols);
            float v[N];
#pragma unroll
            for (int c0 = 0; c0 < N; c0 += 32) {
                float acc[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = 0.f;
                for (int a = 0; a < na; ++a) {
                    float hi[32], lo[32];
                    tmem_ld_32x32(taddr + a * Cfg::acc_cols + c0, hi);
                    tmem_ld_32x32(taddr + a * Cfg::acc_cols + N + c0, lo);
#pragma unroll
                    for (int i = 0; i < 32; ++i) acc[i] += hi[i] + lo[i];
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) v[c0 + i] = acc[i] + __ldg(p.bias + c0 + i);
            }
            if (h == 1) {
                tc_fence_before();
                mbar_arrive(&tempty[buf]);  // accumulators are in registers: the MMA warp may reuse the buffer
            }
            const long long rw = row + h * kFcHM + q * 32;  // first row of this warp's 32
            const long long r = rw + lane;
            if (p.logits) {
                // The warp's 32 rows x C logits are one contiguous block of memory.  One thread storing its row as C
                // scalars makes every store instruction touch 32 different lines (7680 L1 wavefronts per tile on the
                // load/store pipe the converters share: a 7 us stall at every tile boundary in the clock64 trace);
                // transposed through shuffles, lane c holds column c and every store instruction writes ONE row.
#pragma unroll
                for (int c0 = 0; c0 < N; c0 += 32) {
                    float w[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) w[i] = v[c0 + i];
                    warp_transpose32(w, lane);
                    const int c = c0 + lane;
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c < p.C && rw + i < r1) p.logits[(rw + i) * p.C + c] = w[i];
                }
            }
            if (r >= r1) continue;  // not this CTA's row (or past the end): never written
            if (p.y == nullptr && p.path == nullptr) continue;
            const int t = (int)(r / p.B), b = (int)(r - (long long)t * p.B);
            if (p.act_lens && t >= min(max(p.act_lens[b], 0), p.T)) continue;  // padding frame: never read
            float m = -INFINITY;
            int arg = 0;
#pragma unroll
            for (int c = 0; c < N; ++c)
                if (c < p.C && v[c] > m) { m = v[c]; arg = c; }  // strict >: the first maximum wins (torch.max)
            if (p.path) p.path[(size_t)b * p.T + t] = arg;
            if (p.y == nullptr) continue;
            // the softmax row exactly as ctc_prep_kernel forms it: expf, one division, clamp to >= 2^-126, planar words
            float sum = 0.f;
#pragma unroll
            for (int c = 0; c < N; ++c) {
                v[c] = (c < p.C) ? expf(v[c] - m) : 0.f;
                sum += v[c];
            }
            const float inv = 1.0f / sum;
            const bool bad = sum != sum;
            uint4 *yd = reinterpret_cast<uint4 *>(p.y + ((size_t)b * p.T + t) * kYP * N);
#pragma unroll
            for (int c = 0; c < N; c += 4) {
                uint4 hw, l;
                float e = fmaxf(v[c] * inv, 1.17549435e-38f);     hw.x = y_hi_word(e); l.x = y_lo_word(e);
                e = fmaxf(v[c + 1] * inv, 1.17549435e-38f);       hw.y = y_hi_word(e); l.y = y_lo_word(e);
                e = fmaxf(v[c + 2] * inv, 1.17549435e-38f);       hw.z = y_hi_word(e); l.z = y_lo_word(e);
                e = fmaxf(v[c + 3] * inv, 1.17549435e-38f);       hw.w = y_hi_word(e); l.w = y_lo_word(e);
                if (bad) { hw = make_uint4(0x7ff80000u, 0x7ff80000u, 0x7ff80000u, 0x7ff80000u); l = make_uint4(0, 0, 0, 0); }
                yd[c / 4] = hw;
                if (kYP == 2) yd[N / 4 + c / 4] = l;
            }
          }
        }
    }
    // ---- teardown: everybody is done with TMEM before its owner frees it
    tc_fence_before();
    __syncthreads();
    if (p.trace && tid == 0) p.trace[kFcTraceK * 8 + 2 * blockIdx.x + 1] = (long long)global_ns();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::tmem_cols);
    }
}


}  // namespace ctcb200
