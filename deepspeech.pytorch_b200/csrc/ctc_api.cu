// csrc/ctc_api.cu -- C ABI of libctc_b200.so (see include/ctc_b200.h for the contract and the reference
// interfaces each entry point replaces: warp-ctc's get_workspace_size / compute_ctc_loss / ctcGetStatusString /
// get_warpctc_version reached from /root/reference/train.py:12,904,600,352,457 through warpctc_pytorch).
//
// Host side only plans the launch (shape -> lattice warps, rows per checkpoint block, shared memory, workspace
// carving, length-bucket slices) and enqueues; it never synchronises the device.  State kept per device: a
// "max dynamic smem attribute already set" cache and a small pool of internal side streams + events on which the
// per-slice kernels are forked from and joined back to the caller's stream.
#include <cuda_runtime.h>
#include <chrono>
#include <stdio.h>
#include <mutex>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <new>

#include "../../include/ctc_b200.h"
#include "ctc_kernels.cuh"
#include "ctc_metrics.cuh"
#include "fc_kernels.cuh"

using namespace ctcb200;

// Communicator of the fused loss all-reduce (see ctc_sum_costs_kernel): this rank's slot buffer + the peers' buffers
// mapped through CUDA IPC (or given as raw pointers when all ranks live in one process).
struct ctcB200Comm {
    int rank = 0, world = 1;
    unsigned long long *local = nullptr;       // cudaMalloc: [2][kCommMaxWorld] slots + counter
    unsigned long long **peers_dev = nullptr;  // device array [world]
    void *opened[kCommMaxWorld] = {};
    bool ready = false;
    unsigned long long timeout_ns = 120ull * 1000000000ull;  // wall-clock bound of one exchange
    int *timed_out_host = nullptr;             // pinned, host-mapped: set by the kernel when a peer never arrived
    int *timed_out_dev = nullptr;              // the same word as the device sees it
};

namespace {

constexpr int kK = 8;               // states per lattice thread
constexpr size_t kSmemLimit = 227 * 1024;
thread_local int g_last_launches = 0;
thread_local unsigned long long g_last_h2d = 0, g_last_d2h = 0;  // bytes the last ctc_b200_compute_host copied

// where the summed loss goes, and (multi-GPU) the communicator its kernel all-reduces through
struct LossOut {
    float *loss_sum = nullptr;  // this rank's sum
    float *loss_all = nullptr;  // sum over all ranks of the communicator
    CommArgs comm = {nullptr, nullptr, 0, 1, 0ull, nullptr};
    explicit operator bool() const { return loss_sum != nullptr || loss_all != nullptr; }
};

// ---- length-bucket slices (see sort_by_length in ctc_kernels.cuh)
constexpr int kMaxSlices = 16;     // upper bound (CTC_B200_SLICES)
constexpr int kDefaultSlices = 8;
constexpr int kMinSliceB = 8;  // utterances per slice below which slicing stops paying
struct SidePool {
    bool ok = false;
    cudaStream_t side[kMaxSlices], side2[kMaxSlices], aux;  // side2: the small warp class of a slice's gradient work
    cudaEvent_t fork, join[kMaxSlices], join2[kMaxSlices], eng_done[kMaxSlices], eng_done2[kMaxSlices], aux_done;
};
SidePool g_pool[64];
std::mutex g_pool_mu;  // held while a sliced launch is being enqueued: the pool's events are shared by all callers
std::mutex g_attr_mu;  // guards the "max dynamic shared memory already raised on this device" caches below

SidePool *side_pool(int dev) {
    if (dev < 0 || dev >= 64) return nullptr;
    SidePool &sp = g_pool[dev];
    if (sp.ok) return &sp;
    if (cudaEventCreateWithFlags(&sp.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&sp.aux_done, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaStreamCreateWithFlags(&sp.aux, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    for (int k = 0; k < kMaxSlices; ++k) {
        if (cudaStreamCreateWithFlags(&sp.side[k], cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        if (cudaStreamCreateWithFlags(&sp.side2[k], cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&sp.join[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&sp.join2[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&sp.eng_done2[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&sp.eng_done[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
    }
    sp.ok = true;
    return &sp;
}

int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// Two forms of the recursion kernel.  The second (ctc_engine2_kernel) has the shorter dependent chain per group and
// wins whenever an engine CTA has an SM to itself (B=64: 137 against 157 us; B=32,T=4000: 346 against 394); the first
// issues fewer instructions per step (its warp-edge accesses are branches, not predicated instructions on all lanes)
// and wins once three or more engine CTAs share an SM and the kernel is bound by issue slots instead of latency (B=256:
// 259 against 270 us, B=1024: 731 against 790).  CTC_B200_ENGINE=1 / 2 forces one form (A/B measurements); 0 = by batch.
int engine_version() { return env_int("CTC_B200_ENGINE", 0); }  // read per call: tests switch it inside one process
inline bool engine_first_form(int B) {
    const int v = engine_version();
    // measured crossover (T=1500, C=29, us, first / second form): B=96 192 / 154, B=128 192 / 164, B=160 188 / 168,
    // B=192 236 / 207, B=256 259 / 270, B=1024 731 / 790
    return v == 1 || (v == 0 && B > 224);
}

struct Plan {
    int nwl;   // lattice warps per CTA
    int R;     // rows per checkpoint block (16 or 8)
    int nblk;  // ceil(T / R)
    int Lmax;
    int Cp;    // emission-row pitch in floats (C rounded up to a multiple of 4, >= 32 words for C <= 32: rows are
               // 16-byte multiples for TMA and gathers are bank-conflict-free)
    size_t smem_prep, smem_engine, smem_grad;
    // workspace byte offsets
    int nqmax;   // grad kernel warps per CTA
    int nq_small;        // > 0: utterances that need <= nq_small warps get their own launch with CTAs of that size
    int nsplit_small;    // ... its CTAs per utterance
    size_t smem_grad_small;
    int nsplit;  // grad kernel CTAs per utterance
    int big;     // large-alphabet path (C > kMaxSmallC)
    size_t o_ctab, o_bord;
    size_t smem_rows;  // large alphabets: dynamic shared memory of the row kernel
    int nslices;           // length-bucket slices the engine / grad launches are cut into (1 = unsliced)
    size_t smem_engine_req;  // dynamic shared memory REQUESTED for an engine CTA (>= smem_engine, see make_plan)
    size_t o_y, o_ckA, o_ckB, o_ckAE, o_ckBE, o_Pm, o_Pe, o_feas, o_laboff, o_tlab, o_labels, o_llen, o_alen;
    size_t o_Pm2, o_Pe2, o_est;
    int two_part;          // 1: recursions run in two launches with the early gradient blocks in between
    int cluster;           // 1: engines run as 2-CTA clusters (more than 4 lattice warps, see ctc_engine2c_kernel)
    size_t smem_engine_c;  // dynamic shared memory of one CTA of such a cluster
    size_t total;
};

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Shape -> launch plan.  Returns false if the shape is outside what this build supports.
bool make_plan(int T, int B, int C, int Lmax, Plan *pl) {
    if (T < 0 || B < 1 || C < 2 || Lmax < 0) return false;
    pl->big = C > kMaxSmallC;
    if (pl->big && Lmax > 511) return false;  // position-indexed emission rows: 1 + L <= 512 (this build)
    const int S = 2 * Lmax + 1;
    const int nwl = (S + 32 * kK - 1) / (32 * kK);
    if (nwl > kMaxLatticeWarps) return false;  // label length > 1023
    pl->nwl = nwl;
    pl->Lmax = Lmax;
    pl->Cp = pl->big ? (Lmax <= 255 ? 256 : 512) : (C <= 32 ? 32 : (C <= 64 ? 64 : 128));
    const int NT = nwl * 32;
    // one checkpoint per group of 8 time steps: the grad kernel then never renormalises or re-settles exponents
    const int R = kGroup;
    pl->nqmax = (S + kK * kCoreThreads - 1) / (kK * kCoreThreads);
    if (pl->nqmax > kMaxGradWarps) return false;
    if (grad_smem(C, pl->Cp, Lmax, pl->nqmax, kK, pl->big != 0).total > kSmemLimit) return false;
    pl->R = R;
    pl->nblk = T > 0 ? (T + R - 1) / R : 1;
    pl->smem_prep = pl->big ? prep_big_smem(C, Lmax) : prep_smem(C, Lmax);
    pl->smem_rows = pl->big ? prep_big_row_smem(C) : 0;
    pl->smem_engine = engine_smem(pl->Cp, nwl).total;
    pl->smem_grad = grad_smem(C, pl->Cp, Lmax, pl->nqmax, kK, pl->big != 0).total;
    {
        // deal the time groups of an utterance round-robin to enough CTAs for ~7 waves of resident CTAs: utterances
        // differ in length and lattice width, so many small CTAs balance better than a few long ones (measured on
        // the headline shape: 10 CTAs/utterance 154 us, 19: 141, 32: 139, 63: 141)
        // resident CTAs per SM: shared memory, and 128 registers per thread (the kernel's __maxnreg__)
        const int by_smem = (int)(kSmemLimit / pl->smem_grad) > 0 ? (int)(kSmemLimit / pl->smem_grad) : 1;
        const int by_regs = 65536 / (128 * 32 * pl->nqmax) > 0 ? 65536 / (128 * 32 * pl->nqmax) : 1;
        const int resident = by_smem < by_regs ? by_smem : by_regs;
        int ns = (7 * 148 * resident + B - 1) / B;
        if (ns > pl->nblk / 2) ns = pl->nblk / 2;  // >= 2 groups per CTA (measured 2: 0.233, 4: 0.235, 6: 0.240 ms/step)
        ns = env_int("CTC_B200_NSPLIT", ns);       // tuning experiments
        if (ns > pl->nblk) ns = pl->nblk;
        if (ns < 1) ns = 1;
        pl->nsplit = ns;
        // Warp classes (small alphabets, >= 2 gradient warps): an utterance whose lattice fits half of the warps would
        // leave the other half of a full-size CTA idle -- and this kernel's throughput follows the number of warps
        // that have work -- so those utterances get a launch of their own with half-size CTAs (half the registers,
        // about half the shared memory: twice as many resident).  Gradient kernel alone (tools/ubench/engine_bench):
        // B=1024 1085 -> 944 us.  Every class launch covers all utterances of its slice (the host does not see
        // device-side lengths), so the CTAs of the other class's utterances start and leave at once -- but a
        // full-size CTA needs a full-size slot even for that: with few utterances per slice (B=64, 8 slices of 8) the
        // nearly empty full-size launches of the short slices queue behind resident work and hold the step's join
        // back.  Whole step, one launch against two classes (tools/grad_classes_sweep.py): B=64 179 / 194 us, B=128
        // 269 / 277, B=192 380 / 369, B=256 478 / 471, B=1024 1704 / 1601.  Hence by batch size;
        // CTC_B200_GRAD_CLASSES=0 / 1 forces one launch for all / the two classes.
        const int cls = env_int("CTC_B200_GRAD_CLASSES", -1);
        pl->nq_small = (!pl->big && pl->nqmax >= 2 && (cls < 0 ? B >= 192 : cls != 0)) ? pl->nqmax / 2 : 0;
        pl->smem_grad_small = 0;
        pl->nsplit_small = ns;
        if (pl->nq_small) {
            pl->smem_grad_small =
                grad_smem(C, pl->Cp, grad_class_lmax(pl->nq_small, Lmax), pl->nq_small, kK, false).total;
            int res = (int)(kSmemLimit / pl->smem_grad_small) > 0 ? (int)(kSmemLimit / pl->smem_grad_small) : 1;
            const int res_regs = 65536 / (128 * 32 * pl->nq_small);
            if (res > res_regs) res = res_regs;
            int n2 = (7 * 148 * res + B - 1) / B;
            if (n2 > pl->nblk / 2) n2 = pl->nblk / 2;
            pl->nsplit_small = n2 < 1 ? 1 : n2;
        }
    }
    if (pl->smem_engine > kSmemLimit || pl->smem_prep > kSmemLimit) return false;
    // more than 4 lattice warps (L > 511): split them over the two CTAs of a cluster so that each has an SMSP to
    // itself.  CTC_B200_CLUSTER=0 keeps everything in one CTA (A/B measurements).
    pl->cluster = (nwl > 4 && !engine_first_form(B) && env_int("CTC_B200_CLUSTER", 1) != 0) ? 1 : 0;
    pl->smem_engine_c = engine_smem(pl->Cp, (nwl + 1) / 2).total;
    // Two-part recursion (engine_phase_range; second engine form, one CTA per utterance and direction): OFF by default.
    // CTC_B200_TWO_PART=2 runs the two parts back to back (bit-identical results: test_two_part_recursion_...).
    // Measured on the headline: the split costs 8.7 us of the recursion (pipeline drain + refill + launch: 132.2 ->
    // 140.9 us), and a work-conserving model of the slices on this batch says that starting the gradient of the middle
    // time blocks at 3/4 of each recursion can bring the end of the step forward by at most ~10 us (185 -> 175 us
    // modelled, before the part-2 engine CTAs have to queue behind resident gradient CTAs for registers) -- so the
    // mid-point P kernel, the gradient kernel's block filter and the second set of streams were not built.
    pl->two_part = (!engine_first_form(B) && !pl->cluster && pl->nblk >= 16) ? env_int("CTC_B200_TWO_PART", 0) : 0;
    {
        // Length-bucket slices: up to kMaxSlices, each at least kMinSliceB utterances.  CTC_B200_SLICES overrides.
        int ns = B / kMinSliceB;
        ns = ns < 1 ? 1 : (ns > kDefaultSlices ? kDefaultSlices : ns);
        ns = env_int("CTC_B200_SLICES", ns);
        ns = ns < 1 ? 1 : (ns > kMaxSlices ? kMaxSlices : ns);
        if (ns > B || B > kMaxSortB) ns = 1;
        pl->nslices = ns;
        if (ns > 1 && pl->smem_prep < sizeof(int) * (size_t)B) pl->smem_prep = sizeof(int) * (size_t)B;
        // (Padding the engine's shared-memory request to keep gradient CTAs of other slices off its SM was measured:
        // no effect at 0 / 120 / 180 KB, so engines simply request what they use.  Likewise one engine CTA per SM
        // when 2B exceeds the SM count: B=128 recursion 155 -> 174 us, B=160 163 -> 218: two CTAs share an SM better
        // than the second one waits for it -- tools/batch_size_probe.py.)
        const size_t req = 0;
        pl->smem_engine_req = req > pl->smem_engine ? req : pl->smem_engine;
    }
    size_t o = 0;
    const size_t ck = sizeof(int) * (size_t)B * pl->nblk * NT * kK;
    const size_t ckE = sizeof(int) * (size_t)B * pl->nblk * NT;
    pl->o_y = o;      o = align256(o + sizeof(unsigned) * (size_t)T * B * kYP * pl->Cp);
    pl->o_ckA = o;    o = align256(o + ck);
    pl->o_ckB = o;    o = align256(o + ck);
    pl->o_ckAE = o;   o = align256(o + ckE);
    pl->o_ckBE = o;   o = align256(o + ckE);
    pl->o_est = o;    o = align256(o + sizeof(double) * (size_t)B * 2 * kEngStateFields * NT);
    pl->o_Pm2 = o;    o = align256(o + sizeof(double) * 2 * B);
    pl->o_Pe2 = o;    o = align256(o + sizeof(int) * 2 * B);
    pl->o_Pm = o;     o = align256(o + sizeof(double) * B);
    pl->o_Pe = o;     o = align256(o + sizeof(int) * B);
    pl->o_feas = o;   o = align256(o + sizeof(int) * B);
    pl->o_laboff = o; o = align256(o + sizeof(int) * B);
    pl->o_tlab = o;   o = align256(o + sizeof(uint4) * (size_t)B * NT);
    pl->o_ctab = o;    if (pl->big) o = align256(o + sizeof(int2) * (size_t)B * (Lmax > 0 ? Lmax : 1));
    pl->o_bord = o;   o = align256(o + sizeof(int) * (size_t)B);
    pl->o_labels = o; o = align256(o + sizeof(int) * (size_t)B * (Lmax > 0 ? Lmax : 1));
    pl->o_llen = o;   o = align256(o + sizeof(int) * B);
    pl->o_alen = o;   o = align256(o + sizeof(int) * B);
    pl->total = o;
    return true;
}

// Engine + gradient launches, cut into length-bucket slices on the pool's side streams (forked from and joined back
// to the caller's stream with events, so the caller sees ordinary stream semantics and graph capture still works).
// All engines are enqueued before any gradient kernel: the longest slice (rank 0..) starts first.
// Slices are equal rank ranges.  Measured and rejected (headline shape): TRIANGULAR slice sizes (few longest
// utterances in slice 0, its time groups spread over one full wave of CTAs) -- 0.225 against 0.223 ms/step: the end
// of the step is bound by the total gradient work that only becomes available late (CTC_B200_TRACE shows every
// slice's gradient kernel finishing ~30 us after its engine), not by the size of the last slice.
inline int slice_begin(const Plan &pl, int B, int k) {
    const int ns = pl.nslices;
    return k <= 0 ? 0 : (k >= ns ? B : (int)((long long)B * k / ns));
}

// The collapse of the best path (decode by-product of the prep kernel) is independent of the lattice kernels: it is
// enqueued behind the engine of the LAST slice (the shortest utterances, far from the critical path).
inline void launch_collapse(const CtcParams &p, cudaStream_t s, int *launches) {
    if (!p.path || !p.dec_ids) return;
    ctc_collapse_kernel<<<p.B, kCollapseThreads, 0, s>>>(p.path, p.act_lens, p.T, p.B, p.blank, p.dec_ids, p.dec_off,
                                                         p.dec_len);
    ++*launches;
}

template <typename EngineLaunch, typename GradLaunch>
cudaError_t launch_sliced(const CtcParams &p, const Plan &pl, cudaStream_t st, bool do_engine, bool do_grad,
                          int *launches, cudaEvent_t costs_ready, const LossOut &loss_sum, EngineLaunch eng, GradLaunch grd) {
    if (pl.nslices <= 1 || (!do_engine && !do_grad)) {
        if (do_engine) eng(p, p.B, st);
        if (loss_sum && do_engine) {
            ctc_sum_costs_kernel<<<1, 32, 0, st>>>(p.costs, p.B, loss_sum.loss_sum, loss_sum.loss_all, loss_sum.comm);
            ++*launches;
        }
        if (costs_ready && cudaEventRecord(costs_ready, st) != cudaSuccess) return cudaErrorUnknown;
        if (do_grad) grd(p, p.B, st, st);
        launch_collapse(p, st, launches);
        return cudaGetLastError();
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(g_pool_mu);
    SidePool *sp = side_pool(dev);
    if (!sp) return cudaErrorUnknown;
    const int ns = pl.nslices;
    // CTC_B200_TRACE=1 (diagnostic; synchronises the caller's stream): device timeline of the sliced launch, per
    // slice the time its engine and its gradient kernel finished, in us after the fork point (= end of the prep kernel)
    static const bool trace = env_int("CTC_B200_TRACE", 0) != 0;
    cudaEvent_t tr0 = nullptr, tr[kMaxSlices][2] = {};
    if (trace) {
        cudaEventCreate(&tr0);
        for (int k = 0; k < ns; ++k) { cudaEventCreate(&tr[k][0]); cudaEventCreate(&tr[k][1]); }
        cudaEventRecord(tr0, st);
    }
    if ((e = cudaEventRecord(sp->fork, st)) != cudaSuccess) return e;
    for (int k = 0; k < ns; ++k)
        if ((e = cudaStreamWaitEvent(sp->side[k], sp->fork, 0)) != cudaSuccess) return e;
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 0 ? !do_engine : !do_grad) continue;
        for (int k = 0; k < ns; ++k) {
            CtcParams pk = p;
            pk.slice0 = slice_begin(pl, p.B, k);
            const int nb = slice_begin(pl, p.B, k + 1) - pk.slice0;
            if (pass == 0) {
                eng(pk, nb, sp->side[k]);
            } else if (pl.nq_small) {
                // the slice's small warp class runs beside its full-size one, on a stream of its own
                if ((e = cudaEventRecord(sp->eng_done2[k], sp->side[k])) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(sp->side2[k], sp->eng_done2[k], 0)) != cudaSuccess) return e;
                grd(pk, nb, sp->side[k], sp->side2[k]);
                if ((e = cudaEventRecord(sp->join2[k], sp->side2[k])) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(st, sp->join2[k], 0)) != cudaSuccess) return e;
            } else {
                grd(pk, nb, sp->side[k], sp->side[k]);
            }
            if (trace) cudaEventRecord(tr[k][pass], sp->side[k]);
        }
        if (pass == 0) {
            if (costs_ready || loss_sum) {
                // costs[] is final once every slice's recursion is done: gather those points on the aux stream, sum
                // the costs and record the caller's event there; the gradient kernels are enqueued behind this and
                // run on
                for (int k = 0; k < ns; ++k) {
                    if ((e = cudaEventRecord(sp->eng_done[k], sp->side[k])) != cudaSuccess) return e;
                    if ((e = cudaStreamWaitEvent(sp->aux, sp->eng_done[k], 0)) != cudaSuccess) return e;
                }
                if (loss_sum) {
                    ctc_sum_costs_kernel<<<1, 32, 0, sp->aux>>>(p.costs, p.B, loss_sum.loss_sum, loss_sum.loss_all, loss_sum.comm);
                    ++*launches;
                }
                if (costs_ready && (e = cudaEventRecord(costs_ready, sp->aux)) != cudaSuccess) return e;
                if ((e = cudaEventRecord(sp->aux_done, sp->aux)) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(st, sp->aux_done, 0)) != cudaSuccess) return e;
            }
            launch_collapse(p, sp->side[ns - 1], launches);
        }
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    for (int k = 0; k < ns; ++k) {
        if ((e = cudaEventRecord(sp->join[k], sp->side[k])) != cudaSuccess) return e;
        if ((e = cudaStreamWaitEvent(st, sp->join[k], 0)) != cudaSuccess) return e;
    }
    if (trace) {
        cudaStreamSynchronize(st);
        fprintf(stderr, "[ctc_b200 trace]");
        for (int k = 0; k < ns; ++k) {
            float a = 0, b = 0;
            if (do_engine) cudaEventElapsedTime(&a, tr0, tr[k][0]);
            if (do_grad) cudaEventElapsedTime(&b, tr0, tr[k][1]);
            fprintf(stderr, " s%d[%d..%d) eng %.0f grad %.0f;", k, slice_begin(pl, p.B, k), slice_begin(pl, p.B, k + 1),
                    a * 1e3, b * 1e3);
            cudaEventDestroy(tr[k][0]);
            cudaEventDestroy(tr[k][1]);
        }
        fprintf(stderr, "\n");
        cudaEventDestroy(tr0);
    }
    return cudaSuccess;
}

template <int CP>
cudaError_t launch_all_cp(const CtcParams &p, const Plan &pl, cudaStream_t st, int phase_mask, int *launches,
                          cudaEvent_t costs_ready, const LossOut &loss_sum) {
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    std::unique_lock<std::mutex> attr_lock(g_attr_mu);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(ctc_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_engine_kernel<kK, CP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_engine2_kernel<kK, CP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_engine2c_kernel<kK, CP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_grad_kernel<kK, CP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    attr_lock.unlock();
    cudaError_t e = cudaSuccess;
    if (!(phase_mask & 4) && p.T > 0) {
        const long long rows = (long long)p.T * p.B;
        const int nrb = (int)((rows + kPrepRowsCta - 1) / kPrepRowsCta);
        static const int prep_dbg = env_int("CTC_B200_PREP_DBG", 0);  // timing experiment: 1 = table blocks only,
        CtcParams pp = p;                                             // 2 = softmax row blocks only
        pp.prep_dbg = prep_dbg;
        ctc_prep_kernel<<<nrb + p.B + (pl.nslices > 1 ? 1 : 0), kPrepRows, pl.smem_prep, st>>>(pp);
        ++*launches;
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    const bool do_engine = !(phase_mask & 1);
    const bool do_grad = !p.score_only && p.T > 0 && !(phase_mask & 2);
    return launch_sliced(
        p, pl, st, do_engine, do_grad, launches, costs_ready, loss_sum,
        [&](const CtcParams &pk, int nb, cudaStream_t s) {
            if (engine_first_form(pk.B))
                ctc_engine_kernel<kK, CP><<<pk.score_only ? nb : 2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(pk);
            else if (pl.cluster)
                ctc_engine2c_kernel<kK, CP><<<2 * (pk.score_only ? nb : 2 * nb), ((pl.nwl + 1) / 2 + 1) * 32,
                                              pl.smem_engine_c, s>>>(pk);
            else if (pl.two_part && !pk.score_only) {
                CtcParams p1 = pk, p2 = pk;
                p1.phase = 1;
                p2.phase = 2;
                ctc_engine2_kernel<kK, CP><<<2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(p1);
                ctc_engine2_kernel<kK, CP><<<2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(p2);
                ++*launches;
            } else
                ctc_engine2_kernel<kK, CP><<<pk.score_only ? nb : 2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(pk);
            ++*launches;
        },
        [&](const CtcParams &pk, int nb, cudaStream_t s, cudaStream_t s_small) {
            CtcParams pf = pk;
            if (pl.nq_small) {
                CtcParams ps = pk;  // utterances with 1 .. nq_small lattice warps
                ps.nqmax = pl.nq_small;
                ps.gLmax = grad_class_lmax(pl.nq_small, pl.Lmax);
                ps.nsplit = pl.nsplit_small;
                ctc_grad_kernel<kK, CP, false><<<dim3(ps.nsplit, nb), pl.nq_small * 32, pl.smem_grad_small, s_small>>>(ps);
                ++*launches;
                pf.gq_lo = pl.nq_small + 1;
            }
            ctc_grad_kernel<kK, CP, false><<<dim3(pf.nsplit, nb), pl.nqmax * 32, pl.smem_grad, s>>>(pf);
            ++*launches;
        });
}

// Large alphabets: ONE row kernel over the whole batch (softmax, dense gradient part, emission gather -- the
// HBM-bound part, every logit read once and every gradient element written once), then the sliced lattice launches
// of the small path; the gradient kernel subtracts the occupancies in place.
// Measured and rejected on B200 (B=64, T=1000, C=5000; 0.593 ms as shipped): a per-slice pipeline (row kernel of
// slice k+1 on its own stream under the lattice kernels of slice k; 2 / 3 / 4 / 8 slices, with and without stream
// priorities, gradient kernels held back until all rows are done or not): 0.586 - 0.66 ms.  The row kernel needs its
// full occupancy (6 CTAs/SM: 0.357 ms, 5: 0.41, 4: 0.46), so either the lattice kernels wait for free slots until the
// row kernel drains, or (with priorities) they take the slots and the rows slow down by more than the recursion hides;
// the gradient kernel's scattered reductions also disturb the streaming.
template <int CP>
cudaError_t launch_big(const CtcParams &p, const Plan &pl, cudaStream_t st, int phase_mask, int *launches,
                       cudaEvent_t costs_ready, const LossOut &loss_sum) {
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::unique_lock<std::mutex> attr_lock(g_attr_mu);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        e = cudaFuncSetAttribute(ctc_prep_utt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_prep_big_kernel<CP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_prep_big_reg_kernel<CP, 16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_prep_big_reg_kernel<CP, 16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_engine_kernel<kK, CP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_engine2_kernel<kK, CP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_engine2c_kernel<kK, CP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ctc_grad_kernel<kK, CP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    attr_lock.unlock();
    const bool do_rows = !(phase_mask & 4) && p.T > 0;
    const bool do_engine = !(phase_mask & 1);
    const bool do_grad = !p.score_only && p.T > 0 && !(phase_mask & 2);
    if (do_rows) {
        ctc_prep_utt_kernel<<<p.B + (pl.nslices > 1 ? 1 : 0), kPrepRows, pl.smem_prep, st>>>(p);
        ++*launches;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    // rows that are 16-byte aligned and fit 5 / 8 / 16 float4 per thread stay in registers
    const bool aligned = (p.C % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.acts) & 15) == 0) &&
                         (!p.grads || (reinterpret_cast<uintptr_t>(p.grads) & 15) == 0);
    auto rows_at = [&](CtcParams pk, int nb, cudaStream_t s) {
        pk.slice_nb = nb;
        const unsigned grid = (unsigned)((long long)p.T * nb);
        const bool ntz = p.nan_to_zero != 0;
        const size_t smem_reg = 256 + sizeof(float) * (size_t)p.C;
#define CTC_ROWS_REG(NV)                                                                              \
        do {                                                                                              \
            if (ntz) ctc_prep_big_reg_kernel<CP, NV, true><<<grid, kBigThreads, smem_reg, s>>>(pk);            \
            else ctc_prep_big_reg_kernel<CP, NV, false><<<grid, kBigThreads, smem_reg, s>>>(pk);               \
        } while (0)
        if (aligned && p.C <= 1024 * 5) CTC_ROWS_REG(5);
        else if (aligned && p.C <= 1024 * 8) CTC_ROWS_REG(8);
        else if (aligned && p.C <= 1024 * 16) CTC_ROWS_REG(16);
        else ctc_prep_big_kernel<CP><<<grid, kBigThreads, pl.smem_rows, s>>>(pk);
#undef CTC_ROWS_REG
        ++*launches;
    };
    auto rows = [&](CtcParams pk, int nb, cudaStream_t s) {
        pk.bord = nullptr;  // whole batch in memory order
        pk.slice0 = 0;
        rows_at(pk, nb, s);
    };
    auto eng = [&](const CtcParams &pk, int nb, cudaStream_t s) {
        if (CP < 512 && engine_first_form(pk.B))  // (the first form's producer keeps the fixed tile look-ahead)
            ctc_engine_kernel<kK, CP><<<pk.score_only ? nb : 2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(pk);
        else if (pl.cluster)
            ctc_engine2c_kernel<kK, CP><<<2 * (pk.score_only ? nb : 2 * nb), ((pl.nwl + 1) / 2 + 1) * 32,
                                          pl.smem_engine_c, s>>>(pk);
        else if (pl.two_part && !pk.score_only) {
            CtcParams p1 = pk, p2 = pk;
            p1.phase = 1;
            p2.phase = 2;
            ctc_engine2_kernel<kK, CP><<<2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(p1);
            ctc_engine2_kernel<kK, CP><<<2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(p2);
            ++*launches;
        } else
            ctc_engine2_kernel<kK, CP><<<pk.score_only ? nb : 2 * nb, (pl.nwl + 1) * 32, pl.smem_engine_req, s>>>(pk);
        ++*launches;
    };
    auto grd = [&](const CtcParams &pk, int nb, cudaStream_t s, cudaStream_t) {
        ctc_grad_kernel<kK, CP, true><<<dim3(pk.nsplit, nb), pl.nqmax * 32, pl.smem_grad, s>>>(pk);
        ++*launches;
    };
    if ((long long)p.T * p.B >= (1ll << 31)) return cudaErrorInvalidValue;
    if (do_rows) rows(p, p.B, st);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    return launch_sliced(p, pl, st, do_engine, do_grad, launches, costs_ready, loss_sum, eng, grd);
}

cudaError_t launch_all(const CtcParams &p, const Plan &pl, cudaStream_t st, int phase_mask, int *launches,
                       cudaEvent_t costs_ready, const LossOut &loss_sum) {
    if (pl.big)
        return pl.Cp == 256 ? launch_big<256>(p, pl, st, phase_mask, launches, costs_ready, loss_sum)
                            : launch_big<512>(p, pl, st, phase_mask, launches, costs_ready, loss_sum);
    switch (pl.Cp) {
        case 32: return launch_all_cp<32>(p, pl, st, phase_mask, launches, costs_ready, loss_sum);
        case 64: return launch_all_cp<64>(p, pl, st, phase_mask, launches, costs_ready, loss_sum);
        default: return launch_all_cp<128>(p, pl, st, phase_mask, launches, costs_ready, loss_sum);
    }
}

}  // namespace

extern "C" {

int ctc_b200_get_version(void) { return 100; }  // 0.1.0

const char *ctc_b200_status_string(ctcB200Status s) {
    switch (s) {
        case CTC_B200_STATUS_SUCCESS: return "no error";
        case CTC_B200_STATUS_MEMOPS_FAILED: return "cuda memcpy or memset failed";
        case CTC_B200_STATUS_INVALID_VALUE: return "invalid value";
        case CTC_B200_STATUS_EXECUTION_FAILED: return "execution failed";
        default: return "unknown error";
    }
}

void ctc_b200_default_options(ctcB200Options *o) {
    if (!o) return;
    memset(o, 0, sizeof(*o));
    o->blank = 0;
    o->zero_infinity = 1;
    o->grad_scale = 1.0f;
}

ctcB200Status ctc_b200_workspace_size(int T, int B, int C, int max_label_len, const ctcB200Options *opts,
                                      size_t *bytes) {
    (void)opts;
    if (!bytes) return CTC_B200_STATUS_INVALID_VALUE;
    Plan pl;
    if (!make_plan(T, B, C, max_label_len, &pl)) return CTC_B200_STATUS_INVALID_VALUE;
    *bytes = pl.total;
    return CTC_B200_STATUS_SUCCESS;
}

int ctc_b200_last_launch_count(void) { return g_last_launches; }

void ctc_b200_last_copy_bytes(unsigned long long *h2d, unsigned long long *d2h) {
    if (h2d) *h2d = g_last_h2d;
    if (d2h) *d2h = g_last_d2h;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------ compute
struct DecodeOut {
    int *path, *ids, *offsets, *out_lens;
};

static ctcB200Status compute_impl(const float *acts, float *grads, const int *labels, const int *label_lens,
                                  const int *act_lens, int T, int B, int C, float *costs, void *workspace,
                                  size_t workspace_bytes, const ctcB200Options *opts_in, bool allow_slices,
                                  const DecodeOut *dec = nullptr) {
    g_last_launches = 0;
    ctcB200Options o;
    if (opts_in) o = *opts_in; else ctc_b200_default_options(&o);
    if (!acts || !costs || !label_lens || !act_lens || !workspace) return CTC_B200_STATUS_INVALID_VALUE;
    if (B < 1 || T < 0 || C < 2 || o.blank < 0 || o.blank >= C) return CTC_B200_STATUS_INVALID_VALUE;
    if (((uintptr_t)workspace & 255) != 0) return CTC_B200_STATUS_INVALID_VALUE;
    cudaStream_t st = (cudaStream_t)o.stream;

    int Lmax = o.max_label_len;
    long long total_labels = 0;
    if (!o.lengths_on_device) {
        // host lengths (what train.py:600 passes): validate like warp-ctc would and derive the launch bound
        int mx = 0;
        for (int b = 0; b < B; ++b) {
            if (label_lens[b] < 0 || act_lens[b] < 0 || act_lens[b] > T) return CTC_B200_STATUS_INVALID_VALUE;
            mx = label_lens[b] > mx ? label_lens[b] : mx;
            total_labels += label_lens[b];
        }
        if (total_labels > 0 && !labels) return CTC_B200_STATUS_INVALID_VALUE;
        for (long long i = 0; i < total_labels; ++i)
            if (labels[i] < 0 || labels[i] >= C || labels[i] == o.blank) return CTC_B200_STATUS_INVALID_VALUE;
        Lmax = mx;
    } else if (Lmax <= 0) {
        return CTC_B200_STATUS_INVALID_VALUE;  // device lengths need an explicit bound
    }
    Plan pl;
    if (!make_plan(T, B, C, Lmax, &pl)) return CTC_B200_STATUS_INVALID_VALUE;
    if (workspace_bytes < pl.total) return CTC_B200_STATUS_INVALID_VALUE;
    if (!allow_slices) pl.nslices = 1;  // the host-buffer path pipelines whole utterance slices itself
    // the length ranking (bord) is written by the prep launch: without it (T == 0, or the prep phase masked out on
    // a call that was not preceded by a full one) the engines must not index through it
    if (T == 0) pl.nslices = 1;

    unsigned char *ws = static_cast<unsigned char *>(workspace);
    CtcParams p;
    memset(&p, 0, sizeof(p));
    p.acts = acts;
    p.grads = grads;
    p.costs = costs;
    if (o.lengths_on_device) {
        p.labels = labels;
        p.label_lens = label_lens;
        p.act_lens = act_lens;
    } else {
        int *dl = reinterpret_cast<int *>(ws + pl.o_labels);
        int *dll = reinterpret_cast<int *>(ws + pl.o_llen);
        int *dal = reinterpret_cast<int *>(ws + pl.o_alen);
        if (total_labels > 0 &&
            cudaMemcpyAsync(dl, labels, sizeof(int) * (size_t)total_labels, cudaMemcpyHostToDevice, st) != cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        if (cudaMemcpyAsync(dll, label_lens, sizeof(int) * B, cudaMemcpyHostToDevice, st) != cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        if (cudaMemcpyAsync(dal, act_lens, sizeof(int) * B, cudaMemcpyHostToDevice, st) != cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        p.labels = dl;
        p.label_lens = dll;
        p.act_lens = dal;
    }
    p.T = T; p.B = B; p.C = C;
    p.blank = o.blank;
    p.zero_infinity = o.zero_infinity;
    p.grad_scale = o.grad_scale;
    p.Cp = pl.Cp;
    p.nwl = pl.nwl;
    p.R = pl.R;
    p.nblk = pl.nblk;
    p.Lmax = pl.Lmax;
    p.score_only = grads ? 0 : 1;
    p.y = reinterpret_cast<unsigned *>(ws + pl.o_y);
    p.ckA = reinterpret_cast<int *>(ws + pl.o_ckA);
    p.ckB = reinterpret_cast<int *>(ws + pl.o_ckB);
    p.ckAE = reinterpret_cast<int *>(ws + pl.o_ckAE);
    p.ckBE = reinterpret_cast<int *>(ws + pl.o_ckBE);
    p.est = reinterpret_cast<double *>(ws + pl.o_est);
    p.Pm2 = reinterpret_cast<double *>(ws + pl.o_Pm2);
    p.Pe2 = reinterpret_cast<int *>(ws + pl.o_Pe2);
    p.Pm = reinterpret_cast<double *>(ws + pl.o_Pm);
    p.Pe = reinterpret_cast<int *>(ws + pl.o_Pe);
    p.feas = reinterpret_cast<int *>(ws + pl.o_feas);
    p.laboff = reinterpret_cast<int *>(ws + pl.o_laboff);
    p.tlab = reinterpret_cast<uint4 *>(ws + pl.o_tlab);
    p.big = pl.big;
    p.bpos = pl.big ? 0 : o.blank;
    p.ctab = reinterpret_cast<int2 *>(ws + pl.o_ctab);
    p.nsplit = pl.nsplit;
    p.nqmax = pl.nqmax;
    p.gq_lo = 1;
    p.gLmax = pl.Lmax;
    if (dec) {
        if (!dec->path || !dec->ids || !dec->out_lens) return CTC_B200_STATUS_INVALID_VALUE;
        p.path = dec->path;
        p.dec_ids = dec->ids;
        p.dec_off = dec->offsets;
        p.dec_len = dec->out_lens;
    }
    p.bord = pl.nslices > 1 ? reinterpret_cast<int *>(ws + pl.o_bord) : nullptr;
    p.slice0 = 0;
    p.slice_nb = B;
    if (o.emissions_ready && (pl.big || o.nan_to_zero || dec)) return CTC_B200_STATUS_INVALID_VALUE;
    p.y_ready = o.emissions_ready ? 1 : 0;
    p.nan_to_zero = o.nan_to_zero ? 1 : 0;
    p.nan_count = o.nan_to_zero ? o.nan_count : nullptr;

    int launches = 0;
    LossOut lo;
    lo.loss_sum = o.loss_sum;
    lo.loss_all = o.loss_sum_all;
    if (o.comm) {
        const ctcB200Comm *cm = static_cast<const ctcB200Comm *>(o.comm);
        if (!o.loss_sum_all || !cm->ready) return CTC_B200_STATUS_INVALID_VALUE;
        // a peer missed an earlier exchange (it skipped a call, or stalled past the timeout): the ranks are out of
        // step and every further exchange would be wrong -- refuse instead of running the protocol on
        if (cm->timed_out_host && *reinterpret_cast<volatile int *>(cm->timed_out_host))
            return CTC_B200_STATUS_EXECUTION_FAILED;
        lo.comm = CommArgs{cm->local, cm->peers_dev, cm->rank, cm->world, cm->timeout_ns, cm->timed_out_dev};
    }
    cudaError_t e = launch_all(p, pl, st, o.phase_mask, &launches, (cudaEvent_t)o.costs_ready, lo);
    g_last_launches = launches;
    return e == cudaSuccess ? CTC_B200_STATUS_SUCCESS : CTC_B200_STATUS_EXECUTION_FAILED;
}

extern "C" ctcB200Status ctc_b200_compute(const float *acts, float *grads, const int *labels, const int *label_lens,
                                          const int *act_lens, int T, int B, int C, float *costs, void *workspace,
                                          size_t workspace_bytes, const ctcB200Options *opts_in) {
    return compute_impl(acts, grads, labels, label_lens, act_lens, T, B, C, costs, workspace, workspace_bytes, opts_in,
                        true);
}

extern "C" ctcB200Status ctc_b200_compute_decode(const float *acts, float *grads, const int *labels,
                                                 const int *label_lens, const int *act_lens, int T, int B, int C,
                                                 float *costs, void *workspace, size_t workspace_bytes, int *path,
                                                 int *ids, int *offsets, int *out_lens, const ctcB200Options *opts_in) {
    const DecodeOut dec = {path, ids, offsets, out_lens};
    return compute_impl(acts, grads, labels, label_lens, act_lens, T, B, C, costs, workspace, workspace_bytes, opts_in,
                        true, &dec);
}

extern "C" ctcB200Status ctc_b200_greedy_decode(const float *acts, long long stride_t, long long stride_b,
                                                const int *act_lens_dev, int T, int B, int C, int *path, int *ids,
                                                int *offsets, int *out_lens, const ctcB200Options *opts_in) {
    g_last_launches = 0;
    ctcB200Options o;
    if (opts_in) o = *opts_in; else ctc_b200_default_options(&o);
    if (!acts || !path || !ids || !out_lens) return CTC_B200_STATUS_INVALID_VALUE;
    if (B < 1 || T < 0 || C < 2 || o.blank < 0 || o.blank >= C) return CTC_B200_STATUS_INVALID_VALUE;
    cudaStream_t st = (cudaStream_t)o.stream;
    const long long rows = (long long)T * B;
    if (rows > 0) {
        ctc_argmax_kernel<<<(unsigned)((rows + 15) / 16), 128, 0, st>>>(acts, stride_t, stride_b, act_lens_dev, T, B, C,
                                                                        path);
        ++g_last_launches;
    }
    ctc_collapse_kernel<<<B, kCollapseThreads, 0, st>>>(path, act_lens_dev, T, B, o.blank, ids, offsets, out_lens);
    ++g_last_launches;
    return cudaGetLastError() == cudaSuccess ? CTC_B200_STATUS_SUCCESS : CTC_B200_STATUS_EXECUTION_FAILED;
}

// ------------------------------------------------------------------------------------------------ fc layer (f3)
namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// the driver's tensor-map encoder, fetched through the runtime (the library does not link libcuda)
EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return reinterpret_cast<EncodeTiledFn>(f);
    }();
    return fn;
}
// 2-D fp32 tensor [rows][cols] (row pitch `pitch_bytes`, a multiple of 16), box [box_rows][32 columns], 128-byte
// swizzle, out-of-range elements read as zero
bool make_map(CUtensorMap *m, const void *base, unsigned long long rows, unsigned long long cols,
              unsigned long long pitch_bytes, unsigned box_rows) {
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return false;
    const cuuint64_t gdim[2] = {cols, rows}, gstr[1] = {pitch_bytes};
    const cuuint32_t box[2] = {(cuuint32_t)kFcBK, box_rows}, estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(base), gdim, gstr, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
struct FcPlan {
    int N;                       // classes padded to 32 / 64
    size_t o_w2, o_bias, o_mean, o_var, o_part, total;
    int nstat;                   // row blocks of the batch-statistics pass
};
bool fc_plan(int T, int B, int H, int C, FcPlan *pl) {
    if (T < 1 || B < 1 || H < 4 || (H & 3) || C < 2 || C > 64) return false;
    if ((long long)T * B > (1ll << 31) - kFcBM) return false;
    pl->N = C <= 32 ? 32 : 64;
    const long long rows = (long long)T * B;
    pl->nstat = (int)((rows + kFcStatRows - 1) / kFcStatRows);
    size_t o = 0;
    pl->o_w2 = o;   o = align256(o + sizeof(float) * 2 * pl->N * (size_t)H);
    pl->o_bias = o; o = align256(o + sizeof(float) * pl->N);
    pl->o_mean = o; o = align256(o + sizeof(float) * H);
    pl->o_var = o;  o = align256(o + sizeof(float) * H);
    pl->o_part = o; o = align256(o + sizeof(double) * 2 * (size_t)pl->nstat * H);
    pl->total = o;
    return true;
}
template <int N>
cudaError_t fc_launch(const CUtensorMap &mx, const CUtensorMap &mx32, const CUtensorMap &mw, const FcParams &fp, int grid,
                      cudaStream_t st) {
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    {
        std::lock_guard<std::mutex> lock(g_attr_mu);
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            cudaError_t e = cudaFuncSetAttribute(fc_logits_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)FcCfg<N>::smem);
            if (e != cudaSuccess) return e;
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
    }
    fc_logits_kernel<N><<<grid, kFcThreads, FcCfg<N>::smem, st>>>(mx, mx32, mw, fp);
    return cudaGetLastError();
}
}  // namespace

extern "C" ctcB200Status ctc_b200_fc_workspace_size(int T, int B, int H, int C, size_t *bytes) {
    FcPlan pl;
    if (!bytes || !fc_plan(T, B, H, C, &pl)) return CTC_B200_STATUS_INVALID_VALUE;
    *bytes = pl.total;
    return CTC_B200_STATUS_SUCCESS;
}

extern "C" ctcB200Status ctc_b200_fc_forward(const float *x, int T, int B, int H, int C, const ctcB200FcLayer *layer,
                                             float *logits, void *ctc_workspace, const int *act_lens_dev, int *path,
                                             void *fc_workspace, size_t fc_workspace_bytes,
                                             const ctcB200Options *opts_in) {
    g_last_launches = 0;
    FcPlan pl;
    if (!x || !layer || !layer->weight || !fc_workspace || !fc_plan(T, B, H, C, &pl))
        return CTC_B200_STATUS_INVALID_VALUE;
    if (fc_workspace_bytes < pl.total || ((uintptr_t)fc_workspace & 255) || ((uintptr_t)x & 15))
        return CTC_B200_STATUS_INVALID_VALUE;
    if (!logits && !ctc_workspace && !path) return CTC_B200_STATUS_INVALID_VALUE;
    if (ctc_workspace && ((uintptr_t)ctc_workspace & 255)) return CTC_B200_STATUS_INVALID_VALUE;
    cudaStream_t st = opts_in ? (cudaStream_t)opts_in->stream : nullptr;
    unsigned char *ws = static_cast<unsigned char *>(fc_workspace);
    float *W2 = reinterpret_cast<float *>(ws + pl.o_w2), *bias = reinterpret_cast<float *>(ws + pl.o_bias);
    const long long rows = (long long)T * B;
    const float *mean = layer->bn_mean, *var = layer->bn_var;
    if (layer->bn_batch_stats) {
        // training-mode forward of nn.BatchNorm1d: normalise with the statistics of THIS batch (biased variance)
        float *m = reinterpret_cast<float *>(ws + pl.o_mean), *v = reinterpret_cast<float *>(ws + pl.o_var);
        double *part = reinterpret_cast<double *>(ws + pl.o_part);
        fc_bn_partial_kernel<<<pl.nstat, 256, 0, st>>>(x, rows, H, part);
        fc_bn_combine_kernel<<<(H + 255) / 256, 256, 0, st>>>(part, pl.nstat, rows, H, m, v);
        g_last_launches += 2;
        if (layer->batch_mean_out &&
            cudaMemcpyAsync(layer->batch_mean_out, m, sizeof(float) * H, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        if (layer->batch_var_out &&
            cudaMemcpyAsync(layer->batch_var_out, v, sizeof(float) * H, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        mean = m;
        var = v;
    }
    if (!(layer->reuse_folded && !layer->bn_batch_stats)) {
        fc_fold_kernel<<<pl.N, 256, 0, st>>>(layer->weight, layer->bn_weight, layer->bn_bias, mean, var, layer->bn_eps,
                                             H, C, pl.N, W2, bias);
        ++g_last_launches;
    }
    if (cudaGetLastError() != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
    CUtensorMap mx, mx32, mw;
    if (!make_map(&mx, x, (unsigned long long)rows, (unsigned long long)H, (unsigned long long)H * 4, kFcBM) ||
        !make_map(&mx32, x, (unsigned long long)rows, (unsigned long long)H, (unsigned long long)H * 4, 32) ||
        !make_map(&mw, W2, (unsigned long long)(2 * pl.N), (unsigned long long)H, (unsigned long long)H * 4,
                  (unsigned)(2 * pl.N)))
        return CTC_B200_STATUS_EXECUTION_FAILED;
    FcParams fp;
    fp.x = x;
    fp.trace = nullptr;
    static const bool fc_trace = env_int("CTC_FC_TRACE", 0) != 0;  // diagnostic: synchronises and prints
    static long long *trace_dev = nullptr;
    if (fc_trace) {
        if (!trace_dev) cudaMalloc((void **)&trace_dev, sizeof(long long) * (kFcTraceK * 8 + 2 * 256));
        cudaMemsetAsync(trace_dev, 0, sizeof(long long) * (kFcTraceK * 8 + 2 * 256), st);
        fp.trace = trace_dev;
    }
    fp.rows = rows; fp.T = T; fp.B = B; fp.H = H; fp.C = C;
    fp.bias = bias;
    fp.logits = logits;
    fp.y = static_cast<unsigned *>(ctc_workspace);  // the emission rows are the first array of the CTC workspace
    fp.act_lens = act_lens_dev;
    fp.path = path;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // equal row ranges per SM (multiples of 8 rows), at least one 32-row box each
    long long per = (rows + sms - 1) / sms;
    per = (per + 7) & ~7ll;
    if (per < 32) per = 32;
    const int grid = (int)((rows + per - 1) / per);
    fp.rows_per_cta = (int)per;
    // CTC_FC_TS=1 (C <= 32): the TS form -- A operands written to TMEM by the converters (tcgen05.st) and read from
    // there by the tensor core, half the shared-memory traffic.  Measured equal to the SS form (0.083 against 0.082 ms)
    // with one accumulator per tile half instead of two (error 1.3e-5 against 7e-6), so the SS form is the default.
    static const bool use_ts = env_int("CTC_FC_TS", 0) != 0;
    cudaError_t e;
    if (pl.N == 32 && use_ts) {
        static bool ts_attr[64] = {false};
        {
            std::lock_guard<std::mutex> lock(g_attr_mu);
            if (dev < 0 || dev >= 64 || !ts_attr[dev]) {
                if (cudaFuncSetAttribute(fc_logits_ts_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)FcTsCfg<32>::smem) != cudaSuccess)
                    return CTC_B200_STATUS_EXECUTION_FAILED;
                if (dev >= 0 && dev < 64) ts_attr[dev] = true;
            }
        }
        fc_logits_ts_kernel<32><<<grid, kFcThreads, FcTsCfg<32>::smem, st>>>(mx, mx32, mw, fp);
        e = cudaGetLastError();
    } else {
        e = pl.N == 32 ? fc_launch<32>(mx, mx32, mw, fp, grid, st) : fc_launch<64>(mx, mx32, mw, fp, grid, st);
    }
    ++g_last_launches;
    if (fc_trace && e == cudaSuccess) {
        static long long h[kFcTraceK * 8 + 2 * 256];
        cudaStreamSynchronize(st);
        cudaMemcpy(h, trace_dev, sizeof(h), cudaMemcpyDeviceToHost);
        const long long t0 = h[0];
        fprintf(stderr, "[fc trace] k-block: producer-issue | conv: lo-free tile-landed done | mma: tile-landed lo-ready committed  (clk after first issue)\n");
        {
            long long b0 = h[kFcTraceK * 8], e0 = 0, emin = 1ll << 62;
            for (int c = 0; c < grid && c < 256; ++c) {
                b0 = h[kFcTraceK * 8 + 2 * c] < b0 ? h[kFcTraceK * 8 + 2 * c] : b0;
                e0 = h[kFcTraceK * 8 + 2 * c + 1] > e0 ? h[kFcTraceK * 8 + 2 * c + 1] : e0;
                emin = h[kFcTraceK * 8 + 2 * c + 1] < emin ? h[kFcTraceK * 8 + 2 * c + 1] : emin;
            }
            fprintf(stderr, "[fc trace] CTAs: first start 0, first end %.1f us, last end %.1f us; per CTA (start,end us):",
                    (emin - b0) * 1e-3, (e0 - b0) * 1e-3);
            for (int c = 0; c < grid && c < 256; c += 9)
                fprintf(stderr, " %d:(%.1f,%.1f)", c, (h[kFcTraceK * 8 + 2 * c] - b0) * 1e-3,
                        (h[kFcTraceK * 8 + 2 * c + 1] - b0) * 1e-3);
            fprintf(stderr, "\n");
        }
        for (int g = 0; g < kFcTraceK && h[g * 8]; ++g)
            fprintf(stderr, "[fc trace] %3d: %7lld | %7lld %7lld %7lld | %7lld %7lld %7lld\n", g, h[g * 8] - t0,
                    h[g * 8 + 1] - t0, h[g * 8 + 2] - t0, h[g * 8 + 3] - t0, h[g * 8 + 4] - t0, h[g * 8 + 5] - t0,
                    h[g * 8 + 6] - t0);
    }
    return e == cudaSuccess ? CTC_B200_STATUS_SUCCESS : CTC_B200_STATUS_EXECUTION_FAILED;
}

// ------------------------------------------------------------------------------------------------ no-allocation entry
// Scratch of ctc_b200_compute_static: a file-scope device array (every GPU that loads the library holds these
// CTC_B200_STATIC_WORKSPACE_BYTES until the process exits).  In HOST code the name `g_static_ws` is only a stand-in
// for the symbol, never a device address: the address is taken once per device with cudaGetSymbolAddress (that
// device current) and cached.
static __device__ __align__(256) unsigned char g_static_ws[CTC_B200_STATIC_WORKSPACE_BYTES];

namespace {
std::mutex g_static_mu;
void *g_static_ptr[64] = {};

// device address of g_static_ws on the CURRENT device (nullptr on failure)
void *static_workspace() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void *ptr = nullptr;
        if (cudaGetSymbolAddress(&ptr, g_static_ws) != cudaSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        g_static_ptr[dev] = ptr;
    }
    return g_static_ptr[dev];
}
}  // namespace

extern "C" ctcB200Status ctc_b200_compute_static(const float *acts, float *grads, const int *labels,
                                                 const int *label_lens, const int *act_lens, int T, int B, int C,
                                                 float *costs, const ctcB200Options *opts_in) {
    g_last_launches = 0;
    ctcB200Options o;
    if (opts_in) o = *opts_in; else ctc_b200_default_options(&o);
    if (!label_lens || B < 1) return CTC_B200_STATUS_INVALID_VALUE;
    // the shape's scratch must fit the static array: same bound the workspace query reports
    int Lmax = o.max_label_len;
    if (!o.lengths_on_device) {
        Lmax = 0;
        for (int b = 0; b < B; ++b) Lmax = label_lens[b] > Lmax ? label_lens[b] : Lmax;
    }
    size_t need = 0;
    if (ctc_b200_workspace_size(T, B, C, Lmax, &o, &need) != CTC_B200_STATUS_SUCCESS)
        return CTC_B200_STATUS_INVALID_VALUE;
    if (need > (size_t)CTC_B200_STATIC_WORKSPACE_BYTES) {
        fprintf(stderr,
                "ctc_b200_compute_static: T=%d B=%d C=%d max_label_len=%d needs %zu bytes of scratch, the static "
                "array holds %zu (CTC_B200_STATIC_WORKSPACE_BYTES); use ctc_b200_compute with a caller-owned workspace\n",
                T, B, C, Lmax, need, (size_t)CTC_B200_STATIC_WORKSPACE_BYTES);
        return CTC_B200_STATUS_INVALID_VALUE;
    }
    void *ws = static_workspace();
    if (!ws) return CTC_B200_STATUS_EXECUTION_FAILED;
    // unsliced: every kernel goes to opts->stream itself, so not even the side-stream pool is created
    return compute_impl(acts, grads, labels, label_lens, act_lens, T, B, C, costs, ws,
                        (size_t)CTC_B200_STATIC_WORKSPACE_BYTES, &o, false);
}

// ------------------------------------------------------------------------------------------------ loss all-reduce
constexpr size_t kCommBytes = sizeof(unsigned long long) * 2 * kCommMaxWorld + 64;

extern "C" ctcB200Status ctc_b200_comm_create(ctcB200Comm **out, int rank, int world) {
    if (!out || world < 1 || world > kCommMaxWorld || rank < 0 || rank >= world) return CTC_B200_STATUS_INVALID_VALUE;
    ctcB200Comm *c = new (std::nothrow) ctcB200Comm();
    if (!c) return CTC_B200_STATUS_MEMOPS_FAILED;
    c->rank = rank;
    c->world = world;
    if (cudaMalloc((void **)&c->local, kCommBytes) != cudaSuccess ||
        cudaMemset(c->local, 0, kCommBytes) != cudaSuccess ||
        cudaMalloc((void **)&c->peers_dev, sizeof(void *) * kCommMaxWorld) != cudaSuccess ||
        cudaHostAlloc((void **)&c->timed_out_host, sizeof(int), cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer((void **)&c->timed_out_dev, c->timed_out_host, 0) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess) {
        cudaFree(c->local);
        cudaFree(c->peers_dev);
        if (c->timed_out_host) cudaFreeHost(c->timed_out_host);
        delete c;
        return CTC_B200_STATUS_MEMOPS_FAILED;
    }
    *c->timed_out_host = 0;
    if (const char *v = getenv("CTC_B200_COMM_TIMEOUT_MS"))
        if (*v) c->timeout_ns = (unsigned long long)atoll(v) * 1000000ull;
    *out = c;
    return CTC_B200_STATUS_SUCCESS;
}

extern "C" ctcB200Status ctc_b200_comm_set_timeout(ctcB200Comm *c, long long milliseconds) {
    if (!c || milliseconds <= 0) return CTC_B200_STATUS_INVALID_VALUE;
    c->timeout_ns = (unsigned long long)milliseconds * 1000000ull;
    return CTC_B200_STATUS_SUCCESS;
}

// 1 once an exchange of this communicator has timed out (sticky; the communicator is then unusable: destroy it and
// build a new one on every rank, or continue with the caller's own collective)
extern "C" int ctc_b200_comm_timed_out(const ctcB200Comm *c) {
    return (c && c->timed_out_host) ? *reinterpret_cast<volatile int *>(c->timed_out_host) : 0;
}

extern "C" ctcB200Status ctc_b200_comm_handle(ctcB200Comm *c, void *handle64) {
    if (!c || !handle64) return CTC_B200_STATUS_INVALID_VALUE;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the ABI hands the IPC handle around as 64 opaque bytes");
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, c->local) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
    memcpy(handle64, &h, 64);
    return CTC_B200_STATUS_SUCCESS;
}

extern "C" void *ctc_b200_comm_local(ctcB200Comm *c) { return c ? (void *)c->local : nullptr; }

static ctcB200Status comm_finish(ctcB200Comm *c, unsigned long long **ptrs) {
    ptrs[c->rank] = c->local;
    if (cudaMemcpy(c->peers_dev, ptrs, sizeof(void *) * c->world, cudaMemcpyHostToDevice) != cudaSuccess)
        return CTC_B200_STATUS_MEMOPS_FAILED;
    c->ready = true;
    return CTC_B200_STATUS_SUCCESS;
}

// handles: world x 64 bytes, the ctc_b200_comm_handle of every rank in rank order (this rank's own entry is ignored)
extern "C" ctcB200Status ctc_b200_comm_open(ctcB200Comm *c, const void *handles) {
    if (!c || !handles || c->ready) return CTC_B200_STATUS_INVALID_VALUE;
    unsigned long long *ptrs[kCommMaxWorld] = {};
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char *>(handles) + 64 * (size_t)r, 64);
        if (cudaIpcOpenMemHandle(&c->opened[r], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            return CTC_B200_STATUS_EXECUTION_FAILED;
        }
        ptrs[r] = static_cast<unsigned long long *>(c->opened[r]);
    }
    return comm_finish(c, ptrs);
}

// all ranks in ONE process (one device, or devices with peer access enabled by the caller): raw slot-buffer pointers
// (ctc_b200_comm_local of every rank, rank order)
extern "C" ctcB200Status ctc_b200_comm_set_peers(ctcB200Comm *c, void *const *locals) {
    if (!c || !locals || c->ready) return CTC_B200_STATUS_INVALID_VALUE;
    unsigned long long *ptrs[kCommMaxWorld] = {};
    for (int r = 0; r < c->world; ++r) ptrs[r] = static_cast<unsigned long long *>(locals[r]);
    return comm_finish(c, ptrs);
}

extern "C" ctcB200Status ctc_b200_comm_destroy(ctcB200Comm *c) {
    if (!c) return CTC_B200_STATUS_SUCCESS;
    for (int r = 0; r < kCommMaxWorld; ++r)
        if (c->opened[r]) cudaIpcCloseMemHandle(c->opened[r]);
    cudaFree(c->local);
    cudaFree(c->peers_dev);
    if (c->timed_out_host) cudaFreeHost(c->timed_out_host);
    delete c;
    return CTC_B200_STATUS_SUCCESS;
}

// ------------------------------------------------------------------------------------------------ backward scaling
extern "C" ctcB200Status ctc_b200_scale_grads(float *grads, size_t count, const float *scale_dev,
                                              const ctcB200Options *opts_in) {
    g_last_launches = 0;
    if (!grads || !scale_dev) return CTC_B200_STATUS_INVALID_VALUE;
    if (count == 0) return CTC_B200_STATUS_SUCCESS;
    cudaStream_t st = opts_in ? (cudaStream_t)opts_in->stream : nullptr;
    // enough CTAs to saturate HBM when the pass is needed (8 resident CTAs of 256 threads on each of the 148 SMs),
    // never more than one float4 per thread would need
    size_t want = (count / 4 + 255) / 256;
    if (want < 1) want = 1;
    const unsigned grid = (unsigned)(want < 148 * 8 ? want : 148 * 8);
    ctc_scale_grads_kernel<<<grid, 256, 0, st>>>(grads, count, scale_dev);
    ++g_last_launches;
    return cudaGetLastError() == cudaSuccess ? CTC_B200_STATUS_SUCCESS : CTC_B200_STATUS_EXECUTION_FAILED;
}

// ------------------------------------------------------------------------------------------------ WER / CER counts
extern "C" ctcB200Status ctc_b200_error_counts(const int *hyp_ids, int hyp_pitch, const int *hyp_lens,
                                               const int *ref_labels, const int *ref_lens, int B, int max_hyp_len,
                                               int max_ref_len, int space_index, int *counts,
                                               const ctcB200Options *opts_in) {
    g_last_launches = 0;
    ctcB200Options o;
    if (opts_in) o = *opts_in; else ctc_b200_default_options(&o);
    if (!hyp_ids || !hyp_lens || !ref_lens || !counts) return CTC_B200_STATUS_INVALID_VALUE;
    if (B < 1 || max_hyp_len < 0 || max_ref_len < 0 || hyp_pitch < max_hyp_len) return CTC_B200_STATUS_INVALID_VALUE;
    if (max_ref_len > 0 && !ref_labels) return CTC_B200_STATUS_INVALID_VALUE;
    const size_t smem = metrics_smem(max_hyp_len, max_ref_len);
    if (smem > kSmemLimit) return CTC_B200_STATUS_INVALID_VALUE;  // sequences too long for one CTA's shared memory
    static bool attr_set[64] = {false};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
    std::unique_lock<std::mutex> attr_lock(g_attr_mu);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        if (cudaFuncSetAttribute(ctc_error_counts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)kSmemLimit) != cudaSuccess)
            return CTC_B200_STATUS_EXECUTION_FAILED;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    attr_lock.unlock();
    ctc_error_counts_kernel<<<B, kMetricThreads, smem, (cudaStream_t)o.stream>>>(
        hyp_ids, hyp_pitch, hyp_lens, ref_labels, ref_lens, B, max_hyp_len, max_ref_len, space_index,
        reinterpret_cast<int4 *>(counts));
    ++g_last_launches;
    return cudaGetLastError() == cudaSuccess ? CTC_B200_STATUS_SUCCESS : CTC_B200_STATUS_EXECUTION_FAILED;
}

// ------------------------------------------------------------------------------------------------ host-buffer path
// The batch is cut into up to kHostChunks slices of utterances, each with its own stream, compact device buffers and
// workspace: the host->device copy of slice k+1, the kernels of slice k and the device->host copy of slice k-1
// overlap (the kernels are bound by the T-step recursion, not by the SM count, so slices also run side by side).
constexpr int kHostChunks = 4;
constexpr int kHostMinChunk = 8;  // utterances per slice below which slicing stops paying

struct ctcB200Context {
    float *d_acts = nullptr, *d_grads = nullptr, *d_costs = nullptr;
    void *d_ws = nullptr;
    int *d_meta = nullptr, *h_meta = nullptr;  // [labels | label_lens | act_lens]: device copy, pinned staging
    size_t n_acts = 0, n_grads = 0, n_costs = 0, n_ws = 0, n_meta = 0, n_hmeta = 0;
    cudaStream_t side[kHostChunks] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t done[kHostChunks] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t uploaded[kHostChunks] = {nullptr, nullptr, nullptr, nullptr};
    cudaStream_t up = nullptr;  // all uploads, one after the other: the first slice gets the whole link
    cudaEvent_t start = nullptr;
    bool streams_ok = false;
};

static bool grow(void **ptr, size_t *have, size_t need) {
    if (need <= *have) return true;
    if (*ptr) cudaFree(*ptr);
    *ptr = nullptr;
    *have = 0;
    if (cudaMalloc(ptr, need) != cudaSuccess) return false;
    *have = need;
    return true;
}

static bool make_streams(ctcB200Context *c) {
    if (c->streams_ok) return true;
    if (cudaEventCreateWithFlags(&c->start, cudaEventDisableTiming) != cudaSuccess) return false;
    if (cudaStreamCreateWithFlags(&c->up, cudaStreamNonBlocking) != cudaSuccess) return false;
    for (int k = 0; k < kHostChunks; ++k) {
        if (cudaStreamCreateWithFlags(&c->side[k], cudaStreamNonBlocking) != cudaSuccess) return false;
        if (cudaEventCreateWithFlags(&c->done[k], cudaEventDisableTiming) != cudaSuccess) return false;
        if (cudaEventCreateWithFlags(&c->uploaded[k], cudaEventDisableTiming) != cudaSuccess) return false;
    }
    c->streams_ok = true;
    return true;
}

extern "C" ctcB200Status ctc_b200_create(ctcB200Context **ctx) {
    if (!ctx) return CTC_B200_STATUS_INVALID_VALUE;
    *ctx = new (std::nothrow) ctcB200Context();
    return *ctx ? CTC_B200_STATUS_SUCCESS : CTC_B200_STATUS_MEMOPS_FAILED;
}

extern "C" ctcB200Status ctc_b200_destroy(ctcB200Context *c) {
    if (!c) return CTC_B200_STATUS_SUCCESS;
    cudaFree(c->d_acts);
    cudaFree(c->d_grads);
    cudaFree(c->d_costs);
    cudaFree(c->d_ws);
    cudaFree(c->d_meta);
    if (c->h_meta) cudaFreeHost(c->h_meta);
    for (int k = 0; k < kHostChunks; ++k) {
        if (c->side[k]) cudaStreamDestroy(c->side[k]);
        if (c->done[k]) cudaEventDestroy(c->done[k]);
        if (c->uploaded[k]) cudaEventDestroy(c->uploaded[k]);
    }
    if (c->up) cudaStreamDestroy(c->up);
    if (c->start) cudaEventDestroy(c->start);
    delete c;
    return CTC_B200_STATUS_SUCCESS;
}

extern "C" ctcB200Status ctc_b200_compute_host(ctcB200Context *c, const float *acts_host, float *grads_host,
                                               const int *labels, const int *label_lens, const int *act_lens, int T,
                                               int B, int C, float *costs_host, double *loss_sum,
                                               const ctcB200Options *opts_in) {
    if (!c || !acts_host || !costs_host || !label_lens || !act_lens) return CTC_B200_STATUS_INVALID_VALUE;
    if (B < 1 || T < 0 || C < 2) return CTC_B200_STATUS_INVALID_VALUE;
    ctcB200Options o;
    if (opts_in) o = *opts_in; else ctc_b200_default_options(&o);
    if (o.lengths_on_device) return CTC_B200_STATUS_INVALID_VALUE;
    if (o.blank < 0 || o.blank >= C) return CTC_B200_STATUS_INVALID_VALUE;
    // validate once for the whole batch (what ctc_b200_compute does for host lengths), then hand the kernels DEVICE
    // copies of labels and lengths: one pinned staging copy instead of three pageable ones per slice
    size_t total_labels = 0;
    for (int b = 0; b < B; ++b) {
        if (label_lens[b] < 0 || act_lens[b] < 0 || act_lens[b] > T) return CTC_B200_STATUS_INVALID_VALUE;
        total_labels += (size_t)label_lens[b];
    }
    if (total_labels > 0 && !labels) return CTC_B200_STATUS_INVALID_VALUE;
    for (size_t i = 0; i < total_labels; ++i)
        if (labels[i] < 0 || labels[i] >= C || labels[i] == o.blank) return CTC_B200_STATUS_INVALID_VALUE;

    // ---- slices of utterances [b0, b0 + nb)
    // Slice widths are kept at multiples of 64 bytes where the batch allows it: the strided 2-D copies run at the full
    // 50 GB/s of the link for 64-byte-multiple rows and at about half of it otherwise (measured: 16 x 29 floats =
    // 1856 B rows 60 us per slice, 19 x 29 floats = 2204 B rows 116 us).  g = utterances per 64-byte granule.
    int g = 16;
    for (int c2 = C; (c2 & 1) == 0 && g > 1; c2 >>= 1) g >>= 1;
    int b0s[kHostChunks + 1], nbs[kHostChunks];
    int nch;
    if (B / g >= 2 && g >= kHostMinChunk / 2) {
        const int units = B / g;
        nch = units > kHostChunks ? kHostChunks : units;
        for (int k = 0; k < nch; ++k) nbs[k] = g * (units / nch + (k < units % nch ? 1 : 0));
        nbs[nch - 1] += B % g;  // ragged remainder rides with the last slice
    } else {
        nch = B / kHostMinChunk;
        nch = nch < 1 ? 1 : (nch > kHostChunks ? kHostChunks : nch);
        for (int k = 0; k < nch; ++k) nbs[k] = B / nch + (k < B % nch ? 1 : 0);
        if (g > 1 && g < kHostMinChunk)  // small granule (C a multiple of 4 or 8): round the cuts to it
            for (int k = 0; k + 1 < nch; ++k) {
                const int r = nbs[k] % g;
                nbs[k] -= r;
                nbs[k + 1] += r;
            }
    }
    size_t ws_off[kHostChunks + 1], lab_off[kHostChunks + 1];
    int mxs[kHostChunks], Ts[kHostChunks];
    b0s[0] = 0; ws_off[0] = 0; lab_off[0] = 0;
    for (int k = 0; k < nch; ++k) {
        const int nb = nbs[k];
        b0s[k + 1] = b0s[k] + nb;
        int mx = 0, mt = 0;
        size_t nl = 0;
        for (int b = b0s[k]; b < b0s[k + 1]; ++b) {
            mx = label_lens[b] > mx ? label_lens[b] : mx;
            mt = act_lens[b] > mt ? act_lens[b] : mt;
            nl += (size_t)label_lens[b];
        }
        lab_off[k + 1] = lab_off[k] + nl;
        mxs[k] = mx;
        // Frames past the slice's longest utterance are padding: they are neither uploaded nor computed nor downloaded
        // (their gradient rows are zero-filled on the host while the GPU works).  The reference's collate function sorts
        // every batch by length (data_loader_aug.py:527), so on its batches the later slices are much shorter than T.
        Ts[k] = mt;
        size_t wsb = 0;
        ctcB200Status s = ctc_b200_workspace_size(mt, nb, C, mx > 0 ? mx : 1, &o, &wsb);
        if (s != CTC_B200_STATUS_SUCCESS) return s;
        ws_off[k + 1] = ws_off[k] + align256(wsb);
    }
    const size_t nel = (size_t)T * B * C;
    {
        void *pa = c->d_acts, *pg = c->d_grads, *pc = c->d_costs;
        if (!grow(&pa, &c->n_acts, nel * sizeof(float))) { c->d_acts = nullptr; return CTC_B200_STATUS_MEMOPS_FAILED; }
        c->d_acts = (float *)pa;
        if (grads_host) {
            if (!grow(&pg, &c->n_grads, nel * sizeof(float))) { c->d_grads = nullptr; return CTC_B200_STATUS_MEMOPS_FAILED; }
            c->d_grads = (float *)pg;
        }
        if (!grow(&pc, &c->n_costs, sizeof(float) * B)) { c->d_costs = nullptr; return CTC_B200_STATUS_MEMOPS_FAILED; }
        c->d_costs = (float *)pc;
        if (!grow(&c->d_ws, &c->n_ws, ws_off[nch])) return CTC_B200_STATUS_MEMOPS_FAILED;
        void *pm = c->d_meta;
        const size_t meta_bytes = sizeof(int) * (total_labels + 2 * (size_t)B);
        if (!grow(&pm, &c->n_meta, meta_bytes)) { c->d_meta = nullptr; return CTC_B200_STATUS_MEMOPS_FAILED; }
        c->d_meta = (int *)pm;
        if (meta_bytes > c->n_hmeta) {
            if (c->h_meta) cudaFreeHost(c->h_meta);
            c->h_meta = nullptr;
            c->n_hmeta = 0;
            if (cudaHostAlloc((void **)&c->h_meta, meta_bytes, cudaHostAllocDefault) != cudaSuccess)
                return CTC_B200_STATUS_MEMOPS_FAILED;
            c->n_hmeta = meta_bytes;
        }
    }
    cudaStream_t user = (cudaStream_t)o.stream;
    // the previous call synchronised `user` before returning, so the staging buffer is free to overwrite
    if (total_labels) memcpy(c->h_meta, labels, sizeof(int) * total_labels);
    memcpy(c->h_meta + total_labels, label_lens, sizeof(int) * (size_t)B);
    memcpy(c->h_meta + total_labels + B, act_lens, sizeof(int) * (size_t)B);
    if (cudaMemcpyAsync(c->d_meta, c->h_meta, sizeof(int) * (total_labels + 2 * (size_t)B), cudaMemcpyHostToDevice,
                        user) != cudaSuccess)
        return CTC_B200_STATUS_MEMOPS_FAILED;
    const int *d_labels = c->d_meta, *d_llen = c->d_meta + total_labels, *d_alen = c->d_meta + total_labels + B;
    if (nch > 1) {
        if (!make_streams(c)) return CTC_B200_STATUS_MEMOPS_FAILED;
        if (cudaEventRecord(c->start, user) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
    }
    int launches = 0;
    unsigned long long h2d = sizeof(int) * (total_labels + 2 * (size_t)B), d2h = 0;
    const size_t host_pitch = (size_t)B * C * sizeof(float);
    // CTC_B200_HOST_TRACE=1: print the device timeline of this call (per slice: upload done, kernels done, download
    // done, in us after the call's first enqueue) and the host-side enqueue / wait times.  A diagnostic: the events
    // are created per call.
    static const bool trace = env_int("CTC_B200_HOST_TRACE", 0) != 0;
    cudaEvent_t tr0 = nullptr, tr[kHostChunks][3] = {};
    const auto wall0 = std::chrono::steady_clock::now();
    if (trace) {
        cudaEventCreate(&tr0);
        for (int k = 0; k < nch; ++k)
            for (int i = 0; i < 3; ++i) cudaEventCreate(&tr[k][i]);
        cudaEventRecord(tr0, user);
    }
    // Chunks are issued in batch order.  On the reference's length-sorted batches that is longest first, i.e. the
    // critical recursion starts first.  Measured and rejected: longest first, then the others shortest first (so that
    // their downloads would not queue behind the long chunk's) -- uploads and downloads contend on the link and the
    // call got slower (headline batch: 1.66e8 against 1.95e8 frames/s); with the serial upload stream below the two
    // orders measure the same (1.99e8 / 2.01e8): the slice uploaded last then is the second longest, and the call ends
    // with ITS recursion instead of with the queue of downloads.
    // Uploads go through ONE stream, slice after slice: issued on the slices' own streams they share the link, and the
    // first slice -- the longest utterances on the reference's length-sorted batches, i.e. the critical recursion --
    // would have its logits only when nearly everything has arrived.  CTC_B200_HOST_SERIAL_UPLOAD=0: the old way.
    static const bool serial_up = env_int("CTC_B200_HOST_SERIAL_UPLOAD", 1) != 0;
    if (nch > 1 && serial_up && cudaStreamWaitEvent(c->up, c->start, 0) != cudaSuccess)
        return CTC_B200_STATUS_EXECUTION_FAILED;
    for (int k = 0; k < nch; ++k) {
        const int b0 = b0s[k], nb = b0s[k + 1] - b0;
        cudaStream_t st = nch > 1 ? c->side[k] : user;
        cudaStream_t sup = (nch > 1 && serial_up) ? c->up : st;
        if (nch > 1 && cudaStreamWaitEvent(st, c->start, 0) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
        // compact device copies [T][nb][C] of this slice (rows of the host tensor are B*C floats apart)
        float *da = c->d_acts + (size_t)T * b0 * C;
        float *dg = grads_host ? c->d_grads + (size_t)T * b0 * C : nullptr;
        const size_t width = (size_t)nb * C * sizeof(float);
        const int Tk = Ts[k];
        if (Tk > 0 && cudaMemcpy2DAsync(da, width, acts_host + (size_t)b0 * C, host_pitch, width, (size_t)Tk,
                                        cudaMemcpyHostToDevice, sup) != cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        if (sup != st) {
            if (cudaEventRecord(c->uploaded[k], sup) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
            if (cudaStreamWaitEvent(st, c->uploaded[k], 0) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
        }
        h2d += (unsigned long long)width * Tk;
        d2h += (dg ? (unsigned long long)width * Tk : 0ull) + sizeof(float) * (size_t)nb;
        if (trace) cudaEventRecord(tr[k][0], st);
        ctcB200Options ok = o;
        ok.stream = (void *)st;
        ok.lengths_on_device = 1;
        ok.costs_ready = nullptr;  // meaningless here: the call synchronises before it returns
        ok.loss_sum = nullptr;     // the host entry sums the downloaded costs itself
        ok.comm = nullptr;
        ok.loss_sum_all = nullptr;
        ok.max_label_len = mxs[k] > 0 ? mxs[k] : 1;
        ctcB200Status s = compute_impl(da, dg, d_labels + lab_off[k], d_llen + b0, d_alen + b0, Tk, nb, C,
                                       c->d_costs + b0, static_cast<unsigned char *>(c->d_ws) + ws_off[k],
                                       ws_off[k + 1] - ws_off[k], &ok, false);
        if (s != CTC_B200_STATUS_SUCCESS) return s;
        launches += g_last_launches;
        if (trace) cudaEventRecord(tr[k][1], st);
        if (dg && Tk > 0 && cudaMemcpy2DAsync(grads_host + (size_t)b0 * C, host_pitch, dg, width, width, (size_t)Tk,
                                              cudaMemcpyDeviceToHost, st) != cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        if (cudaMemcpyAsync(costs_host + b0, c->d_costs + b0, sizeof(float) * nb, cudaMemcpyDeviceToHost, st) !=
            cudaSuccess)
            return CTC_B200_STATUS_MEMOPS_FAILED;
        if (trace) cudaEventRecord(tr[k][2], st);
        if (nch > 1) {
            if (cudaEventRecord(c->done[k], st) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
            if (cudaStreamWaitEvent(user, c->done[k], 0) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
        }
    }
    g_last_launches = launches;
    g_last_h2d = h2d;
    g_last_d2h = d2h;
    if (grads_host)  // padding frames past each slice's longest utterance (disjoint from what the downloads write)
        for (int k = 0; k < nch; ++k) {
            const size_t width = (size_t)(b0s[k + 1] - b0s[k]) * C * sizeof(float);
            for (int t = Ts[k]; t < T; ++t)
                memset(reinterpret_cast<char *>(grads_host) + (size_t)t * host_pitch + (size_t)b0s[k] * C * sizeof(float),
                       0, width);
        }
    const auto wall1 = std::chrono::steady_clock::now();
    if (cudaStreamSynchronize(user) != cudaSuccess) return CTC_B200_STATUS_EXECUTION_FAILED;
    if (trace) {
        const auto wall2 = std::chrono::steady_clock::now();
        fprintf(stderr, "[ctc_b200 host trace] enqueue %.1f us, wait %.1f us;",
                std::chrono::duration<double, std::micro>(wall1 - wall0).count(),
                std::chrono::duration<double, std::micro>(wall2 - wall1).count());
        for (int k = 0; k < nch; ++k) {
            float a = 0, b = 0, d = 0;
            cudaEventElapsedTime(&a, tr0, tr[k][0]);
            cudaEventElapsedTime(&b, tr0, tr[k][1]);
            cudaEventElapsedTime(&d, tr0, tr[k][2]);
            fprintf(stderr, " slice %d: up %.0f, kernels %.0f, down %.0f us;", k, a * 1e3, b * 1e3, d * 1e3);
            for (int i = 0; i < 3; ++i) cudaEventDestroy(tr[k][i]);
        }
        fprintf(stderr, "\n");
        cudaEventDestroy(tr0);
    }
    if (loss_sum) {
        double acc = 0.0;
        for (int b = 0; b < B; ++b) acc += (double)costs_host[b];
        *loss_sum = acc;
    }
    return CTC_B200_STATUS_SUCCESS;
}
