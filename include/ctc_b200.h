/*
 * include/ctc_b200.h -- C ABI of the B200-native CTC loss library (libctc_b200.so).
 *
 * Drop-in boundary for the ONE hot path of vadimkantorov/deepspeech.pytorch: the CTC loss forward+backward
 *     loss = criterion(logits, targets, output_sizes.cpu(), target_sizes)        /root/reference/train.py:600
 *     (same call under no_grad at train.py:352 and train.py:457; criterion = CTCLoss() at train.py:904,
 *      imported from the third-party `warpctc_pytorch` at train.py:12).
 * The reference reaches native code for this path through warpctc_pytorch's binding over warp-ctc's C API
 * (get_workspace_size / compute_ctc_loss / ctcGetStatusString / get_warpctc_version -- external, un-vendored,
 * described from its published interface in SURVEY.md section 8(b)).  The entry points below are what a binding
 * for this path binds instead; each names the interface it replaces.  Plain pointers and sizes only -- no torch
 * types cross this boundary.  The reference-side stub is shown in INTEGRATION.md.
 *
 * Conventions
 *   acts   : DEVICE fp32 [T,B,C] contiguous, unnormalised logits (train.py:593 "TxNxH")
 *   grads  : DEVICE fp32 [T,B,C] contiguous, receives grad_scale * d(sum_b cost_b)/d acts; every element is
 *            written exactly once (rows t >= act_lens[b] and infeasible utterances get zeros).  NULL => score only
 *            (the eval call sites train.py:352,457 use only the cost).
 *   labels : int32 flat concatenation of the B targets in batch order (data/data_loader_aug.py:546-547),
 *            values in [0,C) and != blank
 *   label_lens, act_lens : int32 [B]            (data_loader_aug.py:545, model.py:382-393 via train.py:600)
 *   costs  : DEVICE fp32 [B], -log p(l_b | x_b); the caller sums (train.py:601 then divides by B)
 *   lengths/labels may live on the host (what the reference passes) or on the device (lets the caller drop the
 *            `.cpu()` sync at train.py:600) -- see ctcB200Options.lengths_on_device
 *   All work is ordered on options->stream: the caller sees ordinary stream semantics (later work on that stream
 *   sees the results; stream capture into a CUDA graph works).  Internally the recursion and gradient kernels are
 *   launched per LENGTH-BUCKET SLICE of the batch (utterances ranked by frames on the device) on a small per-device
 *   pool of side streams that are forked from and joined back to options->stream with events, so that the gradient
 *   work of short utterances overlaps the recursion of long ones.  The pool (streams + events) is the only state
 *   the library keeps; it is created on first use and shared by all callers (enqueue is serialised by a mutex).
 *   Environment: CTC_B200_SLICES=n overrides the slice count (1 = everything on options->stream itself).
 *   ctc_b200_compute never synchronises the device.  Errors are returned as ctcB200Status, never thrown or aborted.
 */
#ifndef CTC_B200_H_
#define CTC_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* replaces warp-ctc's ctcStatus_t (same five states) */
typedef enum {
    CTC_B200_STATUS_SUCCESS = 0,
    CTC_B200_STATUS_MEMOPS_FAILED = 1,
    CTC_B200_STATUS_INVALID_VALUE = 2,
    CTC_B200_STATUS_EXECUTION_FAILED = 3,
    CTC_B200_STATUS_UNKNOWN_ERROR = 4
} ctcB200Status;

/* replaces warp-ctc's ctcOptions {loc, stream|num_threads, blank_label}; there is no CPU location here */
typedef struct {
    void *stream;          /* cudaStream_t; NULL = legacy default stream */
    int blank;             /* blank label index; the reference uses 0 (labels.json "_", train.py:904 defaults) */
    int zero_infinity;     /* 1: infeasible utterance (L + repeats > T) -> cost 0 ; 0: cost +inf.  Its grads are 0
                              in both modes (never NaN). */
    float grad_scale;      /* multiplies every gradient (fuses train.py:601 `loss / B` into the kernel); 1.0 = raw */
    int lengths_on_device; /* 0: labels/label_lens/act_lens are HOST pointers (reference behaviour)
                              1: they are DEVICE pointers; max_label_len must then be given */
    int max_label_len;     /* upper bound on label_lens[b]; <=0 means "derive it" (host lengths only) */
    int phase_mask;        /* profiling aid, 0 in production: bit0 = skip the recursion (engine) kernel, bit1 = skip
                              the gradient kernel, bit2 = skip the softmax/prep kernel; a skipped phase reuses what
                              the previous full call left in the workspace (bench.py times each kernel alone with
                              CUDA events this way) */
    int nan_to_zero;       /* 1: a NaN logit is read as 0 by every pass -- the training loop's guard
                              (/root/reference/train.py:595-598, `logits[torch.isnan(logits)] = 0` ahead of the
                              criterion) folded into the softmax pass instead of two extra sweeps over the logits.
                              `acts` itself is not modified.  0 (default): logits are used as given */
    void *costs_ready;     /* cudaEvent_t or NULL.  If set, the event is recorded (on an internal stream) as soon as
                              costs[] is final, i.e. when the recursion kernels of every slice have finished -- well
                              before the gradient kernels do.  A caller that all-reduces the summed loss (the path's
                              one collective) waits on it from its communication stream and overlaps the exchange
                              with the gradient pass instead of serialising it behind the call. */
    int *nan_count;        /* DEVICE int32 or NULL; with nan_to_zero, the number of NaN logits found in valid frames
                              (t < act_lens[b]) is ADDED to it (the caller zeroes it; reading it back is what the
                              reference's `torch.isnan(logits).any()` warning needs) */
    float *loss_sum;       /* DEVICE float or NULL; receives sum_b costs[b] -- the value the reference's criterion
                              returns (train.py:600: one element, summed over the batch).  Summed in index order in
                              fp64 by one small kernel that runs as soon as the recursions are done, beside the
                              gradient kernels (the same point costs_ready marks), so the caller needs no reduction
                              launch of its own after the call. */
    void *comm;            /* ctcB200Comm* or NULL.  With a communicator (multi-GPU, one process per GPU)
                              loss_sum_all receives the sum over ALL ranks: the same one-warp kernel exchanges the
                              per-rank sums over NVLink peer memory (see ctc_b200_comm_create).  Requires
                              loss_sum_all; every rank of the communicator must make the same sequence of calls. */
    float *loss_sum_all;   /* DEVICE float or NULL; sum of loss_sum over the ranks of `comm` (without a communicator:
                              the same value as loss_sum) -- what reduce_tensor (data/utils.py:40-44) all-reduces */
    int emissions_ready;   /* 1: the emission rows of `workspace` were already written by ctc_b200_fc_forward for these
                              logits (same T, B, C): the softmax pass is skipped and `acts` is not read (C <= 64) */
} ctcB200Options;

/* replaces get_warpctc_version() */
int ctc_b200_get_version(void);

/* replaces ctcGetStatusString() */
const char *ctc_b200_status_string(ctcB200Status status);

/* fills `opts` with the reference's defaults: blank 0, zero_infinity 1, grad_scale 1, host lengths */
void ctc_b200_default_options(ctcB200Options *opts);

/*
 * replaces get_workspace_size(label_lengths, input_lengths, alphabet_size, minibatch, options, &bytes).
 * Shape-only query (no pointers are dereferenced): T = max time steps, B = minibatch, C = alphabet size,
 * max_label_len = upper bound on any label length.  Workspace is caller-owned scratch; contents are
 * meaningless between calls.
 */
ctcB200Status ctc_b200_workspace_size(int T, int B, int C, int max_label_len, const ctcB200Options *opts,
                                      size_t *bytes);

/*
 * replaces compute_ctc_loss(activations, gradients, flat_labels, label_lengths, input_lengths, alphabet_size,
 *                           minibatch, costs, workspace, options).
 * Enqueues the whole forward+backward on opts->stream and returns.  workspace must be a DEVICE buffer of at
 * least the size reported by ctc_b200_workspace_size for the same (T,B,C,max_label_len), 256-byte aligned.
 */
ctcB200Status ctc_b200_compute(const float *acts, float *grads, const int *labels, const int *label_lens,
                               const int *act_lens, int T, int B, int C, float *costs, void *workspace,
                               size_t workspace_bytes, const ctcB200Options *opts);

/*
 * Host-buffer convenience path (what a non-torch caller of the reference's CPU location would use, and what
 * bench.py's end-to-end leg times): acts/grads/costs are HOST pointers (pinned memory recommended); the call
 * stages them through device buffers owned by `ctx`, runs the kernels and copies results back, ordered on
 * opts->stream, then synchronises that stream.  The batch is pipelined in up to 4 slices of utterances on streams
 * owned by `ctx`: the host->device copy of slice k+1, the kernels of slice k and the device->host copy of slice k-1
 * overlap.  *loss_sum receives sum_b costs[b].
 */
typedef struct ctcB200Context ctcB200Context;
ctcB200Status ctc_b200_create(ctcB200Context **ctx);
ctcB200Status ctc_b200_destroy(ctcB200Context *ctx);
ctcB200Status ctc_b200_compute_host(ctcB200Context *ctx, const float *acts_host, float *grads_host,
                                    const int *labels, const int *label_lens, const int *act_lens, int T, int B,
                                    int C, float *costs_host, double *loss_sum, const ctcB200Options *opts);

/*
 * Greedy (best-path) decoding of the same logits -- the first "next" row of the scope table.
 * Replaces GreedyDecoder.decode (/root/reference/decoder.py:182-197: argmax over the classes of every frame) and
 * process_string (decoder.py:165-180: drop blanks and frames equal to the previous frame; the offset of a kept
 * character is its frame index), which the reference runs in Python on every training batch (train.py:575) and
 * eval batch (train.py:367,472).  Integer results, bit-exact against oracle/greedy_oracle.py; ties in a frame go to
 * the lowest class index (torch.max).
 *   path     : DEVICE int32 [B][T], receives the argmax class of every valid frame (the reference's `max_probs`)
 *   ids      : DEVICE int32 [B][T], receives the collapsed sequence of utterance b in ids[b*T .. b*T + out_lens[b])
 *   offsets  : DEVICE int32 [B][T] or NULL, frame index of every kept character
 *   out_lens : DEVICE int32 [B]
 * ctc_b200_compute_decode = ctc_b200_compute + the decode outputs, produced from the softmax pass that reads the
 * logits anyway (no second read).  ctc_b200_greedy_decode is the stand-alone form for eval: element (t,b,c) of
 * `acts` is read at acts[t*stride_t + b*stride_b + c], so both the loss layout [T,B,C] (stride_t = B*C, stride_b = C)
 * and the reference decoder's batch-first [B,T,C] (stride_t = C, stride_b = T*C) work in place; act_lens is a
 * DEVICE int32 [B] or NULL (= T for every utterance).  Both enqueue on opts->stream and return.
 */
ctcB200Status ctc_b200_compute_decode(const float *acts, float *grads, const int *labels, const int *label_lens,
                                      const int *act_lens, int T, int B, int C, float *costs, void *workspace,
                                      size_t workspace_bytes, int *path, int *ids, int *offsets, int *out_lens,
                                      const ctcB200Options *opts);
ctcB200Status ctc_b200_greedy_decode(const float *acts, long long stride_t, long long stride_b,
                                     const int *act_lens_dev, int T, int B, int C, int *path, int *ids, int *offsets,
                                     int *out_lens, const ctcB200Options *opts);

/*
 * The path's one exchange step, the all-reduce of the summed loss -- replaces reduce_tensor
 * (/root/reference/data/utils.py:40-44: dist.all_reduce of the loss, gloo by default, train.py:99,356,461,606).
 * Fused into the kernel that sums the costs: every rank stores its partial sum straight into every peer's slot buffer
 * over NVLink (CUDA IPC peer mappings), polls its own buffer until all partials of this call have arrived and adds
 * them in rank order, so every rank obtains the same bits and no collective kernel is launched.  Set-up (once):
 *   ctc_b200_comm_create(&c, rank, world)        on every rank, its CUDA device current (world <= 32)
 *   ctc_b200_comm_handle(c, handle64)            64 opaque bytes (a cudaIpcMemHandle_t); exchange them by any means
 *                                                (the Python side uses torch.distributed.all_gather_object)
 *   ctc_b200_comm_open(c, handles)               handles = world x 64 bytes in rank order
 * then pass `comm` + `loss_sum_all` in the options of ctc_b200_compute.  ctc_b200_comm_set_peers is the same for ranks
 * that share one process (raw ctc_b200_comm_local pointers instead of IPC handles).
 */
typedef struct ctcB200Comm ctcB200Comm;
ctcB200Status ctc_b200_comm_create(ctcB200Comm **comm, int rank, int world);
ctcB200Status ctc_b200_comm_handle(ctcB200Comm *comm, void *handle64);
void *ctc_b200_comm_local(ctcB200Comm *comm);
ctcB200Status ctc_b200_comm_open(ctcB200Comm *comm, const void *handles);
ctcB200Status ctc_b200_comm_set_peers(ctcB200Comm *comm, void *const *locals);
ctcB200Status ctc_b200_comm_destroy(ctcB200Comm *comm);
/*
 * A late peer (checkpoint save, validation, a data-loader stall) is WAITED for, as reduce_tensor's collective
 * would; only after `milliseconds` of wall-clock time (default 120 000, or the environment variable
 * CTC_B200_COMM_TIMEOUT_MS at creation) is an exchange given up: loss_sum_all then receives NaN, the communicator's
 * sticky flag is set (ctc_b200_comm_timed_out returns 1) and every later ctc_b200_compute that names this
 * communicator returns CTC_B200_STATUS_EXECUTION_FAILED without enqueueing anything -- ranks that are out of step
 * (one of them skipped a call) are reported, not silently summed.  Recover by destroying the communicator on every
 * rank and creating a new one, or by all-reducing loss_sum with the caller's own collective.
 */
ctcB200Status ctc_b200_comm_set_timeout(ctcB200Comm *comm, long long milliseconds);
int ctc_b200_comm_timed_out(const ctcB200Comm *comm);

/*
 * grads[i] *= *scale_dev for i in [0, count), in place, enqueued on opts->stream -- the autograd backward of the
 * reference binding (`cached_grads * grad_output`, reached from loss.backward() at /root/reference/train.py:620
 * with grad_output = 1/B from train.py:601).  The scalar is read on the DEVICE (no host synchronisation); when it
 * is exactly 1.0 the kernel returns without touching the gradients.
 */
ctcB200Status ctc_b200_scale_grads(float *grads, size_t count, const float *scale_dev, const ctcB200Options *opts);

/*
 * Word / character error counts of decoded transcripts against the targets -- SURVEY.md 8(f) rank 4.  Replaces, per
 * utterance, get_cer_wer (/root/reference/data/utils.py:47-57) over Decoder.wer / Decoder.cer
 * (/root/reference/decoder.py:44-73), which the reference evaluates in Python (python-Levenshtein) for every
 * utterance of every training and eval batch (train.py:369-385, 577-591):
 *   counts[4*b + 0] = wer     : edit distance between the WORD sequences (words = maximal runs of non-space ids)
 *   counts[4*b + 1] = cer     : edit distance between the id sequences with the spaces removed
 *   counts[4*b + 2] = wer_ref : number of reference words, or 1 if there is none
 *   counts[4*b + 3] = cer_ref : number of reference non-space ids, or 1 if there is none
 * (the reference's strings map 1:1 to label ids, decoder.py:165-180, so the counts are computed on ids).
 *   hyp_ids   : DEVICE int32 [B][hyp_pitch], utterance b in hyp_ids[b*hyp_pitch .. + hyp_lens[b]) -- the `ids` /
 *               `out_lens` outputs of ctc_b200_compute_decode / ctc_b200_greedy_decode (hyp_pitch = T)
 *   ref_labels: DEVICE int32 flat targets in batch order, ref_lens: DEVICE int32 [B] -- the criterion's own inputs
 *   max_hyp_len / max_ref_len: upper bounds on hyp_lens / ref_lens (size the CTA's shared memory; longer sequences
 *               are truncated); INVALID_VALUE if they do not fit one CTA (about 6*max_hyp + 3*max_ref ints)
 *   counts    : DEVICE int32 [B][4]
 * Integer results, bit-exact against oracle/wer_oracle.py.  Enqueues one kernel on opts->stream and returns.
 */
ctcB200Status ctc_b200_error_counts(const int *hyp_ids, int hyp_pitch, const int *hyp_lens, const int *ref_labels,
                                    const int *ref_lens, int B, int max_hyp_len, int max_ref_len, int space_index,
                                    int *counts, const ctcB200Options *opts);

/*
 * The logits producer -- SURVEY.md 8(f) rank 3.  Replaces, in the forward direction, the reference's final layer
 *     fully_connected = nn.Sequential(nn.BatchNorm1d(H), nn.Linear(H, C, bias=False)); SequenceWise(fully_connected)
 * (/root/reference/model.py:335-341, applied to x [T,B,H] at model.py:373) and, optionally, the softmax the CTC path
 * applies to its output (the criterion's own pass over the logits; model.py:377): one TMA + tcgen05 (tf32, split for
 * fp32-level accuracy) kernel reads x [T*B, H] once and writes
 *   logits        DEVICE fp32 [T,B,C] (nullable), what model.py:373 returns before the transposes at :375 / train.py:593;
 *   ctc_workspace (nullable) the emission rows of a CTC workspace sized by ctc_b200_workspace_size for the SAME T, B, C:
 *                 a following ctc_b200_compute with opts->emissions_ready = 1 then skips its own softmax pass and never
 *                 reads the logits (rows t >= act_lens_dev[b] are skipped when act_lens_dev is given);
 *   path          (nullable) DEVICE int32 [B][T], argmax class per frame: the input of ctc_b200 greedy collapse.
 * BatchNorm is folded into the weights per call: with bn_batch_stats = 0 it uses bn_mean / bn_var (module.eval(), the
 * reference's validation path train.py:329-431); with 1 it first computes this batch's mean and biased variance
 * (training-mode forward; written to batch_mean_out / batch_var_out if given, so the caller can update running stats).
 * Any of bn_weight / bn_bias / bn_mean / bn_var may be NULL (1 / 0 / 0 / 1).  Forward only: gradients of the layer stay
 * with the caller's autograd.  Limits: C <= 64, H a multiple of 4, x 16-byte aligned.
 */
typedef struct {
    const float *weight;     /* DEVICE fp32 [C,H]: nn.Linear(H, C, bias=False).weight */
    const float *bn_weight;  /* DEVICE fp32 [H] or NULL: BatchNorm1d.weight (gamma) */
    const float *bn_bias;    /* DEVICE fp32 [H] or NULL: BatchNorm1d.bias (beta) */
    const float *bn_mean;    /* DEVICE fp32 [H] or NULL: running_mean (bn_batch_stats = 0) */
    const float *bn_var;     /* DEVICE fp32 [H] or NULL: running_var  (bn_batch_stats = 0) */
    float bn_eps;            /* BatchNorm1d.eps (1e-5) */
    int bn_batch_stats;      /* 1: normalise with the statistics of x itself (module.train()) */
    float *batch_mean_out;   /* DEVICE fp32 [H] or NULL: receives the batch mean (bn_batch_stats = 1) */
    float *batch_var_out;    /* DEVICE fp32 [H] or NULL: receives the biased batch variance */
    int reuse_folded;        /* 1: fc_workspace still holds the folded weights of an earlier call with the SAME layer
                                (weights and statistics unchanged, bn_batch_stats = 0): the fold kernel is skipped --
                                what an evaluation loop wants after its first batch */
} ctcB200FcLayer;
ctcB200Status ctc_b200_fc_workspace_size(int T, int B, int H, int C, size_t *bytes);
ctcB200Status ctc_b200_fc_forward(const float *x, int T, int B, int H, int C, const ctcB200FcLayer *layer,
                                  float *logits, void *ctc_workspace, const int *act_lens_dev, int *path,
                                  void *fc_workspace, size_t fc_workspace_bytes, const ctcB200Options *opts);

/*
 * ctc_b200_compute without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and results as ctc_b200_compute, minus the workspace: the scratch is a file-scope
 * `static __device__` array of CTC_B200_STATIC_WORKSPACE_BYTES inside the library (held by every GPU that loads the
 * library until the process exits).  The call performs no cudaMalloc / cudaFree and calls no allocating library
 * routine; every kernel is enqueued on opts->stream itself (no length-bucket slices, no internal streams).
 *   Limit: shapes whose ctc_b200_workspace_size exceeds CTC_B200_STATIC_WORKSPACE_BYTES are rejected with
 *   CTC_B200_STATUS_INVALID_VALUE and a message on stderr naming the bytes needed (B=16,T=500,C=29 needs 4.2 MiB,
 *   B=32,T=500 8.4 MiB; the B=64,T=1500 headline needs 129 MiB and must use ctc_b200_compute).
 *   Sharing: ALL calls on a device use the same array, and it keeps the previous call's contents.  Nothing in it has
 *   to start at zero (every kernel writes what it later reads), but two calls that overlap on the GPU -- different
 *   streams, or a captured graph replayed beside an eager call -- RACE on it: serialise them.
 */
#define CTC_B200_STATIC_WORKSPACE_BYTES (16u << 20)
ctcB200Status ctc_b200_compute_static(const float *acts, float *grads, const int *labels, const int *label_lens,
                                      const int *act_lens, int T, int B, int C, float *costs,
                                      const ctcB200Options *opts);

/* number of kernel launches the last ctc_b200_compute on this thread enqueued (bench.py's gpu_launches claim) */
int ctc_b200_last_launch_count(void);
/* bytes the last ctc_b200_compute_host on this thread actually moved over the bus, host->device and device->host
   (frames past a slice's longest utterance are neither uploaded nor downloaded); bench.py's e2e byte counts */
void ctc_b200_last_copy_bytes(unsigned long long *h2d, unsigned long long *d2h);

#ifdef __cplusplus
}
#endif
#endif /* CTC_B200_H_ */
