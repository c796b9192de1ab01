/*
 * rafiki_b200 -- C ABI of the B200-native Rafiki serving hot path.
 *
 * The reference (vivansxu/rafiki) is pure Python and has no FFI of its own: its hot
 * path is a Python plugin API (rafiki/model/model.py:20-127) whose arithmetic runs
 * inside tensorflow==1.12.0 and NumPy.  This header is therefore the *new* boundary
 * the Python drop-ins (rafiki_b200/model, rafiki_b200/predictor, rafiki_b200/worker)
 * bind with ctypes; each entry point names the reference code it replaces.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++/torch types.
 *   - Every call returns 0 on success, else an RFK_ERR_* code; rfk_last_error() then
 *     holds a thread-local message.  Nothing throws.
 *   - Data buffers are caller-owned DEVICE memory unless a parameter says HOST.
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued on it and the call
 *     returns without synchronising, unless stated.
 *   - sm_100a only.  There is no CPU fallback: without a B200 every compute entry
 *     point fails with RFK_ERR_CUDA / RFK_ERR_UNSUPPORTED.
 */
#ifndef RAFIKI_B200_H
#define RAFIKI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RFK_VERSION 100

#define RFK_OK 0
#define RFK_ERR_INVALID 1     /* bad argument */
#define RFK_ERR_CUDA 2        /* CUDA runtime/driver error */
#define RFK_ERR_UNSUPPORTED 3 /* shape or device outside what the kernels cover */
#define RFK_ERR_COMM 4        /* NCCL / peer-memory error */

#define RFK_F32 0  /* fp32 CUDA-core arithmetic (parity path, <= 1e-5 of the oracle) */
#define RFK_BF16 1 /* bf16 operands on tcgen05 tensor cores, fp32 accumulate */
#define RFK_F16 2  /* fp16 operands on tcgen05 tensor cores, fp32 accumulate: same speed as
                      RFK_BF16, 8x finer operand rounding (11-bit significand); weights must
                      stay below 65504 in magnitude, hidden activations saturate there */

#define RFK_MAX_TRIALS 32  /* most trials one ensemble call accepts */
#define RFK_MAX_CLASSES 16 /* most classes the fused kernels accept */

typedef struct rfk_mlp rfk_mlp_t; /* one trial: packed TfFeedForward weights on one GPU */

int rfk_version(void);
const char* rfk_last_error(void);

/* Launch accounting (no counterpart in the reference): how many kernels this library has launched
 * in this process whose name contains `name_part` (NULL or "": all of them), e.g. "fused",
 * "resize", "peer_reduce".  Lets a caller, a test or bench.py's `gpu_launches` count the CUDA
 * work a call really did instead of assuming it. */
unsigned long long rfk_launch_count(const char* name_part);

/* 0 when `device` is a compute-capability 10.x GPU this library can run on. */
int rfk_device_check(int device);

/*
 * Kernel 1 -- query decode / normalise.
 * Replaces: `X.astype(float32)` feed + Flatten + BatchNormalization(inference) of
 * TfFeedForward (examples/models/image_classification/TfFeedForward.py:95,147-148).
 *   y[i, j] = x[i, j] * scale[j] + shift[j]      (scale/shift NULL: plain cast)
 * x: uint8 [n, d] row-major.  y: out_dtype [n, ld_out] row-major, ld_out >= d, columns
 * d..ld_out-1 are written as zero.  scale, shift: float [d] device arrays.
 */
int rfk_decode_u8(const uint8_t* x, void* y, int64_t n, int d, int ld_out, const float* scale,
                  const float* shift, int out_dtype, void* stream);

/*
 * Follow-on row of SURVEY.md 8(f)3 -- Pillow NEAREST resize folded into the decode.
 * Replaces: ModelDatasetUtils.resize_as_images (rafiki/model/dataset.py:68-78) + the
 * decode above.  x: uint8 [n, s_in, s_in, ch]; y: [n, ld_out] with d = s_out*s_out*ch.
 * Source index = floor((dst + 0.5) * s_in / s_out) per axis (Pillow 5.3.0 default).
 * out_dtype may also be -1 for a uint8 result (resize only; scale/shift ignored).
 */
int rfk_resize_decode_u8(const uint8_t* x, void* y, int64_t n, int s_in, int s_out, int ch,
                         int ld_out, const float* scale, const float* shift, int out_dtype,
                         void* stream);

/*
 * Kernel 2 -- the trial network.
 * Replaces: the Keras graph of TfFeedForward._build_model run by Model.predict
 * (TfFeedForward.py:95,141-157) and its load_parameters (TfFeedForward.py:118-135).
 *
 * rfk_mlp_create packs one trial on the CURRENT device.  All array arguments are HOST
 * pointers.  kernels[i] is the Keras Dense kernel of layer i, row-major [in, out];
 * layers 0..n_hidden-1 are ReLU layers of units[i] outputs, layer n_hidden is the
 * softmax layer with n_classes outputs.  bn_* are the BatchNormalization vectors [d].
 * compute_dtype selects the arithmetic (RFK_F32 / RFK_BF16 / RFK_F16).  Synchronous.
 */
int rfk_mlp_create(rfk_mlp_t** out, int d, int n_hidden, const int* units, int n_classes,
                   const float* const* kernels, const float* const* biases, const float* bn_gamma,
                   const float* bn_beta, const float* bn_mean, const float* bn_var, float bn_eps,
                   int compute_dtype);
int rfk_mlp_destroy(rfk_mlp_t* h);

/* Describe a packed trial: out[0..5] = d, n_hidden, n_classes, compute_dtype,
 * 1 if the single-kernel fused path applies, widest hidden layer. */
int rfk_mlp_info(const rfk_mlp_t* h, int* out6);

/* Scratch bytes rfk_mlp_forward needs for a batch of n queries (0 when the fused
 * single-kernel path applies). */
int rfk_mlp_workspace_bytes(const rfk_mlp_t* h, int64_t n, size_t* out);

/*
 * probs[n, n_classes] (float, row-major) = softmax(MLP(BN(x))) for x: uint8 [n, d].
 * `workspace` is caller-owned device scratch of at least rfk_mlp_workspace_bytes(h, n)
 * bytes (may be NULL when that is 0), 256-byte aligned.
 */
int rfk_mlp_forward(rfk_mlp_t* h, const uint8_t* x, int64_t n, float* probs, void* workspace,
                    size_t workspace_bytes, void* stream);

/*
 * The same forward pass for callers that cannot allocate.  (No counterpart in the reference;
 * it serves hosts that run inside a memory budget fixed at start-up, or that receive raw
 * device pointers from a framework whose allocator they may not call.)  RFK_F32 trials only.
 * The intermediate activations go to a block of RFK_STATIC_SCRATCH_BYTES that is part of the
 * library image on each GPU, so the call takes no workspace and calls no allocator.  A batch
 * whose rfk_mlp_workspace_bytes exceeds that block is refused with RFK_ERR_INVALID.  The
 * first call on a device resolves the block's address (and, under CUDA's lazy module loading, may
 * be the moment the driver places the block on the GPU): make it before any stream capture.  The
 * address is cached per device and does not survive cudaDeviceReset.
 * All calls on a GPU share the block: do not let two of them overlap in time.
 */
#define RFK_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int rfk_mlp_forward_static(rfk_mlp_t* h, const uint8_t* x, int64_t n, float* probs, void* stream);

/*
 * Kernel 3 -- cross-trial probability averaging.
 * Replaces: ensemble_predictions, IMAGE_CLASSIFICATION branch
 * (rafiki/predictor/ensemble.py:10-14): out[i, :] = mean over trials of probs_w[i, :].
 * probs: HOST array of w DEVICE pointers, each float [n, c] (pointers may be peer-GPU
 * memory mapped into this process).  The sum runs in float64 in trial order and is
 * rounded once to float, so out equals float32(reference mean) bit for bit.
 */
int rfk_ensemble_mean(const float* const* probs, int w, int64_t n, int c, float* out, void* stream);
/* Same, without the division: out = float32(sum in float64) -- adds up the partial means of ranks. */
int rfk_ensemble_sum(const float* const* probs, int w, int64_t n, int c, float* out, void* stream);

/*
 * Kernels 1+2+3 for trials that live on one GPU: every trial's forward pass and the
 * averaging, x read once.  Replaces Predictor.predict's fan-out / gather / ensemble
 * (rafiki/predictor/predictor.py:35-70) when the workers share a device.
 * trials: HOST array of w handles created on the current device with the same d and
 * n_classes.  out = sum over the w given trials of probs_w / total_trials: a complete
 * ensemble passes total_trials = w (the mean); a rank that holds only part of a larger
 * ensemble passes the full trial count so that partial results add up across ranks.
 */
int rfk_ensemble_workspace_bytes(rfk_mlp_t* const* trials, int w, int64_t n, size_t* out);
int rfk_ensemble_forward(rfk_mlp_t* const* trials, int w, int total_trials, const uint8_t* x,
                         int64_t n, float* out, void* workspace, size_t workspace_bytes,
                         void* stream);

/*
 * Kernel 3 across the GPUs of one box: the mean over ranks fused with the collective.
 * Replaces: the Predictor's gather of one answer per worker + ensemble_predictions
 * (rafiki/predictor/predictor.py:46-70, rafiki/predictor/ensemble.py:10-14) for
 * one-process-per-GPU ranks.  probs[j] / flags[j]: rank j's partial-mean buffer (count floats,
 * as written by rfk_ensemble_forward with total_trials = all trials) and flag block
 * (2 * world + 1 zero-initialised uint32), mapped into this process (CUDA IPC); HOST arrays of
 * `world` DEVICE pointers.  Every rank calls rfk_peer_reduce with the same step (1, 2, ...),
 * count and grid_blocks; rank r sums slice r of the elements over all ranks (float64, rank
 * order, one rounding) and stores it into root_out, the ROOT's result buffer.  The root then
 * calls rfk_peer_wait on its own flag block; after it, root_out holds the full mean.
 * Callers alternate two probs / result buffers by step parity.
 */
/* Set-up of the peer path: rfk_peer_alloc makes one zero-filled device block other processes of the
 * box can map (handle64: 64 bytes to send them); rfk_peer_open maps such a block for kernels of the
 * current device; rfk_peer_close / rfk_peer_free undo them.  Set-up only: they allocate and synchronise. */
int rfk_peer_alloc(size_t bytes, void** dptr, void* handle64);
int rfk_peer_open(const void* handle64, void** dptr);
int rfk_peer_close(void* dptr);
int rfk_peer_free(void* dptr);
/* Let kernels on the current device load/store the memory of `peer_device` (once per pair). */
int rfk_peer_enable(int peer_device);
int rfk_peer_reduce(const float* const* probs, unsigned int* const* flags, int world, int rank,
                    int root, unsigned int step, int64_t count, float* root_out, int grid_blocks,
                    void* stream);
int rfk_peer_wait(const unsigned int* flags_of_this_rank, int world, unsigned int step, void* stream);
/*
 * Gather form of the same collective, fused into the trial kernel: a rank passes rfk_ensemble_forward
 * an `out` pointer that lies in the ROOT's memory (its slot of the root's rfk_peer_alloc block), so the
 * kernel's epilogue stores the partial means across NVLink tile by tile while it computes.
 * rfk_peer_signal, stream-ordered after that kernel, publishes "step s is in place" by storing `step`
 * into `count` flags in peer memory (HOST array of DEVICE pointers); rfk_peer_wait_flags holds a stream until `count` consecutive flags have
 * all reached `step`.  The root then adds its slots with rfk_ensemble_sum.
 */
int rfk_peer_signal(unsigned int* const* flags, int count, unsigned int step, void* stream);
int rfk_peer_wait_flags(const unsigned int* flags, int count, unsigned int step, void* stream);

/*
 * One call per serving step of the trial-per-GPU ensemble (BASELINE configs[2]).
 * Replaces: one Predictor.predict round -- the same queries pushed to every worker
 * (rafiki/predictor/predictor.py:39-41), one answer gathered per worker and averaged
 * (predictor.py:46-70) -- for one-process-per-GPU ranks of one box.
 *
 * rfk_peer_t holds every rank's mapped block, the block layout and the step counters.  Set-up:
 * each rank allocates rfk_peer_block_bytes(...) with rfk_peer_alloc, the 64-byte handles are
 * exchanged by the host (torch.distributed in the Python layer), every rank maps the others with
 * rfk_peer_open and passes the `world` base pointers (its own at [rank]) to rfk_peer_create.
 *
 * Queries, in place of the reference's per-worker RPUSH of the same JSON query
 * (rafiki/cache/cache.py:36-45): rank r uploads only rows rfk_peer_slice_rows(n, world, r) of the
 * uint8 batch over its own PCIe link, to the address rfk_peer_upload_slot returns, and
 * rfk_peer_allgather_u8 (every rank, same n, stream-ordered after the upload) pulls the other
 * slices over NVLink; *gathered is then the whole batch on this device.
 *
 * rfk_peer_ensemble_step (every rank, same n) enqueues: this rank's fused trial kernel on `x`
 * (w may be 0: the rank contributes zeros) -> the peer-memory reduce of the partial means ->
 * on the root, the wait for all slices.  *result is the root's mean for this step (valid on
 * the root once `stream` has passed the call).  Buffers alternate by step parity: a result
 * must be consumed (on `stream`) before the step after next is enqueued.  All ranks must make
 * the same sequence of calls; every wait inside is bounded (a dead rank makes the others trap).
 */
typedef struct rfk_peer rfk_peer_t;
int rfk_peer_block_bytes(int world, int64_t max_n, int d, int n_classes, size_t* out);
int rfk_peer_slice_rows(int64_t n, int world, int rank, int64_t* row_lo, int64_t* row_hi);
int rfk_peer_create(rfk_peer_t** out, int world, int rank, int root, void* const* bases,
                    int64_t max_n, int d, int n_classes);
int rfk_peer_destroy(rfk_peer_t* p);
int rfk_peer_upload_slot(rfk_peer_t* p, int64_t n, uint8_t** dst, int64_t* row_lo, int64_t* row_hi);
int rfk_peer_allgather_u8(rfk_peer_t* p, int64_t n, const uint8_t** gathered, void* stream);
int rfk_peer_ensemble_step(rfk_peer_t* p, rfk_mlp_t* const* trials, int w, int total_trials,
                           const uint8_t* x, int64_t n, void* workspace, size_t workspace_bytes,
                           float** result, void* stream);
/*
 * Pipelined form of the same step for back-to-back batches (a serving loop): with one trial per GPU
 * the cross-rank sum of batch s is carried inside the trial kernel of batch s + 1 (an idle warp of
 * every CTA adds this GPU's share of the rows while the tensor cores work on the new batch), so a
 * step is ONE kernel per rank and no communication kernel stands between two trial kernels.
 * *prev_result is the result buffer of the PREVIOUS call's batch (NULL after the first call); on the
 * root it is complete once `stream` has passed this call.  rfk_peer_flush completes the last batch
 * and returns its buffer.  Other configurations run rfk_peer_ensemble_step underneath and hand each
 * result out one call later, so a caller sees one behaviour.  Replaces the same reference code as
 * rfk_peer_ensemble_step.
 */
int rfk_peer_ensemble_step_pipelined(rfk_peer_t* p, rfk_mlp_t* const* trials, int w, int total_trials,
                                     const uint8_t* x, int64_t n, void* workspace,
                                     size_t workspace_bytes, float** prev_result, void* stream);
int rfk_peer_flush(rfk_peer_t* p, float** result, void* stream);
/*
 * A promise by the caller about rfk_peer_ensemble_step_pipelined: the query batches handed to it are complete
 * before the call BEFORE it was enqueued (resident batches, or batches gathered two steps ahead).  The trial
 * kernel of a step then starts while the previous step's kernel is still finishing (it is launched as a
 * programmatic dependent and, with this promise, never waits for the previous grid: every other dependency
 * between steps already goes through the flag words).  Default off: the kernel waits for the previous kernel
 * of its stream before it reads the queries.  No counterpart in the reference (its workers never overlap).
 */
int rfk_peer_set_inputs_resident(rfk_peer_t* p, int on);

#ifdef __cplusplus
}
#endif
#endif /* RAFIKI_B200_H */
