// Internal entry points shared between the translation units of librafiki_b200.
#pragma once
#include "rfk_common.cuh"

namespace rfk {

// mlp_f32.cu -- fp32 CUDA-core path
size_t mlp_f32_workspace_bytes(const rfk_mlp* h, int64_t n);
int mlp_f32_forward(rfk_mlp* h, const uint8_t* x, int64_t n, float* probs, void* ws, size_t ws_bytes,
                    cudaStream_t st);

// gemm_tc.cu -- layered tcgen05 path (decode -> GEMM+bias+ReLU ... -> GEMM+bias+softmax)
int tc_prepare(rfk_mlp* h);
void tc_release(rfk_mlp* h);
size_t tc_workspace_bytes(const rfk_mlp* h, int64_t n);
int tc_forward(rfk_mlp* h, const uint8_t* x, int64_t n, float* probs, void* ws, size_t ws_bytes,
               cudaStream_t st);

// fused_trial.cu -- single-kernel tcgen05 path for TfFeedForward-shaped trials
bool fused_applicable(const rfk_mlp* h);
int fused_prepare(rfk_mlp* h);
void fused_release(rfk_mlp* h);
// out = sum_w probs_w / total_trials over the w trials given (w == total_trials: the mean)
// optional scratch of fused_workspace_bytes() enables the one-item-per-SM schedule for small batches
size_t fused_workspace_bytes(int w, int n_classes, int64_t n, int d);
int fused_forward(rfk_mlp* const* trials, int w, int total_trials, const uint8_t* x, int64_t n,
                  float* out, void* ws, size_t ws_bytes, cudaStream_t st);

// device images of a trial packed for the fused kernels (owned by the handle)
struct FusedImages {
  const uint8_t* w1;
  const uint8_t* w2;
  const uint8_t* w3;
  const float* bias;
  int h1_pad, h2_pad, n_hidden;
};
bool fused_images(const rfk_mlp* h, FusedImages* out);

// Optional hand-shake of a forward call whose output lies in ANOTHER GPU's memory (peer.cu: a rank's
// kernel stores its partial means straight into the root's slot over NVLink).  Before its first output
// store the kernel waits until *wait_flag >= wait_value (the root has consumed what the slot held); after
// its last store it publishes signal_value in *signal_flag at system scope.  `counter` is a zero-initialised
// word in local memory that the kernel leaves at zero.  Only the pair kernel's single-trial path takes
// the hook inside the kernel; peer.cu brackets every other path with its own tiny kernels.
constexpr int kPeerMaxRanks = 16;
struct PeerHook {
  // before the first output store: every one of the n_wait words at wait_flags is >= wait_value
  const unsigned int* wait_flags;
  int n_wait;
  unsigned int wait_value;
  // after the last output store: signal_value is published (system scope) in each of the n_signal words
  unsigned int* signal_flags[kPeerMaxRanks];
  int n_signal;
  unsigned int signal_value;
  unsigned int* counter;
  // scatter form: the rows of pair tile pt go to out_owner[j] (a full-size [n, C] buffer in GPU j's memory,
  // indexed by global row) for owner_bound[j] <= pt < owner_bound[j + 1]; n_owner == 0: the call's `out`
  float* out_owner[kPeerMaxRanks];
  int owner_bound[kPeerMaxRanks + 1];
  int n_owner;
  // Carried reduction: the sum of the PREVIOUS step's slots that this GPU owns, done by an otherwise idle warp
  // of every CTA while the tensor cores work on this step (n_carry == 0: nothing to carry).  carry_slot[j] is
  // rank j's slot in THIS GPU's memory; rows [carry_v_lo, carry_v_hi) in float4 units; the sums go to
  // carry_out (the root's result buffer of that step), then carry_done (root) and carry_consumed[j] (every
  // rank) receive carry_step.
  const float* carry_slot[kPeerMaxRanks];
  unsigned int* carry_consumed[kPeerMaxRanks];
  const unsigned int* carry_arrived;
  unsigned int* carry_done;
  unsigned int* carry_counter;
  float* carry_out;
  long long carry_v_lo, carry_v_hi;
  unsigned int carry_step;
  int n_carry;
  // The caller guarantees that nothing this launch reads (the queries) was produced by earlier work of its
  // stream and every cross-launch dependency goes through the flags above: the kernel, launched as a
  // programmatic dependent, then does not wait for the previous grid at all (its first CTAs run while the last
  // CTAs of the previous step finish and publish their flags).
  int no_grid_wait;
  // Several trials on this GPU: the trial kernel writes per-trial probabilities to the scratch and carries the
  // previous step's sum only (no owners, no flags of its own); the caller runs the scatter-mean kernel next,
  // so the library's own kernel 3 is skipped.
  int skip_mean;
};
void set_peer_hook(const PeerHook* hook);     // thread-local; consumed (and cleared) by the next pair_forward with w == 1
bool peer_hook_pending();

// fused_pair.cu -- CTA-pair kernel for large batches (weights resident in shared memory, trial-outer order)
bool pair_applicable(int d, int64_t n);
size_t pair_workspace_bytes(int w, int n_classes, int64_t n);
int pair_forward(rfk_mlp* const* trials, int w, int total_trials, const uint8_t* x, int64_t n, float* out, void* ws,
                 size_t ws_bytes, cudaStream_t st);

// ensemble.cu
int scale_inplace(float* p, int64_t count, double divisor, cudaStream_t st);
int ensemble_partial_mean(const float* const* probs, int w, int64_t n, int c, float* out, double divisor,
                          cudaStream_t st);

}  // namespace rfk
