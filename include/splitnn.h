/* libsplitnn — C ABI of the B200-native split-NN hot path.
 *
 * The reference (OniReimu/SplitLearning) has no FFI: its boundary is a Python surface whose
 * arithmetic is delegated to PyTorch-CPU operators.  Each entry point below replaces the
 * operator calls at the cited reference lines (paths relative to the reference repo root).
 * The Python host mirror (splitlearning_b200/{models,data_entities*}.py) binds these through
 * ctypes (splitlearning_b200/_lib.py); INTEGRATION.md shows the stub a maintainer would add.
 *
 * Conventions
 *   - plain pointers and sizes only; every `const float*`/`float*` is a DEVICE pointer unless
 *     the parameter is documented as host memory.  The library never frees caller memory.
 *   - every call returns 0 on success, non-zero on error; `sn_last_error()` (thread-local) then
 *     holds a message.  Nothing falls back to the CPU.
 *   - calls that take `stream` only enqueue work on that cudaStream_t (pass it as void*;
 *     NULL = legacy default stream) and return without synchronising.
 *   - a handle lives on one device; calls switch to that device themselves.  The kernels of a
 *     step overlap their prologues with the previous launch (programmatic dependent launch) and
 *     the library tracks, per DEVICE, whether that previous launch may be writing the weights the
 *     next one wants to prefetch - which assumes the library's work on one device is enqueued on
 *     one stream at a time.  Order different streams on one device with events yourself.
 *   - weights are [out,in] row-major fp32, exactly PyTorch's nn.Linear / nn.Conv2d layout, and
 *     are OWNED BY THE CALLER (bound by pointer, updated in place); optimizer state,
 *     activations kept for the backward pass and split-K scratch are owned by the handle.
 */
#ifndef SPLITNN_H
#define SPLITNN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sn_bob sn_bob;     /* a server-side MLP (also used for Alice's U-shape tail) */
typedef struct sn_alice sn_alice; /* a client-side conv block */

enum { SN_OPT_NONE = 0, SN_OPT_SGD_MOMENTUM = 1, SN_OPT_ADAM = 2 };

int sn_version(void);
const char* sn_last_error(void);
/* Arithmetic of the tensor-core matrix products (forward of batches > 64, backward of batches > 16);
 * process-wide.  All three run on the same tcgen05 kernel except mode 2.
 *   0 (default) 3xTF32: each fp32 operand is split into hi + lo TF32 halves on the fly and the
 *               product is formed as hi*hi + lo*hi + hi*lo with fp32 accumulation: ~1e-5 relative,
 *               inside the 1e-4 parity bound, at one third of the TF32 rate.
 *   1           TF32, one pass: ~5e-4 relative on this model's logits, outside the parity bound;
 *               for throughput runs (BASELINE.json config 5 sweep).
 *   2           plain fp32 FMA (a simple tiled SGEMM): slow, kept as an independent check.
 * The forward of batches <= 64 and the whole step of batches <= 16 always use the fp32
 * weight-streaming kernels.  (No counterpart in the reference: one CPU fp32 path.) */
int sn_set_matmul_precision(int mode);
int sn_get_matmul_precision(void);
/* Route the weight stream of big layers at batch <= 16 through the tensor cores (1) or the CUDA
 * cores (0, default).  The tensor-core form (3xTF32 with the weight operand in tensor memory,
 * gemm_tc.cuh: skinny_fwd_x3_kernel) is fp32-accurate and passes the same parity tests, but an
 * M=128 N=16 K=8 MMA costs ~120 cycles whatever N is, so at batch 16 it is slower than the FMA
 * kernel (36 vs 28 us on fc1); kept selectable for measurement.  Ignored in precision mode 2.
 * (No counterpart in the reference.) */
int sn_set_skinny_tensor_forward(int on);
int sn_get_skinny_tensor_forward(void);
/* Run the small middle of a batch <= 16 training step of a three-layer server (first layer's
 * epilogue, second layer forward, its epilogue, last layer forward, loss, last layer backward, second
 * layer backward + update, its input-gradient reduction) as ONE persistent kernel whose phases meet
 * at grid-wide barriers (1), or as seven dependent launches (0, default).  Results are bit-identical.
 * north_star asked for "a persistent whole-MLP kernel"; built and MEASURED SLOWER on B200: 76-83 us
 * against 51 us for the seven launches (profiles/r2e_trace_vanilla_tail.txt).  Kept
 * selectable, with its parity tests, as that measured answer.  (No counterpart in the reference.) */
int sn_set_persistent_tail(int on);
int sn_get_persistent_tail(void);
/* Batches of 17..rows rows run their backward as fp32 weight-stream traversals in tiles of 16 with a
 * gradient buffer (the round-1 form); larger ones as tcgen05 GEMMs with the optimizer in the dW GEMM's
 * epilogue.  Default 16: the GEMMs are faster from 17 rows up and each is exact to ~1e-6 (the forward
 * of batches <= 64 stays on the fp32 FMA kernels either way, so no ReLU gate can flip).  rows in
 * [16, 64]; SN_BWD_TILED_MAX sets the initial value.  (No counterpart in the reference.) */
int sn_set_bwd_tiled_max(int rows);
int sn_get_bwd_tiled_max(void);
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
uint64_t sn_launch_count(void);

/* ------------------------------------------------------------------ stateless operators */

/* conv3x3(1->32)+bias -> ReLU -> maxpool2x2 -> flatten.  x [B,1,28,28] -> act [B,5408]
 * (index c*169+h*13+w).  Replaces model1_sisa.forward / model1.forward, models.py:27-30,11-14. */
int sn_conv_pool_forward(const float* x, const float* w, const float* b, int B, float* act,
                         void* stream);

/* dst <- src for one conv block's parameters (w [32,1,3,3], b [32]; device pointers on the current
 * device, or peer-accessible): the Alice -> Alice weight hand-off of the vanilla / U-shape protocol,
 * `self.model.load_state_dict(last_alice_rref.rpc_sync().give_weights())`, data_entities_vanilla.py:59-64,
 * as one small launch in the stream's kernel chain. */
int sn_conv_weights_copy(float* dst_w, float* dst_b, const float* src_w, const float* src_b, void* stream);

/* Weight/bias gradient of the block above given dL/dact [B,5408]; no input gradient (the input
 * is data).  dw [32*9], db [32].  `scratch` >= sn_conv_pool_backward_scratch_bytes(B) bytes.
 * Replaces autograd's conv/relu/pool backward under loss.backward(),
 * data_entities_vanilla_sisa.py:68 and dist_autograd.backward, data_entities_vanilla.py:231. */
size_t sn_conv_pool_backward_scratch_bytes(int B);
int sn_conv_pool_backward(const float* x, const float* w, const float* b, const float* dact, int B,
                          float* dw, float* db, void* scratch, void* stream);

/* y[B,N] = act(x[B,K] . W[N,K]^T + bias) (* keep_mask * mask_scale).  relu: 0/1.  keep_mask
 * (nullable) is [B,N] bytes, 1 = keep.  Split-K partial sums go through `scratch`
 * (>= sn_linear_forward_scratch_bytes(B,K,N) bytes).  Replaces F.linear/F.relu/F.dropout in
 * model2_sisa.forward, models.py:55-63 (and model2 :39-44, model3 :92-94). */
size_t sn_linear_forward_scratch_bytes(int B, int K, int N);
int sn_linear_forward(const float* x, const float* W, const float* bias, int B, int K, int N,
                      int relu, const uint8_t* keep_mask, float mask_scale, float* y,
                      void* scratch, size_t scratch_bytes, void* stream);

/* sn_linear_forward for callers that cannot allocate GPU memory and have no scratch buffer to
 * lend.  (No counterpart in the reference, which never manages device memory.)  Who that is: a
 * process whose HBM is already owned by another allocator (a serving stack's memory pool, a
 * fixed partition per tenant), where a cudaMalloc from inside a library can fail or fragment the
 * pool; and latency-critical loops, where cudaMalloc/cudaFree stall every stream on the device.
 * Same arguments and result as sn_linear_forward minus the scratch pair: the split-K partial sums
 * go to a 4 MiB array built into the library (sn_linear_forward_noalloc_max_scratch_bytes()),
 * which each GPU that loads the library keeps for the life of the process.  Shapes whose partials
 * do not fit (check with sn_linear_forward_scratch_bytes) are refused with an error; nothing is
 * truncated.  The call itself performs no cudaMalloc/cudaFree or other allocating call.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *     with that device current); make that first call before starting any stream capture.
 *   - All calls on a device share the one array and it keeps whatever the previous call left in
 *     it.  Nothing in it needs clearing between calls (every value read was written earlier in
 *     the same call), but two calls whose kernels overlap on the GPU — different streams with no
 *     dependency between them — race on it.  Serialise them on one stream or with events. */
size_t sn_linear_forward_noalloc_max_scratch_bytes(void);
int sn_linear_forward_noalloc(const float* x, const float* W, const float* bias, int B, int K,
                              int N, int relu, const uint8_t* keep_mask, float mask_scale,
                              float* y, void* stream);

/* Mean cross-entropy over B rows of C logits with int64 labels, and its gradient
 * (softmax - onehot)/B.  loss: device scalar.  dlogits nullable.  Replaces
 * nn.CrossEntropyLoss, data_entities_vanilla.py:224,230; data_entities_vanilla_sisa.py:65,308;
 * data_entities.py:76. */
int sn_cross_entropy(const float* logits, const int64_t* labels, int B, int C, float* loss,
                     float* dlogits, void* stream);

/* counters6 += {correct, total, correct&label==omit, label==omit, correct&label!=omit,
 * label!=omit} with prediction = first argmax over C.  counters6: device int64[6].  Replaces
 * the torch.max/compare/sum block of alice.eval_breakdown, data_entities_vanilla.py:181-194
 * (and alice.eval :93-95 via counters 0,1). */
int sn_eval_counts(const float* logits, const int64_t* labels, int B, int C, int64_t omit_label,
                   int64_t* counters6, void* stream);

/* ------------------------------------------------------------------ Bob: the server MLP */

/* dims[0..n_layers] = layer widths, e.g. {5408,5000,1000,100} (model2_sisa, models.py:46-53),
 * {5408,1000,100} with relu_last=1 (model2, :33-37), {5408K,5000,1000,100K} (concat, :66-73),
 * {100,10} (model3, :87-90).  Hidden layers get ReLU and, when dropout_p>0 and the call is in
 * train mode, inverted dropout.  W[l] [dims[l+1],dims[l]] and b[l] are caller-owned device
 * buffers updated in place.  n_opt_slots independent optimizer states are kept: the reference
 * builds one DistributedOptimizer per Alice over Bob's parameters
 * (data_entities_vanilla.py:37-42), so Bob holds one momentum set per Alice.
 * optimizer: SN_OPT_SGD_MOMENTUM (lr, momentum; F.sgd) or SN_OPT_ADAM (lr, beta1, beta2, eps,
 * coupled-L2 weight_decay; _single_tensor_adam).  Hyper-parameters are doubles because torch keeps
 * them as python floats and derives Adam's bias corrections in double before rounding to fp32. */
int sn_bob_create(sn_bob** out, int device, const int* dims, int n_layers, int relu_last,
                  double dropout_p, int optimizer, double lr, double momentum, double beta1,
                  double beta2, double eps, double weight_decay, int n_opt_slots, int max_batch,
                  float* const* W, float* const* b);
int sn_bob_destroy(sn_bob* h);

/* Fresh optimizer state in `slot` (zeroed moments, step 0): alice.unlearn builds a new
 * optimizer, data_entities_vanilla.py:132-137; data_entities_vanilla_sisa.py:145. */
int sn_bob_reset_opt_slot(sn_bob* h, int slot, void* stream);
/* Optimizer-state access for checkpointing / teacher-forced tests.  which: 0 = momentum buffer
 * or Adam exp_avg, 1 = Adam exp_avg_sq.  param: 2*l = W[l], 2*l+1 = b[l]. */
int sn_bob_opt_state(sn_bob* h, int slot, int which, int param, float** dptr, size_t* count);
int sn_bob_get_opt_step(sn_bob* h, int slot, int64_t* step);
int sn_bob_set_opt_step(sn_bob* h, int slot, int64_t step);
int sn_bob_set_lr(sn_bob* h, double lr);
/* Treat the last layer's dims[L] outputs as `heads` independent classifiers of dims[L]/heads
 * classes each: the loss becomes the mean CE over B*heads rows and `labels` is [B,heads].
 * heads=1 (default) is the reference's single CE.  Used by the --concat server, whose fc3 emits
 * 100 logits per client (models.py:73); the reference's own concat loop crashes before any loss
 * is computed (data_entities_sisa_concat.py:270-276), so these semantics are defined here. */
int sn_bob_set_ce_heads(sn_bob* h, int heads);
/* Evaluation over a whole test shard in a few launches.  By default forward batches above 64 rows
 * take the tcgen05 GEMM (3xTF32: 1e-5 off the 16-row path, enough to flip a near-tie argmax).  After
 * this call, sn_bob_forward calls of up to `rows` rows (<= max_batch) run the fp32
 * weight-stream kernels in row tiles of 16 inside one launch per layer: each row's arithmetic is then
 * exactly what a batch-of-16 loop computes, so the counters of alice.eval_breakdown
 * (data_entities_vanilla.py:159-202; sisa :93-137) come out identical with ~7 launches per shard
 * instead of 7 per 16 samples.  The split-K scratch grows with `rows`; when max_scratch_bytes > 0 the
 * row count is halved until the scratch fits, and *rows_granted (nullable) returns what was set.
 * Synchronises the device (it swaps the scratch buffer). */
int sn_bob_set_stream_rows(sn_bob* h, int rows, size_t max_scratch_bytes, int* rows_granted);
/* Measurement aid (no counterpart in the reference): while enabled, every launch of the two
 * weight-streaming kernels is bracketed by CUDA events on the launching stream.
 * sn_bob_profile_read sums them for kind 0 = forward (layer l), 1 = backward+update (layer l),
 * synchronising on the recorded events.  Enabling or disabling clears earlier records. */
int sn_bob_profile(sn_bob* h, int enable);
int sn_bob_profile_read(sn_bob* h, int kind, int layer, double* total_ms, int* launches);

/* One training batch on Bob: forward (train mode), mean CE over the last layer's logits,
 * backward, optimizer update of `slot`; cut_grad [B,dims[0]] = dL/dact (nullable: SISA's input
 * is detached).  drop_masks: nullable host array of (n_layers-1) device pointers to [B,dims[l+1]]
 * keep-bytes (teacher-forced / torch-CPU-stream parity mode); when NULL masks come from Philox
 * keyed by (seed, offset).  logits_out [B,dims[L]] nullable.  loss_out: one float, nullable - device
 * memory, or pinned HOST memory (device-visible under UVA): the loss kernel then stores the value over
 * PCIe itself and no D2H copy is needed in the stream (valid once the stream is synchronised).
 * Batches of 1..16 rows take the fused weight-stream kernels, 17..64 rows the same kernels in tiles of 16
 * with gradient accumulation, larger ones the tcgen05 GEMMs (sn_set_matmul_precision).
 * Replaces bob.train_and_backward + the Bob half of dist_optimizer.step,
 * data_entities_vanilla.py:226-232,76 ; and the batch body of SISA's server loop,
 * data_entities_vanilla_sisa.py:305-313. */
int sn_bob_train_step(sn_bob* h, int slot, const float* act, const int64_t* labels, int B,
                      const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset,
                      float* cut_grad, float* logits_out, float* loss_out, void* stream);

/* Forward only; out [B,dims[L]].  train_mode=1 applies dropout (the reference's --vanilla never
 * calls .eval(), so bob.inference is stochastic there) and keeps activations for a following
 * sn_bob_backward_update.  Replaces bob.inference, data_entities_vanilla.py:234-235;
 * data_entities.py:166-167. */
int sn_bob_forward(sn_bob* h, const float* act, int B, int train_mode,
                   const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset, float* out,
                   void* stream);

/* Backward + update from dL/dout [B,dims[L]] of the preceding train-mode sn_bob_forward;
 * cut_grad nullable.  Replaces Bob's share of dist_autograd.backward + dist_optimizer.step in
 * the U-shape step, data_entities.py:79,81. */
int sn_bob_backward_update(sn_bob* h, int slot, const float* act, const float* dout, int B,
                           float* cut_grad, void* stream);

/* SISA server epoch: for i in [0,n_batches): one sn_bob_train_step (no cut_grad) on rows
 * [batch_offsets[i], batch_offsets[i+1]) of the HBM-resident cache [rows,dims[0]] / labels[rows].
 * batch_offsets: HOST int64[n_batches+1].  drop_masks: nullable host array of device pointers to
 * [rows,dims[l+1]] keep-bytes.  loss_per_step: device float[n_batches], nullable.  Replaces the
 * inner loop of bob.train_and_backward, data_entities_vanilla_sisa.py:304-313. */
int sn_bob_server_epoch(sn_bob* h, int slot, const float* cache, const int64_t* labels,
                        const int64_t* batch_offsets, int n_batches,
                        const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset,
                        float* loss_per_step, void* stream);

/* ------------------------------------------------------------------ Alice: the client conv block */

/* w [32,1,3,3], b [32] caller-owned, updated in place.  optimizer as for Bob (Alice's half of
 * the per-Alice DistributedOptimizer, data_entities_vanilla.py:37-42; local Adam wd=1e-5,
 * data_entities_vanilla_sisa.py:48). */
int sn_alice_create(sn_alice** out, int device, int optimizer, double lr, double momentum,
                    double beta1, double beta2, double eps, double weight_decay, int max_batch,
                    float* w, float* b);
int sn_alice_destroy(sn_alice* h);
int sn_alice_reset_opt(sn_alice* h, void* stream);
int sn_alice_opt_state(sn_alice* h, int which, int param, float** dptr, size_t* count);
int sn_alice_get_opt_step(sn_alice* h, int64_t* step);
int sn_alice_set_opt_step(sn_alice* h, int64_t step);

/* act [B,5408] may live on another device (peer-mapped) — e.g. directly in Bob's HBM. */
int sn_alice_forward(sn_alice* h, const float* x, int B, float* act, void* stream);
/* The same for a batch that still lies in PINNED HOST memory (cudaHostAlloc / torch pin_memory: device-
 * visible under UVA): the conv kernel reads host_x [B,1,28,28] and host_y [B] over PCIe itself, in its
 * prologue - before its dependency wait, so the transfer overlaps the previous kernel of the stream -
 * and leaves device copies in x_dev [B,784] / y_dev [B] for whoever needs the batch again (the conv
 * backward, the server's loss).  Replaces the DataLoader -> model hand-off of a CPU batch,
 * data_entities_vanilla.py:67-72, without a separate H2D copy, side stream or event in the step.  The
 * caller alternates between two (x_dev, y_dev) pairs from step to step: this step's prologue may
 * overwrite them while the previous step's backward still reads its own pair.  host_y/y_dev may both
 * be NULL.  B < 512. */
int sn_alice_forward_host(sn_alice* h, const float* host_x, const int64_t* host_y, int B, float* act,
                          float* x_dev, int64_t* y_dev, void* stream);
/* The same with a row stride: sample b's 5408 activations go to act + b*ld_act (ld_act >= 5408, in
 * floats).  With ld_act = 5408*K and act pointing at column 5408*k of the --concat server's
 * [rows, 5408*K] cache - normally a peer pointer into Bob's GPU - one launch per Alice fills her slot
 * of every concatenated row in place: the gather of alice.give_activation_and_labels,
 * data_entities_vanilla_sisa.py:196-211, feeding model2_sisa_concat's input, models.py:75-76, without
 * a torch.cat.  From 512 samples up both forms run a streaming kernel (same numbers). */
int sn_alice_forward_ld(sn_alice* h, const float* x, int B, float* act, size_t ld_act, void* stream);
/* conv weight gradient from the cut-layer gradient + optimizer update.  x must be the batch given
 * to the matching forward.  Replaces Alice's share of dist_autograd.backward +
 * dist_optimizer.step, data_entities_vanilla.py:231,76. */
int sn_alice_backward_update(sn_alice* h, const float* x, const float* cut_grad, int B,
                             void* stream);
/* SISA phase-1 / unlearn-retrain batch: forward, CE over the 5408 activations, backward, update.
 * Replaces data_entities_vanilla_sisa.py:62-70 and :160-168. */
int sn_alice_local_step(sn_alice* h, const float* x, const int64_t* labels, int B,
                        float* loss_out, void* stream);

/* One pass of sn_alice_local_step over n device-resident samples in batches of `batch` (the last
 * one may be short): the whole inner loop of alice.train, data_entities_vanilla_sisa.py:59-70, in
 * one call.  loss_per_step: device float[ceil(n/batch)], nullable. */
int sn_alice_local_epoch(sn_alice* h, const float* x, const int64_t* labels, int64_t n, int batch,
                         float* loss_per_step, void* stream);

/* One whole vanilla split step on one device in one call: sn_alice_forward -> sn_bob_train_step
 * (Philox dropout) -> sn_alice_backward_update, i.e. the body of alice.train's batch loop,
 * data_entities_vanilla.py:70-76.  act and cut_grad are caller scratch [B,5408]. */
int sn_vanilla_step(sn_alice* alice, sn_bob* bob, int slot, const float* x, const int64_t* labels,
                    int B, uint64_t seed, uint64_t offset, float* act, float* cut_grad,
                    float* loss_out, void* stream);

/* sn_vanilla_step with the batch taken from pinned host memory (sn_alice_forward_host): the whole
 * end-to-end step - host batch in, loss out (loss_out may be pinned host memory too) - is one call and
 * one unbroken chain of dependent launches. */
int sn_vanilla_step_host(sn_alice* alice, sn_bob* bob, int slot, const float* host_x, const int64_t* host_y,
                         int B, uint64_t seed, uint64_t offset, float* x_dev, int64_t* y_dev, float* act,
                         float* cut_grad, float* loss_out, void* stream);

/* ------------------------------------------------------------------ NVLink plumbing */
int sn_p2p_enable_all(int n_devices);
int sn_p2p_copy(void* dst, int dst_device, const void* src, int src_device, size_t bytes,
                void* stream);

/* Peer-memory mailboxes between the one-process-per-GPU ranks: what carries the cut-layer
 * activations, labels and the cut-layer gradient in place of the reference's rpc_sync /
 * dist_autograd round trips (data_entities_vanilla.py:63,73 and :231) - and in place of an NCCL
 * send/recv pair, whose ~20 us latency per message is a quarter of the whole step at batch 16.
 * A rank allocates a mailbox on its own GPU (sn_ipc_alloc: plain cudaMalloc, zero-filled, plus
 * its 64-byte CUDA IPC handle), ranks exchange the handles out of band (torch.distributed), and
 * sn_ipc_open maps a peer's mailbox (another PROCESS's allocation; NVLink peer access is enabled
 * by the mapping).  The producing kernel then writes its output straight into the consumer's
 * mailbox over NVLink (the conv kernel's `act`, the backward's `cut_grad` are just peer
 * pointers), and a 4-byte flag in the same mailbox orders the two streams:
 *   sn_stream_signal(flag, v, stream)  after all prior work of `stream`: *flag = v
 *   sn_stream_wait(flag, v, stream)    work enqueued after it starts once *flag >= v
 * (cuStreamWriteValue32 / cuStreamWaitValue32: no kernel spins, no host involvement).
 * Flags must be 4-byte aligned device memory of a mailbox; values must grow monotonically. */
int sn_ipc_alloc(size_t bytes, void** dptr, void* handle64);
int sn_ipc_open(const void* handle64, void** dptr);
int sn_ipc_close(void* dptr);
int sn_ipc_free(void* dptr);
int sn_stream_signal(void* flag, uint32_t value, void* stream);
int sn_stream_wait(void* flag, uint32_t value, void* stream);
/* The same two operations as one-CTA gate kernels (ld.acquire.sys polling / st.release.sys), which
 * is what the multi-GPU schedule uses: a stream-ordered value wait costs ~20 us per hop on this
 * system, a gate kernel ~3 us, and the launch queued behind a waiting gate overlaps its prologue
 * with the wait.  The waiter polls memory of its OWN GPU that a kernel on the PEER GPU writes.
 * sn_flag_signal can first copy `copy_bytes` (a multiple of 8, <= 64 KiB; 0 = none) from
 * `copy_src` to `copy_dst` (the labels, into the peer's mailbox). */
int sn_flag_signal(void* flag, uint32_t value, void* copy_dst, const void* copy_src, size_t copy_bytes,
                   void* stream);
int sn_flag_wait(const void* flag, uint32_t value, void* stream);
/* One gate launch that waits for `count` (<= 32) flags, `stride_bytes` apart, all to reach `value`
 * (the head GPU waiting for every shard's slice). */
int sn_flag_wait_many(const void* flags, int count, int stride_bytes, uint32_t value, void* stream);
/* One gate launch that sets `count` (<= 128) flags whose addresses are in a DEVICE array (one flag
 * in each consumer's own memory). */
int sn_flag_signal_many(const void* flag_table_dev, int count, uint32_t value, void* stream);
/* A gate kernel does not poll for ever: after SN_FLAG_TIMEOUT_MS (environment, default 20000; 0 = no
 * limit) it gives up, records which flag it was waiting on in pinned host memory and lets the stream
 * run on.  sn_flag_status() returns non-zero (message in sn_last_error()) once that has happened in
 * this process - results computed after the timed-out wait are invalid - and every later
 * sn_flag_wait* call fails the same way until sn_flag_status_clear().  The reference's analogue is
 * the 60 s RPC timeout of TensorPipe raising at the caller. */
int sn_flag_status(void);
int sn_flag_status_clear(void);

/* ------------------------------------------------------------------ tensor-parallel first layer
 * the server step is a weight stream, and the first layer is 84 % of it, so its
 * rows are split over G GPUs: each GPU streams 1/G of W1 and of its optimizer state.  An sn_shard is
 * rows [n_off, n_off+n_rows) of a [n_total, K] Linear (+ReLU, dropout) with that slice's optimizer
 * state; it borrows W_slice [n_rows,K] and b_slice [n_rows] (updated in place), batch <= 16.
 *   sn_shard_forward          H_out[b*ld_out + n_off + j] = dropout(relu(X W_slice^T + b))  -  H_out is
 *                             the FULL-width activation [B, n_total], normally a peer pointer into the
 *                             head GPU's mailbox; the dropout decision of an element is that of the
 *                             unsharded layer (mask bytes [B,n_total] or Philox on the global index).
 *   sn_shard_backward_update  dH [B, n_total] (row stride ld_dh; normally read from the head GPU over
 *                             NVLink) is the gradient w.r.t. the layer's OUTPUT as the head's
 *                             backward returns it (its cut gradient, not yet gated); the slice gates
 *                             it with its own activation, forms dW, updates W/b and the optimizer
 *                             state in one traversal, and - when X_next is given (SISA: the next
 *                             cached batch) - also multiplies the next batch through the updated
 *                             rows, so the following sn_shard_forward is only an epilogue.
 * No input gradient is produced (SISA server epochs detach the cached activations,
 * data_entities_vanilla_sisa.py:307).  No counterpart in the reference (one CPU process per role). */
typedef struct sn_shard sn_shard;
int sn_shard_create(sn_shard** out, int device, int K, int n_total, int n_off, int n_rows, int relu,
                    double dropout_p, int optimizer, double lr, double momentum, double beta1,
                    double beta2, double eps, double weight_decay, float* W_slice, float* b_slice);
int sn_shard_destroy(sn_shard* h);
int sn_shard_reset_opt(sn_shard* h, void* stream);
/* Optimizer-state access of a slice, for checkpointing and teacher-forced tests: which 0 = momentum
 * / exp_avg, 1 = exp_avg_sq; param 0 = the weight slice [n_rows,K], 1 = the bias slice [n_rows].
 * Setting the step counter also drops any fused next-forward partials (the caller has rewritten
 * weights or state). */
int sn_shard_opt_state(sn_shard* h, int which, int param, float** dptr, size_t* count);
int sn_shard_get_opt_step(sn_shard* h, int64_t* step);
int sn_shard_set_opt_step(sn_shard* h, int64_t step);
int sn_shard_forward(sn_shard* h, const float* X, int B, int train_mode, const uint8_t* keep_mask,
                     uint64_t seed, uint64_t offset, float* H_out, int ld_out, void* stream);
int sn_shard_backward_update(sn_shard* h, const float* X, const float* dH, int ld_dh, int B,
                             const float* X_next, int B_next, void* stream);
/* One pass over cached batches (HOST batch_offsets[n_batches+1], batches of <= 16 rows) with the first
 * layer sharded over `world` ranks: the tensor-parallel form of sn_bob_server_epoch, i.e. of the
 * epoch body of bob.train_and_backward, data_entities_vanilla_sisa.py:298-313.  Every rank calls it
 * with its shard and the same cache (replicated); rank 0 also passes `head` (an sn_bob over the
 * remaining layers, input width = the sharded layer's width) and the labels; rank 0 of a multi-rank
 * run may come WITHOUT a shard (shard = NULL, cache unused: it is then the head only).  Column
 * ownership of H1/dH1: rank 0's shard rows first (possibly none), then equal ranges for ranks 1.. .
 * The head's backward pushes every rank's columns of the input gradient into that rank's memory and
 * then releases the flags.  (SN_TP_EARLY_CUT=1: the gradient is formed by a read-only pass and handed
 * out BEFORE the head's own first-layer update; measured slower from 4 GPUs up, off by default.)
 * H1 is the [16, n_total]
 * activation buffer in rank 0's mailbox (a peer pointer elsewhere) and h_ready_flags the base of the
 * per-rank flags next to it (rank r's at +128*r bytes); dH1 [16, n_total] and dh_ready_flag are THIS
 * rank's own gradient buffer and flag.  Rank 0 also passes two DEVICE arrays: dH1_table_dev[world]
 * (every rank's dH1 buffer - the head pushes rank r's column range into dH1_table_dev[r]) and
 * dh_flag_table_dev[world-1] (the dh_ready flags of ranks 1..world-1).  Flag values used are
 * seq0+1 .. seq0+n_batches.  drop_masks: nullable [first layer (n,n_total),
 * then the head's layers] keep bytes; otherwise Philox keyed by (seed, offset + step). */
int sn_tp_server_epoch(sn_shard* shard, sn_bob* head, int rank, int world, const float* cache,
                       const int64_t* labels, const int64_t* batch_offsets, int n_batches,
                       const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset, float* H1,
                       float* dH1, void* h_ready_flags, void* dh_ready_flag,
                       float* const* dH1_table_dev, const void* dh_flag_table_dev, uint32_t seq0,
                       float* loss_per_step, void* stream);

/* ---- host-side data preparation (no GPU work; these two run on the calling thread) ----
 * LDA (Dirichlet) non-IID partition of n labelled samples over client_num clients: the classification
 * branch of non_iid_partition_with_dirichlet_distribution, /root/reference/data/noniid_partition.py:6-91,
 * as one native call that consumes numpy's LEGACY MT19937 stream exactly as the reference's
 * np.random.shuffle / np.random.dirichlet calls do.  The caller passes the generator state
 * (np.random.get_state(): key[624], pos, has_gauss, cached_gaussian); it is advanced in place, so handing
 * it back with np.random.set_state() leaves the stream where the reference would have left it.
 * out_idx[n] receives the clients' sample ids back to back, out_offsets[client_num+1] the boundaries
 * (samples whose label is outside [0, classes) belong to nobody, as in the reference).  Returns 0, or 2
 * on a NULL/invalid argument. */
int sn_lda_partition(const int64_t* labels, int64_t n, int client_num, int classes, double alpha,
                     uint32_t* mt_key624, int* mt_pos, int* has_gauss, double* cached_gauss,
                     int64_t* out_idx, int64_t* out_offsets);
/* Python's random.shuffle(ids) for the state of random.getstate() (key[624], pos): the shuffle in front
 * of the 80/20 train/test split, /root/reference/data/relational_table_preprocessor.py:83. */
int sn_py_random_shuffle(int64_t* ids, int64_t n, uint32_t* mt_key624, int* mt_pos);

#ifdef __cplusplus
}
#endif
#endif /* SPLITNN_H */
