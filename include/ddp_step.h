/*
 * ddp_step.h — C-ABI of libddp_step.so: the B200-native data-parallel training step.
 *
 * This is the drop-in boundary.  Every entry point is
 * `extern "C"`, takes plain device/host pointers, element counts and a
 * `cudaStream_t` (passed as void*), and returns an int:
 *      0  success
 *     >0  a cudaError_t
 *     <0  -(ncclResult_t) for collective errors, or DDP_EINVAL / DDP_ENOSYS
 * No C++ exception crosses the boundary; `ddp_last_error()` gives the message
 * of the last failing call on the calling thread.
 *
 * Memory: every buffer is owned by the caller (PyTorch's caching allocator in
 * the Python host layer); the library allocates nothing on the device, with one
 * stated exception: ddp_peer_create, whose NVLink receive buffers are
 * communicator resources.  All launches are asynchronous on the given stream
 * and are CUDA-graph capturable.
 *
 * "ref:" citations point at the reference interface each entry point replaces
 * (/root/reference = ArnauGabrielAtienza/pytorch_distributed_examples);
 * "torch:" citations point at the installed torch the reference scripts call into.
 */
#ifndef DDP_STEP_H_
#define DDP_STEP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDP_ABI_VERSION 1
#define DDP_EINVAL (-1000)
#define DDP_ENOSYS (-1001)

/* ---- flat parameter arena of the MNIST CNN `Net` --------------------------
 * ref: horovod/mnist_horovod.py:9-16 (Net.__init__), horovod_mnist_elastic.py:16-23.
 * Eight tensors in state_dict order, each padded to a multiple of 4 floats so
 * every tensor starts 16-byte aligned:
 *   conv1.weight[10,1,5,5] conv1.bias[10] conv2.weight[20,10,5,5] conv2.bias[20]
 *   fc1.weight[50,320] fc1.bias[50] fc2.weight[10,50] fc2.bias[10]
 * The gradient bucket and optimizer state arenas use the same layout.       */
#define DDP_NET_NUM_TENSORS 8
#define DDP_NET_ARENA_ELEMS 21848   /* 21,840 parameters + 8 pad floats */
#define DDP_NET_PARAM_COUNT 21840

int ddp_abi_version(void);
const char* ddp_last_error(void);
/* Binds the calling thread to a device (cudaSetDevice).  Call once per host thread. */
int ddp_set_device(int device);

/* offsets[8], sizes[8] (unpadded element counts) of the arena tensors. */
int ddp_net_param_layout(int64_t* offsets, int64_t* sizes);
/* Bytes of the saved-activation workspace forward() fills and backward() reads. */
int64_t ddp_net_workspace_bytes(int64_t batch);
/* Bytes of scratch (per-CTA partial gradients, inter-kernel gradients, loss partials). */
int64_t ddp_net_scratch_bytes(int64_t batch);

/* ---- Net.forward  (+ F.nll_loss, + argmax for test()) ----------------------
 * ref: horovod/mnist_horovod.py:18-25 (forward), :62 (F.nll_loss);
 *      horovod_mnist_elastic.py:97-99 (torch.max(outputs,1) in test()).
 * params   [DDP_NET_ARENA_ELEMS] f32 arena
 * x        [B,1,28,28] f32, 16-byte aligned
 * target   [B] i64 or NULL (needed for out_loss / out_correct)
 * mask2d   [B,20] f32 in {0,1} or NULL: Dropout2d keep mask (injected for parity)
 * mask1d   [B,50] f32 in {0,1} or NULL: F.dropout keep mask
 *          With training!=0 and a NULL mask the kernel draws the mask from
 *          Philox4x32-10 keyed (seed, offset, element).  training==0: no dropout.
 * offset_dev [1] i64 or NULL: added to `offset` on the device, so a captured CUDA graph draws
 *          fresh masks on every replay (point it at the optimizer's device step counter).
 * workspace  ddp_net_workspace_bytes(B), always required (activations between kernels)
 * out_logp [B,10] f32 log-probabilities (the tensor Net.forward returns)
 * out_loss [1] f32 mean NLL or NULL;  out_argmax [B] i64 or NULL
 * out_correct [1] i64 or NULL: += number of argmax==target (integer atomics)
 * scratch  ddp_net_scratch_bytes(B); required when out_loss != NULL           */
int ddp_net_forward(const float* params, const float* x, const int64_t* target,
                    const float* mask2d, const float* mask1d,
                    uint64_t seed, uint64_t offset, const int64_t* offset_dev,
                    int training, int64_t batch,
                    void* workspace, float* out_logp, float* out_loss,
                    int64_t* out_argmax, int64_t* out_correct, void* scratch,
                    void* stream);

/* ---- loss.backward() through Net ------------------------------------------
 * ref: horovod/mnist_horovod.py:63 ; horovod_mnist_elastic.py:68.
 * dlogp [B,10] f32 = dL/d(log-probs) from autograd, or NULL: then the kernel
 * uses the F.nll_loss(mean) gradient  dlogp[b,c] = -loss_scale/B * [c==target[b]].
 * Writes all 8 gradients into grad_bucket[DDP_NET_ARENA_ELEMS] (overwrite, so
 * optimizer.zero_grad() — ref :60 — costs nothing).  Deterministic: per-CTA
 * partial sums in scratch, reduced in fixed order; no float atomics.
 * Launch structure: the kernels are chained with programmatic dependent launch
 * (environment DDP_NO_PDL=1 turns that off).  For batch <= 1024 the conv2 weight
 * gradient is enqueued on a side stream owned by the library (one per device,
 * created on the first such call made outside a stream capture — until then the
 * kernels simply run one after the other; captured calls record the fork/join
 * in the graph); all
 * work is ordered after `stream` at entry and `stream` waits for it before the
 * final reduction, so the caller sees ordinary stream semantics.              */
int ddp_net_backward(const float* params, const float* x, const float* logp,
                     const float* dlogp, const int64_t* target, float loss_scale,
                     int training, int64_t batch, const void* workspace,
                     void* scratch, float* grad_bucket, void* stream);

/* Forward + backward of ONE training step in a single call: what the reference loop does between zero_grad() and
 * optimizer.step() when the loss is F.nll_loss(output, target) --
 * ref: horovod/mnist_horovod.py:61-63 (output = model(data); loss = F.nll_loss(output, target); loss.backward());
 *      pytorch_elastic/mnist_ddp_elastic.py:72-74 (Trainer._run_batch).
 * Same results as ddp_net_forward(training = 1, out_loss) followed by ddp_net_backward(dlogp = NULL, target), but the
 * head (fc1 .. NLL and its backward) runs as ONE kernel: fc1.weight and the a2 tile are staged once, h / logp are not
 * re-read, one launch less.  loss_scale multiplies the loss gradient (1.0 for the plain mean loss).
 * out_argmax / out_correct / mask2d / mask1d / offset_dev may be NULL. */
int ddp_net_train_fwdbwd(const float* params, const float* x, const int64_t* target,
                         const float* mask2d, const float* mask1d, uint64_t seed, uint64_t offset,
                         const int64_t* offset_dev, float loss_scale, int64_t batch, void* workspace,
                         float* out_logp, float* out_loss, int64_t* out_argmax, int64_t* out_correct,
                         void* scratch, float* grad_bucket, void* stream);

/* ---- flatten-into-bucket (DDP Reducer copy / Horovod fusion buffer) ---------
 * torch: nn/parallel/distributed.py:1255-1280 (Reducer buckets);
 * ref: pytorch_elastic/mnist_ddp_elastic.py:58 (DDP wrap), horovod/mnist_horovod.py:53.
 * A byte-granular gather/scatter: item i copies nbytes[i] between tensor
 * ptrs[i] and bucket + offsets[i].  The three tables live in DEVICE memory
 * (built once per bucket layout).  chunk_item/chunk_start map each CTA to a
 * (item, byte start) so large and small tensors balance.  direction 0 =
 * tensors -> bucket (flatten), 1 = bucket -> tensors (unflatten).             */
#define DDP_BUCKET_CHUNK_BYTES 16384
int ddp_bucket_copy(const uint64_t* ptrs, const int64_t* offsets, const int64_t* nbytes,
                    const int32_t* chunk_item, const int64_t* chunk_start, int64_t nchunks,
                    void* bucket, int direction, void* stream);

/* ---- fused (1/N)*g + optimizer update, in place -----------------------------
 * ref: horovod/mnist_horovod.py:50,64 (SGD lr=.01 + step);
 *      horovod_mnist_elastic.py:41,69 (AdamW); pytorch_elastic/mnist_ddp_elastic.py:173 (Adam).
 * torch: optim/sgd.py:343-380, optim/adam.py:414-547, optim/adadelta.py.
 * One pass over n elements: g = grad[i]*inv_world (+ weight decay), then the
 * update of kind `opt`.  state1/state2: momentum_buffer | exp_avg,exp_avg_sq |
 * square_avg,acc_delta (NULL when unused).  grad_dtype 0 = f32, 1 = bf16
 * (params and state are always f32 master copies).
 * step: 1-based step count used for Adam bias correction / SGD first-step
 * momentum; if step_dev != NULL the kernel instead reads *step_dev + 1 from
 * the device (use ddp_step_increment after it) so a captured CUDA graph can be
 * replayed without re-baking the step.                                        */
enum { DDP_OPT_SGD = 0, DDP_OPT_ADAM = 1, DDP_OPT_ADAMW = 2, DDP_OPT_ADADELTA = 3 };
typedef struct {
  /* doubles, like the Python floats torch.optim holds; derived constants
   * (1-beta1, lr/(1-beta1^t), ...) are formed in double and rounded once. */
  double lr, momentum, dampening, weight_decay;
  double beta1, beta2, eps, rho;
  int nesterov;
  int write_back_grad; /* 1: store the averaged gradient back into grad (f32 only) */
  int bump_step_dev;   /* 1 (with step_dev != NULL): the kernel's last CTA also does *step_dev += 1,
                          saving the separate ddp_step_increment launch; calls on different streams
                          must then not overlap (one library-internal ticket counter) */
} ddp_hyper_t;
int ddp_fused_update(int opt, const void* grad, int grad_dtype, float* params,
                     float* state1, float* state2, int64_t n, float inv_world,
                     const ddp_hyper_t* hyper, int64_t step, const int64_t* step_dev,
                     void* stream);
int ddp_step_increment(int64_t* step_dev, void* stream);
/* Scale a bucket in place: buf[i] *= s (the DDP allreduce hook's 1/N). */
int ddp_scale(void* buf, int dtype, int64_t n, float s, void* stream);

/* ---- DistributedSampler shard indices (bit-exact) ---------------------------
 * ref: horovod/mnist_horovod.py:41-42 ; pytorch_elastic/mnist_ddp_elastic.py:188,89.
 * torch: utils/data/distributed.py:102-134 (g=mt19937(seed+epoch); randperm;
 * pad by wrap-around to ceil(N/W)*W, or truncate when drop_last; take [rank::W]).
 * perm_scratch: device buffer of n int32 (the whole permutation).
 * out: [num_samples] i64, num_samples = ddp_sampler_num_samples(...).          */
int64_t ddp_sampler_num_samples(int64_t n, int world, int drop_last);
int ddp_sampler_indices(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                        uint64_t epoch, int drop_last, int32_t* perm_scratch,
                        int64_t* out, void* stream);
/* Same result as ddp_sampler_indices, for hosts that have no device allocator to take perm_scratch
 * from (all device memory already handed to a fixed pool, a plain-C caller, or a loop that cannot
 * afford cudaMalloc's device-wide synchronisation).  It has no counterpart in the reference: it
 * exists only because of that constraint.  The permutation scratch is a 256 KiB array built into
 * the library (resident on every GPU that loads it, until exit), so n is limited to
 * DDP_SAMPLER_STATIC_MAX_N; larger n is rejected with DDP_EINVAL and a message.  It allocates and
 * frees nothing.  The first call on each device resolves the array's device address — issue it
 * before any stream capture.  All calls share the array: do not let two calls overlap on the GPU. */
#define DDP_SAMPLER_STATIC_MAX_N 65536
int ddp_sampler_indices_static(int64_t n, int world, int rank, int shuffle, uint64_t seed,
                               uint64_t epoch, int drop_last, int64_t* out, void* stream);

/* ---- on-device synthetic 28x28 batch ----------------------------------------
 * replaces ref: horovod/mnist_horovod.py:34-38 (datasets.MNIST + ToTensor + Normalize),
 * which needs a network download.  Sample i of the virtual dataset: pixels u8 ~
 * Philox4x32-10(key=seed, counter=(i, block)), x=(u8/255-0.1307)/0.3081; label =
 * philox % 10.  indices [B] i64 (dataset indices from the sampler) or NULL for
 * base_index + b.                                                             */
int ddp_synth_batch(uint64_t seed, const int64_t* indices, int64_t base_index,
                    int64_t batch, float* out_x, int64_t* out_y, void* stream);
/* The same samples as raw uint8 pixels [B,784] (MNIST's on-disk format, what datasets.MNIST holds
 * before its transform runs; ref: horovod/mnist_horovod.py:38).                                   */
int ddp_synth_batch_u8(uint64_t seed, const int64_t* indices, int64_t base_index,
                       int64_t batch, uint8_t* out_px, int64_t* out_y, void* stream);
/* transforms.ToTensor() + transforms.Normalize((mean,), (std,)) on the device
 * (ref: horovod/mnist_horovod.py:34-37; horovod_mnist_elastic.py:45-48; mnist_ddp_elastic.py:166-169):
 * out[i] = (in[i]/255 - mean)/std in IEEE fp32, bit-identical to torchvision's op order.  Lets a
 * host feed uint8 batches (4x fewer PCIe bytes than fp32) — in/out 16-byte aligned.               */
int ddp_normalize_u8(const uint8_t* in, float* out, int64_t n, float mean, float stdv, void* stream);

/* ---- collectives over NCCL / NVLink -----------------------------------------
 * ref: pytorch_elastic/mnist_ddp_elastic.py:26,58 (gloo process group + DDP allreduce);
 *      horovod/mnist_horovod.py:28,53,56 (hvd.init, allreduce, broadcast_parameters).
 * uid: 128 bytes (ncclUniqueId) created on rank 0 and distributed by the host
 * layer through the torchrun TCP store.  One communicator per elastic generation. */
#define DDP_UID_BYTES 128
int ddp_comm_unique_id(void* uid_out);
int ddp_comm_init(void** comm_out, const void* uid, int world, int rank);
int ddp_comm_destroy(void* comm);
int ddp_comm_abort(void* comm);
/* Failure detection (ref: horovod/horovod_mnist_elastic.py:55 -- @hvd.elastic.run retries after a failed collective):
 * *nccl_result_out <- ncclCommGetAsyncError (0 = healthy, 7 = still in progress, anything else = the ring is broken).
 * The host layer polls it while it waits for a collective with a deadline; a broken or timed-out ring is torn down
 * with ddp_comm_abort (never ddp_comm_destroy, which would wait for the stuck kernel). */
int ddp_comm_async_error(void* comm, int* nccl_result_out);
/* dtype 0 = f32, 1 = bf16, 2 = i64, 3 = u8 */
int ddp_allreduce_sum(void* comm, void* buf, int64_t count, int dtype, void* stream);
int ddp_allreduce_max(void* comm, void* buf, int64_t count, int dtype, void* stream);
int ddp_broadcast(void* comm, void* buf, int64_t nbytes, int root, void* stream);

/* ---- one-shot allreduce over NVLink peer memory, fused with the update ---------
 * For latency-bound buckets on ONE node (world <= 8): replaces ncclAllReduce + ddp_fused_update for the
 * gradient averaging hvd.DistributedOptimizer.step() / DDP perform every step
 * (ref: horovod/mnist_horovod.py:53,64 ; pytorch_elastic/mnist_ddp_elastic.py:58).  Each rank pushes its
 * bucket into every peer's receive slot through cudaIpc-mapped NVLink addresses as 64-bit (value, epoch)
 * words, polls its own slots until every word carries this step's epoch, sums the W contributions in rank
 * order (replicas stay bit-identical) and applies the optimizer rule — one kernel, no fences, one one-way
 * NVLink latency exposed.  (Environment DDP_PEER_LL=0 selects the first protocol: plain data + a separate
 * epoch flag behind system-scope fences.)  The receive buffers are communicator resources allocated by
 * ddp_peer_create (the one place this library calls cudaMalloc); handles travel through the host's
 * rendezvous store: create -> local_handle -> (exchange) -> connect.  n % 4 == 0, n <= max_elems.
 * Every rank must launch the call for the same step; do not overlap two calls on one communicator.      */
#define DDP_IPC_HANDLE_BYTES 64
int ddp_peer_create(void** peer_out, int world, int rank, int64_t max_elems);
int ddp_peer_local_handle(void* peer, void* handle_out /* DDP_IPC_HANDLE_BYTES */);
int ddp_peer_connect(void* peer, const void* all_handles /* world x DDP_IPC_HANDLE_BYTES, rank order */);
int ddp_peer_destroy(void* peer);
/* in-place sum over ranks */
int ddp_peer_allreduce(void* peer, float* bucket, int64_t n, void* stream);
/* sum over ranks, 1/world, optimizer update (same arguments as ddp_fused_update; hyper->write_back_grad
 * stores the averaged gradient back into bucket) */
int ddp_peer_allreduce_update(void* peer, float* bucket, int64_t n, int opt, float* params,
                              float* state1, float* state2, const ddp_hyper_t* hyper, int64_t step,
                              const int64_t* step_dev, void* stream);
/* The WHOLE training step of one rank in one call -- forward, F.nll_loss, backward, gradient exchange, averaged
 * optimizer update -- ref: horovod/mnist_horovod.py:59-64 (zero_grad / model(data) / F.nll_loss / backward /
 * hvd.DistributedOptimizer.step()).  Same arithmetic as ddp_net_forward + ddp_net_backward(target) +
 * ddp_peer_allreduce_update, but the backward's per-CTA partial gradient sums are reduced straight into the peer
 * exchange: the fc / conv2 part is pushed over NVLink by a side-stream kernel WHILE conv1's weight gradient is still
 * being computed (batch > 1024), so only conv1's 260 gradients and one flag round trip stay exposed.
 * params / state1 / state2 are updated in place; grad_bucket (may be NULL unless hyper->write_back_grad) receives the
 * averaged gradient.  fused_head != 0 runs the head as one kernel (see ddp_net_train_fwdbwd).  `peer` as above. */
int ddp_net_train_step_peer(float* params, const float* x, const int64_t* target, const float* mask2d,
                            const float* mask1d, uint64_t seed, uint64_t offset, const int64_t* offset_dev,
                            int64_t batch, void* workspace, float* out_logp, float* out_loss, void* scratch,
                            float* grad_bucket, void* peer, int opt, float* state1, float* state2,
                            const ddp_hyper_t* hyper, int64_t step, const int64_t* step_dev, int fused_head,
                            void* stream);

/* ---- two-shot allreduce for large buckets in symmetric memory (NVSwitch multicast / peer load-store) ----
 * For bandwidth-bound buckets on ONE node (ResNet-50: 102 MB fp32 / 51 MB bf16): replaces the per-bucket
 * allreduce + 1/N of DDP (ref: pytorch_elastic/mnist_ddp_elastic.py:58; torch: distributed/algorithms/
 * ddp_comm_hooks/default_hooks.py:35-54) and of hvd.DistributedOptimizer (ref: horovod/mnist_horovod.py:53).
 * The bucket lives in a SYMMETRIC buffer: the same allocation size on every rank, every rank's replica mapped
 * into this process (replicas[q], q = rank order; replicas[rank] is the local one) and, optionally, all
 * replicas bound to one NVSwitch multicast object mapped at `multicast` (0 = none).  The host layer gets such
 * memory from torch.distributed._symmetric_memory (plumbing) — the library allocates nothing here.
 * The last DDP_SYM_FLAG_BYTES of the buffer, starting at flag_offset (128-byte aligned, zeroed once before
 * the first call and never touched by the caller again), hold the synchronisation flags.
 * In place: elements [byte_offset, byte_offset + n*elt) of every replica are replaced by scale * sum over
 * ranks; rank r reduces the r-th W-th of the range (in the switch with multimem.ld_reduce when `multicast`
 * is given, else by loading the W replicas in rank order) and writes the result into every replica, so all
 * replicas end bit-identical.  dtype 0 = f32, 1 = bf16 (fp32 accumulation); n*elt % 16 == 0.
 * max_ctas <= 0 lets the library size the grid; unroll = 16-byte vectors in flight per thread (0 = the
 * library's measured default; 1, 2, 4, 8 are tuning values).  Every rank must launch the call for the same step with
 * the same arguments; calls on one buffer must not overlap.                                            */
#define DDP_SYM_FLAG_BYTES 4096
int64_t ddp_sym_flag_bytes(void);
int ddp_sym_allreduce(const uint64_t* replicas /* host array [world] of device addresses */, uint64_t multicast,
                      int64_t flag_offset, int world, int rank, int64_t byte_offset, int64_t n, int dtype,
                      float scale, int max_ctas, int unroll, void* stream);

/* ---- the DDP script's own model: MLP + cross-entropy -----------------------------
 * ref: pytorch_elastic/mnist_ddp_elastic.py:133-159 `Model(hidden_layers=5, features=1024)` built at :172,
 *      nn.CrossEntropyLoss at :174, the loop body `_run_batch` at :70-75.
 * Linear(in,F) -> ReLU -> hidden_layers x [Linear(F,F) -> ReLU] -> Linear(F,out).  The wide layers run on the
 * tensor cores as 3xTF32 GEMMs (tcgen05 + TMA; fp32-accurate), the classifier/loss/bias sums as small kernels.
 * Parameter arena: reference state_dict order (input_layer.weight, input_layer.bias, hidden_layers.i.weight,
 * hidden_layers.i.bias, ..., final_layer.weight, final_layer.bias), every tensor padded to 4 floats;
 * ddp_mlp_param_layout returns the number of tensors and fills offsets/sizes (in floats).
 * Constraints: in_features % 4 == 0, features % 4 == 0, hidden_layers <= 5, out_features <= 16,
 * batch % 4 == 0.  The workspace (ddp_mlp_workspace_bytes, caller-allocated, 256-byte aligned) carries the
 * saved activations from forward to backward.
 * forward : out_logits [batch][out] (optional), out_loss = mean cross-entropy (optional, needs target).
 * backward: gradients of loss_scale * loss into grad_bucket (same layout as the arena).  Either `dlogits`
 *           [batch][out] (what autograd hands to the module's backward; then target may be NULL) or, with
 *           dlogits NULL, the fused cross-entropy gradient (softmax - onehot)/batch from `target`.         */
typedef struct ddp_mlp_shape {
  int in_features, features, hidden_layers, out_features;
} ddp_mlp_shape_t;
int64_t ddp_mlp_param_count(const ddp_mlp_shape_t* shape);
int ddp_mlp_param_layout(const ddp_mlp_shape_t* shape, int64_t* offsets, int64_t* sizes, int max_tensors);
int64_t ddp_mlp_workspace_bytes(const ddp_mlp_shape_t* shape, int64_t batch);
int ddp_mlp_forward(const ddp_mlp_shape_t* shape, const float* params, const float* x, const int64_t* target,
                    int64_t batch, void* workspace, float* out_logits, float* out_loss, void* stream);
int ddp_mlp_backward(const ddp_mlp_shape_t* shape, const float* params, const float* dlogits, const int64_t* target,
                     float loss_scale, int64_t batch, void* workspace, float* grad_bucket, void* stream);
/* Building blocks of the above (no counterpart in the reference).  ddp_split_tf32: src [rows][cols] -> two
 * tf32-exact planes hi = rna_tf32(v), lo = rna_tf32(v - hi), row-major and (when hiT/loT are given)
 * transposed [cols][rows].  ddp_gemm_3xtf32: c[M][N] (row pitch ldc) = A . B^T from such planes with three
 * tensor-core products per k-step (lo*hi + hi*lo + hi*hi); the tensor-core accumulator is drained into
 * round-to-nearest fp32 sums every 128 k values.  Operand layouts: a_mn_major = 0: A is [M][K] (K contiguous),
 * 1: A is [K][M]; b_mn_major = 0: B is [N][K], 1: B is [K][N].  The contiguous dimension of each operand and
 * ldc must be multiples of 4, all bases 16-byte aligned.  With an MN-major operand whose row count is not a
 * multiple of 32 the kernel may read (never use) up to 124 bytes past the end of that operand's last row.   */
int ddp_split_tf32(const float* src, int64_t rows, int64_t cols, float* hi, float* lo, float* hiT, float* loT,
                   void* stream);
int ddp_gemm_3xtf32(const float* a_hi, const float* a_lo, int a_mn_major, const float* b_hi, const float* b_lo,
                    int b_mn_major, int M, int N, int K, float* c, int ldc, void* stream);

/* ---- elastic state snapshot -------------------------------------------------
 * ref: horovod_mnist_elastic.py:71-72 (state.commit), :104 (TorchState);
 *      pytorch_elastic/mnist_ddp_elastic.py:95-104 (snapshot.pt).
 * Async copies between a device arena and caller-owned pinned host memory.    */
int ddp_state_snapshot(const void* device_arena, void* pinned_host, int64_t nbytes, void* stream);
int ddp_state_restore(void* device_arena, const void* pinned_host, int64_t nbytes, void* stream);

/* ---- per-kernel timing aid (no counterpart in the reference; used by bench.py's roofline leg) ----
 * When enabled, ddp_net_forward/backward bracket every kernel they launch with CUDA events on the
 * launch stream.  ddp_prof_read synchronises on the last event and returns the duration in ms of
 * the kernels of the most recent forward and backward call, in this order:
 *   conv1_fwd, conv2_fwd, head_fwd, head_bwd, conv2_dgrad, conv2_wgrad, conv1_wgrad, reduce_partials
 * Do not enable during CUDA-graph capture.                                                        */
#define DDP_PROF_KERNELS 8
int ddp_prof_enable(int on);
int ddp_prof_read(float* ms_out /* [DDP_PROF_KERNELS] */);

#ifdef __cplusplus
}
#endif
#endif /* DDP_STEP_H_ */
